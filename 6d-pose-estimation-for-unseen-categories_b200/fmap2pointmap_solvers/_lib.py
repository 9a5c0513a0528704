"""ctypes binding of libf2p_sm100.so (C-ABI in include/f2p.h) plus the torch-side plumbing: device memory,
streams and a per-device workspace cache.  There is no CPU fallback: if the CUDA library is missing or no CUDA
device is present every solver call raises.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_NAME = "libf2p_sm100.so"
LIB_PATH = os.path.join(_HERE, LIB_NAME)

F2P_INT16_COMPAT = 1
F2P_MAX_K = 256

_c_f32p = ctypes.c_void_p  # device pointers travel as integers
_c_vp = ctypes.c_void_p


class F2PPairDesc(ctypes.Structure):
    """Mirror of `F2PPairDesc` in include/f2p.h (offsets are in elements from the batch base pointers)."""
    _fields_ = [
        ("evecs_x_off", ctypes.c_int64), ("evecs_y_off", ctypes.c_int64), ("c_off", ctypes.c_int64),
        ("cad_off", ctypes.c_int64), ("pc_off", ctypes.c_int64), ("out_off", ctypes.c_int64),
        ("n1", ctypes.c_int32), ("n2", ctypes.c_int32), ("ldx", ctypes.c_int32), ("ldy", ctypes.c_int32),
        ("ncad", ctypes.c_int32), ("npc", ctypes.c_int32), ("thr", ctypes.c_float * 4),
    ]


class F2PLstsqDesc(ctypes.Structure):
    """Mirror of `F2PLstsqDesc` in include/f2p.h."""
    _fields_ = [
        ("p_off", ctypes.c_int64), ("npos", ctypes.c_int64), ("evecs1_off", ctypes.c_int64),
        ("evecs2_off", ctypes.c_int64), ("n1", ctypes.c_int32), ("n2", ctypes.c_int32),
    ]


F2P_LSTSQ_MAX_K = 64

# every symbol include/f2p.h declares: (restype, argtypes)
_SIGNATURES = {
    "f2p_version": (ctypes.c_int, []),
    "f2p_last_error": (ctypes.c_char_p, []),
    "f2p_device_info": (ctypes.c_int, [ctypes.POINTER(ctypes.c_int)] * 3 + [ctypes.POINTER(ctypes.c_size_t)]),
    "f2p_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(F2PPairDesc), ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "f2p_naive_batched": (ctypes.c_int, [_c_vp, _c_vp, _c_vp, ctypes.POINTER(F2PPairDesc), ctypes.c_int, ctypes.c_int,
                                         ctypes.c_uint, _c_vp, _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_spatial_batched": (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, ctypes.POINTER(F2PPairDesc), ctypes.c_int,
                                           ctypes.c_int, ctypes.c_uint, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp,
                                           ctypes.c_size_t, _c_vp]),
    "f2p_project": (ctypes.c_int, [_c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, _c_vp, _c_vp, _c_vp, ctypes.c_int,
                                   _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_nn_argmin": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, ctypes.c_int, _c_vp, _c_vp, ctypes.c_int, ctypes.c_int,
                                     ctypes.c_int, _c_vp, _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_nn_argmin_static": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, ctypes.c_int, _c_vp, _c_vp, ctypes.c_int, ctypes.c_int,
                                            ctypes.c_int, _c_vp, _c_vp]),
    "f2p_nn_argmin_keys": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, _c_vp, _c_vp, ctypes.c_int,
                                          ctypes.c_int, ctypes.c_int, _c_vp, _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_keys_to_map": (ctypes.c_int, [_c_vp, ctypes.c_int, ctypes.c_uint, _c_vp, _c_vp]),
    "f2p_nn_top5": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, ctypes.c_int, _c_vp, _c_vp, ctypes.c_int, ctypes.c_int,
                                   ctypes.c_int, _c_vp, _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_spatial_filter": (ctypes.c_int, [_c_vp, ctypes.c_int, _c_vp, ctypes.c_int, _c_vp, ctypes.c_int,
                                          ctypes.POINTER(ctypes.c_float), _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp,
                                          ctypes.c_size_t, _c_vp]),
    "f2p_project_workspace_bytes": (ctypes.c_size_t, []),
    "f2p_nn_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "f2p_filter_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int]),
    "f2p_naive_host": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_uint, _c_vp]),
    "f2p_spatial_host": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.c_int, _c_vp, ctypes.c_int,
                                        _c_vp, ctypes.c_int, ctypes.POINTER(ctypes.c_float), ctypes.c_uint, _c_vp,
                                        ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    "f2p_pair_stream_create": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(_c_vp)]),
    "f2p_pair_stream_destroy": (ctypes.c_int, [_c_vp]),
    "f2p_pair_stream_submit_naive": (ctypes.c_int, [_c_vp, _c_vp, _c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.c_int,
                                                    ctypes.c_uint, _c_vp, ctypes.POINTER(ctypes.c_longlong)]),
    "f2p_pair_stream_submit_spatial": (ctypes.c_int, [_c_vp, _c_vp, _c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.c_int,
                                                      _c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.POINTER(ctypes.c_float),
                                                      ctypes.c_uint, _c_vp, _c_vp, _c_vp, ctypes.POINTER(ctypes.c_longlong)]),
    "f2p_pair_stream_flush": (ctypes.c_int, [_c_vp, ctypes.POINTER(ctypes.c_longlong)]),
    "f2p_inlier_ratio_batched": (ctypes.c_int, [_c_vp, _c_vp, _c_vp, ctypes.POINTER(F2PPairDesc), ctypes.c_int, ctypes.c_int,
                                                _c_vp, ctypes.POINTER(ctypes.c_float), _c_vp, _c_vp, _c_vp,
                                                ctypes.c_size_t, _c_vp]),
    "f2p_inlier_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int]),
    "f2p_find_positives_count": (ctypes.c_int, [_c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                                _c_vp, _c_vp, _c_vp, _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_find_positives_write": (ctypes.c_int, [_c_vp, ctypes.c_int, _c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                                _c_vp, ctypes.c_int64, _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_positives_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int]),
    "f2p_c_from_sparse_p_batched": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int64, _c_vp, ctypes.c_int64,
                                                   ctypes.POINTER(F2PLstsqDesc), ctypes.c_int, ctypes.c_int, _c_vp, _c_vp,
                                                   _c_vp, ctypes.c_size_t, _c_vp]),
    "f2p_lstsq_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(F2PLstsqDesc), ctypes.c_int, ctypes.c_int]),
    "f2p_tim_graph_count": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, ctypes.c_double, _c_vp, _c_vp, _c_vp, _c_vp,
                                           ctypes.c_size_t, _c_vp]),
    "f2p_tim_graph_write": (ctypes.c_int, [_c_vp, _c_vp, ctypes.c_int, ctypes.c_double, _c_vp, ctypes.c_int64, _c_vp,
                                           ctypes.c_size_t, _c_vp]),
    "f2p_tim_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int]),
    "f2p_compute_tims": (ctypes.c_int, [_c_vp, ctypes.c_int, _c_vp, _c_vp, _c_vp]),
    "f2p_measure_fp32_peak": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), _c_vp]),
    "f2p_test_vote_band": (ctypes.c_int, [ctypes.c_double]),
    "f2p_measure_tf32_peak": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), _c_vp]),
    "f2p_measure_tmem_ld": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), _c_vp]),
    "f2p_tc_last_stats": (ctypes.c_int, [ctypes.POINTER(ctypes.c_ulonglong)]),
    "f2p_test_sqrt_approx_error": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.POINTER(ctypes.c_double), _c_vp]),
    "f2p_stage_timing": (ctypes.c_int, [ctypes.c_int]),
    "f2p_stage_filter_executed": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
    "f2p_stage_times": (ctypes.c_int, [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
}

_lib = None
_lock = threading.Lock()


def exported_symbols():
    return sorted(_SIGNATURES)


def load():
    """Loads the CUDA library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        f"{LIB_NAME} not found at {LIB_PATH}: build it with `python -c 'import __graft_entry__ as g; "
                        "g.build()'` or `make -C 6d-pose-estimation-for-unseen-categories_b200/csrc`. "
                        "This package has no CPU fallback.")
                L = ctypes.CDLL(LIB_PATH)
                for name, (res, args) in _SIGNATURES.items():
                    fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
                    fn.restype = res
                    fn.argtypes = args
                _lib = L
    return _lib


def check(rc, what):
    if rc == 0:
        return
    msg = load().f2p_last_error().decode("utf-8", "replace")
    if rc in (-1, -3):
        raise ValueError(f"{what}: {msg} (code {rc})")
    raise RuntimeError(f"{what}: {msg} (code {rc})")


def cuda_device_for(*tensors):
    """The CUDA device the call runs on: the first CUDA input's device, else the current device.  Raises when no
    CUDA device exists — the arithmetic only exists as sm_100a kernels."""
    for t in tensors:
        if isinstance(t, torch.Tensor) and t.is_cuda:
            return t.device
    if not torch.cuda.is_available():
        raise RuntimeError("fmap2pointmap_solvers (B200 build) needs a CUDA device: there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


_workspaces = {}


def workspace(device, nbytes):
    """Grow-only workspace per (device, current stream) — include/f2p.h forbids sharing a workspace between concurrent
    calls on different streams, so two torch streams (or a DataLoader thread next to the main thread) get their own.
    On growth the old buffer is handed to the caching allocator with `record_stream`, so kernels of that stream that
    are still running keep their memory until they finish."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    stream = torch.cuda.current_stream(device)
    key = (idx, stream.cuda_stream)
    with _lock:
        ws = _workspaces.get(key)
        if ws is None or ws.numel() < nbytes:
            if ws is not None:
                ws.record_stream(stream)
                del _workspaces[key]
                del ws
            ws = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=device)
            _workspaces[key] = ws
        return ws


def release_workspaces():
    with _lock:
        _workspaces.clear()


def stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def dev_f32(t, device):
    """detach -> device -> float32, keeping a row-major view (unit inner stride) without copying when possible."""
    t = t.detach()
    if t.dtype != torch.float32 or t.device != device:
        t = t.to(device=device, dtype=torch.float32, non_blocking=True)
    return t


def rowmajor_2d(t):
    if t.dim() != 2:
        raise ValueError(f"expected a 2-D tensor, got shape {tuple(t.shape)}")
    if t.shape[0] > 1 and (t.stride(1) != 1 or t.stride(0) < t.shape[1]):
        t = t.contiguous()
    elif t.shape[0] <= 1 and t.shape[1] > 1 and t.stride(1) != 1:
        t = t.contiguous()
    return t
