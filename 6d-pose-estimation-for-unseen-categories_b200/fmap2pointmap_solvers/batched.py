"""Batched host side of the B200 fmap->point-map path: packs a list of object/scene pairs into one C-ABI call
(`f2p_naive_batched` / `f2p_spatial_batched`, include/f2p.h).  The per-sample Python loop of the reference
(`scripts/eval.py:83-91`) becomes one launch sequence over all pairs; single-pair solver calls are batches of one.
"""
import ctypes
import os
from typing import List, Sequence

import torch

from . import _lib

_FACTORS = (0.3, 0.15, 0.055, 0.065)  # spacial_filtering.py:54,58,68-73


def thresholds(diam_cad):
    """The four `c * diam_cad` scalars exactly as the reference's comparison sees them: the product is formed by the
    host language with `diam_cad`'s own type (Python/NumPy float -> FP64 product; FP32 tensor -> FP32 product), then
    torch casts it to the FP32 compare type."""
    import numpy as np
    out = []
    for c in _FACTORS:
        v = c * diam_cad          # the reference's own expression (spacial_filtering.py:54): host-side type rules apply
        if isinstance(v, torch.Tensor):
            if v.numel() != 1:
                raise ValueError("diam_cad must be a scalar")
            out.append(float(v.detach().to(torch.float32).cpu().item()))   # 0-d tensors are cast to the FP32 compare type
        else:
            out.append(float(np.float32(float(v))))                       # wrapped number -> FP32 compare type
    return out


def _squeeze_c(C12):
    # naive.py:17-18 / spacial_filtering.py:16-17
    if C12.dim() == 3:
        C12 = C12.squeeze(0)
    if C12.dim() != 2 or C12.shape[0] != C12.shape[1]:
        raise RuntimeError(f"C12 must be [K,K] or [1,K,K], got {tuple(C12.shape)}")
    return C12


class _Packed:
    """Device views of a batch plus the descriptor array (offsets relative to per-operand base pointers)."""

    def __init__(self, C_list, ex_list, ey_list, device, cad_list=None, pc_list=None, diam_list=None):
        n = len(ex_list)
        if not (len(C_list) == len(ey_list) == n):
            raise ValueError("C12 / evecs_x / evecs_y lists differ in length")
        self.n = n
        self.keep = []  # keeps converted tensors alive until the call has been enqueued
        Cs, exs, eys, cads, pcs = [], [], [], [], []
        k = None
        for i in range(n):
            C = _lib.dev_f32(_squeeze_c(C_list[i]), device).contiguous()
            ex = _lib.rowmajor_2d(_lib.dev_f32(ex_list[i], device))
            ey = _lib.rowmajor_2d(_lib.dev_f32(ey_list[i], device))
            if k is None:
                k = int(C.shape[0])
            if C.shape[0] != k or ex.shape[1] != k or ey.shape[1] != k:
                raise RuntimeError(f"pair {i}: shapes {tuple(ex.shape)} @ {tuple(C.shape)}^T vs {tuple(ey.shape)} "
                                   f"are inconsistent (all pairs of a batch must share K={k})")
            Cs.append(C); exs.append(ex); eys.append(ey)
            if cad_list is not None:
                cad = _lib.dev_f32(cad_list[i], device).contiguous()
                pc = _lib.dev_f32(pc_list[i], device).contiguous()
                if cad.dim() != 2 or cad.shape[1] != 3 or pc.dim() != 2 or pc.shape[1] != 3:
                    raise RuntimeError("CAD / PC must be [N,3]")
                cads.append(cad); pcs.append(pc)
        self.k = k if k is not None else 1
        self.keep = [Cs, exs, eys, cads, pcs]
        dummy = torch.zeros(4, dtype=torch.float32, device=device)
        self.keep.append(dummy)

        def base_of(ts):
            # empty tensors have a null data_ptr: they contribute no base and get offset 0
            return min((t.data_ptr() for t in ts if t.numel() > 0), default=dummy.data_ptr())

        self.C_base, self.ex_base, self.ey_base = base_of(Cs), base_of(exs), base_of(eys)
        self.cad_base, self.pc_base = base_of(cads), base_of(pcs)
        self.descs = (_lib.F2PPairDesc * max(n, 1))()
        self.out_sizes = []
        out_off = 0
        for i in range(n):
            d = self.descs[i]
            ex, ey = exs[i], eys[i]
            d.evecs_x_off = (ex.data_ptr() - self.ex_base) // 4 if ex.numel() else 0
            d.evecs_y_off = (ey.data_ptr() - self.ey_base) // 4 if ey.numel() else 0
            d.c_off = (Cs[i].data_ptr() - self.C_base) // 4
            d.n1, d.n2 = int(ex.shape[0]), int(ey.shape[0])
            d.ldx = int(ex.stride(0)) if ex.shape[0] > 1 else self.k
            d.ldy = int(ey.stride(0)) if ey.shape[0] > 1 else self.k
            if cads:
                d.cad_off = (cads[i].data_ptr() - self.cad_base) // 4 if cads[i].numel() else 0
                d.pc_off = (pcs[i].data_ptr() - self.pc_base) // 4 if pcs[i].numel() else 0
                d.ncad, d.npc = int(cads[i].shape[0]), int(pcs[i].shape[0])
                thr = thresholds(diam_list[i])
                for q in range(4):
                    d.thr[q] = thr[q]
                size = 10 * d.n2
            else:
                d.ncad, d.npc = d.n1, d.n2
                size = 2 * d.n2
            d.out_off = out_off
            self.out_sizes.append(size)
            out_off += size
        self.out_total = out_off


def thresholds_many(diam_cad, n):
    """`thresholds` for n pairs at once ([n,4] float32), same type rules: FP32 tensor -> FP32 products, Python / NumPy
    numbers -> FP64 products cast to FP32."""
    import numpy as np
    if isinstance(diam_cad, torch.Tensor):
        d = diam_cad.detach().reshape(-1)
        if d.numel() == 1:
            d = d.expand(n)
        if d.numel() != n:
            raise ValueError("diam_cad must have one entry per pair")
        return torch.stack([(c * d).to(torch.float32) for c in _FACTORS], 1).cpu().numpy()
    if isinstance(diam_cad, (int, float)) or (hasattr(diam_cad, "dtype") and getattr(diam_cad, "ndim", 1) == 0):
        diam_cad = [diam_cad] * n
    if isinstance(diam_cad, np.ndarray) and diam_cad.dtype in (np.float64, np.float32) and diam_cad.shape == (n,):
        return np.stack([(c * diam_cad).astype(np.float32) for c in _FACTORS], 1)
    if len(diam_cad) != n:
        raise ValueError("diam_cad must have one entry per pair")
    if all(type(v) is float for v in diam_cad):
        d = np.asarray(diam_cad, np.float64)
        return np.stack([(c * d).astype(np.float32) for c in _FACTORS], 1)
    return np.asarray([thresholds(v) for v in diam_cad], np.float32)


def _is_stacked(x):
    return isinstance(x, torch.Tensor) and x.dim() == 3


class _PackedStacked:
    """A uniform batch given as stacked tensors (C [P,k,k] or [k,k] shared, evecs_x [P,n1,k], evecs_y [P,n2,k],
    CAD [P,ncad,3], PC [P,npc,3]): one device view per operand and a vectorised descriptor fill — no per-pair Python
    work, and CPU inputs cost one H2D copy per operand.
    The CAD-side operands may be SHARED by runs of consecutive pairs (many scenes of one object, the usual shape of
    an evaluation set): evecs_x [G,n1,k] and CAD [G,ncad,3] with G dividing P serve pairs i*P/G .. (i+1)*P/G - 1; every
    pair still has its own C, evecs_y and PC."""

    def __init__(self, C12, evecs_x, evecs_y, device, CAD=None, PC=None, diam_cad=None):
        import numpy as np
        ex = _lib.dev_f32(evecs_x, device).contiguous()
        ey = _lib.dev_f32(evecs_y, device).contiguous()
        gx, n1, k = (int(v) for v in ex.shape)
        n, n2 = int(ey.shape[0]), int(ey.shape[1])
        if gx < 1 or n % gx:
            raise RuntimeError(f"stacked batch: evecs_x has {gx} entries for {n} pairs (must divide the pair count)")
        rep_x = n // gx
        C = _lib.dev_f32(C12, device)
        if C.dim() == 2:
            C = C.unsqueeze(0)
        if C.dim() == 4 and C.shape[1] == 1:
            C = C.squeeze(1)
        C = C.contiguous()
        if C.dim() != 3 or C.shape[1] != k or C.shape[2] != k or C.shape[0] not in (1, n) or ey.shape[0] != n or ey.shape[2] != k:
            raise RuntimeError(f"stacked batch: shapes C{tuple(C.shape)} evecs_x{tuple(ex.shape)} evecs_y{tuple(ey.shape)} are inconsistent")
        self.n, self.k, self.n1, self.n2 = n, k, n1, n2
        self.keep = [C, ex, ey]
        self.C_base, self.ex_base, self.ey_base = C.data_ptr(), ex.data_ptr(), ey.data_ptr()
        dt = np.dtype([("evecs_x_off", "<i8"), ("evecs_y_off", "<i8"), ("c_off", "<i8"), ("cad_off", "<i8"),
                       ("pc_off", "<i8"), ("out_off", "<i8"), ("n1", "<i4"), ("n2", "<i4"), ("ldx", "<i4"),
                       ("ldy", "<i4"), ("ncad", "<i4"), ("npc", "<i4"), ("thr", "<f4", (4,))])
        assert dt.itemsize == ctypes.sizeof(_lib.F2PPairDesc)
        d = np.zeros(max(n, 1), dt)
        i = np.arange(max(n, 1), dtype=np.int64)
        d["evecs_x_off"] = (i // rep_x) * (n1 * k); d["evecs_y_off"] = i * (n2 * k)
        d["c_off"] = 0 if C.shape[0] == 1 else i * (k * k)
        d["n1"], d["n2"], d["ldx"], d["ldy"] = n1, n2, k, k
        self.cad_base = self.pc_base = ex.data_ptr()
        if CAD is not None:
            cad = _lib.dev_f32(CAD, device).contiguous()
            pc = _lib.dev_f32(PC, device).contiguous()
            if cad.dim() == 2:
                cad = cad.unsqueeze(0)
            if cad.dim() != 3 or cad.shape[2] != 3 or pc.dim() != 3 or pc.shape[2] != 3 or pc.shape[0] != n or cad.shape[0] < 1 or n % cad.shape[0]:
                raise RuntimeError("stacked batch: CAD must be [P,ncad,3] (or [G,ncad,3] shared by runs of P/G pairs) and PC [P,npc,3]")
            self.keep += [cad, pc]
            self.cad_base, self.pc_base = cad.data_ptr(), pc.data_ptr()
            d["cad_off"] = (i // (n // int(cad.shape[0]))) * (int(cad.shape[1]) * 3)
            d["pc_off"] = i * (int(pc.shape[1]) * 3)
            d["ncad"], d["npc"] = int(cad.shape[1]), int(pc.shape[1])
            d["thr"] = thresholds_many(diam_cad, n)
            size = 10 * n2
        else:
            d["ncad"], d["npc"] = n1, n2
            size = 2 * n2
        d["out_off"] = i * size
        self.desc_np = d
        self.descs = d.ctypes.data_as(ctypes.POINTER(_lib.F2PPairDesc))
        self.size = size
        self.out_total = n * size


# ---- single pair: the reference's real call shape (scripts/eval.py:87, one sample per loop iteration) -----------------
# No lists, no descriptor arrays: one cached ctypes descriptor and workspace size per shape, each operand is its own base.
_single_desc = {}


def _single_entry(lib, spatial, n1, n2, k, ldx, ldy, ncad, npc):
    # the workspace size depends on the kernel-2 path the library will pick (F2P_NN_PATH, include/f2p.h)
    key = (spatial, n1, n2, k, ldx, ldy, ncad, npc, os.environ.get("F2P_NN_PATH"))
    ent = _single_desc.get(key)
    if ent is None:
        d = _lib.F2PPairDesc()
        d.n1, d.n2, d.ldx, d.ldy, d.ncad, d.npc = n1, n2, ldx, ldy, ncad, npc
        nbytes = lib.f2p_workspace_bytes(ctypes.byref(d), 1, k, spatial)
        if len(_single_desc) > 256:
            _single_desc.clear()
        ent = _single_desc[key] = (d, nbytes)
    return ent


def _single_operands(C12, evecs_x, evecs_y, device):
    C = _lib.dev_f32(_squeeze_c(C12), device).contiguous()
    ex = _lib.rowmajor_2d(_lib.dev_f32(evecs_x, device))
    ey = _lib.rowmajor_2d(_lib.dev_f32(evecs_y, device))
    k = int(C.shape[0])
    if ex.shape[1] != k or ey.shape[1] != k:
        raise RuntimeError(f"pair 0: shapes {tuple(ex.shape)} @ {tuple(C.shape)}^T vs {tuple(ey.shape)} "
                           f"are inconsistent (all pairs of a batch must share K={k})")
    return C, ex, ey, k


def _naive_single_host(lib, device, C12, evecs_x, evecs_y, int16_compat):
    """All three operands in host memory: one `f2p_naive_host` call (include/f2p.h) — the library uploads from the
    tensors' own memory into its per-thread staging, runs the kernels and writes the map into a host tensor; no torch
    device tensors, allocations or stream work on the way."""
    def host(t):
        t = t.detach()
        if t.dtype != torch.float32:
            t = t.float()
        return t if t.is_contiguous() else t.contiguous()
    C, ex, ey = host(_squeeze_c(C12)), host(evecs_x), host(evecs_y)
    if ex.dim() != 2 or ey.dim() != 2:
        raise ValueError(f"expected 2-D tensors, got shapes {tuple(ex.shape)} and {tuple(ey.shape)}")
    k = int(C.shape[0])
    if ex.shape[1] != k or ey.shape[1] != k:
        raise RuntimeError(f"pair 0: shapes {tuple(ex.shape)} @ {tuple(C.shape)}^T vs {tuple(ey.shape)} "
                           f"are inconsistent (all pairs of a batch must share K={k})")
    n1, n2 = int(ex.shape[0]), int(ey.shape[0])
    if n1 < 1:
        raise RuntimeError("pair 0: evecs_x has no rows (argmin of an empty dimension)")
    if float(n1) * n2 >= 2e9:
        # one huge pair: its upload takes about as long as its kernels, so it goes through the blocked pipeline (row /
        # column blocks, uploads behind the kernels); the result is the same map
        from .pipeline import naive_host_blocked
        with torch.cuda.device(device):
            return naive_host_blocked(C, ex, ey, int16_compat=int16_compat).clone()
    out = torch.empty((2, n2), dtype=torch.int64)
    if n2 > 0:
        with torch.cuda.device(device):
            rc = lib.f2p_naive_host(C.data_ptr(), ex.data_ptr(), n1, ey.data_ptr(), n2, k,
                                    _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr())
        _lib.check(rc, "f2p_naive_host")
    return out


def naive_single(C12, evecs_x, evecs_y, int16_compat=False):
    """`naive_fmap2pointmap` for ONE pair: int64 [2,V2] on the device of `evecs_x`."""
    lib = _lib.load()
    device = _lib.cuda_device_for(evecs_x, evecs_y, C12)
    if not (evecs_x.is_cuda or evecs_y.is_cuda or C12.is_cuda):
        return _naive_single_host(lib, device, C12, evecs_x, evecs_y, int16_compat)
    with torch.cuda.device(device):
        C, ex, ey, k = _single_operands(C12, evecs_x, evecs_y, device)
        n1, n2 = int(ex.shape[0]), int(ey.shape[0])
        if n1 < 1:
            raise RuntimeError("pair 0: evecs_x has no rows (argmin of an empty dimension)")
        out = torch.empty((2, n2), dtype=torch.int64, device=device)
        if n2 > 0:
            d, nbytes = _single_entry(lib, 0, n1, n2, k, int(ex.stride(0)) if n1 > 1 else k,
                                      int(ey.stride(0)) if n2 > 1 else k, n1, n2)
            if nbytes == 0:
                _lib.check(-1 if k <= _lib.F2P_MAX_K else -3, "f2p_workspace_bytes")
            ws = _lib.workspace(device, nbytes)
            rc = lib.f2p_naive_batched(ex.data_ptr(), ey.data_ptr(), C.data_ptr(), ctypes.byref(d), 1, k,
                                       _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr(), ws.data_ptr(),
                                       ws.numel(), _lib.stream_ptr(device))
            _lib.check(rc, "f2p_naive_batched")
        return out if evecs_x.is_cuda else out.to(evecs_x.device)


def naive_batched(C12: Sequence[torch.Tensor], evecs_x: Sequence[torch.Tensor], evecs_y: Sequence[torch.Tensor],
                  int16_compat: bool = False) -> List[torch.Tensor]:
    """`naive_fmap2pointmap` (naive.py:5-21) for a list of pairs sharing K: one int64 [2,V2] tensor per pair, on the
    device of that pair's `evecs_x` (CPU inputs are copied to the current CUDA device and the result copied back)."""
    lib = _lib.load()
    if _is_stacked(evecs_x):
        if evecs_x.shape[0] == 1 and evecs_y.shape[0] == 1 and C12.dim() in (2, 3):
            C1 = C12 if C12.dim() == 2 else C12[0]
            return naive_single(C1, evecs_x[0], evecs_y[0], int16_compat).unsqueeze(0)
        return _naive_stacked(lib, C12, evecs_x, evecs_y, int16_compat)
    n = len(evecs_x)
    if n == 0:
        return []
    if n == 1 and len(C12) == 1 and len(evecs_y) == 1:
        return [naive_single(C12[0], evecs_x[0], evecs_y[0], int16_compat)]
    device = _lib.cuda_device_for(*evecs_x, *evecs_y, *C12)
    with torch.cuda.device(device):
        pk = _Packed(C12, evecs_x, evecs_y, device)
        for i in range(n):
            if pk.descs[i].n1 < 1:
                raise RuntimeError(f"pair {i}: evecs_x has no rows (argmin of an empty dimension)")
        out = torch.empty(max(pk.out_total, 1), dtype=torch.int64, device=device)
        nbytes = lib.f2p_workspace_bytes(pk.descs, n, pk.k, 0)
        if nbytes == 0:
            _lib.check(-1 if pk.k <= _lib.F2P_MAX_K else -3, "f2p_workspace_bytes")
        ws = _lib.workspace(device, nbytes)
        rc = lib.f2p_naive_batched(pk.ex_base, pk.ey_base, pk.C_base, pk.descs, n, pk.k,
                                   _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr(), ws.data_ptr(),
                                   ws.numel(), _lib.stream_ptr(device))
        _lib.check(rc, "f2p_naive_batched")
        res = []
        for i in range(n):
            d = pk.descs[i]
            r = out[d.out_off:d.out_off + 2 * d.n2].view(2, d.n2)
            src = evecs_x[i]
            res.append(r if src.is_cuda else r.to(src.device))
        return res


def spatial_batched(C12: Sequence[torch.Tensor], evecs_x: Sequence[torch.Tensor], evecs_y: Sequence[torch.Tensor],
                    CAD: Sequence[torch.Tensor], PC: Sequence[torch.Tensor], diam_cad: Sequence,
                    int16_compat: bool = True, return_info: bool = False, verbose: bool = True, packed: bool = False):
    """`spacial_filtering_fmap2pointmap` (spacial_filtering.py:4-75) for a list of pairs sharing K.
    Stacked 3-D tensors ([P,n,k] ...) select the uniform-batch fast path; there `packed=True` returns
    (flat int64 [P, 2*5*V2], counts int32 [P], status int32 [P]) on the device WITHOUT a host sync: pair i's dense
    [2,M'_i] map is `flat[i, :2*counts[i]].view(2, -1)`.
    Returns one int64 [2,M'] tensor per pair (and, with return_info, per-pair dicts with the pre-filter candidates,
    the "skipped" flag and the status word).  The one host sync is the read of the survivor counts."""
    lib = _lib.load()
    if _is_stacked(evecs_x):
        return _spatial_stacked(lib, C12, evecs_x, evecs_y, CAD, PC, diam_cad, int16_compat, return_info, verbose, packed)
    if packed:
        raise ValueError("packed=True needs a uniform batch given as stacked tensors")
    n = len(evecs_x)
    if n == 0:
        return ([], []) if return_info else []
    device = _lib.cuda_device_for(*evecs_x, *evecs_y, *C12, *CAD, *PC)
    with torch.cuda.device(device):
        pk = _Packed(C12, evecs_x, evecs_y, device, CAD, PC, diam_cad)
        out = torch.empty(max(pk.out_total, 1), dtype=torch.int64, device=device)
        cand = torch.empty(max(pk.out_total, 1), dtype=torch.int64, device=device) if return_info else None
        meta = torch.zeros(2 * n, dtype=torch.int32, device=device)
        nbytes = lib.f2p_workspace_bytes(pk.descs, n, pk.k, 1)
        if nbytes == 0:
            # re-run the checks through the real entry point to get the precise message
            rc = lib.f2p_spatial_batched(pk.ex_base, pk.ey_base, pk.C_base, pk.cad_base, pk.pc_base, pk.descs, n, pk.k, 0,
                                         out.data_ptr(), meta.data_ptr(), meta.data_ptr() + 4 * n, None, out.data_ptr(),
                                         0, _lib.stream_ptr(device))
            _lib.check(rc if rc else -1, "f2p_spatial_batched")
        ws = _lib.workspace(device, nbytes)
        rc = lib.f2p_spatial_batched(pk.ex_base, pk.ey_base, pk.C_base, pk.cad_base, pk.pc_base, pk.descs, n, pk.k,
                                     _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr(), meta.data_ptr(),
                                     meta.data_ptr() + 4 * n, cand.data_ptr() if cand is not None else None,
                                     ws.data_ptr(), ws.numel(), _lib.stream_ptr(device))
        _lib.check(rc, "f2p_spatial_batched")
        host = meta.cpu()  # the unavoidable sync: M' sizes the returned tensors (spacial_filtering.py:68 syncs too)
        res, infos = [], []
        for i in range(n):
            d = pk.descs[i]
            m, status = int(host[i]), int(host[n + i])
            if status & 2:
                raise IndexError(f"pair {i}: candidate index out of range of CAD/PC (CAD rows {d.ncad}, PC rows {d.npc})")
            M = 5 * d.n2
            r = out[d.out_off:d.out_off + 2 * m].view(2, m)   # the library writes the survivors dense: a view, no copy
            if status & 1 and verbose:
                print("skipped")  # spacial_filtering.py:70-71
            src = evecs_x[i]
            res.append(r if src.is_cuda else r.to(src.device))
            if return_info:
                c = cand[d.out_off:d.out_off + 2 * M].view(2, M)
                infos.append({"candidates": c if src.is_cuda else c.to(src.device), "skipped": bool(status & 1),
                              "status": status})
        return (res, infos) if return_info else res


def _naive_stacked(lib, C12, evecs_x, evecs_y, int16_compat):
    """Uniform batch as stacked tensors: returns ONE int64 [P,2,V2] tensor (index [i] = pair i's [2,V2] map)."""
    device = _lib.cuda_device_for(evecs_x, evecs_y, C12)
    with torch.cuda.device(device):
        pk = _PackedStacked(C12, evecs_x, evecs_y, device)
        if pk.n == 0 or pk.n2 == 0:
            return torch.empty((pk.n, 2, pk.n2), dtype=torch.int64, device=evecs_x.device)
        if pk.n1 < 1:
            raise RuntimeError("evecs_x has no rows (argmin of an empty dimension)")
        out = torch.empty(pk.out_total, dtype=torch.int64, device=device)
        nbytes = lib.f2p_workspace_bytes(pk.descs, pk.n, pk.k, 0)
        if nbytes == 0:
            _lib.check(-1 if pk.k <= _lib.F2P_MAX_K else -3, "f2p_workspace_bytes")
        ws = _lib.workspace(device, nbytes)
        rc = lib.f2p_naive_batched(pk.ex_base, pk.ey_base, pk.C_base, pk.descs, pk.n, pk.k,
                                   _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr(), ws.data_ptr(),
                                   ws.numel(), _lib.stream_ptr(device))
        _lib.check(rc, "f2p_naive_batched")
        res = out.view(pk.n, 2, pk.n2)
        return res if evecs_x.is_cuda else res.to(evecs_x.device)


def _spatial_stacked(lib, C12, evecs_x, evecs_y, CAD, PC, diam_cad, int16_compat, return_info, verbose, packed=False):
    """Uniform batch as stacked tensors.  Returns the list of per-pair int64 [2,M'] maps; each is a VIEW into one
    flat result buffer (no per-pair copy kernels), on the input device."""
    device = _lib.cuda_device_for(evecs_x, evecs_y, C12, CAD, PC)
    with torch.cuda.device(device):
        pk = _PackedStacked(C12, evecs_x, evecs_y, device, CAD, PC, diam_cad)
        n = pk.n
        if n == 0:
            if packed:
                z = torch.zeros(0, dtype=torch.int32, device=device)
                return torch.empty((0, pk.size), dtype=torch.int64, device=device), z, z
            return ([], []) if return_info else []
        out = torch.empty(max(pk.out_total, 1), dtype=torch.int64, device=device)
        cand = torch.empty(max(pk.out_total, 1), dtype=torch.int64, device=device) if return_info else None
        meta = torch.zeros(2 * n, dtype=torch.int32, device=device)
        nbytes = lib.f2p_workspace_bytes(pk.descs, n, pk.k, 1)
        if nbytes == 0:
            rc = lib.f2p_spatial_batched(pk.ex_base, pk.ey_base, pk.C_base, pk.cad_base, pk.pc_base, pk.descs, n, pk.k, 0,
                                         out.data_ptr(), meta.data_ptr(), meta.data_ptr() + 4 * n, None, out.data_ptr(),
                                         0, _lib.stream_ptr(device))
            _lib.check(rc if rc else -1, "f2p_spatial_batched")
        ws = _lib.workspace(device, nbytes)
        rc = lib.f2p_spatial_batched(pk.ex_base, pk.ey_base, pk.C_base, pk.cad_base, pk.pc_base, pk.descs, n, pk.k,
                                     _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr(), meta.data_ptr(),
                                     meta.data_ptr() + 4 * n, cand.data_ptr() if cand is not None else None,
                                     ws.data_ptr(), ws.numel(), _lib.stream_ptr(device))
        _lib.check(rc, "f2p_spatial_batched")
        if packed:
            return out.view(n, pk.size), meta[:n], meta[n:]
        host = meta.cpu().numpy()     # the one host sync: M' sizes the returned views (spacial_filtering.py:68 syncs too)
        counts, status = host[:n], host[n:]
        if (status & 2).any():
            bad = int((status & 2).nonzero()[0][0])
            raise IndexError(f"pair {bad}: candidate index out of range of CAD/PC")
        on_dev = evecs_x.is_cuda
        flat = out if on_dev else out.to(evecs_x.device)
        size, M = pk.size, 5 * pk.n2
        res, infos = [], []
        for i in range(n):
            m = int(counts[i])
            res.append(flat[i * size:i * size + 2 * m].view(2, m))
            if status[i] & 1 and verbose:
                print("skipped")  # spacial_filtering.py:70-71
        if return_info:
            cflat = cand if on_dev else cand.to(evecs_x.device)
            for i in range(n):
                infos.append({"candidates": cflat[i * size:(i + 1) * size].view(2, M), "skipped": bool(status[i] & 1),
                              "status": int(status[i])})
        return (res, infos) if return_info else res
