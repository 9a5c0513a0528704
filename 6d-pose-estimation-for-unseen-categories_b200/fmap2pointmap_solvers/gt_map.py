"""Ground-truth functional map of a sample on the device (second half).

`C_from_sparse_P(P, evecs1, evecs2)` keeps the reference's signature and result (`utils/utils.py:67-80`): the float32
`[k,k]` least-squares solution of `evecs2[P[:,1]] @ C = evecs1[P[:,0]]` that `torch.linalg.lstsq(...)[0][:k]` returns
(minimum-norm when the system is rank deficient), on the inputs' device.  The callers slice the bases first
(`cad[:,:30]`, `scripts/eval.py:74`, `scripts/train.py:101`); such column slices are read in place through their row
stride, nothing is copied.  `C_from_sparse_P_batched` solves a whole batch (the `torch.stack([...])` of the same lines)
with one launch pair.  The arithmetic — FP64 Gram accumulation + FP64 solve, rounded once to FP32 — runs in
csrc/lstsq.cu through the C-ABI; there is no CPU path.  Parity with the reference is by tolerance
(tests/test_gpu_parity.py::test_c_from_sparse_p_*).
"""
import torch

from . import _lib


def _basis(t, device):
    """float32 2-D view with unit inner stride on `device` (a column slice keeps its row stride)."""
    t = _lib.dev_f32(t, device)
    if t.dim() != 2:
        raise ValueError(f"eigenbases must be 2-D, got {tuple(t.shape)}")
    if t.shape[1] > 1 and t.stride(1) != 1:
        t = t.contiguous()
    if t.shape[0] > 1 and t.stride(0) < t.shape[1]:
        t = t.contiguous()
    return t


def _positives(P, device):
    P = P.detach()
    if P.dim() != 2 or P.shape[1] != 2:
        raise ValueError(f"P must be [N,2], got {tuple(P.shape)}")
    if P.dtype.is_floating_point or P.dtype == torch.bool:
        raise IndexError("tensors used as indices must be long, int, byte or bool tensors")
    return P.to(device=device, dtype=torch.int64).contiguous()


def C_from_sparse_P_batched(P_list, evecs1_list, evecs2_list, return_status=False):
    """[B,k,k] float32 ground-truth maps of B samples (ragged positives and point counts, shared k).

    P_list[b] int [N_b,2]; evecs1_list[b] [n1_b,k]; evecs2_list[b] [n2_b,k] (views with a row stride are fine).
    `return_status=True` returns (C, status) without a host sync: int32 words on the device, bit0 = rank deficient
    (minimum-norm solution), bit1 = an index of P out of range (that sample's C is NaN)."""
    lib = _lib.load()
    B = len(P_list)
    if not (B == len(evecs1_list) == len(evecs2_list)):
        raise ValueError("P_list, evecs1_list and evecs2_list must have the same length")
    if B == 0:
        raise ValueError("empty batch")
    device = _lib.cuda_device_for(*evecs1_list, *evecs2_list, *P_list)
    with torch.cuda.device(device):
        e1 = [_basis(t, device) for t in evecs1_list]
        e2 = [_basis(t, device) for t in evecs2_list]
        k = int(e1[0].shape[1])
        if any(int(t.shape[1]) != k for t in e1 + e2):
            raise RuntimeError("all eigenbases of a batch must have the same number of columns k")
        if k < 1 or k > _lib.F2P_LSTSQ_MAX_K:
            raise ValueError(f"k={k} outside [1, {_lib.F2P_LSTSQ_MAX_K}]")
        # one row stride per side: re-pack the (rare) odd ones
        ld1 = max(int(t.stride(0)) if t.shape[0] > 1 else k for t in e1)
        ld2 = max(int(t.stride(0)) if t.shape[0] > 1 else k for t in e2)
        e1 = [t if t.shape[0] <= 1 or t.stride(0) == ld1 else _restride(t, ld1) for t in e1]
        e2 = [t if t.shape[0] <= 1 or t.stride(0) == ld2 else _restride(t, ld2) for t in e2]
        Ps = [_positives(P, device) for P in P_list]
        Pcat = torch.cat(Ps, 0) if B > 1 else Ps[0]
        d = (_lib.F2PLstsqDesc * B)()
        base1, base2 = min(t.data_ptr() for t in e1), min(t.data_ptr() for t in e2)
        off = 0
        for b in range(B):
            if e1[b].shape[0] < 1 or e2[b].shape[0] < 1:
                raise IndexError("index out of range: empty eigenbasis")
            d[b].p_off, d[b].npos = 2 * off, int(Ps[b].shape[0])
            d[b].evecs1_off, d[b].evecs2_off = (e1[b].data_ptr() - base1) // 4, (e2[b].data_ptr() - base2) // 4
            d[b].n1, d[b].n2 = int(e1[b].shape[0]), int(e2[b].shape[0])
            off += int(Ps[b].shape[0])
        C = torch.empty((B, k, k), dtype=torch.float32, device=device)
        status = torch.empty(B, dtype=torch.int32, device=device)
        ws = _lib.workspace(device, lib.f2p_lstsq_workspace_bytes(d, B, k))
        rc = lib.f2p_c_from_sparse_p_batched(Pcat.data_ptr() if off else None, base1, ld1, base2, ld2, d, B, k,
                                             C.data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(),
                                             _lib.stream_ptr(device))
        _lib.check(rc, "f2p_c_from_sparse_p_batched")
        if return_status:                      # asynchronous: the caller inspects the words (bit1 = bad index, C = NaN)
            return C, status
        st = status.cpu()                      # host sync: torch raises on a bad index, so this does too
        if bool((st & 2).any()):
            bad = int((st & 2).nonzero()[0][0])
            raise IndexError(f"sample {bad}: an index of P is out of range of its eigenbasis")
        return C


def _restride(t, ld):
    buf = torch.empty((t.shape[0], ld), dtype=t.dtype, device=t.device)
    buf[:, :t.shape[1]] = t
    return buf[:, :t.shape[1]]


def C_from_sparse_P(P, evecs1, evecs2):
    """Drop-in for `utils.C_from_sparse_P` (utils/utils.py:67-80)."""
    C = C_from_sparse_P_batched([P], [evecs1], [evecs2])[0]
    src = evecs1.device if isinstance(evecs1, torch.Tensor) else C.device
    return C if src.type == "cuda" else C.to(src)
