"""Host-resident batches through the solvers with the copies hidden behind the kernels.

`scripts/eval.py:62-87` receives every batch from the DataLoader in host memory and moves it to the device before the
solver loop.  For a uniform batch given as STACKED host tensors these entry points cut the batch into pair groups and
run three stream-ordered stages per group — H2D of the group's operands (copy stream), the solver kernels (the
caller's current stream), D2H of the group's result slots (a second copy stream) — so that group g+1 is uploaded and
group g-1 downloaded while group g computes.  Pairs are independent (no exchange between groups), so the result is
bit-identical to the one-shot call.  Inputs should be pinned (`tensor.pin_memory()`): pageable memory still works, but
the copies then serialise with the host.

    flat, counts, status = spatial_host_pipelined(C, evecs_x, evecs_y, CAD, PC, diam)    # host tensors in and out
    maps                 = naive_host_pipelined(C, evecs_x, evecs_y)                     # int64 [P,2,V2] on the host

The returned host tensors are pinned buffers cached per shape: consume (or clone) them before the next call of the
same shape.  All arithmetic runs in the CUDA library; there is no CPU path.
"""
import torch

from . import _lib
from .batched import naive_batched, spatial_batched

_streams = {}
_host_out = {}


def _copy_streams(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _streams:
        _streams[key] = (torch.cuda.Stream(device), torch.cuda.Stream(device))
    return _streams[key]


def _pinned(tag, shape, dtype):
    key = (tag, tuple(shape), dtype)
    buf = _host_out.get(key)
    if buf is None:
        buf = torch.empty(shape, dtype=dtype, pin_memory=True)
        _host_out[key] = buf
    return buf


def _groups(n, chunks, align=1):
    """Pair groups [lo, hi): `chunks` equal groups, the first one split again into a quarter and three quarters — the
    first upload is the only one no kernel can hide, so it is kept short (a quarter of a group).  Boundaries are
    multiples of `align` (the number of consecutive pairs that share a CAD-side operand)."""
    chunks = max(1, min(int(chunks), n))
    step = (n + chunks - 1) // chunks
    step = (step + align - 1) // align * align
    groups = [(lo, min(n, lo + step)) for lo in range(0, n, step)]
    lo, hi = groups[0]
    head = (hi - lo) // 4 // align * align
    if len(groups) > 1 and head >= 8:
        groups[0:1] = [(lo, lo + head), (lo + head, hi)]
    return groups


def _check_host(*ts):
    for t in ts:
        if not isinstance(t, torch.Tensor) or t.is_cuda or t.dim() != 3:
            raise ValueError("the pipelined entry points take stacked 3-D HOST tensors ([P,n,k] ...)")


def _share_align(n, *ops):
    """Pairs per shared entry of the most widely shared operand (1 when every pair has its own operands)."""
    align = 1
    for t in ops:
        g = int(t.shape[0])
        if g < 1 or n % g:
            raise ValueError(f"an operand with {g} entries cannot be shared by {n} pairs (must divide the pair count)")
        if g > 1 or n == 1:
            align = max(align, n // g)
    return align


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("fmap2pointmap_solvers (B200 build) needs a CUDA device: there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _sl(t, lo, hi, n):
    """Pair-group slice of a stacked operand with t.shape[0] entries for n pairs (an entry serves n / t.shape[0]
    consecutive pairs; group boundaries are aligned to that, see _groups)."""
    rep = n // t.shape[0]
    return t[lo // rep:(hi + rep - 1) // rep]


def _upload(h2d, comp, device, host_ops, lo, hi, dev_ops, n, wait=True):
    """Enqueues the H2D copies of pair group [lo, hi) on the copy stream; returns the event that marks their end (and
    makes the compute stream wait for it unless wait=False)."""
    with torch.cuda.stream(h2d):
        for src, dst in zip(host_ops, dev_ops):
            if src.shape[0] == 1 and lo > 0:
                continue                      # shared by the whole batch: uploaded with the first group
            rep = n // src.shape[0]
            a, b = lo // rep, (hi + rep - 1) // rep
            dst[a:b].copy_(src[a:b], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(h2d)
    if wait:
        comp.wait_event(ev)
    return ev


def spatial_host_pipelined(C12, evecs_x, evecs_y, CAD, PC, diam_cad, chunks=4, int16_compat=True, prefix_only=True):
    """(flat int64 [P, 2*5*V2], counts int32 [P], status int32 [P]) as pinned HOST tensors; pair i's dense [2,M'_i] map
    is `flat[i, :2*counts[i]].view(2, -1)` (the layout of `spatial_batched(..., packed=True)`).  Only those prefixes are
    copied back: the slots behind them are unspecified."""
    _lib.load()
    _check_host(C12 if C12.dim() == 3 else C12.reshape(-1, C12.shape[-2], C12.shape[-1]), evecs_x, evecs_y, CAD, PC)
    device = _device()
    n, n2 = int(evecs_y.shape[0]), int(evecs_y.shape[1])
    if len(diam_cad) != n:
        raise ValueError("diam_cad must have one entry per pair")
    align = _share_align(n, C12 if C12.dim() == 3 else C12.reshape(-1, C12.shape[-2], C12.shape[-1]), evecs_x, CAD, PC)
    flat_h = _pinned("sp_flat", (n, 10 * n2), torch.int64)
    meta_h = _pinned("sp_meta", (2, n), torch.int32)
    if n == 0:
        return flat_h, meta_h[0], meta_h[1]
    with torch.cuda.device(device):
        comp = torch.cuda.current_stream(device)
        h2d, d2h = _copy_streams(device)
        h2d.wait_stream(comp)                    # the staging tensors below may reuse memory of work still in flight
        C3 = C12 if C12.dim() == 3 else C12.reshape(-1, C12.shape[-2], C12.shape[-1])
        host_ops = [C3.float(), evecs_x.float(), evecs_y.float(), CAD.float(), PC.float()]   # no-ops for float32
        dev_ops = [torch.empty(t.shape, dtype=torch.float32, device=device) for t in host_ops]
        for t in dev_ops:
            t.record_stream(h2d)
        diam = list(diam_cad)
        pending = []        # (lo, hi, flat, event after the group's counts reached the host), oldest first

        def drain(p):
            # the result buffer is sized for the worst case (10*V2 slots per pair) and nearly empty behind the dense
            # [2,M'] prefixes: once the group's counts are on the host, only the first 2*max(M') slots are copied back
            plo, phi, pflat, pev = p
            if prefix_only:
                pev.synchronize()
                width = min(int(pflat.shape[1]), 2 * int(meta_h[0, plo:phi].max())) if phi > plo else 0
            else:
                width = int(pflat.shape[1])
            if width > 0:
                with torch.cuda.stream(d2h):
                    flat_h[plo:phi, :width].copy_(pflat[:, :width], non_blocking=True)

        groups = _groups(n, chunks, align)
        # every group's upload is queued up front (the copy stream runs them back to back); the host may then block on a
        # group's survivor counts without delaying any upload
        up = [_upload(h2d, comp, device, host_ops, lo, hi, dev_ops, n, wait=False) for lo, hi in groups]
        for gi, (lo, hi) in enumerate(groups):
            comp.wait_event(up[gi])
            flat, counts, status = spatial_batched(*[_sl(t, lo, hi, n) for t in dev_ops], diam[lo:hi], int16_compat=int16_compat,
                                                   verbose=False, packed=True)
            ev = torch.cuda.Event()
            ev.record(comp)
            for t in (flat, counts, status):
                t.record_stream(d2h)
            with torch.cuda.stream(d2h):
                d2h.wait_event(ev)
                meta_h[0, lo:hi].copy_(counts, non_blocking=True)
                meta_h[1, lo:hi].copy_(status, non_blocking=True)
                cev = torch.cuda.Event()
                cev.record(d2h)
            # maps are fetched two groups behind: that group's counts are on the host by now (the group in between is
            # still running), so the host does not block while there is work left to queue
            pending.append((lo, hi, flat, cev))
            if len(pending) > 2:
                drain(pending.pop(0))
        for p in pending:
            drain(p)
        d2h.synchronize()
    if bool((meta_h[1] & 2).any()):
        bad = int((meta_h[1] & 2).nonzero()[0][0])
        raise IndexError(f"pair {bad}: candidate index out of range of CAD/PC")
    return flat_h, meta_h[0], meta_h[1]


def naive_host_pipelined(C12, evecs_x, evecs_y, chunks=4, int16_compat=False):
    """int64 [P,2,V2] index maps as one pinned HOST tensor."""
    _lib.load()
    C3 = C12 if C12.dim() == 3 else C12.reshape(-1, C12.shape[-2], C12.shape[-1])
    _check_host(C3, evecs_x, evecs_y)
    device = _device()
    n, n2 = int(evecs_y.shape[0]), int(evecs_y.shape[1])
    align = _share_align(n, C3, evecs_x)
    out_h = _pinned("nv", (n, 2, n2), torch.int64)
    if n == 0 or n2 == 0:
        return out_h
    with torch.cuda.device(device):
        comp = torch.cuda.current_stream(device)
        h2d, d2h = _copy_streams(device)
        h2d.wait_stream(comp)
        host_ops = [C3.float(), evecs_x.float(), evecs_y.float()]
        dev_ops = [torch.empty(t.shape, dtype=torch.float32, device=device) for t in host_ops]
        for t in dev_ops:
            t.record_stream(h2d)
        for lo, hi in _groups(n, chunks, align):
            _upload(h2d, comp, device, host_ops, lo, hi, dev_ops, n)
            maps = naive_batched(*[_sl(t, lo, hi, n) for t in dev_ops], int16_compat=int16_compat)
            ev = torch.cuda.Event()
            ev.record(comp)
            maps.record_stream(d2h)
            with torch.cuda.stream(d2h):
                d2h.wait_event(ev)
                out_h[lo:hi].copy_(maps, non_blocking=True)
        d2h.synchronize()
    return out_h

