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
    first upload is the only one no kernel can hide, so it is kept short (a quarter of a group, at most 32 pairs).
    Boundaries are multiples of `align` (the number of consecutive pairs that share a CAD-side operand).
    How many groups: fewer groups cost fewer launch sequences (cfg3 on one GPU: 2 groups 24.8 ms, 4 groups 25.7 ms per 512
    pairs), more groups leave a shorter tail after the last upload when the host link is the limit (eight ranks sharing
    one host); 4 is the default for both."""
    chunks = max(1, min(int(chunks), n))
    step = (n + chunks - 1) // chunks
    step = (step + align - 1) // align * align
    groups = [(lo, min(n, lo + step)) for lo in range(0, n, step)]
    lo, hi = groups[0]
    head = min((hi - lo) // 4, max(32, align)) // align * align       # ... and never more than 32 pairs
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



def naive_host_blocked(C12, evecs_x, evecs_y, xb=2, yb=2, int16_compat=False):
    """`naive_fmap2pointmap` for ONE large host-resident pair with the upload hidden behind the kernels.

    A single pair has no pair groups to pipeline, and its kernels (3.4 ms at 100k x 100k x 128) now take little longer than
    its upload (102 MB, ~2 ms), so the pair is cut into xb row blocks of the CAD basis and yb column blocks of the scene
    basis: block (i, j) is an independent arg-min whose exact (distance, global row) keys are merged into the scene
    block's persistent keys (`f2p_nn_argmin_keys`), so the result is bit-identical to the one-shot call.  Pieces are
    uploaded in the order the blocks need them (x0, y0, x1, ..., y1, ...) on the copy stream; only the first two are
    exposed.  Every piece is projected / transposed once (`f2p_project`).  Returns int64 [2, V2] as a pinned host tensor."""
    import ctypes
    lib = _lib.load()
    C = C12.squeeze(0) if C12.dim() == 3 else C12
    ex = evecs_x.squeeze(0) if evecs_x.dim() == 3 else evecs_x
    ey = evecs_y.squeeze(0) if evecs_y.dim() == 3 else evecs_y
    for t in (C, ex, ey):
        if not isinstance(t, torch.Tensor) or t.is_cuda or t.dim() != 2:
            raise ValueError("naive_host_blocked takes 2-D HOST tensors (one pair)")
    n1, k = (int(v) for v in ex.shape)
    n2 = int(ey.shape[0])
    if C.shape[0] != k or C.shape[1] != k or ey.shape[1] != k:
        raise RuntimeError(f"shapes {tuple(ex.shape)} @ {tuple(C.shape)}^T vs {tuple(ey.shape)} are inconsistent")
    if n1 < 1:
        raise RuntimeError("evecs_x has no rows (argmin of an empty dimension)")
    out_h = _pinned("nvb", (2, n2), torch.int64)
    if n2 == 0:
        return out_h
    device = _device()
    C, ex, ey = C.float().contiguous(), ex.float().contiguous(), ey.float().contiguous()

    def cuts(n, parts):      # block boundaries, multiples of 128
        step = ((n + parts - 1) // parts + 127) // 128 * 128
        return [(lo, min(n, lo + step)) for lo in range(0, n, step)]
    xcut, ycut = cuts(n1, max(1, int(xb))), cuts(n2, max(1, int(yb)))
    with torch.cuda.device(device):
        comp = torch.cuda.current_stream(device)
        h2d, d2h = _copy_streams(device)
        h2d.wait_stream(comp)
        st = _lib.stream_ptr(device)
        ld1 = [((hi - lo) + 127) // 128 * 128 for lo, hi in xcut]
        ld2 = [((hi - lo) + 127) // 128 * 128 for lo, hi in ycut]
        C_d = torch.empty((k, k), dtype=torch.float32, device=device)
        ex_d = torch.empty((n1, k), dtype=torch.float32, device=device)
        ey_d = torch.empty((n2, k), dtype=torch.float32, device=device)
        xT = [torch.empty((k, l), dtype=torch.float32, device=device) for l in ld1]
        nx = [torch.empty((l,), dtype=torch.float32, device=device) for l in ld1]
        yT = [torch.empty((k, l), dtype=torch.float32, device=device) for l in ld2]
        ny = [torch.empty((l,), dtype=torch.float32, device=device) for l in ld2]
        keys = [torch.full((l,), -1, dtype=torch.int64, device=device) for l in ld2]          # every byte 0xFF
        out_d = torch.empty((2, n2), dtype=torch.int64, device=device)
        wsp = _lib.workspace(device, lib.f2p_project_workspace_bytes() +
                             max(lib.f2p_nn_workspace_bytes(b - a, d - c, k, 0) for a, b in xcut for c, d in ycut) + 1024)
        pbytes = (lib.f2p_project_workspace_bytes() + 255) // 256 * 256
        ws_nn = wsp[pbytes:]
        for t in (C_d, ex_d, ey_d):
            t.record_stream(h2d)
        # uploads, in the order the blocks need them
        ev_x, ev_y = [None] * len(xcut), [None] * len(ycut)

        def up(dst, src, lo, hi):
            with torch.cuda.stream(h2d):
                dst[lo:hi].copy_(src[lo:hi], non_blocking=True)
                e = torch.cuda.Event(); e.record(h2d)
            return e
        with torch.cuda.stream(h2d):
            C_d.copy_(C, non_blocking=True)
        order = [("x", 0), ("y", 0)] + [("x", i) for i in range(1, len(xcut))] + [("y", j) for j in range(1, len(ycut))]
        for kind, i in order:
            if kind == "x":
                ev_x[i] = up(ex_d, ex, *xcut[i])
            else:
                ev_y[i] = up(ey_d, ey, *ycut[i])
        done_x, done_y = [False] * len(xcut), [False] * len(ycut)
        for j, (c0, c1) in enumerate(ycut):
            if not done_y[j]:
                comp.wait_event(ev_y[j])
                _lib.check(lib.f2p_project(ey_d[c0:c1].data_ptr(), c1 - c0, k, k, None, yT[j].data_ptr(), ny[j].data_ptr(), ld2[j],
                                           wsp.data_ptr(), pbytes, st), "f2p_project")
                done_y[j] = True
            for i, (r0, r1) in enumerate(xcut):
                if not done_x[i]:
                    comp.wait_event(ev_x[i])
                    _lib.check(lib.f2p_project(ex_d[r0:r1].data_ptr(), r1 - r0, k, k, C_d.data_ptr(), xT[i].data_ptr(),
                                               nx[i].data_ptr(), ld1[i], wsp.data_ptr(), pbytes, st), "f2p_project")
                    done_x[i] = True
                _lib.check(lib.f2p_nn_argmin_keys(xT[i].data_ptr(), nx[i].data_ptr(), r1 - r0, ld1[i], r0, yT[j].data_ptr(),
                                                  ny[j].data_ptr(), c1 - c0, ld2[j], k, keys[j].data_ptr(), ws_nn.data_ptr(),
                                                  ws_nn.numel(), st), "f2p_nn_argmin_keys")
        # keys -> map, block by block (row 1 is the scene point index of the whole pair)
        flags = _lib.F2P_INT16_COMPAT if int16_compat else 0
        tmp = torch.empty((2, max(ld2)), dtype=torch.int64, device=device)
        for j, (c0, c1) in enumerate(ycut):
            w = c1 - c0
            _lib.check(lib.f2p_keys_to_map(keys[j].data_ptr(), w, 0, tmp.data_ptr(), st), "f2p_keys_to_map")
            out_d[0, c0:c1].copy_(tmp.view(-1)[:w])
        idx = torch.arange(n2, dtype=torch.int64, device=device)
        out_d[1].copy_(((idx + 32768) % 65536) - 32768 if int16_compat else idx)      # naive.py:21's int16 wrap, exactly
        ev = torch.cuda.Event(); ev.record(comp)
        out_d.record_stream(d2h)
        with torch.cuda.stream(d2h):
            d2h.wait_event(ev)
            out_h.copy_(out_d, non_blocking=True)
        d2h.synchronize()
    return out_h


def _host_f32(t, what):
    if not isinstance(t, torch.Tensor) or t.is_cuda:
        raise ValueError(f"{what}: the pair streams take HOST tensors")
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class _NativePairStream:
    """ctypes face of F2PPairStream (include/f2p.h): submissions complete in order; `pending` holds, per submission in
    flight, the tensors the C side still reads from or writes to."""

    def __init__(self, depth):
        import collections
        import ctypes
        self.ct = ctypes
        self.lib = _lib.load()
        self.h = ctypes.c_void_p()
        _lib.check(self.lib.f2p_pair_stream_create(int(depth), ctypes.byref(self.h)), "f2p_pair_stream_create")
        self.pending = collections.deque()
        self.completed = 0
        self.popped = 0
        self.done = ctypes.c_longlong(0)

    def ready(self):
        """Submissions whose results have arrived since the last call, oldest first."""
        out = []
        while self.popped < self.done.value:
            out.append(self.pending.popleft())
            self.popped += 1
        return out

    def flush(self):
        _lib.check(self.lib.f2p_pair_stream_flush(self.h, self.ct.byref(self.done)), "f2p_pair_stream_flush")
        return self.ready()

    def close(self):
        if self.h:
            self.lib.f2p_pair_stream_destroy(self.h)
            self.h = self.ct.c_void_p()

    # Creating the object costs streams, events and (on first use) pinned and device allocations — far more than a
    # pair — so finished generators hand their object back and the next stream on this thread / device / depth reuses it.
    _idle = {}

    @classmethod
    def take(cls, device, depth):
        import threading
        key = (device.index, int(depth), threading.get_ident())
        ps = cls._idle.pop(key, None)
        if ps is None:
            ps = cls(depth)
        ps.key = key
        return ps

    def give_back(self, clean):
        """Back to the idle pool after a complete run; an abandoned or failed run is flushed and destroyed."""
        if clean and not self.pending and self.key not in self._idle:
            self._idle[self.key] = self
        else:
            self.close()
            self.pending.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def naive_pair_stream(pairs, depth=3, int16_compat=False):
    """`naive_fmap2pointmap` over a stream of host-resident pairs — the per-sample loop of `scripts/eval.py:83-91` with
    `depth` pairs in flight (F2PPairStream, include/f2p.h): pair i+1 is uploaded and pair i-1 downloaded while pair i
    runs.  `pairs` is any iterable of (C12, evecs_x, evecs_y) host tensors ([K,K] or [1,K,K]; shapes may differ from
    pair to pair); results come back in order, one int64 [2,V2] host tensor per pair, each bit-identical to the single
    call.  Pinned inputs keep the uploads asynchronous.

        for p_pred in naive_pair_stream((s["C12"], s["evecs_x"], s["evecs_y"]) for s in samples): ...
    """
    from .batched import _squeeze_c
    device = _device()
    flags = _lib.F2P_INT16_COMPAT if int16_compat else 0
    with torch.cuda.device(device):
        ps, clean = _NativePairStream.take(device, depth), False
        try:
            for C12, ex, ey in pairs:
                C, ex, ey = _host_f32(_squeeze_c(C12), "C12"), _host_f32(ex, "evecs_x"), _host_f32(ey, "evecs_y")
                if ex.dim() != 2 or ey.dim() != 2 or ex.shape[1] != C.shape[0] or ey.shape[1] != C.shape[0]:
                    raise RuntimeError(f"shapes {tuple(ex.shape)} @ {tuple(C.shape)}^T vs {tuple(ey.shape)} are inconsistent")
                if ex.shape[0] < 1:
                    raise RuntimeError("evecs_x has no rows (argmin of an empty dimension)")
                out = torch.empty((2, ey.shape[0]), dtype=torch.int64)
                _lib.check(ps.lib.f2p_pair_stream_submit_naive(ps.h, C.data_ptr(), ex.data_ptr(), ex.shape[0], ey.data_ptr(),
                                                               ey.shape[0], C.shape[0], flags, out.data_ptr(),
                                                               ps.ct.byref(ps.done)), "f2p_pair_stream_submit_naive")
                ps.pending.append((out, C, ex, ey))
                for item in ps.ready():
                    yield item[0]
            for item in ps.flush():
                yield item[0]
            clean = True
        finally:
            ps.give_back(clean)


def spatial_pair_stream(samples, depth=3, int16_compat=True, verbose=True):
    """`spacial_filtering_fmap2pointmap` over a stream of host-resident samples (C12, evecs_x, evecs_y, CAD, PC,
    diam_cad), `depth` in flight; yields, in order, the int64 [2,M'] host map of every sample, bit-identical to the single
    call — including its side effects: "skipped" is printed and [2,0] returned when the filter leaves nothing
    (spacial_filtering.py:70-71), and a candidate index outside CAD / PC raises IndexError."""
    import ctypes
    from .batched import _squeeze_c, thresholds
    device = _device()
    flags = _lib.F2P_INT16_COMPAT if int16_compat else 0

    def result(item):
        out, meta = item[0], item[1]
        m, status = int(meta[0]), int(meta[1])
        if status & 2:
            raise IndexError(f"candidate index out of range of CAD/PC (CAD rows {item[5].shape[0]}, PC rows {item[6].shape[0]})")
        if status & 1 and verbose:
            print("skipped")
        return out[:2 * m].view(2, m).clone() if m else torch.empty((2, 0), dtype=torch.int64)

    with torch.cuda.device(device):
        ps, clean = _NativePairStream.take(device, depth), False
        try:
            for C12, ex, ey, cad, pc, diam in samples:
                C, ex, ey = _host_f32(_squeeze_c(C12), "C12"), _host_f32(ex, "evecs_x"), _host_f32(ey, "evecs_y")
                cad, pc = _host_f32(cad, "CAD"), _host_f32(pc, "PC")
                if ex.dim() != 2 or ey.dim() != 2 or ex.shape[1] != C.shape[0] or ey.shape[1] != C.shape[0]:
                    raise RuntimeError(f"shapes {tuple(ex.shape)} @ {tuple(C.shape)}^T vs {tuple(ey.shape)} are inconsistent")
                if cad.dim() != 2 or cad.shape[1] != 3 or pc.dim() != 2 or pc.shape[1] != 3:
                    raise RuntimeError("CAD / PC must be [N,3]")
                thr = (ctypes.c_float * 4)(*thresholds(diam))
                out = torch.empty(max(10 * ey.shape[0], 1), dtype=torch.int64)
                meta = torch.zeros(2, dtype=torch.int32)
                _lib.check(ps.lib.f2p_pair_stream_submit_spatial(
                    ps.h, C.data_ptr(), ex.data_ptr(), ex.shape[0], ey.data_ptr(), ey.shape[0], C.shape[0], cad.data_ptr(),
                    cad.shape[0], pc.data_ptr(), pc.shape[0], thr, flags, out.data_ptr(), meta.data_ptr(),
                    meta.data_ptr() + 4, ps.ct.byref(ps.done)), "f2p_pair_stream_submit_spatial")
                ps.pending.append((out, meta, C, ex, ey, cad, pc))
                for item in ps.ready():
                    yield result(item)
            for item in ps.flush():
                yield result(item)
            clean = True
        finally:
            ps.give_back(clean)
