#!/usr/bin/env python
"""Tensor-pipe screen (csrc/tc.cu) against the FP32 gated kernels on the same operands: indices must be identical.
python tools/tc_check.py [n1 n2 k top5]...   (default: a ladder of shapes); prints times and the screen's counters."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch
from fmap2pointmap_solvers import _lib
from fmap2pointmap_solvers.synth import make_pair

L = _lib.load(); dev = torch.device("cuda", 0)
st = _lib.stream_ptr(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def stats():
    a = (ctypes.c_ulonglong * 8)()
    _lib.check(L.f2p_tc_last_stats(a), "stats")
    return list(a)[:4]


def run(n1, n2, k, top5, reps=5):
    P = make_pair(n1, n2, k, seed=99)
    ex, ey, C = P.evecs_x.to(dev), P.evecs_y.to(dev), P.C12.to(dev)
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    xT = torch.empty(k, ld1, device=dev); nx = torch.empty(ld1, device=dev)
    yT = torch.empty(k, ld2, device=dev); ny = torch.empty(ld2, device=dev)
    wsp = torch.empty(L.f2p_project_workspace_bytes() + 256, dtype=torch.uint8, device=dev)
    os.environ["F2P_NN_PATH"] = "tc"
    wsn = torch.empty(L.f2p_nn_workspace_bytes(n1, n2, k, top5) + 256, dtype=torch.uint8, device=dev)
    out = torch.empty(n2 * (5 if top5 else 1), dtype=torch.int64, device=dev)
    _lib.check(L.f2p_project(ex.data_ptr(), n1, k, k, C.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, wsp.data_ptr(), wsp.numel(), st), "p")
    _lib.check(L.f2p_project(ey.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, wsp.data_ptr(), wsp.numel(), st), "p")
    fn = L.f2p_nn_top5 if top5 else L.f2p_nn_argmin
    res = {}
    for path in ("gated", "tc"):
        os.environ["F2P_NN_PATH"] = path
        ts = []
        for it in range(reps):
            flush.fill_(1)
            out.fill_(-7)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(fn(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k, out.data_ptr(), wsn.data_ptr(), wsn.numel(), st), "nn")
            e1.record(); e1.synchronize()
            if it >= 2: ts.append(e0.elapsed_time(e1))
        res[path] = (out.clone(), min(ts))
    same = bool((res["gated"][0] == res["tc"][0]).all())
    nbad = int((res["gated"][0] != res["tc"][0]).sum())
    s = stats()
    fl = 2.0 * n1 * n2 * k
    print(f"{n1}x{n2}x{k} top5={top5}: gated {res['gated'][1]:.4f} ms ({fl/res['gated'][1]/1e9:.1f} TF)  tc {res['tc'][1]:.4f} ms "
          f"({fl/res['tc'][1]/1e9:.1f} TF)  same={same} bad={nbad}  recorded={s[0]} ovf={s[1]} verified={s[2]} "
          f"({s[2]/max(n2,1):.2f}/col) violations={s[3]}", flush=True)
    if not same:
        g, t = res["gated"][0].cpu(), res["tc"][0].cpu()
        idx = (g != t).nonzero().flatten()[:10]
        print("   first mismatches (pos, gated, tc):", [(int(i), int(g[i]), int(t[i])) for i in idx])
    return same


if __name__ == "__main__":
    pk = ctypes.c_double()
    _lib.check(L.f2p_measure_tf32_peak(ctypes.byref(pk), st), "peak")
    print(f"tf32 tensor-pipe peak (M128 N256 K8, smem operands): {pk.value:.1f} TFLOP/s", flush=True)
    bpc, gbs = ctypes.c_double(), ctypes.c_double()
    _lib.check(L.f2p_measure_tmem_ld(ctypes.byref(bpc), ctypes.byref(gbs), st), "tmem")
    print(f"TMEM -> register (tcgen05.ld 32x32b.x32, 8 warps/SM): {bpc.value:.1f} B/clk/SM at the nominal clock, {gbs.value:.0f} GB/s", flush=True)
    a = [int(v) for v in sys.argv[1:]]
    shapes = [tuple(a[i:i + 4]) for i in range(0, len(a), 4)] or [
        (256, 128, 26, 0), (1000, 300, 30, 0), (5000, 2000, 50, 0), (5000, 2000, 50, 1), (10000, 4000, 128, 0),
        (50000, 20000, 64, 0), (50000, 20000, 64, 1), (100000, 100000, 128, 0)]
    ok = True
    for sh in shapes:
        ok = run(*sh) and ok
    print("ALL SAME" if ok else "MISMATCH")
    sys.exit(0 if ok else 1)
