#!/usr/bin/env python
"""Times kernel 2 alone (f2p_nn_argmin / f2p_nn_top5) under experiment switches: python tools/nn_time.py n1 n2 k top5 VAR..."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch
from fmap2pointmap_solvers import _lib
from fmap2pointmap_solvers.synth import make_pair
n1, n2, k, top5 = (int(v) for v in sys.argv[1:5])
variants = sys.argv[5:] or ["0"]
L = _lib.load(); dev = torch.device("cuda", 0)
P = make_pair(n1, n2, k, seed=99)
ex, ey, C = P.evecs_x.to(dev), P.evecs_y.to(dev), P.C12.to(dev)
ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
xT = torch.empty(k, ld1, device=dev); nx = torch.empty(ld1, device=dev)
yT = torch.empty(k, ld2, device=dev); ny = torch.empty(ld2, device=dev)
wsp = torch.empty(L.f2p_project_workspace_bytes() + 256, dtype=torch.uint8, device=dev)
wsn = torch.empty(L.f2p_nn_workspace_bytes(n1, n2, k, top5) + 256, dtype=torch.uint8, device=dev)
out = torch.empty(n2 * (5 if top5 else 1), dtype=torch.int64, device=dev)
st = _lib.stream_ptr(dev)
_lib.check(L.f2p_project(ex.data_ptr(), n1, k, k, C.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, wsp.data_ptr(), wsp.numel(), st), "p")
_lib.check(L.f2p_project(ey.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, wsp.data_ptr(), wsp.numel(), st), "p")
fn = L.f2p_nn_top5 if top5 else L.f2p_nn_argmin
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ref = None
for v in variants:
    if "=" in v:
        for kk in list(os.environ):
            if kk.startswith("F2P_"): del os.environ[kk]
        kk, vv = v.split("=", 1); os.environ[kk] = vv; tag = v
    else:
        for kk in list(os.environ):
            if kk.startswith("F2P_"): del os.environ[kk]
        tag = "default"
    ts = []
    for it in range(8):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(fn(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k, out.data_ptr(), wsn.data_ptr(), wsn.numel(), st), "nn")
        e1.record(); e1.synchronize()
        if it >= 3: ts.append(e0.elapsed_time(e1))
    o = out.clone()
    same = True if ref is None else bool((o == ref).all())
    if ref is None: ref = o
    ms = sum(ts) / len(ts)
    print(f"{tag:28s} {ms:9.4f} ms  {2.0*n1*n2*k/ms/1e9:7.2f} TFLOP/s  same={same}", flush=True)
