#!/usr/bin/env python
"""Resident single-pair calls (cfg2: 5k x 2k x 50, naive solver): the command to put under
`ncu --metrics gpu__time_duration.sum` for the per-kernel split of one call.  python tools/pair_launches.py [n1 n2 k] [calls]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch
from fmap2pointmap_solvers import naive_fmap2pointmap
from fmap2pointmap_solvers.synth import make_pair
a = [int(v) for v in sys.argv[1:]]
n1, n2, k = a[:3] if len(a) >= 3 else (5000, 2000, 50)
calls = a[3] if len(a) > 3 else 4
P = make_pair(n1, n2, k, seed=1)
C, ex, ey = P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for it in range(calls):
    flush.fill_(it)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = naive_fmap2pointmap(C, ex, ey); e1.record(); e1.synchronize()
    print(f"call {it}: {1e3 * e0.elapsed_time(e1):.1f} us", flush=True)
