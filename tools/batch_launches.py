#!/usr/bin/env python
"""Resident cfg3 steps only (512 pairs of 5k x 2k x 50, spatial solver, one CAD basis per 16 scenes): the command to put
under `ncu --metrics gpu__time_duration.sum` for a per-kernel split of the batch step.  python tools/batch_launches.py [pairs] [calls]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch
from fmap2pointmap_solvers.batched import spatial_batched
from fmap2pointmap_solvers.synth import make_pair
ppg = int(sys.argv[1]) if len(sys.argv) > 1 else 512
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 3
share, n1, n2, k = 16, 5000, 2000, 50
base = [make_pair(n1, n2, k, seed=1234 + i) for i in range(16)]
objs = [base[i % 16] for i in range(ppg // share)]
pairs = [objs[i // share] for i in range(ppg)]
dev = torch.device("cuda", 0)
C = torch.stack([p.C12 for p in pairs]).to(dev); ex = torch.stack([o.evecs_x for o in objs]).to(dev)
ey = torch.stack([p.evecs_y for p in pairs]).to(dev); cad = torch.stack([o.CAD for o in objs]).to(dev)
pc = torch.stack([p.PC for p in pairs]).to(dev); diam = [p.diam_cad for p in pairs]
for it in range(calls):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    flat, counts, status = spatial_batched(C, ex, ey, cad, pc, diam, verbose=False, packed=True)
    b.record(); b.synchronize()
    print(f"call {it}: {a.elapsed_time(b):.3f} ms, survivors {int(counts.sum())}", flush=True)
