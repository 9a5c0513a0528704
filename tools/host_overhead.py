"""Host-side cost of one single-pair call (enqueue time without synchronisation) next to its GPU time, and a cProfile
of the Python part.  Side measurement for the launch-bound cfg2 workload."""
import cProfile
import pstats
import sys
import time

import torch

sys.path.insert(0, "6d-pose-estimation-for-unseen-categories_b200")
from fmap2pointmap_solvers import naive_fmap2pointmap  # noqa: E402
from fmap2pointmap_solvers.batched import naive_batched  # noqa: E402
from fmap2pointmap_solvers.synth import make_pair  # noqa: E402

P = make_pair(5000, 2000, 50, seed=1)
C, ex, ey = P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()
Cs, exs, eys = C.unsqueeze(0), ex.unsqueeze(0), ey.unsqueeze(0)
for fn, name in ((lambda: naive_fmap2pointmap(C, ex, ey), "naive_fmap2pointmap"), (lambda: naive_batched(Cs, exs, eys), "naive_batched[stacked]")):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    N = 300
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(N):
        fn()
    b.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"{name}: enqueue {1e6 * (t1 - t0) / N:.1f} us/call, GPU span {1e3 * a.elapsed_time(b) / N:.1f} us/call")
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    naive_batched(Cs, exs, eys)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
