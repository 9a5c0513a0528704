"""Host side of the single-pair call shape (cfg2: 5k x 2k x 50, naive solver; scripts/eval.py:87):
  1. device-resident call: enqueue time (no synchronisation) next to its GPU span;
  2. host tensors in / host tensor out, one synchronous call at a time (f2p_naive_host under naive_fmap2pointmap);
  3. the same pairs as a stream with 1 / 2 / 3 / 6 in flight (naive_pair_stream over F2PPairStream);
  4. a cProfile of the stream's Python part.
python tools/host_overhead.py > profiles/rNN_host_overhead.txt"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch  # noqa: E402
from fmap2pointmap_solvers import naive_fmap2pointmap  # noqa: E402
from fmap2pointmap_solvers.pipeline import naive_pair_stream  # noqa: E402
from fmap2pointmap_solvers.synth import make_pair  # noqa: E402

P = make_pair(5000, 2000, 50, seed=1)
C, ex, ey = P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()
for _ in range(20):
    naive_fmap2pointmap(C, ex, ey)
torch.cuda.synchronize()
N = 300
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
a.record()
for _ in range(N):
    naive_fmap2pointmap(C, ex, ey)
b.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"device-resident call: enqueue {1e6 * (t1 - t0) / N:.1f} us/call, GPU span {1e3 * a.elapsed_time(b) / N:.1f} us/call (L2 warm)")

for pinned in (True, False):
    Ch, exh, eyh = (t.pin_memory() if pinned else t for t in (P.C12, P.evecs_x, P.evecs_y))
    for _ in range(20):
        naive_fmap2pointmap(Ch, exh, eyh)
    t0 = time.perf_counter()
    for _ in range(N):
        naive_fmap2pointmap(Ch, exh, eyh)
    dt = (time.perf_counter() - t0) / N
    print(f"host tensors ({'pinned' if pinned else 'pageable'}), synchronous call: {1e6 * dt:.1f} us/pair = {1 / dt:.0f} pairs/s")
    feed = lambda m: ((Ch, exh, eyh) for _ in range(m))      # noqa: E731
    for depth in (1, 2, 3, 6):
        for _ in naive_pair_stream(feed(32), depth=depth):
            pass
        t0 = time.perf_counter()
        n = sum(1 for _ in naive_pair_stream(feed(1024), depth=depth))
        dt = (time.perf_counter() - t0) / n
        print(f"host tensors ({'pinned' if pinned else 'pageable'}), stream with {depth} in flight: {1e6 * dt:.1f} us/pair = {1 / dt:.0f} pairs/s")

Ch, exh, eyh = P.C12.pin_memory(), P.evecs_x.pin_memory(), P.evecs_y.pin_memory()
pr = cProfile.Profile()
pr.enable()
for _ in naive_pair_stream(((Ch, exh, eyh) for _ in range(512)), depth=3):
    pass
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(10)
