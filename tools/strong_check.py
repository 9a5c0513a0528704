"""One pair with its CAD rows split over the ranks (distributed.naive_row_sharded) against the unsplit call, under torchrun:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/strong_check.py
Checks bit-identical maps (incl. duplicated rows across the split and a shape where the last rank gets no rows) and prints
the time per pair of the cfg4 shape, exchange included (CUDA events, max over ranks)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from fmap2pointmap_solvers import naive_fmap2pointmap  # noqa: E402
from fmap2pointmap_solvers.distributed import naive_row_sharded  # noqa: E402
from fmap2pointmap_solvers.synth import make_pair  # noqa: E402

rank, local, world = (int(os.environ.get(v, d)) for v, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for n1, n2, k, seed in ((20000, 8000, 64, 1), (5000, 2000, 50, 2), (100, 300, 30, 3), (50000, 20000, 128, 4)):
    P = make_pair(n1, n2, k, seed=seed)
    ex = P.evecs_x.clone()
    if n1 >= 5000:
        ex[n1 - 7] = ex[11]; ex[n1 // 2 + 3] = ex[n1 // 2 - 300]       # exact duplicates across the split
    C, ex, ey = P.C12.to(dev), ex.to(dev), P.evecs_y.to(dev)
    for compat in (False, True):
        got = naive_row_sharded(C, ex, ey, int16_compat=compat)
        want = naive_fmap2pointmap(C, ex, ey, int16_compat=compat)
        ok = ok and bool(torch.equal(got, want))
flag = torch.tensor([0.0 if ok else 1.0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MAX)
P = make_pair(100000, 100000, 128, seed=1234)
C, ex, ey = P.C12.to(dev), P.evecs_x.to(dev), P.evecs_y.to(dev)
for _ in range(3):
    naive_row_sharded(C, ex, ey, grid=(world, 1))
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out = naive_row_sharded(C, ex, ey, grid=(world, 1))
e1.record(); e1.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
for R in (1, 2, 4, 8):          # every grid of this world: which split of rows x scene points is fastest
    if world % R:
        continue
    g = (R, world // R)
    same = bool(torch.equal(naive_row_sharded(C, ex, ey, grid=g), out))
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        naive_row_sharded(C, ex, ey, grid=g)
    e1.record(); e1.synchronize()
    tg = torch.tensor([e0.elapsed_time(e1) / 10, 0.0 if same else 1.0], device=dev)
    dist.all_reduce(tg, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"world {world}: grid {g[0]} row shards x {g[1]} scene shards: {tg[0].item():.3f} ms, identical {tg[1].item() == 0}", flush=True)
from fmap2pointmap_solvers.distributed import _local_keys_cuda, row_cuts  # noqa: E402
lo, hi = row_cuts(100000, world)[rank]          # the all-rows split, for the exchange share
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
e0.record()
for _ in range(10):
    _local_keys_cuda(C, ex[lo:hi], ey, lo, dev)
e1.record(); e1.synchronize()
tl = torch.tensor([e0.elapsed_time(e1) / 10], device=dev)
dist.all_reduce(tl, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"world {world}: local kernels only (projection of the row slice and of all scene points, packing, screen, "
          f"verification, key merge) {tl.item():.3f} ms; exchange + map {t.item() - tl.item():.3f} ms", flush=True)
if rank == 0:
    print(f"world {world}: maps identical on every rank: {flag.item() == 0}; cfg4 pair {t.item():.3f} ms "
          f"({1e10 / (t.item() * 1e-3) / 1e9:.0f} G point-queries/s, exchange included)", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if flag.item() == 0 else 1)
