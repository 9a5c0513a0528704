"""BASELINE cfg5 sweep (configs[4]): n1 in {2k,10k,50k} (n2 = 0.4 n1) x k in {30,64,128}, both solvers, as BATCHES of pairs
sharded over the ranks of a torchrun world (weak scaling: every rank converts its own pairs; 1 process = 1 GPU).

    python tools/sweep.py [--cpu]                                                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 tools/sweep.py

Per shape and solver: device-resident and end-to-end (pinned host in, host out, pipelined) pairs/s of the whole job
(CUDA events / wall clock around synchronised regions, max over ranks, L2 flushed between resident steps).  --cpu (rank
0): the reference's own solvers on the host cores next to it — naive on all CAD rows x the first 2048 scene points
(exact by column-chunk invariance, scaled to the pair); spatial on one whole pair where the reference's [M,M,3]
temporaries fit (n1 <= 10k), stated as infeasible above that.  One JSON line per (shape, solver) on rank 0."""
import argparse
import contextlib
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))
import torch  # noqa: E402
from fmap2pointmap_solvers.batched import naive_batched, spatial_batched  # noqa: E402
from fmap2pointmap_solvers.pipeline import naive_host_pipelined, spatial_host_pipelined  # noqa: E402
from fmap2pointmap_solvers.synth import make_pair  # noqa: E402

PAIRS = {2000: 128, 10000: 64, 50000: 4}       # pairs per GPU


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    rank, local, world = (int(os.environ.get(v, d)) for v, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def job_max(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for n1 in (2000, 10000, 50000):
        n2 = int(n1 * 0.4)
        ppg = PAIRS[n1]
        for k in (30, 64, 128):
            base = [make_pair(n1, n2, k, seed=n1 + k + 7 * rank + i) for i in range(2)]
            pairs = [base[i % 2] for i in range(ppg)]
            host = {a: torch.stack([getattr(p, b) for p in pairs]).pin_memory()
                    for a, b in (("C", "C12"), ("ex", "evecs_x"), ("ey", "evecs_y"), ("cad", "CAD"), ("pc", "PC"))}
            devt = {a: v.to(dev) for a, v in host.items()}
            diam = [p.diam_cad for p in pairs]
            for solver in ("naive", "spatial"):
                if solver == "naive":
                    res = lambda: naive_batched(devt["C"], devt["ex"], devt["ey"])                      # noqa: E731
                    e2e = lambda: naive_host_pipelined(host["C"], host["ex"], host["ey"])               # noqa: E731
                else:
                    res = lambda: spatial_batched(devt["C"], devt["ex"], devt["ey"], devt["cad"], devt["pc"], diam,  # noqa: E731
                                                  verbose=False, packed=True)
                    e2e = lambda: spatial_host_pipelined(host["C"], host["ex"], host["ey"], host["cad"], host["pc"], diam)  # noqa: E731
                for _ in range(2):
                    res()
                torch.cuda.synchronize()
                if dist is not None:
                    dist.barrier()
                ms = 0.0
                for _ in range(args.steps):
                    flush.fill_(1)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(); res(); b.record(); b.synchronize()
                    ms += a.elapsed_time(b)
                ms = job_max(ms)
                e2e()
                torch.cuda.synchronize()
                if dist is not None:
                    dist.barrier()
                t0 = time.perf_counter()
                for _ in range(3):
                    e2e()
                torch.cuda.synchronize()
                es = job_max(time.perf_counter() - t0)
                row = {"n1": n1, "n2": n2, "k": k, "solver": solver, "pairs_per_gpu": ppg, "n_gpus": world,
                       "ms_per_step": ms / args.steps, "pairs_per_s": ppg * world * args.steps / (ms * 1e-3),
                       "e2e_pairs_per_s": ppg * world * 3 / es,
                       "g_point_queries_per_s": float(n1) * n2 * ppg * world * args.steps / (ms * 1e-3) / 1e9}
                if args.cpu and rank == 0:
                    row.update(cpu_row(base[0], n1, n2, k, solver))
                if rank == 0:
                    print(json.dumps(row), flush=True)
            del devt, host
            torch.cuda.empty_cache()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def cpu_row(P, n1, n2, k, solver):
    sys.path.insert(0, ROOT)
    import bench
    ref = bench.reference_solvers()
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    if solver == "naive":
        cols = min(2048, n2)
        ey = P.evecs_y[:cols].contiguous()
        ref["naive"](P.C12, P.evecs_x, ey)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); ref["naive"](P.C12, P.evecs_x, ey); ts.append(time.perf_counter() - t0)
        s = sorted(ts)[1] * n2 / cols
        return {"cpu_pairs_per_s": 1.0 / s, "cpu_ms_per_pair": 1e3 * s, "cpu_kind": ref["kind"], "cpu_cores": threads,
                "cpu_sample": f"all {n1} rows x first {cols} of {n2} scene points, median of 3, scaled to the pair"}
    if n1 > 10000:
        M = 5 * n2
        return {"cpu_pairs_per_s": None, "cpu_kind": ref["kind"], "cpu_cores": threads,
                "cpu_sample": f"infeasible on the host: the reference materialises 2 x [M,M,3] + 3 x [M,M] FP32 temporaries per "
                              f"round at M = {M} ({(9 * M * M * 4) / 2**30:.0f} GiB)"}
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        ref["spatial"](P.C12, P.evecs_x, P.evecs_y, CAD=P.CAD, PC=P.PC, diam_cad=P.diam_cad)
    s = time.perf_counter() - t0
    return {"cpu_pairs_per_s": 1.0 / s, "cpu_ms_per_pair": 1e3 * s, "cpu_kind": ref["kind"], "cpu_cores": threads,
            "cpu_sample": "one whole pair, one run"}


if __name__ == "__main__":
    main()
