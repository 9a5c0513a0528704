#!/usr/bin/env python
"""bench.py — fmap->point-map throughput on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload large|pair|batch] [--impl reference]

Workloads (BASELINE.json configs):
    large  cfg4: one pair n1=n2=100k, k=128, naive solver   (default; the config the roofline target is quoted on)
    pair   cfg2: one pair 5k x 2k, k=50, naive solver
    batch  cfg3: pairs of 5k x 2k, k=50 with spatial-consistency filtering, 512 pairs per GPU (8 GPUs = the 4096 pairs
           of configs[2]); one CAD eigenbasis / model per 16 scenes, every pair its own C, evecs_y, PC
One "step" = one pass of the hot path over the rank's pair(s).  Multi-GPU (torchrun): every rank converts its own
pair slice (weak scaling) and the index maps are returned to every rank with an NCCL all-gather per step (int32 on the
wire, issued on a side stream so that it overlaps the next step).

Prints ONE JSON line (rank 0).  `value` = device-resident throughput of the headline workload; `e2e` = the same through
the public Python API from pinned HOST buffers including H2D/D2H; `roofline` = kernel 2 (tensor-pipe screen + exact
verification) against the measured TF32 peak; `cpu_baseline` = the reference's own solvers on the host cores (bounded
sample); `other_workloads` = cfg2 and cfg3 measured in the same process group (resident and e2e pairs/s, step split).
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))

import torch  # noqa: E402

NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.5: 148 SMs x 128 lanes x 2 flop x 1.965 GHz
NOMINAL_TF32_TFLOPS = 148 * 2048 * 2 * 1.965e9 / 1e12  # 1191: 148 SMs x 2048 TF32 MAC/clk x 2 flop x 1.965 GHz

WORKLOADS = {
    "large": dict(n1=100_000, n2=100_000, k=128, solver="naive", pairs_per_gpu=1, baseline_cols=4096, share=1),
    "pair": dict(n1=5_000, n2=2_000, k=50, solver="naive", pairs_per_gpu=1, baseline_cols=2000, share=1),
    "batch": dict(n1=5_000, n2=2_000, k=50, solver="spatial", pairs_per_gpu=512, baseline_cols=2000, share=16),
}
METRIC = "fmap->p2p G point-queries/s"
UNIT = "G point-queries/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="large", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pairs-per-gpu", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the cfg2 / cfg3 side measurements of the default run")
    return ap.parse_args()


def make_pairs(n1, n2, k, seed0, count):
    from fmap2pointmap_solvers.synth import make_pair
    return [make_pair(n1, n2, k, seed=seed0 + i) for i in range(count)]


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        clocks, mx, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                clocks.append(float(f[0])); mx = float(f[1]); power.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v == "Active":
                    reasons.add(nm)
        clocks.sort()
        load = clocks[len(clocks) // 2:] if clocks else []   # upper half: idle samples before/after are excluded
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(clocks),
                "power_w_max": max(power) if power else None}


# ------------------------------------------------------------------------------------------------ CPU arm
_REF = {}


def reference_solvers():
    """The reference's own solver module (imported UNMODIFIED from /root/reference in the build container, or from the
    copy `__graft_entry__.build()` stages into git-ignored oracle/_ref/ for the GPU box) -> kind "reference";
    the ATen op-for-op port (oracle/aten_port.py) when neither exists -> kind "port"."""
    if not _REF:
        from oracle import ref_loader
        if ref_loader.reference_available():
            mod = ref_loader.load_reference()
            _REF.update(kind="reference", naive=mod.naive_fmap2pointmap, spatial=mod.spacial_filtering_fmap2pointmap,
                        where=ref_loader.reference_root())
        else:
            from oracle import aten_port
            _REF.update(kind="port", naive=lambda C, ex, ey: aten_port.naive(C, ex, ey),
                        spatial=lambda C, ex, ey, CAD, PC, diam_cad: aten_port.spatial(C, ex, ey, CAD, PC, diam_cad),
                        where="oracle/aten_port.py")
    return _REF


def cpu_reference_sample(wl, P, threads, reps):
    """Times the reference solver on a bounded sample of the workload on `threads` host threads.
    naive: all V1 rows against the first `baseline_cols` scene points (cdist is bitwise invariant to column chunking,
    SURVEY 8(a).1-3, and its cost is linear in the column count).  spatial: one whole pair.
    Returns (G point-queries/s of the median run, median seconds, sample text, queries of the sample, last result)."""
    import contextlib
    import io
    ref = reference_solvers()
    torch.set_num_threads(threads)
    cols = min(wl["baseline_cols"], wl["n2"])
    times, out = [], None
    if wl["solver"] == "naive":
        ey = P.evecs_y[:cols].contiguous()
        for _ in range(reps):
            t0 = time.perf_counter()
            out = ref["naive"](P.C12, P.evecs_x, ey)
            times.append(time.perf_counter() - t0)
        queries = wl["n1"] * cols
        sample = f"all {wl['n1']} CAD rows x first {cols} of {wl['n2']} scene points, k={wl['k']}, median of {reps}"
    else:
        for _ in range(reps):
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                out = ref["spatial"](P.C12, P.evecs_x, P.evecs_y, CAD=P.CAD, PC=P.PC, diam_cad=P.diam_cad)
            times.append(time.perf_counter() - t0)
        queries = wl["n1"] * wl["n2"]
        sample = f"1 whole pair {wl['n1']}x{wl['n2']}, k={wl['k']}, spatial filter included, median of {reps}"
    med = statistics.median(times)
    return queries / med / 1e9, med, sample, queries, out


def cpu_baseline_block(wl, P, reps_all, reps_one):
    """All-core and 1-thread figures of the reference solver on the bounded sample."""
    threads = os.cpu_count() or 1
    cpu_reference_sample(wl, P, threads, 1)                       # warm-up (MKL thread pool, page faults)
    v, s, sample, queries, out = cpu_reference_sample(wl, P, threads, reps_all)
    v1, s1, _, _, _ = cpu_reference_sample(wl, P, 1, reps_one)
    torch.set_num_threads(threads)
    full = float(wl["n1"]) * wl["n2"]
    ref = reference_solvers()
    return {"value": v, "unit": UNIT, "cores": threads, "kind": ref["kind"], "sample": sample, "seconds": s,
            "seconds_extrapolated_to_one_pair": s * full / queries,
            "one_thread": {"value": v1, "seconds": s1, "runs": reps_one},
            "note": "%s (%s) on the box's host cores" % (
                "the reference's own fmap2pointmap_solvers, imported unmodified" if ref["kind"] == "reference"
                else "ATen op-for-op port of the reference", ref["where"])}, out


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    P = make_pairs(wl["n1"], wl["n2"], wl["k"], 1234, 1)[0]
    threads = os.cpu_count() or 1
    spatial = wl["solver"] == "spatial"
    for _ in range(max(args.warmup, 1) if not spatial else 1):
        cpu_reference_sample(wl, P, threads, 1)
    secs, queries, sample = [], 0, ""
    for _ in range(args.steps if not spatial else min(args.steps, 3)):
        _, s, sample, queries, _ = cpu_reference_sample(wl, P, threads, 1)
        secs.append(s)
    steps = len(secs)
    value = queries * steps / sum(secs) / 1e9
    _, s1, _, _, _ = cpu_reference_sample(wl, P, 1, 1)
    torch.set_num_threads(threads)
    full = float(wl["n1"]) * wl["n2"]
    ref = reference_solvers()
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / steps,
        "ms_per_step_extrapolated": 1e3 * sum(secs) / steps * full / queries,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, wl, args.pairs_per_gpu or wl["pairs_per_gpu"], args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": ref["kind"],
                         "sample": sample.replace("median of 1", "one run per step"),
                         "one_thread": {"value": queries / s1 / 1e9, "seconds": s1, "runs": 1},
                         "note": "ms_per_step is the time of the SAMPLE; ms_per_step_extrapolated scales it to one whole "
                                 "step of the workload (the cost is linear in the scene-point count)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(name, wl, ppg, n_gpus):
    cfg = {"workload": f"{name}: {ppg} pair(s)/GPU of n1={wl['n1']} x n2={wl['n2']}, k={wl['k']}, {wl['solver']} solver",
           "n1": wl["n1"], "n2": wl["n2"], "k": wl["k"], "solver": wl["solver"], "pairs_per_gpu": ppg,
           "pairs_total": ppg * n_gpus, "parallelism": f"pair-sharded x{n_gpus}",
           "l2": "flushed between timed steps (256 MiB write)"}
    if wl.get("share", 1) > 1 and ppg >= wl["share"]:
        cfg["cad_sharing"] = wl["share"]
    return cfg


# ------------------------------------------------------------------------------------------------ GPU arm
class Ctx:
    pass


def gpu_context():
    from fmap2pointmap_solvers import _lib
    c = Ctx()
    c.rank = int(os.environ.get("RANK", "0"))
    c.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    c.world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback); use --impl reference for the CPU arm")
    c.pinned_cpus = pin_rank_to_its_numa_node(c.local_rank, c.world)
    torch.cuda.set_device(c.local_rank)
    c.dev = torch.device("cuda", c.local_rank)
    c._lib = _lib
    c.L = _lib.load()
    c.dist = None
    if c.world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=c.dev)
        c.dist = dist
    c.comm = torch.cuda.Stream(c.dev)
    c.flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=c.dev)
    return c


def pin_rank_to_its_numa_node(local_rank, world):
    """Multi-GPU runs: bind the rank (and with it the pinned buffers it allocates afterwards: first touch) to the CPUs the
    kernel lists as local to its GPU (/sys/bus/pci/devices/<bus id>/local_cpulist).  Best effort: a box that reports
    every CPU as local to every GPU (this pool does) is left untouched, and so is a single-GPU run."""
    if world <= 1:
        return None
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        txt = open(f"/sys/bus/pci/devices/{bus}/local_cpulist").read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-"); cpus.update(range(int(a), int(b) + 1))
            elif part.strip().isdigit():
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if len(cpus) >= 2 and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            return sorted(cpus)
    except Exception:
        pass
    return None


def measure_workload(ctx, name, wl, ppg, steps, warmup, e2e_steps):
    """Resident and end-to-end throughput of one workload on this process group."""
    from fmap2pointmap_solvers import naive_fmap2pointmap
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.pipeline import naive_host_blocked, naive_host_pipelined, spatial_host_pipelined
    dev, dist, world, rank = ctx.dev, ctx.dist, ctx.world, ctx.rank
    n1, n2, k = wl["n1"], wl["n2"], wl["k"]
    spatial = wl["solver"] == "spatial"
    share = wl.get("share", 1) if ppg >= wl.get("share", 1) and ppg % wl.get("share", 1) == 0 else 1
    # distinct objects per rank (seed = 1234 + global index); a batch cycles 16 distinct objects; with sharing, every
    # object (CAD eigenbasis + model) serves `share` consecutive scenes that differ in C, evecs_y and PC
    nobj = ppg // share
    distinct = min(nobj, 16)
    base = make_pairs(n1, n2, k, 1234 + rank * nobj, distinct)
    objs = [base[i % distinct] for i in range(nobj)]
    pairs = [objs[i // share] for i in range(ppg)]
    host = {"C": torch.stack([p.C12 for p in pairs], 0), "ex": torch.stack([o.evecs_x for o in objs], 0),
            "ey": torch.stack([p.evecs_y for p in pairs], 0)}
    if spatial:
        host["cad"] = torch.stack([o.CAD for o in objs], 0)
        host["pc"] = torch.stack([p.PC for p in pairs], 0)
    if share > 1:
        # scenes of one object: own functional map (a small rotation of the object's), own eigenbasis / cloud noise
        g = torch.Generator().manual_seed(99 + rank)
        host["ey"] = host["ey"] + (0.02 / n1 ** 0.5) * torch.randn(host["ey"].shape, generator=g)
        if spatial:
            host["pc"] = host["pc"] + 0.02 * torch.randn(host["pc"].shape, generator=g)
    host = {kk: v.contiguous().pin_memory() for kk, v in host.items()}
    devt = {kk: v.to(dev) for kk, v in host.items()}
    diam = [p.diam_cad for p in pairs]

    def step_resident():
        # device-resident, asynchronous form of the call: spatial returns (flat maps, counts, status) with no host sync
        if spatial:
            return spatial_batched(devt["C"], devt["ex"], devt["ey"], devt["cad"], devt["pc"], diam, verbose=False, packed=True)
        return naive_batched(devt["C"], devt["ex"], devt["ey"])

    gbuf = {}

    def gather_maps(out):
        """Index maps to every rank.  naive: int32 on the wire, on the side stream (overlaps the next step's kernels).
        spatial: counts first, then only the used prefix of the result slots (allgather_packed_maps; its read of the
        largest count is a host sync, so it stays on the compute stream)."""
        if dist is None:
            return None
        if spatial:
            from fmap2pointmap_solvers.distributed import allgather_packed_maps
            return allgather_packed_maps(out[0], out[1])
        comp = torch.cuda.current_stream(dev)
        key = tuple(out.shape)
        if key not in gbuf:       # wire / gathered / widened buffers are allocated once per workload, not per step
            gbuf[key] = (torch.empty(key, dtype=torch.int32, device=dev),
                         torch.empty((world,) + key, dtype=torch.int32, device=dev),
                         torch.empty((world,) + key, dtype=torch.int64, device=dev))
        wire, allm, res = gbuf[key]
        comp.wait_stream(ctx.comm)            # the previous step's gather has finished with the buffers
        wire.copy_(out)
        ev = torch.cuda.Event(); ev.record(comp)
        with torch.cuda.stream(ctx.comm):
            ctx.comm.wait_event(ev)
            dist.all_gather_into_tensor(allm, wire)
            res.copy_(allm)
        return res

    d2h_bytes = [0]

    def step_e2e():
        # public API on pinned HOST buffers.  Batches go through the pipelined entry points (pair groups: upload of
        # group g+1 and download of group g-1 overlap the kernels of group g); a single pair has nothing to overlap.
        if ppg > 1:
            if spatial:
                flat, counts, status = spatial_host_pipelined(host["C"], host["ex"], host["ey"], host["cad"], host["pc"], diam)
                d2h_bytes[0] = int(2 * counts.max()) * 8 * flat.shape[0] + counts.numel() * 8   # upper bound: prefix width of the widest group
                return flat
            res = naive_host_pipelined(host["C"], host["ex"], host["ey"])
            d2h_bytes[0] = res.numel() * 8
            return res
        if not spatial and float(n1) * n2 >= 2e9:
            # one huge pair: row / column blocks, uploads behind the kernels (pipeline.naive_host_blocked)
            res = naive_host_blocked(host["C"][0], host["ex"][0], host["ey"][0])
            d2h_bytes[0] = res.numel() * 8
            return res
        if spatial:
            src = {kk: v.to(dev, non_blocking=True) for kk, v in host.items()}
            out = spatial_batched(src["C"], src["ex"], src["ey"], src["cad"], src["pc"], diam, verbose=False)
            res = torch.cat([o.reshape(-1) for o in out]).cpu()
        else:
            # the reference's own call on host tensors (scripts/eval.py:87): upload, kernels and download inside
            res = naive_fmap2pointmap(host["C"][0], host["ex"][0], host["ey"][0])
        d2h_bytes[0] = res.numel() * 8
        return res

    def timed(fn, nsteps, nwarm):
        comp = torch.cuda.current_stream(dev)
        for _ in range(nwarm):
            gather_maps(fn())
        comp.wait_stream(ctx.comm)
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
        total_ms, keep = 0.0, None
        for _ in range(nsteps):
            ctx.flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            keep = gather_maps(fn())
            e1.record()
            e1.synchronize()
            total_ms += e0.elapsed_time(e1)
        # the last step's side-stream gather is the only one nothing overlaps: its exposed tail is part of the job
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(); comp.wait_stream(ctx.comm); t1.record(); t1.synchronize()
        total_ms += t0.elapsed_time(t1)
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        del keep
        return float(t.item())

    ms_total = timed(step_resident, steps, warmup)
    queries_per_step = float(n1) * n2 * ppg * world
    value = queries_per_step * steps / (ms_total * 1e-3) / 1e9

    # share of the all-gather in a step (multi-GPU): the same steps without it
    split = {}
    if dist is not None:
        saved = dist
        dist = None
        ms_nogather = timed(step_resident, max(2, steps // 2), 1)
        dist = saved
        split["kernels_ms_per_step"] = ms_nogather / max(2, steps // 2)
        split["allgather_exposed_ms_per_step"] = max(0.0, ms_total / steps - split["kernels_ms_per_step"])

    # e2e from pinned host buffers (CUDA events cannot bracket the host-side .cpu(): wall clock inside a synced region)
    for _ in range(2):
        step_e2e()
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    torch.cuda.synchronize(dev)
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_s = float(dt.item())
    stream = None
    if ppg == 1 and not spatial:
        # the eval loop's shape (scripts/eval.py:83-91): a STREAM of single-pair conversions with a few pairs in flight
        # (pipeline.naive_pair_stream over F2PPairStream).  Every step's upload from pinned host memory and its result
        # download are inside the timed region; they overlap the neighbouring steps' kernels instead of waiting for them.
        from fmap2pointmap_solvers.pipeline import naive_pair_stream
        depth = 2 if float(n1) * n2 >= 2e9 else 3
        npairs = e2e_steps if float(n1) * n2 >= 2e9 else 64 * max(4, e2e_steps // 8)
        feed = lambda m: ((host["C"][0], host["ex"][0], host["ey"][0]) for _ in range(m))    # noqa: E731
        for _ in naive_pair_stream(feed(2 * depth + 2), depth=depth):
            pass
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        got = sum(int(r.shape[1] == n2) for r in naive_pair_stream(feed(npairs), depth=depth))
        torch.cuda.synchronize(dev)
        dts = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(dts, op=dist.ReduceOp.MAX)
        assert got == npairs
        stream = {"seconds": float(dts.item()), "pairs": npairs, "in_flight": depth,
                  "entry": "naive_pair_stream: host tensors in, host tensors out, results in order"}
    h2d = sum(v.numel() * 4 for v in host.values())
    split["h2d_ms_per_step_at_25GBps"] = h2d / 25e9 * 1e3

    # kernels of this library launched per step
    tiles = ppg * ((n1 + 127) // 128 + (n2 + 127) // 128)
    prep = 1 if tiles <= 8 * 148 else 2
    tc = 2.0 * n1 * n2 * k * ppg >= 8.0e9 and k >= 19
    if spatial:
        k2 = (4 if tc else 0) + 3               # [pack x, pack y, screen, verify] + gated + exact fallback + merge
        launches = prep + k2 + 1 + 3 * 5        # + filter init + 3 x (vote, ordered vote, classify, verify, combine)
    elif tc:
        launches = prep + 4 + 2                 # pack x, pack y, screen, verify + gated (flagged tiles) + finalise
    else:
        launches = prep + (1 if (n1 * n2 <= 10_000_000 and ppg <= 64) else 2)
    res = {
        "name": name, "config": workload_config(name, wl, ppg, world), "value": value, "ms_per_step": ms_total / steps,
        "pairs_per_s": ppg * world * steps / (ms_total * 1e-3),
        "e2e": {"value": queries_per_step * e2e_steps / e2e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h_bytes[0] + (8 * ppg if spatial else 0),
                "pairs_per_s": ppg * world * e2e_steps / e2e_s, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                "mode": "one call per step, the next step starts when the result is on the host",
                "timer": "host wall clock around synchronized region (includes pinned H2D and result D2H)"},
        "gpu_launches_per_step": launches, "split": split, "steps": steps,
    }
    if stream is not None:
        # throughput metric, so the stream is the end-to-end figure; the one-call-at-a-time figure stays beside it
        one = dict(res["e2e"])
        sec, m = stream["seconds"], stream["pairs"]
        res["e2e"] = {"value": float(n1) * n2 * world * m / sec / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": 2 * n2 * 8, "pairs_per_s": world * m / sec, "ms_per_step": 1e3 * sec / m,
                      "mode": f"stream of single-pair calls, {stream['in_flight']} in flight: every step uploads its own "
                              "inputs from pinned host memory and downloads its result inside the timed region, "
                              "overlapped with the neighbouring steps' kernels",
                      "entry": stream["entry"], "steps": m, "timer": one["timer"],
                      "one_call_at_a_time": {kk: one[kk] for kk in ("value", "pairs_per_s", "ms_per_step", "mode")}}
    return res, step_resident, base, devt


def run_b200_arm(args, wl):
    ctx = gpu_context()
    ppg = args.pairs_per_gpu or wl["pairs_per_gpu"]
    sampler = ClockSampler(ctx.local_rank)
    if ctx.rank == 0:
        sampler.start()
    res, step_resident, base, devt = measure_workload(ctx, args.workload, wl, ppg, args.steps, args.warmup,
                                                      max(3, min(args.steps, 10)))
    clocks = sampler.stop() if ctx.rank == 0 else None
    line = {
        "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": ctx.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": res["config"], "pairs_per_s": res["pairs_per_s"], "e2e": res["e2e"], "clocks": clocks,
        "gpu_launches": res["gpu_launches_per_step"] * args.steps,
    }
    if res["split"]:
        line["step_split"] = res["split"]
    if ctx.rank == 0:
        stages = stage_breakdown(ctx.L, ctx._lib, step_resident, ctx.flush, max(2, min(args.steps, 5)))
        line["roofline"] = roofline_section(ctx, wl, ppg, stages)
        if not args.no_cpu_baseline and ctx.world == 1:
            spatial = wl["solver"] == "spatial"
            cb, ref_out = cpu_baseline_block(wl, base[0], 2 if spatial else 5, 1)
            line["cpu_baseline"] = cb
            line["parity"] = parity_block(ctx, wl, base[0], ref_out)
    del step_resident, devt
    torch.cuda.empty_cache()
    # the pairs/s half of BASELINE.json's metric, in the same process group (every rank takes part)
    if args.workload == "large" and not args.no_extra:
        other = {}
        for name, st, wu in (("pair", 50, 5), ("batch", 5, 3)):
            w2 = WORKLOADS[name]
            r2, s2, _, d2 = measure_workload(ctx, name, w2, w2["pairs_per_gpu"], st, wu, 5 if name == "batch" else 20)
            entry = {"workload": r2["config"]["workload"], "config": r2["config"], "pairs_per_s": r2["pairs_per_s"],
                     "ms_per_step": r2["ms_per_step"], "value": r2["value"], "unit": UNIT,
                     "e2e_pairs_per_s": r2["e2e"]["pairs_per_s"], "e2e_ms_per_step": r2["e2e"]["ms_per_step"],
                     "h2d_bytes_per_step": r2["e2e"]["h2d_bytes_per_step"], "step_split": r2["split"]}
            if "one_call_at_a_time" in r2["e2e"]:
                entry["e2e_mode"] = r2["e2e"]["mode"]
                entry["e2e_one_call_at_a_time"] = r2["e2e"]["one_call_at_a_time"]
            if ctx.rank == 0:
                st2 = stage_breakdown(ctx.L, ctx._lib, s2, ctx.flush, 3)
                entry["stages_ms"] = st2["ms"]
                fl = 2.0 * w2["n1"] * w2["n2"] * w2["k"] * w2["pairs_per_gpu"]
                entry["kernel2_tflops"] = fl / (st2["ms"]["kernel2"] * 1e-3) / 1e12 if st2["ms"]["kernel2"] > 0 else None
                entry["tensor_pipe_screen"] = tc_stats(ctx)
            other[name] = entry
            del s2, d2
            torch.cuda.empty_cache()
        if ctx.world > 1:
            other["large_row_sharded"] = strong_scaled_pair(ctx, wl)
        line["other_workloads"] = other
    if ctx.rank == 0:
        print(json.dumps(line), flush=True)
    if ctx.dist is not None:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


def strong_scaled_pair(ctx, wl, steps=10, warmup=3):
    """ONE pair of the headline workload split over all ranks as a grid of row x scene-point blocks
    (distributed.naive_row_sharded): every rank evaluates its block, one NCCL all-reduce(MIN) of the (distance, global row)
    keys merges the row shards, assembles the scene-point shards and returns the pair's arg-min to every rank.  Strong scaling of a single huge pair; the exchange is inside the
    timed region; rank 0 also checks the map against its own unsplit call."""
    from fmap2pointmap_solvers import naive_fmap2pointmap
    from fmap2pointmap_solvers.distributed import block_grid, naive_row_sharded
    dev, dist = ctx.dev, ctx.dist
    n1, n2, k = wl["n1"], wl["n2"], wl["k"]
    grid = block_grid(n1, n2, ctx.world)
    P = make_pairs(n1, n2, k, 1234, 1)[0]                       # the SAME pair on every rank
    C, ex, ey = P.C12.to(dev), P.evecs_x.to(dev), P.evecs_y.to(dev)
    for _ in range(warmup):
        out = naive_row_sharded(C, ex, ey)
    same = bool(torch.equal(out, naive_fmap2pointmap(C, ex, ey)))
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    total = 0.0
    for _ in range(steps):
        ctx.flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = naive_row_sharded(C, ex, ey)
        e1.record()
        e1.synchronize()
        total += e0.elapsed_time(e1)
    t = torch.tensor([total, 0.0 if same else 1.0], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0].item()) / steps
    return {"workload": f"one pair of n1={n1} x n2={n2}, k={k}, naive solver, split over {ctx.world} GPUs as "
                        f"{grid[0]} CAD-row shards x {grid[1]} scene-point shards",
            "scaling": "strong", "ms_per_pair": ms, "value": float(n1) * n2 / (ms * 1e-3) / 1e9, "unit": UNIT,
            "exchange": f"all-reduce(MIN) of {n2} int64 keys per pair (NCCL), inside the timed region",
            "identical_to_unsplit_call_on_every_rank": float(t[1].item()) == 0.0, "steps": steps}


def tc_stats(ctx):
    a = (ctypes.c_ulonglong * 8)()
    if ctx.L.f2p_tc_last_stats(a) != 0:
        return None
    return {"candidates_recorded": a[0], "lists_overflowed": a[1], "candidates_verified": a[2], "lower_bound_violations": a[3]}


def parity_block(ctx, wl, P, ref_out):
    """The GPU result on the cpu_baseline sample against what the reference solver just returned for it, plus the
    at-risk census of SURVEY 8(c)(3) on that sample (naive workloads)."""
    import fmap2pointmap_solvers as S
    if wl["solver"] != "naive" or ref_out is None:
        return None
    cols = min(wl["baseline_cols"], wl["n2"])
    got = S.naive_fmap2pointmap(P.C12.to(ctx.dev), P.evecs_x.to(ctx.dev), P.evecs_y[:cols].contiguous().to(ctx.dev),
                                int16_compat=True).cpu()
    mism = int((got != ref_out).sum())
    out = {"checked_columns": cols, "mismatches": mism,
           "checked_against": "%s solver output on the cpu_baseline sample" % reference_solvers()["kind"]}
    if wl["n1"] * cols <= 500_000_000:
        feat = torch.matmul(P.evecs_x, P.C12.t())
        d = torch.cdist(feat, P.evecs_y[:cols].contiguous())
        t2 = d.topk(2, dim=0, largest=False).values.contiguous().view(torch.int32)
        out["at_risk_columns"] = int(((t2[1] - t2[0]).abs() <= 1).sum())
        out["at_risk_definition"] = "two smallest reference distances within 1 ulp after the reference's sqrt"
    return out


def _event_time(fn, reps, warmup, flush):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return sum(ts) / len(ts), ts[0]


def _traffic(wl, tensor):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture of this workload (None when the
    workload has no capture)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            t = json.load(f)
    except Exception:
        return None
    if wl["n1"] == 100_000 and wl["k"] == 128 and wl["solver"] == "naive":
        return t.get("large_tc" if tensor else "large_fp32", {}).get("bytes")
    return None


STAGE_NAMES = ("projection", "kernel2", "filter_init", "vote0", "combine0", "vote1", "combine1", "vote2", "combine2")


def stage_breakdown(L, _lib, step, flush, reps):
    """Per-stage durations of the timed workload's own step: the library records CUDA events on the launching stream at
    its stage boundaries (f2p_stage_timing / f2p_stage_times, include/f2p.h); L2 flushed before every step."""
    ms = (ctypes.c_float * len(STAGE_NAMES))()
    evals, amb, execd, skipped = ctypes.c_double(0), ctypes.c_double(0), ctypes.c_double(0), ctypes.c_double(0)
    acc = [0.0] * len(STAGE_NAMES)
    _lib.check(L.f2p_stage_timing(1), "f2p_stage_timing")
    try:
        for _ in range(reps):
            flush.fill_(1)
            step()
            _lib.check(L.f2p_stage_times(ms, ctypes.byref(evals), ctypes.byref(amb)), "f2p_stage_times")
            _lib.check(L.f2p_stage_filter_executed(ctypes.byref(execd), ctypes.byref(skipped)), "f2p_stage_filter_executed")
            for i in range(len(acc)):
                acc[i] += ms[i]
    finally:
        L.f2p_stage_timing(0)
    return {"ms": {n: a / reps for n, a in zip(STAGE_NAMES, acc)}, "filter_evals_per_step": evals.value, "ambiguous_columns_per_step": amb.value, "steps": reps,
            "filter_evals_executed_per_step": execd.value, "filter_votes_skipped_per_step": skipped.value,
            "timed": "CUDA events recorded by the library on the launching stream at the stage boundaries of the timed "
                     "step (f2p_stage_times)"}


def roofline_section(ctx, wl, ppg=1, stages=None):
    """Per-kernel rooflines measured live with CUDA events on the launching (current) stream.
    Dominant kernel = kernel 2 (fused distance + argmin/top-5).  Large problems run it as a tensor-pipe (tcgen05, TF32)
    screen + exact FP32 verification: bound = tensor, algorithmic flops 2*n1*n2*k against the TF32 peak measured live with
    the same MMA shape; the FP32-pipe form of the kernel (fallback, small problems) is timed beside it."""
    L, _lib, dev = ctx.L, ctx._lib, ctx.dev
    n1, n2, k = wl["n1"], wl["n2"], wl["k"]
    top5 = wl["solver"] == "spatial"
    P = make_pairs(n1, n2, k, 99, 1)[0]
    ex, ey, C = P.evecs_x.to(dev), P.evecs_y.to(dev), P.C12.to(dev)
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    xT = torch.empty(k, ld1, device=dev); nx = torch.empty(ld1, device=dev)
    yT = torch.empty(k, ld2, device=dev); ny = torch.empty(ld2, device=dev)
    wsp = torch.empty(L.f2p_project_workspace_bytes() + 256, dtype=torch.uint8, device=dev)
    os.environ["F2P_NN_PATH"] = "tc"
    ws_tc = L.f2p_nn_workspace_bytes(n1, n2, k, 1 if top5 else 0)
    del os.environ["F2P_NN_PATH"]
    wsn = torch.empty(max(ws_tc, L.f2p_nn_workspace_bytes(n1, n2, k, 1 if top5 else 0)) + 256, dtype=torch.uint8, device=dev)
    out = torch.empty(n2 * (5 if top5 else 1), dtype=torch.int64, device=dev)
    flush = ctx.flush
    st = _lib.stream_ptr(dev)

    def proj_x():
        _lib.check(L.f2p_project(ex.data_ptr(), n1, k, k, C.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1,
                                 wsp.data_ptr(), wsp.numel(), st), "f2p_project")

    def proj_y():
        _lib.check(L.f2p_project(ey.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2,
                                 wsp.data_ptr(), wsp.numel(), st), "f2p_project")

    def nn():
        fn = L.f2p_nn_top5 if top5 else L.f2p_nn_argmin
        _lib.check(fn(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k, out.data_ptr(),
                      wsn.data_ptr(), wsn.numel(), st), "f2p_nn")

    proj_y()
    pm, pb = _event_time(proj_x, 10, 3, flush)
    flops = 2.0 * n1 * n2 * k
    kp = (k + 6 + 7) // 8 * 8
    tensor = flops >= 8.0e9 and k >= 19          # the library's own path selection (csrc/internal.h TC_MIN_FLOPS)
    nm, nb = _event_time(nn, 10, 3, flush)        # default path
    stats = tc_stats(ctx) if tensor else None
    os.environ["F2P_NN_PATH"] = "gated"
    fm, fb = _event_time(nn, 3 if flops > 1e12 else 10, 1, flush)
    del os.environ["F2P_NN_PATH"]
    packed, scalar, tf32 = ctypes.c_double(0), ctypes.c_double(0), ctypes.c_double(0)
    _lib.check(L.f2p_measure_fp32_peak(ctypes.byref(packed), ctypes.byref(scalar), st), "f2p_measure_fp32_peak")
    _lib.check(L.f2p_measure_tf32_peak(ctypes.byref(tf32), st), "f2p_measure_tf32_peak")
    achieved = flops / (nm * 1e-3) / 1e12
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    proj_bytes = 4.0 * (2.0 * n1 * k + k * k + n1)
    fp32 = {"kernel": "nn_gated_kernel<%d> (FFMA2 fused distance + %s; fallback / small problems)" % (5 if top5 else 1, "top-5" if top5 else "arg-min"),
            "bound": "fp32", "achieved": flops / (fm * 1e-3) / 1e12, "peak": packed.value, "unit": "TFLOP/s",
            "frac": flops / (fm * 1e-3) / 1e12 / packed.value, "frac_nominal": flops / (fm * 1e-3) / 1e12 / NOMINAL_FP32_TFLOPS,
            "peak_nominal": NOMINAL_FP32_TFLOPS, "peak_scalar_ffma": scalar.value, "ms_per_launch": fm, "ms_best": fb,
            "peak_source": "measured live: packed fma.rn.f32x2 8x8 register outer product (f2p_measure_fp32_peak)"}
    if tensor:
        half_bf16 = peaks.get("bf16_tflops", 0.0) / 2.0
        roof = {
            "kernel": "nn_tc_screen_kernel<%d> + nn_tc_verify_kernel (tcgen05 TF32 screen with a proven error band, exact FP32 "
                      "re-evaluation of in-band candidates; %s)" % (5 if top5 else 1, "top-5" if top5 else "arg-min"),
            "bound": "tensor", "achieved": achieved, "peak": tf32.value, "unit": "TFLOP/s", "frac": achieved / tf32.value,
            "peak_source": "measured live: tcgen05.mma kind::tf32 M128 N256 K8 from shared memory, no loads / epilogue "
                           "(f2p_measure_tf32_peak); MEASURED_PEAKS.json has bf16 only",
            "peak_nominal": NOMINAL_TF32_TFLOPS, "frac_nominal": achieved / NOMINAL_TF32_TFLOPS,
            "peak_half_of_measured_bf16": half_bf16 or None, "frac_half_of_measured_bf16": achieved / half_bf16 if half_bf16 else None,
            "algorithmic_flops_per_launch": flops, "tensor_flops_executed_per_launch": 2.0 * (ld1 + (-ld1) % 256) * ld2 * kp,
            "ms_per_launch": nm, "ms_best": nb,
            "timed": "CUDA events on the launching stream around f2p_nn_%s (operand packing + screen + verification + "
                     "FP32 pass over flagged tiles + index finalise)" % ("top5" if top5 else "argmin"),
            "candidates": dict(stats or {}, per_column=(stats["candidates_verified"] / n2) if stats else None),
            "speedup_over_fp32_path": fm / nm, "fp32_path": fp32, "traffic": _traffic(wl, True),
        }
    else:
        roof = dict(fp32, algorithmic_flops_per_launch=flops, traffic=_traffic(wl, False),
                    timed="CUDA events on the launching stream around f2p_nn_%s (key fill + fused kernel + index finalise)" % ("top5" if top5 else "argmin"))
    # projection: 2k flop per 8 bytes moved (row read + k-major row written) — past the FP32 / HBM ridge of this device
    # (~11.5 flop/B) the single-accumulator FP32 chains bound it, not the memory system; both roofs are printed
    proj_gbs, proj_tf = proj_bytes / (pm * 1e-3) / 1e9, 2.0 * n1 * k * k / (pm * 1e-3) / 1e12
    fp32_peak = fp32["peak"]
    proj_fp32_bound = (2.0 * k / 8.0) > (fp32_peak * 1e12) / (hbm_peak * 1e9)
    roof["projection"] = {"bound": "fp32" if proj_fp32_bound else "hbm",
                          "achieved": proj_tf if proj_fp32_bound else proj_gbs,
                          "peak": fp32_peak if proj_fp32_bound else hbm_peak, "unit": "TFLOP/s" if proj_fp32_bound else "GB/s",
                          "frac": proj_tf / fp32_peak if proj_fp32_bound else proj_gbs / hbm_peak, "ms_per_launch": pm,
                          "hbm": {"achieved": proj_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": proj_gbs / hbm_peak},
                          "fp32": {"achieved": proj_tf, "peak": fp32_peak, "unit": "TFLOP/s", "frac": proj_tf / fp32_peak},
                          "flop_per_byte": 2.0 * k / 8.0,
                          "peak_source": ("MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s") +
                                         "; FP32: the live packed-FMA measurement of fp32_path",
                          "flops_per_launch": 2.0 * n1 * k * k}
    if stages is not None:
        sm = stages["ms"]
        roof["stages"] = stages
        if ppg > 1 and sm["kernel2"] > 0:
            # the launch the step really makes: all pairs of the batch in one grid
            # the library picks the path from the flops of the whole launch: a batch of small pairs takes the tensor pipe
            # where one of its pairs alone would not
            batch_tensor = flops * ppg >= 8.0e9 and k >= 19
            peak = tf32.value if batch_tensor else packed.value
            roof["single_pair_launch"] = {"achieved": achieved, "ms_per_launch": nm, "bound": roof["bound"],
                                          "peak": roof["peak"], "frac": roof["frac"], "kernel": roof["kernel"]}
            bat = flops * ppg / (sm["kernel2"] * 1e-3) / 1e12
            if batch_tensor and not tensor:
                roof["kernel"] = ("nn_tc_screen_kernel<%d> + nn_tc_verify_kernel (tcgen05 TF32 screen with a proven error "
                                  "band, exact FP32 re-evaluation of in-band candidates; %s), all pairs of the batch in one "
                                  "grid" % (5 if top5 else 1, "top-5" if top5 else "arg-min"))
                roof["peak_source"] = ("measured live: tcgen05.mma kind::tf32 M128 N256 K8 from shared memory, no loads / "
                                       "epilogue (f2p_measure_tf32_peak)")
                roof["candidates"] = tc_stats(ctx)
            roof.update({"bound": "tensor" if batch_tensor else "fp32", "peak": peak, "achieved": bat, "frac": bat / peak,
                         "algorithmic_flops_per_launch": flops * ppg, "ms_per_launch": sm["kernel2"], "ms_best": None,
                         "timed": stages["timed"] + ": %d pairs per launch (operand packing + screen + verification + "
                                  "FP32 pass over flagged tiles)" % ppg})
            for stale in ("frac_nominal", "peak_nominal", "peak_scalar_ffma"):
                if batch_tensor and not tensor:
                    roof.pop(stale, None)
        if top5 and stages["filter_evals_per_step"] > 0:
            # kernel 3 = 21 flop per (a,b) evaluation of the reference x sum_r M_r^2, against the FP32 pipe;
            # `executed` = what the kernel really does: each unordered pair once (symmetry), nothing for a skipped round
            vote_ms = sm["vote0"] + sm["vote1"] + sm["vote2"]
            comb_ms = sm["combine0"] + sm["combine1"] + sm["combine2"]
            fl = 21.0 * stages["filter_evals_per_step"]
            ach = fl / (vote_ms * 1e-3) / 1e12 if vote_ms > 0 else 0.0
            roof["filter"] = {
                "kernel": "vote_kernel (spatial-consistency vote, 3 rounds)", "bound": "fp32", "achieved": ach,
                "peak": packed.value, "unit": "TFLOP/s", "frac": ach / packed.value,
                "algorithmic_flops_per_step": fl, "evaluations_per_step": stages["filter_evals_per_step"],
                "ambiguous_columns_per_step": stages["ambiguous_columns_per_step"],
                "ms_vote": vote_ms, "ms_combine": comb_ms, "ms_init": sm["filter_init"],
                "note": "reference work (both triangles, B per entry, every round); see `executed` for the work the kernel does"}
            # executed work: the screened tile issues 25 warp instructions per 64 evaluations (ncu source page);
            # issue peak = 4 warp instructions per clock and SM at the clock sampled under load
            ex_ev = stages.get("filter_evals_executed_per_step", 0.0)
            clk = 1.9e9
            roof["filter"]["executed"] = {
                "evaluations_per_step": ex_ev, "votes_skipped_per_step": stages.get("filter_votes_skipped_per_step"),
                "warp_instructions_per_64_evaluations": 25,
                # one MUFU.SQRT per evaluation for A, B's is reused while the scene point repeats (5 candidates per
                # scene point in round 0): 2.3 warp-level MUFU per 64 evaluations on the ncu source page = 1.15 each
                "sqrt_per_evaluation": 1.15,
                "issue_slot_fraction": (ex_ev / 64.0 * 25.0) / (vote_ms * 1e-3 * 4 * 148 * clk) if vote_ms > 0 else None,
                "mufu_fraction_at_16_per_clk_sm": (1.15 * ex_ev) / (vote_ms * 1e-3 * 16 * 148 * clk) if vote_ms > 0 else None,
                "g_evaluations_per_s": ex_ev / (vote_ms * 1e-3) / 1e9 if vote_ms > 0 else None,
                "assumed_sm_clock_hz": clk}
    return roof


def main():
    args = parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl)
    else:
        run_b200_arm(args, wl)


if __name__ == "__main__":
    main()
