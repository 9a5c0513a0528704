#!/usr/bin/env python
"""bench.py — fmap->point-map throughput on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload large|pair|batch] [--impl reference]

Workloads (BASELINE.json configs):
    large  cfg4: one pair n1=n2=100k, k=128, naive solver   (default; the config the FP32-roofline target is quoted on)
    pair   cfg2: one pair 5k x 2k, k=50, naive solver
    batch  cfg3: pairs of 5k x 2k, k=50 with spatial-consistency filtering (256 pairs per GPU by default)
One "step" = one pass of the hot path over the rank's pair(s).  Multi-GPU (torchrun): every rank converts its own
pair slice (weak scaling) and the index maps are returned with one NCCL all-gather per step.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput; `e2e` = the same through the public Python
API from pinned HOST buffers including H2D/D2H; `roofline` = the fused distance+argmin kernel against the FP32 pipe;
`cpu_baseline` = the reference's ATen op sequence (oracle/aten_port.py) on the host cores, bounded sample.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200"))

import torch  # noqa: E402

NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.5: 148 SMs x 128 lanes x 2 flop x 1.965 GHz

WORKLOADS = {
    "large": dict(n1=100_000, n2=100_000, k=128, solver="naive", pairs_per_gpu=1, baseline_cols=4096),
    "pair": dict(n1=5_000, n2=2_000, k=50, solver="naive", pairs_per_gpu=1, baseline_cols=2000),
    "batch": dict(n1=5_000, n2=2_000, k=50, solver="spatial", pairs_per_gpu=256, baseline_cols=2000),
}


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


# ------------------------------------------------------------------------------------------------ inputs
def gen_pair_fast(n1, n2, k, seed, with_xyz):
    """Synthetic pair of SURVEY 8(d) shape (same construction as fmap2pointmap_solvers.synth.make_pair)."""
    from fmap2pointmap_solvers.synth import make_pair
    P = make_pair(n1, n2, k, seed=seed)
    return P


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
        # median over the samples taken under load (upper half: idle samples before/after are excluded)
        load = clocks[len(clocks) // 2:] if clocks else []
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(clocks),
                "power_w_max": max(power) if power else None}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_reference_sample(wl, P, threads, reps):
    """Times the reference's ATen op sequence (oracle/aten_port.py) on a bounded sample of the workload.
    naive: all V1 rows against the first `baseline_cols` scene points (cdist is bitwise invariant to column
    chunking, SURVEY 8(a).1-3, and its cost is linear in the column count).  spatial: one whole pair."""
    from oracle import aten_port
    torch.set_num_threads(threads)
    cols = min(wl["baseline_cols"], wl["n2"])
    times = []
    if wl["solver"] == "naive":
        ey = P.evecs_y[:cols].contiguous()
        for _ in range(reps):
            t0 = time.perf_counter()
            aten_port.naive(P.C12, P.evecs_x, ey)
            times.append(time.perf_counter() - t0)
        queries = wl["n1"] * cols
        sample = f"all {wl['n1']} CAD rows x first {cols} of {wl['n2']} scene points, k={wl['k']}, best of {reps}"
    else:
        for _ in range(reps):
            t0 = time.perf_counter()
            aten_port.spatial(P.C12, P.evecs_x, P.evecs_y, P.CAD, P.PC, P.diam_cad)
            times.append(time.perf_counter() - t0)
        queries = wl["n1"] * wl["n2"]
        sample = f"1 whole pair {wl['n1']}x{wl['n2']}, k={wl['k']}, spatial filter included, best of {reps}"
    best = min(times)
    return queries / best / 1e9, best, sample


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    P = gen_pair_fast(wl["n1"], wl["n2"], wl["k"], 1234, wl["solver"] == "spatial")
    threads = os.cpu_count() or 1
    for _ in range(max(args.warmup, 1) if wl["solver"] == "naive" else 1):
        cpu_reference_sample(wl, P, threads, 1)
    vals, secs = [], []
    for _ in range(args.steps if wl["solver"] == "naive" else min(args.steps, 3)):
        v, s, sample = cpu_reference_sample(wl, P, threads, 1)
        vals.append(v); secs.append(s)
    steps = len(vals)
    value = sum(wl["n1"] * min(wl["baseline_cols"], wl["n2"]) if wl["solver"] == "naive" else wl["n1"] * wl["n2"]
                for _ in vals) / sum(secs) / 1e9
    line = {
        "impl": "reference", "metric": "fmap->p2p G point-queries/s", "value": value, "unit": "G point-queries/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, wl, args.gpus),
        "cpu_baseline": {"value": value, "unit": "G point-queries/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "G point-queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, wl, n_gpus):
    ppg = args.pairs_per_gpu or wl["pairs_per_gpu"]
    return {"workload": f"{args.workload}: {ppg} pair(s)/GPU of n1={wl['n1']} x n2={wl['n2']}, k={wl['k']}, {wl['solver']} solver",
            "n1": wl["n1"], "n2": wl["n2"], "k": wl["k"], "solver": wl["solver"], "pairs_per_gpu": ppg,
            "pairs_total": ppg * n_gpus, "parallelism": f"pair-sharded x{n_gpus}",
            "l2": "flushed between timed steps (256 MiB write)"}


# ------------------------------------------------------------------------------------------------ GPU arm
def run_b200_arm(args, wl):
    import fmap2pointmap_solvers as S
    from fmap2pointmap_solvers import _lib
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.pipeline import naive_host_pipelined, spatial_host_pipelined

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    L = _lib.load()
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    ppg = args.pairs_per_gpu or wl["pairs_per_gpu"]
    n1, n2, k = wl["n1"], wl["n2"], wl["k"]
    spatial = wl["solver"] == "spatial"
    # distinct pairs per rank (seed = 1234 + global pair index); large batches cycle 16 distinct pairs
    distinct = min(ppg, 16)
    base = [gen_pair_fast(n1, n2, k, 1234 + rank * ppg + i, spatial) for i in range(distinct)]
    pairs = [base[i % distinct] for i in range(ppg)]

    # The batch is handed over as STACKED tensors ([P,n,k] ...): one view per operand, no per-pair host work.
    # pinned host copies (e2e) and device-resident copies (value)
    keys = ("C", "ex", "ey", "cad", "pc") if spatial else ("C", "ex", "ey")
    attr = dict(C="C12", ex="evecs_x", ey="evecs_y", cad="CAD", pc="PC")
    host = {kk: torch.stack([getattr(p, attr[kk]) for p in pairs], 0).pin_memory() for kk in keys}
    devt = {kk: v.to(dev) for kk, v in host.items()}
    diam = [p.diam_cad for p in pairs]

    def solve(src):
        if spatial:
            return spatial_batched(src["C"], src["ex"], src["ey"], src["cad"], src["pc"], diam, verbose=False)
        return naive_batched(src["C"], src["ex"], src["ey"])

    def step_resident():
        # device-resident, asynchronous form of the call: spatial returns (flat maps, counts, status) with no host sync
        if spatial:
            return spatial_batched(devt["C"], devt["ex"], devt["ey"], devt["cad"], devt["pc"], diam, verbose=False, packed=True)
        return naive_batched(devt["C"], devt["ex"], devt["ey"])

    def gather_maps(out):
        # return the index maps to every rank: NCCL all-gather over NVLink of the rank's result buffer (+ counts)
        if dist is None:
            return
        if spatial:
            # counts first, then only the used prefix of every pair's result slots, int32 on the wire
            # (fmap2pointmap_solvers.distributed.allgather_packed_maps): the worst-case buffer is 10*n2 int64 slots per
            # pair and nearly empty behind the dense [2,M'] prefix (8 ranks: 328 MB per rank and step before this)
            from fmap2pointmap_solvers.distributed import allgather_packed_maps
            return allgather_packed_maps(out[0], out[1])
        parts = [out]
        res = []
        for t in parts:
            allm = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=dev)
            dist.all_gather_into_tensor(allm, t.contiguous())
            res.append(allm)
        return res

    d2h_bytes = [0]

    def step_e2e():
        # public API on pinned HOST buffers.  Batches go through the pipelined entry points (pair groups: upload of
        # group g+1 and download of group g-1 overlap the kernels of group g); a single pair has nothing to overlap
        # (measured: moving one 100k x 100k pair in column groups costs more in per-launch tails than the copies it hides).
        if ppg > 1:
            if spatial:
                flat, counts, status = spatial_host_pipelined(host["C"], host["ex"], host["ey"], host["cad"], host["pc"], diam)
                d2h_bytes[0] = flat.numel() * 8 + counts.numel() * 8
                return flat
            res = naive_host_pipelined(host["C"], host["ex"], host["ey"])
            d2h_bytes[0] = res.numel() * 8
            return res
        out = solve({kk: v.to(dev, non_blocking=True) for kk, v in host.items()})
        res = torch.cat([o.reshape(-1) for o in out]).cpu() if spatial else out.cpu()
        d2h_bytes[0] = res.numel() * 8
        return res

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def timed(fn, steps, warmup, with_gather):
        for _ in range(warmup):
            o = fn()
            if with_gather:
                gather_maps(o)
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
        total_ms = 0.0
        t_wall0 = time.perf_counter()
        for _ in range(steps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            o = fn()
            if with_gather:
                gather_maps(o)
            e1.record()
            e1.synchronize()
            total_ms += e0.elapsed_time(e1)
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        wall = time.perf_counter() - t_wall0
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), wall

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total, _ = timed(step_resident, args.steps, args.warmup, True)
    clocks = sampler.stop() if rank == 0 else None

    queries_per_step = float(n1) * n2 * ppg * world
    value = queries_per_step * args.steps / (ms_total * 1e-3) / 1e9

    # e2e from pinned host buffers (CUDA events cannot bracket the host-side .cpu(): wall clock inside a synced region)
    def e2e_run(steps, warmup):
        for _ in range(warmup):
            step_e2e()
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            step_e2e()
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    e2e_steps = max(3, min(args.steps, 10))
    e2e_s = e2e_run(e2e_steps, 2)
    e2e_value = queries_per_step * e2e_steps / e2e_s / 1e9
    h2d = sum(v.numel() * 4 for v in host.values())

    line = {
        "metric": "fmap->p2p G point-queries/s", "value": value, "unit": "G point-queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, wl, world),
        "pairs_per_s": ppg * world * args.steps / (ms_total * 1e-3),
        "e2e": {"value": e2e_value, "unit": "G point-queries/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h_bytes[0] + (8 * ppg if spatial else 0), "pairs_per_s": ppg * world * e2e_steps / e2e_s,
                "timer": "host wall clock around synchronized region (includes pinned H2D and result D2H)"},
        "clocks": clocks,
    }
    # kernels of this library launched per step.  Preparation (projection + transpose, fused norms, arg-min state reset):
    # 1 launch up to 8 x 148 tiles, else 2.  Naive: gated arg-min that finalises its own column tiles (1) up to ~8 waves
    # of items, else arg-min + finalise (2).  Spatial: gated top-5 + exact fallback + finalise 3, filter init +
    # 3 x (vote + classify + verify + combine) 13.
    tiles = ppg * ((n1 + 127) // 128 + (n2 + 127) // 128)
    prep_launches = 1 if tiles <= 8 * 148 else 2
    naive_launches = 1 if (n1 * n2 <= 10_000_000 and ppg <= 64) else 2
    line["gpu_launches"] = (prep_launches + (16 if spatial else naive_launches)) * args.steps

    if rank == 0 and world == 1 and args.workload == "large" and not args.no_extra:
        line["other_workloads"] = side_workloads()
    if rank == 0:
        stages = stage_breakdown(L, _lib, step_resident, flush, max(2, min(args.steps, 5)))
        line["roofline"] = roofline_section(L, _lib, wl, dev, ppg, stages)
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            cpu_reference_sample(wl, base[0], threads, 1)
            v, s, sample = cpu_reference_sample(wl, base[0], threads, 2 if spatial else 3)
            line["cpu_baseline"] = {"value": v, "unit": "G point-queries/s", "cores": threads, "kind": "port",
                                    "sample": sample, "seconds": s,
                                    "note": "reference ATen op sequence (oracle/aten_port.py) on the box's host cores"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def side_workloads():
    """The pairs/s half of BASELINE.json's metric: cfg2 (one 5k x 2k x 50 pair, naive) and cfg3 (256 pairs with the
    spatial filter) measured by this same script in a child process each, reported beside the cfg4 headline."""
    out = {}
    for name, extra in (("pair", ["--steps", "50", "--warmup", "5"]), ("batch", ["--steps", "5", "--warmup", "3"])):
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--workload", name, "--no-cpu-baseline"] + extra,
                               capture_output=True, text=True, timeout=600)
            d = json.loads(r.stdout.strip().splitlines()[-1])
            out[name] = {"workload": d["config"]["workload"], "pairs_per_s": d["pairs_per_s"], "ms_per_step": d["ms_per_step"],
                         "value": d["value"], "unit": d["unit"], "e2e_pairs_per_s": d["e2e"]["pairs_per_s"],
                         "kernel2_tflops": d["roofline"]["achieved"]}
        except Exception as e:   # a side measurement must never cost the headline line
            out[name] = {"error": repr(e)[:200]}
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


def _traffic(wl):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture of this workload (None when the
    workload has no capture)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            t = json.load(f)
    except Exception:
        return None
    if wl["n1"] == 100_000 and wl["k"] == 128 and wl["solver"] == "naive":
        return t.get("large", {}).get("bytes")
    return None


STAGE_NAMES = ("projection", "kernel2", "filter_init", "vote0", "combine0", "vote1", "combine1", "vote2", "combine2")


def stage_breakdown(L, _lib, step, flush, reps):
    """Per-stage durations of the timed workload's own step: the library records CUDA events on the launching stream at
    its stage boundaries (f2p_stage_timing / f2p_stage_times, include/f2p.h); L2 flushed before every step."""
    ms = (ctypes.c_float * len(STAGE_NAMES))()
    evals, amb = ctypes.c_double(0), ctypes.c_double(0)
    acc = [0.0] * len(STAGE_NAMES)
    _lib.check(L.f2p_stage_timing(1), "f2p_stage_timing")
    try:
        for _ in range(reps):
            flush.fill_(1)
            step()
            _lib.check(L.f2p_stage_times(ms, ctypes.byref(evals), ctypes.byref(amb)), "f2p_stage_times")
            for i in range(len(acc)):
                acc[i] += ms[i]
    finally:
        L.f2p_stage_timing(0)
    return {"ms": {n: a / reps for n, a in zip(STAGE_NAMES, acc)}, "filter_evals_per_step": evals.value, "ambiguous_columns_per_step": amb.value, "steps": reps,
            "timed": "CUDA events recorded by the library on the launching stream at the stage boundaries of the timed "
                     "step (f2p_stage_times)"}


def roofline_section(L, _lib, wl, dev, ppg=1, stages=None):
    """Per-kernel rooflines measured live with CUDA events on the launching (current) stream.
    Dominant kernel = kernel 2 (fused distance + argmin/top-5): FP32-pipe bound, algorithmic flops 2*n1*n2*k.
    Batches: kernel 2 and the spatial filter are taken from the stage events of the batched step itself."""
    n1, n2, k = wl["n1"], wl["n2"], wl["k"]
    top5 = wl["solver"] == "spatial"
    P = gen_pair_fast(n1, n2, k, 99, False)
    ex, ey, C = P.evecs_x.to(dev), P.evecs_y.to(dev), P.C12.to(dev)
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    xT = torch.empty(k, ld1, device=dev); nx = torch.empty(ld1, device=dev)
    yT = torch.empty(k, ld2, device=dev); ny = torch.empty(ld2, device=dev)
    wsp = torch.empty(L.f2p_project_workspace_bytes() + 256, dtype=torch.uint8, device=dev)
    wsn = torch.empty(L.f2p_nn_workspace_bytes(n1, n2, k, 1 if top5 else 0) + 256, dtype=torch.uint8, device=dev)
    out = torch.empty(n2 * (5 if top5 else 1), dtype=torch.int64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
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
    nm, nb = _event_time(nn, 10, 3, flush)
    packed, scalar = ctypes.c_double(0), ctypes.c_double(0)
    _lib.check(L.f2p_measure_fp32_peak(ctypes.byref(packed), ctypes.byref(scalar), st), "f2p_measure_fp32_peak")
    flops = 2.0 * n1 * n2 * k
    achieved = flops / (nm * 1e-3) / 1e12
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    proj_bytes = 4.0 * (2.0 * n1 * k + k * k + n1)
    roof = {
        "kernel": "nn_kernel<%d> (fused distance + %s)" % (1 if top5 else 0, "top-5" if top5 else "arg-min"),
        "bound": "fp32", "achieved": achieved, "peak": packed.value, "unit": "TFLOP/s", "frac": achieved / packed.value,
        "peak_source": "measured live: packed fma.rn.f32x2 8x8 register outer product (f2p_measure_fp32_peak); "
                       "MEASURED_PEAKS.json has no FP32 figure",
        "peak_nominal": NOMINAL_FP32_TFLOPS, "frac_nominal": achieved / NOMINAL_FP32_TFLOPS,
        "peak_scalar_ffma": scalar.value,
        "algorithmic_flops_per_launch": flops, "ms_per_launch": nm, "ms_best": nb,
        "timed": "CUDA events on the launching stream around f2p_nn_%s (key fill + fused kernel + index finalise)" % ("top5" if top5 else "argmin"),
        "traffic": _traffic(wl),
        "projection": {"bound": "hbm", "achieved": proj_bytes / (pm * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                       "frac": proj_bytes / (pm * 1e-3) / 1e9 / hbm_peak, "ms_per_launch": pm,
                       "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s",
                       "flops_per_launch": 2.0 * n1 * k * k},
    }
    if stages is not None:
        sm = stages["ms"]
        roof["stages"] = stages
        if ppg > 1 and sm["kernel2"] > 0:
            # the launch the step really makes: all pairs of the batch in one grid
            roof["single_pair_launch"] = {"achieved": achieved, "frac": achieved / packed.value, "ms_per_launch": nm}
            bat = flops * ppg / (sm["kernel2"] * 1e-3) / 1e12
            roof.update({"achieved": bat, "frac": bat / packed.value, "frac_nominal": bat / NOMINAL_FP32_TFLOPS,
                         "algorithmic_flops_per_launch": flops * ppg, "ms_per_launch": sm["kernel2"], "ms_best": None,
                         "timed": stages["timed"] + ": %d pairs per launch" % ppg})
        if top5 and stages["filter_evals_per_step"] > 0:
            # kernel 3 = 21 flop per (a,b) evaluation of the reference x sum_r M_r^2, against the FP32 pipe
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
                "note": "reference work (both triangles, B per entry, every round); the kernel evaluates each unordered "
                        "pair once, reuses B across the candidates of one scene point and skips a round whose "
                        "predecessor pruned nothing, so the algorithmic rate may exceed the pipe peak"}
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
