"""BASELINE cfg5 sweep (n1 in {2k,10k,50k} x k in {30,64,128}, both solvers): single-pair latency on the GPU with the
operands resident (CUDA events, 10 calls after 3 warm-ups), and the reference's ATen op sequence on the host cores for
the sizes it finishes in seconds.  A side measurement for DESIGN.md, not a bench line."""
import json
import sys
import time

import torch

sys.path.insert(0, "6d-pose-estimation-for-unseen-categories_b200")
sys.path.insert(0, ".")
from fmap2pointmap_solvers import naive_fmap2pointmap, spacial_filtering_fmap2pointmap  # noqa: E402
from fmap2pointmap_solvers.synth import make_pair  # noqa: E402


def gpu_ms(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    from oracle import aten_port
    rows = []
    for n1 in (2000, 10000, 50000):
        n2 = int(n1 * 0.4)
        for k in (30, 64, 128):
            P = make_pair(n1, n2, k, seed=n1 + k)
            C, ex, ey = P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()
            cad, pc = P.CAD.cuda(), P.PC.cuda()
            r = {"n1": n1, "n2": n2, "k": k}
            r["naive_ms"] = gpu_ms(lambda: naive_fmap2pointmap(C, ex, ey))
            r["spatial_ms"] = gpu_ms(lambda: spacial_filtering_fmap2pointmap(C, ex, ey, CAD=cad, PC=pc, diam_cad=P.diam_cad),
                                     n=5, warm=2)
            if n1 <= 10000:
                t = time.perf_counter(); aten_port.naive(P.C12, P.evecs_x, P.evecs_y); r["cpu_naive_ms"] = (time.perf_counter() - t) * 1e3
            if n1 <= 2000:
                t = time.perf_counter(); aten_port.spatial(P.C12, P.evecs_x, P.evecs_y, P.CAD, P.PC, P.diam_cad); r["cpu_spatial_ms"] = (time.perf_counter() - t) * 1e3
            rows.append(r)
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
