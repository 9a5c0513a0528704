"""Times the ground-truth side kernels (find_positives, C_from_sparse_P) on the GPU box next to their CPU forms
(NumPy broadcast for find_positives as the reference writes it; torch.linalg.lstsq on the host cores).  Not a bench
line — a side measurement for DESIGN.md."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, "6d-pose-estimation-for-unseen-categories_b200")
sys.path.insert(0, "tests")
from fmap2pointmap_solvers import C_from_sparse_P, C_from_sparse_P_batched  # noqa: E402
from fmap2pointmap_solvers.positives import find_positives  # noqa: E402
from golden.lstsq_cases import build_lstsq_inputs  # noqa: E402


def gpu_ms(fn, n=20, warm=3):
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


def cpu_ms(fn, n=5):
    fn()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t) / n * 1e3


def main():
    out = {}
    for name in ("lsq_lm_5002x1369_k30", "lsq_many_20000"):
        P, e1, e2, k = build_lstsq_inputs(name)
        Pt, t1, t2 = torch.from_numpy(P), torch.from_numpy(e1)[:, :k], torch.from_numpy(e2)[:, :k]
        Pg, g1, g2 = Pt.cuda(), t1.cuda(), t2.cuda()
        out[name] = {"gpu_ms": gpu_ms(lambda: C_from_sparse_P(Pg, g1, g2)),
                     "cpu_lstsq_ms": cpu_ms(lambda: torch.linalg.lstsq(t2[Pt[:, 1]], t1[Pt[:, 0]])[0][:k])}
        B = 64
        out[name + "_batch64"] = {"gpu_ms": gpu_ms(lambda: C_from_sparse_P_batched([Pg] * B, [g1] * B, [g2] * B, return_status=True), n=5)}
    rng = np.random.default_rng(0)
    a = rng.random((5002, 3)) * 10
    b = a[rng.integers(0, 5002, 1369)] + rng.normal(0, 0.5, (1369, 3))
    ag, bg = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out["find_positives_5002x1369_f64"] = {
        "gpu_ms": gpu_ms(lambda: find_positives(ag, bg, 0.8625)),
        "cpu_numpy_ms": cpu_ms(lambda: np.argwhere(np.linalg.norm(a[:, np.newaxis] - b, axis=2) <= 0.8625), n=3)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
