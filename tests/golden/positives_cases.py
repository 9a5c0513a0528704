"""Seeded inputs of the find_positives golden cases (outputs in positives_golden.npz, made by
oracle/make_golden_positives.py running the reference's own `find_positives` source)."""
import numpy as np

POS_CASES = {
    "pos_f64_700x400": dict(n1=700, n2=400, dtype="float64", seed=1, r=0.865, jitter=0.4),
    "pos_f32_500x300": dict(n1=500, n2=300, dtype="float32", seed=2, r=0.8, jitter=0.3),
    "pos_mixed_300x200": dict(n1=300, n2=200, dtype="mixed", seed=3, r=1.5, jitter=None),
    "pos_lm_5002x1369": dict(n1=5002, n2=1369, dtype="float64", seed=4, r=0.05 * 17.25, jitter=0.5),
    "pos_none_100x50": dict(n1=100, n2=50, dtype="float64", seed=5, r=1e-9, jitter=None),
    "pos_exact_r_64x64": dict(n1=64, n2=64, dtype="float32", seed=6, r=5.0, jitter="grid"),
}


def build_positives_inputs(name):
    c = POS_CASES[name]
    rng = np.random.default_rng(c["seed"])
    if c["jitter"] == "grid":
        # integer lattice points: many distances are EXACTLY r = 5 (3-4-5 triangles) -> the <= boundary is exercised
        a = rng.integers(0, 9, (c["n1"], 3)).astype(np.float32)
        b = rng.integers(0, 9, (c["n2"], 3)).astype(np.float32)
        return a, b, c["r"]
    a = rng.random((c["n1"], 3)) * 10
    if c["jitter"] is None:
        b = rng.random((c["n2"], 3)) * 10
    else:
        b = a[rng.integers(0, c["n1"], c["n2"])] + rng.normal(0, c["jitter"], (c["n2"], 3))
    if c["dtype"] == "float32":
        a, b = a.astype(np.float32), b.astype(np.float32)
    elif c["dtype"] == "mixed":
        a = a.astype(np.float32)
    return np.ascontiguousarray(a), np.ascontiguousarray(b), c["r"]
