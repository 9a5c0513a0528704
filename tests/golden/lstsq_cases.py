"""Seeded inputs of the C_from_sparse_P golden cases (outputs in lstsq_golden.npz, made by oracle/make_golden_lstsq.py
running the reference's own `utils.C_from_sparse_P`)."""
import numpy as np

LSQ_CASES = {
    # name: rows of the two bases, stored columns K (the callers slice [:, :k]), k, positives, kind
    "lsq_lm_5002x1369_k30": dict(n1=5002, n2=1369, K=64, k=30, npos=2500, seed=11, kind="smooth"),
    "lsq_small_300x200_k30": dict(n1=300, n2=200, K=30, k=30, npos=400, seed=12, kind="smooth"),
    "lsq_k50_2000x900": dict(n1=2000, n2=900, K=50, k=50, npos=3000, seed=13, kind="smooth"),
    "lsq_k7_64x48": dict(n1=64, n2=48, K=9, k=7, npos=33, seed=14, kind="random"),
    "lsq_square_n_eq_k": dict(n1=100, n2=100, K=30, k=30, npos=30, seed=15, kind="random"),
    "lsq_under_n_lt_k": dict(n1=100, n2=100, K=30, k=30, npos=12, seed=16, kind="random"),
    "lsq_dup_rows_rank_def": dict(n1=80, n2=60, K=30, k=30, npos=200, seed=17, kind="dup"),
    "lsq_many_20000": dict(n1=6000, n2=3000, K=40, k=30, npos=20000, seed=18, kind="smooth"),
}


def build_lstsq_inputs(name):
    """(P int64 [N,2], evecs1 float32 [n1,K], evecs2 float32 [n2,K], k).  'smooth' bases are low-frequency cosine
    products of random point coordinates (columns of very different scale, like Laplacian eigenvectors restricted to a
    patch); 'dup' draws every positive from 10 distinct scene points, so A has rank <= 10 < k."""
    c = LSQ_CASES[name]
    rng = np.random.default_rng(c["seed"])
    n1, n2, K = c["n1"], c["n2"], c["K"]
    if c["kind"] == "smooth":
        x1, x2 = rng.random((n1, 3)), rng.random((n2, 3))
        f = rng.integers(0, 5, (K, 3)).astype(np.float64)
        ph = rng.random(K) * 6.28
        e1 = np.cos(x1 @ f.T * 3.1 + ph) / np.sqrt(n1)
        e2 = np.cos(x2 @ f.T * 3.1 + ph) / np.sqrt(n2)
        e1 += rng.normal(0, 0.02 / np.sqrt(n1), e1.shape); e2 += rng.normal(0, 0.02 / np.sqrt(n2), e2.shape)
    else:
        e1, e2 = rng.normal(0, 1, (n1, K)), rng.normal(0, 1, (n2, K))
    if c["kind"] == "dup":
        P = np.stack([rng.integers(0, n1, c["npos"]), rng.integers(0, 10, c["npos"])], 1)
    else:
        P = np.stack([rng.integers(0, n1, c["npos"]), rng.integers(0, n2, c["npos"])], 1)
    return P.astype(np.int64), e1.astype(np.float32), e2.astype(np.float32), c["k"]
