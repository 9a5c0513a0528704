"""ctypes/numpy front-end of the CPU oracle (oracle/f2p_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may import
this module.  The product path (fmap2pointmap_solvers -> libf2p_sm100.so) never does.

Each function mirrors one reference function:
  naive()            <- fmap2pointmap_solvers/naive.py:5-21
  spatial()          <- fmap2pointmap_solvers/spacial_filtering.py:4-21
  nn_argmin()/nn_top5() <- naive.py:23-35 / spacial_filtering.py:22-40
  vote_scores()      <- spacial_filtering.py:42-49,54
  spatial_filter()   <- spacial_filtering.py:51-75
  thresholds()       <- the `< c*diam_cad` scalars of spacial_filtering.py:54,58,68-73
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libf2p_oracle.so")
_lib = None

_f32p = ctypes.POINTER(ctypes.c_float)
_i64p = ctypes.POINTER(ctypes.c_int64)
_i32p = ctypes.POINTER(ctypes.c_int)
_i64 = ctypes.c_int64
_int = ctypes.c_int


def build(force: bool = False) -> str:
    """Compiles oracle/f2p_oracle.c with gcc (oracle/Makefile) if the .so is missing or stale."""
    src = os.path.join(_HERE, "f2p_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libf2p_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        L.f2po_rownorm.argtypes = [_f32p, _i64, _int, _i64, _f32p]
        L.f2po_rownorm.restype = None
        L.f2po_project.argtypes = [_f32p, _i64, _int, _i64, _f32p, _f32p]
        L.f2po_project.restype = None
        L.f2po_nn_argmin.argtypes = [_f32p, _i64, _f32p, _i64, _int, _i64p, _f32p]
        L.f2po_nn_argmin.restype = _int
        L.f2po_nn_top5.argtypes = [_f32p, _i64, _f32p, _i64, _int, _i64p, _f32p, _i32p]
        L.f2po_nn_top5.restype = _int
        L.f2po_pair_dist.argtypes = [_f32p, _i64, _f32p, _i64, _int, _i64p, _i64p, _i64, _f32p, _f32p]
        L.f2po_pair_dist.restype = _int
        L.f2po_vote_scores.argtypes = [_f32p, _f32p, _i64, _f32p]
        L.f2po_vote_scores.restype = None
        L.f2po_spatial_filter.argtypes = [_f32p, _i64, _f32p, _i64, _i64p, _i64, _f32p, _i64p, _i64, _i64p, _i32p]
        L.f2po_spatial_filter.restype = _i64
        L.f2po_naive.argtypes = [_f32p, _f32p, _i64, _f32p, _i64, _int, _int, _i64p]
        L.f2po_naive.restype = _int
        L.f2po_spatial.argtypes = [_f32p, _f32p, _i64, _f32p, _i64, _int, _f32p, _i64, _f32p, _i64, _f32p,
                                   _int, _i64p, _i64p, _i64p, _i32p]
        L.f2po_spatial.restype = _i64
        L.f2po_inlier_count.argtypes = [_f32p, _i64, _f32p, _i64, _i64p, _i64p, _i64, ctypes.c_float]
        L.f2po_inlier_count.restype = _i64
        _lib = L
    return _lib


def _f32(a):
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(_f32p)


def _ip(a):
    return a.ctypes.data_as(_i64p)


def _c2d(C):
    C = _f32(C)
    if C.ndim == 3:
        C = C[0]
    return np.ascontiguousarray(C)


def thresholds(diam_cad) -> np.ndarray:
    """fl32(c*diam) for c in (0.3, 0.15, 0.055, 0.065) formed as torch forms them: in FP64 when `diam_cad`
    is a Python/NumPy scalar (scripts/eval.py:87 passes a list element), in FP32 when it is a 0-d FP32 tensor."""
    out = []
    for c in (0.3, 0.15, 0.055, 0.065):
        v = c * diam_cad
        if hasattr(v, "detach"):
            import torch
            out.append(np.float32(v.detach().to(torch.float32).cpu().item()))
        else:
            out.append(np.float32(float(v)))
    return np.array(out, dtype=np.float32)


def project(evecs, C):
    e = _f32(evecs); C = _c2d(C)
    n, k = e.shape
    out = np.empty((n, k), np.float32)
    lib().f2po_project(_fp(e), n, k, k, _fp(C), _fp(out))
    return out


def rownorm(x):
    x = _f32(x)
    out = np.empty(x.shape[0], np.float32)
    lib().f2po_rownorm(_fp(x), x.shape[0], x.shape[1], x.shape[1], _fp(out))
    return out


def nn_argmin(feat_x, feat_y, return_dist=False):
    fx = _f32(feat_x); fy = _f32(feat_y)
    idx = np.empty(fy.shape[0], np.int64)
    d = np.empty(fy.shape[0], np.float32)
    rc = lib().f2po_nn_argmin(_fp(fx), fx.shape[0], _fp(fy), fy.shape[0], fx.shape[1], _ip(idx), _fp(d))
    if rc:
        raise MemoryError("f2po_nn_argmin")
    return (idx, d) if return_dist else idx


def nn_top5(feat_x, feat_y, return_dist=False):
    fx = _f32(feat_x); fy = _f32(feat_y)
    K = min(5, fx.shape[0])
    idx = np.empty((fy.shape[0], K), np.int64)
    d = np.empty((fy.shape[0], K), np.float32)
    kk = ctypes.c_int(0)
    rc = lib().f2po_nn_top5(_fp(fx), fx.shape[0], _fp(fy), fy.shape[0], fx.shape[1], _ip(idx), _fp(d), ctypes.byref(kk))
    if rc:
        raise MemoryError("f2po_nn_top5")
    return (idx, d) if return_dist else idx


def pair_dist(feat_x, feat_y, ii, jj):
    fx = _f32(feat_x); fy = _f32(feat_y)
    ii = np.ascontiguousarray(ii, np.int64); jj = np.ascontiguousarray(jj, np.int64)
    d2 = np.empty(ii.shape[0], np.float32); d = np.empty(ii.shape[0], np.float32)
    lib().f2po_pair_dist(_fp(fx), fx.shape[0], _fp(fy), fy.shape[0], fx.shape[1], _ip(ii), _ip(jj), ii.shape[0], _fp(d2), _fp(d))
    return d2, d


def vote_scores(X, Y):
    X = _f32(X); Y = _f32(Y)
    s = np.empty(X.shape[0], np.float32)
    lib().f2po_vote_scores(_fp(X), _fp(Y), X.shape[0], _fp(s))
    return s


def spatial_filter(CAD, PC, p_pred, diam_cad, return_info=False):
    CAD = _f32(CAD); PC = _f32(PC)
    p = np.ascontiguousarray(p_pred.detach().cpu().numpy() if hasattr(p_pred, "detach") else p_pred, np.int64)
    M = p.shape[1]
    thr = thresholds(diam_cad)
    out = np.empty((2, max(M, 1)), np.int64)
    rounds = np.zeros(3, np.int64)
    skipped = ctypes.c_int(0)
    m = lib().f2po_spatial_filter(_fp(CAD), CAD.shape[0], _fp(PC), PC.shape[0], _ip(p), M, _fp(thr), _ip(out),
                                  out.shape[1], _ip(rounds), ctypes.byref(skipped))
    if m < 0:
        raise IndexError("f2po_spatial_filter: index out of range" if m == -2 else "f2po_spatial_filter failed")
    res = np.ascontiguousarray(out[:, :m])
    return (res, {"round_counts": rounds, "skipped": bool(skipped.value)}) if return_info else res


def naive(C12, evecs_x, evecs_y, int16_compat=True, **kwargs):
    C = _c2d(C12); ex = _f32(evecs_x); ey = _f32(evecs_y)
    out = np.empty((2, ey.shape[0]), np.int64)
    rc = lib().f2po_naive(_fp(C), _fp(ex), ex.shape[0], _fp(ey), ey.shape[0], ex.shape[1], int(int16_compat), _ip(out))
    if rc:
        raise MemoryError("f2po_naive")
    return out


def spatial(C12, evecs_x, evecs_y, CAD, PC, diam_cad, int16_compat=True, return_info=False):
    C = _c2d(C12); ex = _f32(evecs_x); ey = _f32(evecs_y); CAD = _f32(CAD); PC = _f32(PC)
    n2 = ey.shape[0]
    thr = thresholds(diam_cad)
    out = np.empty((2, max(5 * n2, 1)), np.int64)
    cand = np.empty((2, max(5 * n2, 1)), np.int64)
    rounds = np.zeros(3, np.int64)
    skipped = ctypes.c_int(0)
    m = lib().f2po_spatial(_fp(C), _fp(ex), ex.shape[0], _fp(ey), n2, ex.shape[1], _fp(CAD), CAD.shape[0], _fp(PC),
                           PC.shape[0], _fp(thr), int(int16_compat), _ip(out), _ip(cand), _ip(rounds), ctypes.byref(skipped))
    if m < 0:
        raise IndexError("f2po_spatial failed (%d)" % m)
    res = np.ascontiguousarray(out[:, :m])
    if return_info:
        K = min(5, ex.shape[0])
        return res, {"round_counts": rounds, "skipped": bool(skipped.value),
                     "candidates": np.ascontiguousarray(cand.reshape(-1)[: 2 * K * n2].reshape(2, K * n2))}
    return res


def inlier_ratio(pred_corr, CAD, PC_aligned, threshold):
    """compute_inlier_ratio (utils/utils.py:81-105) on a [M,2] (or [2,M] transposed by the caller) index array:
    returns (ratio as np.float32 — 0 for an empty map —, inlier count)."""
    pc = np.ascontiguousarray(np.asarray(pred_corr, np.int64))
    M = pc.shape[0]
    if M == 0:
        return np.float32(0), 0
    p0 = np.ascontiguousarray(pc[:, 0]); p1 = np.ascontiguousarray(pc[:, 1])
    cad = _f32(CAD); pca = _f32(PC_aligned)
    thr = float(np.float32(float(threshold)))
    n = lib().f2po_inlier_count(_fp(cad), cad.shape[0], _fp(pca), pca.shape[0], _ip(p0), _ip(p1), M, thr)
    if n < 0:
        raise IndexError("correspondence index out of range")
    return np.float32(n) / np.float32(M), int(n)


def find_positives(pc1, pc2, r=0.2, chunk=512):
    """find_positives (dataset/object.py:281-288) restated in NumPy in row chunks (same ufuncs, same dtype rules, same
    np.argwhere order; the reference materialises the whole [n1,n2,3] difference tensor)."""
    pc1 = np.asarray(pc1); pc2 = np.asarray(pc2)
    out = []
    for lo in range(0, pc1.shape[0], chunk):
        d = pc1[lo:lo + chunk, np.newaxis] - pc2
        s = d * d                                            # np.linalg.norm: sqrt(add.reduce((x.conj()*x).real, axis))
        dist = np.sqrt(np.add.reduce(s, axis=2))
        hit = np.argwhere(dist <= r)
        hit[:, 0] += lo
        out.append(hit)
    return np.concatenate(out, 0) if out else np.zeros((0, 2), np.int64)


def c_from_sparse_p(P, evecs1, evecs2):
    """C_from_sparse_P (utils/utils.py:67-80) restated: gather, then the least-squares solution of
    evecs2[P[:,1]] @ C = evecs1[P[:,0]] with torch's rank rule (rcond = eps32 * max(N, k)); minimum-norm when rank
    deficient.  Solved in FLOAT64 from the float32 inputs: this is the exact answer both the reference's FP32 gelsy and
    the device's FP64 normal equations approximate, so tests measure each one's distance to it.  Returns float64 [k,k]."""
    P = np.asarray(P).astype(np.int64)
    e1 = np.asarray(evecs1, dtype=np.float32); e2 = np.asarray(evecs2, dtype=np.float32)
    k = e1.shape[1]
    if P.shape[0] == 0:
        return np.zeros((k, k), np.float64)
    A = e2[P[:, 1]].astype(np.float64); B = e1[P[:, 0]].astype(np.float64)
    rcond = float(np.finfo(np.float32).eps) * max(A.shape[0], k)
    return np.linalg.lstsq(A, B, rcond=rcond)[0]


# ---- TEASER++ front half ---------------------------------------------------------------------------------------
# Pinned at mask level to the reference's own fixture (registration_test/fixed_scale_inliers.csv -> tests/golden/
# teaser_golden.npz, oracle/make_golden_teaser.py).  The reference cannot be compiled here (no Eigen), so the order of
# the 3-term norm sum is an assumption; it matters only for a statistic within one ulp of beta, which the fixture has not.
def compute_tims(v):
    """computeTIMs (TEASER-plusplus/teaser/src/registration.cc:481-519): v [3,N] float64 ->
    (vtilde [3, N(N-1)/2], map [2, N(N-1)/2]); segment i holds v_j - v_i for j = i+1..N-1."""
    v = np.asarray(v, np.float64)
    n = v.shape[1]
    iu, ju = np.triu_indices(n, 1)                       # row-major upper triangle = the reference's segment order
    return v[:, ju] - v[:, iu], np.stack([iu, ju]).astype(np.int32)


def scale_inliers(v1, v2, noise_bound, cbar2=1.0):
    """ScaleInliersSelector::solveForScale (registration.cc:427-445) on two [3,M] measurement matrices: column m is an
    inlier iff | |v1_m| - |v2_m| | <= beta, beta = 2*noise_bound*sqrt(cbar2).  Norms are sqrt((t0^2 + t1^2) + t2^2) —
    Eigen's packet order for a 3-element column sum (see csrc/tims.cu for why this is an assumption)."""
    v1, v2 = np.asarray(v1, np.float64), np.asarray(v2, np.float64)
    n1 = np.sqrt((v1[0] * v1[0] + v1[1] * v1[1]) + v1[2] * v1[2])
    n2 = np.sqrt((v2[0] * v2[0] + v2[1] * v2[1]) + v2[2] * v2[2])
    return np.abs(n1 - n2) <= 2.0 * float(noise_bound) * np.sqrt(float(cbar2))


def scale_inlier_graph(src, dst, noise_bound, cbar2=1.0):
    """scale_inliers on the TIMs of src / dst and the edge loop of registration.cc:582-590.
    Returns (edges int32 [E,2], degree int32 [N], mask bool [T])."""
    ts, mp = compute_tims(src)
    td, _ = compute_tims(dst)
    mask = scale_inliers(ts, td, noise_bound, cbar2)
    edges = mp[:, mask].T.astype(np.int32)
    degree = np.bincount(edges.ravel(), minlength=np.asarray(src).shape[1]).astype(np.int32)
    return edges, degree, mask
