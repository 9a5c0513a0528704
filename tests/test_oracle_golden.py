"""CPU tests: the oracle (oracle/f2p_oracle.c and oracle/aten_port.py) against the golden vectors made from the
unmodified reference, and against the live reference when /root/reference exists (build container only)."""
import os

import numpy as np
import pytest
import torch

from conftest import load_case
from golden.cases import CASES
from oracle import aten_port
from oracle import f2p_oracle as O
from oracle.ref_loader import load_reference, reference_available

NAIVE = [n for n, c in CASES.items() if c["kind"] == "naive"]
SPATIAL = [n for n, c in CASES.items() if c["kind"] == "spatial"]


@pytest.mark.parametrize("name", NAIVE)
def test_c_oracle_naive_matches_reference_golden(golden, name):
    c, inp = load_case(golden, name)
    got = O.naive(inp["C12"], inp["evecs_x"], inp["evecs_y"], int16_compat=True)
    want = golden[name + "/p_pred"]
    assert got.dtype == np.int64 and got.shape == want.shape
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("name", NAIVE)
def test_c_oracle_intermediate_bits(golden, name):
    """Projection and unclamped distance-GEMM bits (naive.py:20,33) are reproduced bit-for-bit."""
    if name + "/feat_x" not in golden.files:
        pytest.skip("intermediates stored for small cases only")
    c, inp = load_case(golden, name)
    feat = O.project(inp["evecs_x"], inp["C12"])
    np.testing.assert_array_equal(feat.view(np.int32), golden[name + "/feat_x"].view(np.int32))
    if name + "/d2_sample" in golden.files:
        g = golden[name + "/d2_sample"]
        ii, jj = np.meshgrid(np.arange(g.shape[0]), np.arange(g.shape[1]), indexing="ij")
        d2, _ = O.pair_dist(feat, inp["evecs_y"], ii.ravel(), jj.ravel())
        np.testing.assert_array_equal(d2.reshape(g.shape).view(np.int32), g.view(np.int32))


@pytest.mark.parametrize("name", SPATIAL)
def test_c_oracle_spatial_matches_reference_golden(golden, name):
    c, inp = load_case(golden, name)
    got, info = O.spatial(inp["C12"], inp["evecs_x"], inp["evecs_y"], inp["CAD"], inp["PC"], inp["diam_cad"],
                          return_info=True)
    np.testing.assert_array_equal(info["candidates"], golden[name + "/candidates"])
    np.testing.assert_array_equal(got, golden[name + "/p_pred"])
    assert info["skipped"] == bool(golden[name + "/skipped"])


@pytest.mark.parametrize("name", ["naive_300x200_k30", "naive_duprows_600x400_k30", "naive_wrap_200x40000_k16",
                                  "spatial_300x200_k30", "spatial_fallback065_600x200_k30",
                                  "spatial_skipped_300x100_k30", "spatial_diamtensor_400x150_k30"])
def test_aten_port_matches_reference_golden(golden, name, capsys):
    c, inp = load_case(golden, name)
    if c["kind"] == "naive":
        got = aten_port.naive(inp["C12"], inp["evecs_x"], inp["evecs_y"])
    else:
        got = aten_port.spatial(inp["C12"], inp["evecs_x"], inp["evecs_y"], inp["CAD"], inp["PC"], inp["diam_cad"])
    assert got.dtype == torch.int64
    np.testing.assert_array_equal(got.numpy(), golden[name + "/p_pred"])
    assert ("skipped" in capsys.readouterr().out) == bool(golden[name + "/skipped"])


def test_duplicate_rows_pick_lowest_index(golden):
    c, inp = load_case(golden, "naive_duprows_600x400_k30")
    got = O.naive(inp["C12"], inp["evecs_x"], inp["evecs_y"])
    assert (got[0] < c["n1"] // 2).all()


def test_int16_wrap_row(golden):
    c, inp = load_case(golden, "naive_wrap_200x40000_k16")
    got = O.naive(inp["C12"], inp["evecs_x"], inp["evecs_y"], int16_compat=True)
    assert got[1, 32767] == 32767 and got[1, 32768] == -32768 and got[1, 39999] == 39999 - 65536
    plain = O.naive(inp["C12"], inp["evecs_x"], inp["evecs_y"], int16_compat=False)
    np.testing.assert_array_equal(plain[1], np.arange(40000))
    np.testing.assert_array_equal(plain[0], got[0])


@pytest.mark.parametrize("M", [1, 5, 31, 32, 33, 255, 256, 257, 1000, 4099])
def test_vote_cascade_order_matches_aten_sum(M):
    """`.mean(0)` summation order (spacial_filtering.py:54): bitwise equal to ATen when ATen is fed IEEE sqrt values
    (torch's own CPU sqrt is MKL VML and is not correctly rounded in ~0.7% of elements)."""
    g = torch.Generator().manual_seed(100 + M)
    X = torch.rand(M, 3, generator=g) * 10
    Y = X @ torch.linalg.qr(torch.randn(3, 3, generator=g)).Q.t() + 0.3 * torch.randn(M, 3, generator=g)

    def ed(t):
        return torch.from_numpy(np.sqrt(torch.sum((t[:, None] - t) ** 2, dim=2).numpy()))

    want = torch.absolute(ed(X) - ed(Y)).mean(0).numpy()
    got = O.vote_scores(X, Y)
    np.testing.assert_array_equal(got.view(np.int32), want.view(np.int32))


@pytest.mark.parametrize("k", [1, 7, 8, 30, 33, 50, 64, 66, 128, 130, 256])
def test_rownorm_order_matches_aten(k):
    g = torch.Generator().manual_seed(k)
    x = torch.randn(257, k, generator=g)
    want = x.pow(2).sum(-1).numpy()
    np.testing.assert_array_equal(O.rownorm(x).view(np.int32), want.view(np.int32))


@pytest.mark.skipif(not reference_available(), reason="/root/reference only exists in the build container")
@pytest.mark.parametrize("shape", [(257, 129, 30, 0), (900, 411, 50, 1), (1500, 64, 128, 2)])
def test_c_oracle_matches_live_reference(shape):
    from fmap2pointmap_solvers.synth import make_pair
    n1, n2, k, seed = shape
    ref = load_reference()
    P = make_pair(n1, n2, k, seed=900 + seed)
    want = ref.naive_fmap2pointmap(P.C12, P.evecs_x, P.evecs_y).numpy()
    np.testing.assert_array_equal(O.naive(P.C12, P.evecs_x, P.evecs_y), want)
    want = ref.spacial_filtering_fmap2pointmap(P.C12, P.evecs_x, P.evecs_y, P.CAD, P.PC, P.diam_cad).numpy()
    np.testing.assert_array_equal(O.spatial(P.C12, P.evecs_x, P.evecs_y, P.CAD, P.PC, P.diam_cad), want)


# ---------------------------------------------------------------------------------------------- inlier ratio
ALL_IR = [n for n in CASES]


@pytest.mark.parametrize("name", ALL_IR)
def test_oracle_inlier_ratio_matches_reference_golden(golden, name):
    """oracle.inlier_ratio vs utils.compute_inlier_ratio of the real reference (stored by oracle/make_golden.py)."""
    from golden.cases import aligned_pc, ir_threshold
    from oracle import f2p_oracle as O
    c, inp = load_case(golden, name)
    p = golden[name + "/p_pred"]
    ratio, _ = O.inlier_ratio(p.T, inp["CAD"], aligned_pc(name, inp), ir_threshold(inp))
    assert np.float32(ratio).view(np.int32) == np.float32(golden[name + "/ir"]).view(np.int32)


@pytest.mark.skipif(not reference_available(), reason="/root/reference only exists in the build container")
def test_oracle_inlier_ratio_matches_live_reference():
    from oracle import f2p_oracle as O
    from oracle.ref_loader import load_reference_utils
    ru = load_reference_utils()
    g = torch.Generator().manual_seed(3)
    CAD = torch.rand(4000, 3, generator=g) * 10
    PCa = torch.rand(3000, 3, generator=g) * 10
    corr = torch.stack([torch.randint(0, 4000, (20000,), generator=g), torch.randint(-3000, 3000, (20000,), generator=g)], 1)
    PCa[corr[:5000, 1]] = CAD[corr[:5000, 0]] + 0.9 * torch.randn(5000, 3, generator=g)
    for thr in (1.73, np.float64(0.931), torch.tensor(2.2)):
        want = ru.compute_inlier_ratio(corr, CAD, PCa, thr)
        got, _ = O.inlier_ratio(corr.numpy(), CAD, PCa, float(thr))
        assert np.float32(float(want)).view(np.int32) == np.float32(got).view(np.int32)
    assert ru.compute_inlier_ratio(corr[:0], CAD, PCa, 1.0) == 0 and O.inlier_ratio(corr[:0].numpy(), CAD, PCa, 1.0)[0] == 0


# ---------------------------------------------------------------------------------------------- find_positives
def _pos_case(name):
    from golden.positives_cases import build_positives_inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "positives_golden.npz"))
    a, b, r = build_positives_inputs(name)
    if float(g[name + "/in_sum"]) != float(a.astype(np.float64).sum() + b.astype(np.float64).sum()):
        pytest.skip("NumPy RNG stream changed: inputs differ from the ones the golden output was made from")
    return a, b, r, g[name + "/P"]


from golden.positives_cases import POS_CASES  # noqa: E402


@pytest.mark.parametrize("name", list(POS_CASES))
def test_oracle_find_positives_matches_reference_golden(name):
    a, b, r, want = _pos_case(name)
    got = O.find_positives(a, b, r, chunk=193)
    assert got.dtype == np.int64
    np.testing.assert_array_equal(got, want)


@pytest.mark.skipif(not reference_available(), reason="/root/reference only exists in the build container")
def test_oracle_find_positives_matches_live_reference():
    from oracle.ref_loader import load_reference_method
    fp = load_reference_method("dataset/object.py", "base_object_dataset", "find_positives")
    rng = np.random.default_rng(9)
    a = rng.random((900, 3)) * 10
    b = a[rng.integers(0, 900, 600)] + rng.normal(0, 0.5, (600, 3))
    for r in (0.3, 0.8625):
        np.testing.assert_array_equal(O.find_positives(a, b, r), fp(None, a, b, r=r))
        np.testing.assert_array_equal(O.find_positives(a.astype(np.float32), b.astype(np.float32), r),
                                      fp(None, a.astype(np.float32), b.astype(np.float32), r=r))


# ---------------------------------------------------------------------------------------------- C_from_sparse_P
from golden.lstsq_cases import LSQ_CASES  # noqa: E402

LSQ_REF_TOL = 2e-5     # relative Frobenius distance of the reference's FP32 gelsy result to the FP64 solution


def _lsq_case(name):
    from golden.lstsq_cases import build_lstsq_inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lstsq_golden.npz"))
    P, e1, e2, k = build_lstsq_inputs(name)
    if float(g[name + "/in_sum"]) != float(e1.astype(np.float64).sum() + e2.astype(np.float64).sum() + P.sum()):
        pytest.skip("NumPy RNG stream changed: inputs differ from the ones the golden output was made from")
    return P, e1, e2, k, g[name + "/C"]


@pytest.mark.parametrize("name", list(LSQ_CASES))
def test_oracle_c_from_sparse_p_matches_reference_golden(name):
    """The FP64 restatement (same rank rule) and the reference's FP32 result agree to the reference's own rounding
    error, in full-rank, square, underdetermined and rank-deficient cases alike."""
    P, e1, e2, k, want = _lsq_case(name)
    got = O.c_from_sparse_p(P, e1[:, :k], e2[:, :k])
    assert got.shape == (k, k)
    rel = np.linalg.norm(got - want.astype(np.float64)) / np.linalg.norm(got)
    assert rel < LSQ_REF_TOL, rel


@pytest.mark.skipif(not reference_available(), reason="/root/reference only exists in the build container")
def test_oracle_c_from_sparse_p_matches_live_reference():
    import torch
    from oracle.ref_loader import load_reference_utils
    ru = load_reference_utils()
    rng = np.random.default_rng(21)
    e1 = rng.normal(0, 1, (400, 40)).astype(np.float32)
    e2 = rng.normal(0, 1, (300, 40)).astype(np.float32)
    for npos in (500, 30, 7):
        P = np.stack([rng.integers(0, 400, npos), rng.integers(0, 300, npos)], 1)
        want = ru.C_from_sparse_P(torch.from_numpy(P).type(torch.int), torch.from_numpy(e1)[:, :30],
                                  torch.from_numpy(e2)[:, :30]).numpy().astype(np.float64)
        got = O.c_from_sparse_p(P, e1[:, :30], e2[:, :30])
        assert np.linalg.norm(got - want) / np.linalg.norm(got) < LSQ_REF_TOL


# ---------------------------------------------------------------------------------------------- TEASER++ front half
def test_oracle_tims_follow_the_reference_segment_layout():
    """The layout rules the reference states in its own comments —
    segment i starts at i*N - i*(i+1)/2, has N-1-i columns, and column c of it is v[:, i+1+c] - v[:, i] with map (i, j)
    (TEASER-plusplus/teaser/src/registration.cc:489-516); edges appear in that order and degrees count both ends."""
    rng = np.random.default_rng(5)
    N = 23
    v = rng.normal(size=(3, N))
    vt, mp = O.compute_tims(v)
    assert vt.shape == (3, N * (N - 1) // 2)
    for i in range(N - 1):
        s = i * N - i * (i + 1) // 2
        np.testing.assert_array_equal(vt[:, s:s + N - 1 - i], v[:, i + 1:] - v[:, i:i + 1])
        assert (mp[0, s:s + N - 1 - i] == i).all() and (mp[1, s:s + N - 1 - i] == np.arange(i + 1, N)).all()
    dst = v + rng.normal(0, 0.05, v.shape)
    edges, degree, mask = O.scale_inlier_graph(v, dst, 0.05)
    assert 0 < len(edges) < mask.size and (edges[:, 0] < edges[:, 1]).all()
    assert (np.diff(edges[:, 0] * N + edges[:, 1]) > 0).all()
    assert degree.sum() == 2 * len(edges)


def _teaser_fixture():
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "teaser_golden.npz"))
    n = g["src"].shape[1]
    iu, ju = np.triu_indices(n, 1)
    return g["src"], g["dst"], float(g["noise_bound"]), float(g["cbar2"]), g["inlier_matrix"], g["inlier_matrix"][iu, ju]


def test_oracle_scale_inliers_match_the_reference_own_fixture():
    """Pin: the reference's own test data for the fixed-scale problem (TEASER-plusplus/test/teaser/data/
    registration_test/fixed_scale_inliers.csv, registration-test.cc:300-362) — 14028 TIMs, 2008 inliers, the closest
    statistic 3e-5 away from beta — reproduced exactly by the restatement."""
    src, dst, nb, cbar2, full, upper = _teaser_fixture()
    edges, degree, mask = O.scale_inlier_graph(src, dst, nb, cbar2)
    np.testing.assert_array_equal(mask, upper)
    assert len(edges) == int(upper.sum()) == 2008
    np.testing.assert_array_equal(degree, full.sum(1).astype(np.int32))


def test_oracle_scale_inliers_known_answers_of_the_reference_test():
    """ScaleSolverTest.FixedScale (TEASER-plusplus/test/teaser/scale-solver-test.cc:71-131) restated on its own data
    (objectIn.csv): identical measurements -> all inliers; 3*v + 10 -> none; first column * 10 -> all but the first."""
    obj = _teaser_fixture()[0]
    assert O.scale_inliers(obj, obj, 1.0, 1).all()
    assert not O.scale_inliers(obj, obj * 3 + 10, 1.0, 1).any()
    shifted = obj.copy()
    shifted[:, 0] *= 10
    m = O.scale_inliers(obj, shifted, 1.0, 1)
    assert not m[0] and m[1:].all()
