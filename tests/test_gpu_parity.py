"""GPU parity tests (run on the B200 box with `-m gpu`): the CUDA path, called through the drop-in Python API and the
C-ABI, against the golden vectors of the real reference and against the CPU oracle on the same seeded inputs.
Bar: bit-exact indices / masks / intermediate FP32 bits."""
import ctypes
import os

import numpy as np
import pytest
import torch

from conftest import load_case
from golden.cases import CASES

pytestmark = pytest.mark.gpu

NAIVE = [n for n, c in CASES.items() if c["kind"] == "naive"]
SPATIAL = [n for n, c in CASES.items() if c["kind"] == "spatial"]


@pytest.fixture(scope="module")
def solvers():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import fmap2pointmap_solvers as S
    from fmap2pointmap_solvers import _lib
    _lib.load()  # must load the in-tree CUDA library: no fallback exists
    return S


@pytest.fixture(scope="module")
def oracle():
    from oracle import f2p_oracle as O
    return O


@pytest.fixture(params=["gated", "exact", "two_phase", "tc"])
def argmin_path(request, monkeypatch):
    """Runs a test through every arg-min path: the threshold-gated kernel (default for small problems), the exact kernel
    alone, the raw pass + re-check of sqrt-ambiguous columns, and the tensor-pipe screen + exact verification (default
    for large problems; forced here at every size whose k fills a swizzle row)."""
    monkeypatch.setenv("F2P_NN_PATH", request.param)
    return request.param


@pytest.fixture(params=["gated", "exact", "tc"])
def top5_path(request, monkeypatch):
    """Runs a spatial test through the threshold-gated top-5 kernel, the exact register-list kernel and the tensor-pipe
    screen + exact verification."""
    monkeypatch.setenv("F2P_NN_PATH", request.param)
    return request.param


def _cuda(inp):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) and k != "diam_cad" else v) for k, v in inp.items()}


# ---------------------------------------------------------------------------------------------- naive
@pytest.mark.parametrize("name", NAIVE)
def test_naive_matches_reference_golden(solvers, golden, name, argmin_path):
    c, inp = load_case(golden, name)
    d = _cuda(inp)
    got = solvers.naive_fmap2pointmap(d["C12"], d["evecs_x"], d["evecs_y"], int16_compat=True)
    assert got.dtype == torch.int64 and got.is_cuda and tuple(got.shape) == (2, c["n2"])
    np.testing.assert_array_equal(got.cpu().numpy(), golden[name + "/p_pred"])


def test_naive_cpu_inputs_roundtrip(solvers, golden):
    """CPU tensors in -> H2D copy -> kernels -> result back on the CPU (the reference returns on the input device)."""
    c, inp = load_case(golden, "naive_300x200_k30")
    got = solvers.naive_fmap2pointmap(inp["C12"], inp["evecs_x"], inp["evecs_y"], CAD=None, PC=None, diam_cad=1.0)
    assert not got.is_cuda
    np.testing.assert_array_equal(got.numpy(), golden["naive_300x200_k30/p_pred"])


def test_naive_default_row1_is_true_index(solvers, golden):
    c, inp = load_case(golden, "naive_wrap_200x40000_k16")
    d = _cuda(inp)
    got = solvers.naive_fmap2pointmap(d["C12"], d["evecs_x"], d["evecs_y"]).cpu().numpy()
    np.testing.assert_array_equal(got[1], np.arange(40000))
    np.testing.assert_array_equal(got[0], golden["naive_wrap_200x40000_k16/p_pred"][0])


def test_naive_requires_grad_and_noncontiguous(solvers, oracle):
    from fmap2pointmap_solvers.synth import make_pair
    P = make_pair(700, 300, 64, seed=77)
    ex64 = torch.randn(700, 64)
    ex64[:, :30] = P.evecs_x[:, :30]
    C = P.C12[:30, :30].clone().cuda().requires_grad_(True)
    ex = ex64.cuda()[:, :30]          # column-sliced view, row stride 64 (scripts/eval.py:87)
    ey = P.evecs_y.cuda()[:, :30]
    got = solvers.naive_fmap2pointmap(C, ex, ey).cpu().numpy()
    want = oracle.naive(P.C12[:30, :30], ex64[:, :30], P.evecs_y[:, :30], int16_compat=False)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("shape", [(1, 1, 1), (5, 1, 3), (127, 129, 17), (128, 128, 16), (129, 127, 15),
                                   (1000, 3, 50), (3, 1000, 50), (777, 555, 100), (2500, 300, 256)])
def test_naive_edge_shapes_vs_oracle(solvers, oracle, shape, argmin_path):
    from fmap2pointmap_solvers.synth import make_pair
    n1, n2, k = shape
    P = make_pair(n1, n2, k, seed=500 + n1 + n2 + k)
    got = solvers.naive_fmap2pointmap(P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.naive(P.C12, P.evecs_x, P.evecs_y, int16_compat=False))


def test_naive_errors(solvers):
    with pytest.raises(RuntimeError):
        solvers.naive_fmap2pointmap(torch.eye(4).cuda(), torch.randn(10, 5).cuda(), torch.randn(7, 4).cuda())
    with pytest.raises(ValueError):
        solvers.naive_fmap2pointmap(torch.eye(300).cuda(), torch.randn(10, 300).cuda(), torch.randn(7, 300).cuda())
    out = solvers.naive_fmap2pointmap(torch.eye(4).cuda(), torch.randn(10, 4).cuda(), torch.zeros(0, 4).cuda())
    assert tuple(out.shape) == (2, 0) and out.dtype == torch.int64


def test_naive_batched_ragged(solvers, oracle, argmin_path):
    from fmap2pointmap_solvers.synth import make_pair
    shapes = [(300, 200), (1000, 64), (129, 500), (5, 5), (2048, 1024)]
    Ps = [make_pair(a, b, 30, seed=40 + i) for i, (a, b) in enumerate(shapes)]
    got = solvers.naive_batched([p.C12.cuda() for p in Ps], [p.evecs_x.cuda() for p in Ps], [p.evecs_y.cuda() for p in Ps])
    for g, p in zip(got, Ps):
        np.testing.assert_array_equal(g.cpu().numpy(), oracle.naive(p.C12, p.evecs_x, p.evecs_y, int16_compat=False))


# ---------------------------------------------------------------------------------------------- spatial
@pytest.mark.parametrize("name", SPATIAL)
def test_spatial_matches_reference_golden(solvers, golden, name, capsys, top5_path):
    from fmap2pointmap_solvers.batched import spatial_batched
    c, inp = load_case(golden, name)
    d = _cuda(inp)
    res, infos = spatial_batched([d["C12"]], [d["evecs_x"]], [d["evecs_y"]], [d["CAD"]], [d["PC"]], [d["diam_cad"]],
                                 return_info=True)
    np.testing.assert_array_equal(infos[0]["candidates"].cpu().numpy(), golden[name + "/candidates"])
    got = solvers.spacial_filtering_fmap2pointmap(d["C12"], d["evecs_x"], d["evecs_y"], CAD=d["CAD"], PC=d["PC"],
                                                  diam_cad=d["diam_cad"])
    assert got.dtype == torch.int64 and got.is_cuda and got.shape[0] == 2
    np.testing.assert_array_equal(got.cpu().numpy(), golden[name + "/p_pred"])
    assert ("skipped" in capsys.readouterr().out) == bool(golden[name + "/skipped"])


@pytest.mark.parametrize("shape", [(5, 1, 4), (40, 7, 8), (300, 33, 30), (600, 257, 30), (900, 410, 50), (3000, 820, 64)])
def test_spatial_edge_shapes_vs_oracle(solvers, oracle, shape, top5_path):
    from fmap2pointmap_solvers.synth import make_pair
    n1, n2, k = shape
    P = make_pair(n1, n2, k, seed=700 + n1 + n2)
    for scale in (1.0, 4.0):
        diam = P.diam_cad * scale
        got = solvers.spacial_filtering_fmap2pointmap(P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda(), P.CAD.cuda(),
                                                      P.PC.cuda(), diam).cpu().numpy()
        np.testing.assert_array_equal(got, oracle.spatial(P.C12, P.evecs_x, P.evecs_y, P.CAD, P.PC, diam))


def test_spatial_batched_ragged(solvers, oracle, top5_path):
    from fmap2pointmap_solvers.batched import spatial_batched
    from fmap2pointmap_solvers.synth import make_pair
    shapes = [(300, 200), (1000, 64), (129, 300), (50, 5), (2048, 700)]
    Ps = [make_pair(a, b, 30, seed=60 + i) for i, (a, b) in enumerate(shapes)]
    got = spatial_batched([p.C12.cuda() for p in Ps], [p.evecs_x.cuda() for p in Ps], [p.evecs_y.cuda() for p in Ps],
                          [p.CAD.cuda() for p in Ps], [p.PC.cuda() for p in Ps], [p.diam_cad * 3 for p in Ps])
    for g, p in zip(got, Ps):
        np.testing.assert_array_equal(g.cpu().numpy(), oracle.spatial(p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC, p.diam_cad * 3))


# ---------------------------------------------------------------------------------------------- C-ABI kernels
def _ws(nbytes):
    return torch.empty(int(nbytes) + 256, dtype=torch.uint8, device="cuda")


@pytest.mark.parametrize("shape", [(300, 30), (1000, 50), (129, 7), (640, 128), (257, 256), (100, 1)])
def test_kernel1_projection_and_norm_bits(solvers, oracle, shape):
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    n, k = shape
    g = torch.Generator().manual_seed(n * 7 + k)
    e = torch.randn(n, k, generator=g) / n ** 0.5
    C = torch.linalg.qr(torch.randn(k, k, generator=g)).Q.contiguous()
    ld = (n + 127) // 128 * 128
    xT = torch.full((k, ld), float("nan"), device="cuda")
    nrm = torch.full((ld,), float("nan"), device="cuda")
    ws = _ws(L.f2p_project_workspace_bytes())
    ed, Cd = e.cuda(), C.cuda()
    rc = L.f2p_project(ed.data_ptr(), n, k, k, Cd.data_ptr(), xT.data_ptr(), nrm.data_ptr(), ld, ws.data_ptr(), ws.numel(), None)
    _lib.check(rc, "f2p_project")
    torch.cuda.synchronize()
    feat = oracle.project(e, C)
    got = (-0.5 * xT[:, :n].t()).contiguous().cpu().numpy()
    np.testing.assert_array_equal(got.view(np.int32), feat.view(np.int32))
    np.testing.assert_array_equal(nrm[:n].cpu().numpy().view(np.int32), oracle.rownorm(feat).view(np.int32))
    assert torch.isinf(nrm[n:]).all() and (xT[:, n:] == 0).all()
    # transpose-only operand (evecs_y): C == NULL
    rc = L.f2p_project(ed.data_ptr(), n, k, k, None, xT.data_ptr(), nrm.data_ptr(), ld, ws.data_ptr(), ws.numel(), None)
    _lib.check(rc, "f2p_project")
    torch.cuda.synchronize()
    np.testing.assert_array_equal(xT[:, :n].t().contiguous().cpu().numpy().view(np.int32), e.numpy().view(np.int32))
    np.testing.assert_array_equal(nrm[:n].cpu().numpy().view(np.int32), oracle.rownorm(e).view(np.int32))
    assert (nrm[n:] == 0).all()


@pytest.mark.parametrize("M", [1, 5, 31, 32, 33, 255, 256, 257, 1000, 4099, 9995])
def test_kernel3_vote_scores_bits_and_mask(solvers, oracle, M):
    from fmap2pointmap_solvers import _lib
    from fmap2pointmap_solvers.batched import thresholds
    L = _lib.load()
    g = torch.Generator().manual_seed(1000 + M)
    ncad, npc = 3 * M + 7, M + 3
    CAD = torch.rand(ncad, 3, generator=g) * 10
    Rm = torch.linalg.qr(torch.randn(3, 3, generator=g)).Q
    p0 = torch.randint(0, ncad, (M,), generator=g)
    p1 = torch.sort(torch.randint(0, npc, (M,), generator=g)).values
    PC = torch.rand(npc, 3, generator=g) * 10
    good = torch.rand(M, generator=g) < 0.6
    PC[p1[good]] = CAD[p0[good]] @ Rm.t() + 0.05 * torch.randn(int(good.sum()), 3, generator=g)
    p = torch.stack([p0, p1], 0).contiguous()
    diam = 17.3
    thr = (ctypes.c_float * 4)(*thresholds(diam))
    out = torch.zeros(2, max(M, 1), dtype=torch.int64, device="cuda")
    meta = torch.zeros(2, dtype=torch.int32, device="cuda")
    mask = torch.zeros(max(M, 1), dtype=torch.uint8, device="cuda")
    scores = torch.full((3, max(M, 1)), float("nan"), device="cuda")
    ws = _ws(L.f2p_filter_workspace_bytes(M))
    cd, pd, pp = CAD.cuda(), PC.cuda(), p.cuda()
    rc = L.f2p_spatial_filter(cd.data_ptr(), ncad, pd.data_ptr(), npc, pp.data_ptr(), M, thr, out.data_ptr(),
                              meta.data_ptr(), meta.data_ptr() + 4, mask.data_ptr(), scores.data_ptr(), ws.data_ptr(),
                              ws.numel(), None)
    _lib.check(rc, "f2p_spatial_filter")
    torch.cuda.synchronize()
    s0 = oracle.vote_scores(CAD[p0], PC[p1])
    np.testing.assert_array_equal(scores[0, :M].cpu().numpy().view(np.int32), s0.view(np.int32))
    want, info = oracle.spatial_filter(CAD, PC, p, diam, return_info=True)
    m = int(meta[0])
    assert m == want.shape[1]
    np.testing.assert_array_equal(out.flatten()[:2 * m].view(2, m).cpu().numpy(), want)   # dense [2,M']
    assert bool(int(meta[1]) & 1) == info["skipped"]
    keep = np.zeros(M, bool)
    # survivors are a subsequence of the original columns
    cols = {(int(a), int(b)) for a, b in want.T}
    assert int(mask.sum()) == m


@pytest.mark.parametrize("scale", [1e-16, 3e-19, 1e-22])
def test_kernel3_vote_tiny_nonzero_and_zero_distances(solvers, oracle, scale):
    """Squared distances that are exactly zero (shared CAD vertex / scene point) stay on the packed fast path; nonzero
    ones below 2^-101 (coordinates ~1e-16 apart and smaller, down to subnormal squares) need the full-range sqrt_rn:
    filter_init flags the chunks holding such coordinates and their tiles run the scalar variant of the row loop.
    Scores are compared bit for bit with the oracle."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    M = 700
    g = torch.Generator().manual_seed(77)
    ncad, npc = 90, 150                       # few distinct vertices: many exact-zero A distances
    CAD = torch.rand(ncad, 3, generator=g) * 10
    CAD[:40] = torch.rand(40, 3, generator=g) * scale          # a cluster of nearly coincident vertices
    CAD[40:50] = CAD[:10]                                      # and exact duplicates of some of them
    PC = torch.rand(npc, 3, generator=g) * 10
    PC[:60] = torch.rand(60, 3, generator=g) * scale * 3
    p0 = torch.randint(0, ncad, (M,), generator=g)
    p1 = torch.sort(torch.randint(0, npc, (M,), generator=g)).values
    p = torch.stack([p0, p1], 0).contiguous()
    thr = (ctypes.c_float * 4)(1e30, 1e30, 1e30, 1e30)
    out = torch.zeros(2, M, dtype=torch.int64, device="cuda")
    meta = torch.zeros(2, dtype=torch.int32, device="cuda")
    scores = torch.full((3, M), float("nan"), device="cuda")
    ws = _ws(L.f2p_filter_workspace_bytes(M))
    cd, pd, pp = CAD.cuda(), PC.cuda(), p.cuda()
    rc = L.f2p_spatial_filter(cd.data_ptr(), ncad, pd.data_ptr(), npc, pp.data_ptr(), M, thr, out.data_ptr(),
                              meta.data_ptr(), meta.data_ptr() + 4, None, scores.data_ptr(), ws.data_ptr(),
                              ws.numel(), None)
    _lib.check(rc, "f2p_spatial_filter")
    torch.cuda.synchronize()
    s0 = oracle.vote_scores(CAD[p0], PC[p1])
    for r in range(3):                        # nothing is pruned, so every round repeats the same scores
        np.testing.assert_array_equal(scores[r].cpu().numpy().view(np.int32), s0.view(np.int32))
    assert int(meta[0]) == M


@pytest.mark.parametrize("bad", [1e-12, -3e-30, 1e19])
def test_kernel3_vote_mixed_fast_and_exact_tiles(solvers, oracle, bad):
    """One 256-candidate chunk references a vertex with a coordinate outside the packed square root's proven range
    (tiny but nonzero, or huge): only the tiles touching that chunk take the scalar variant in round 0, every tile does
    in the later rounds.  Scores of all three rounds and the survivors must match the oracle bit for bit."""
    from fmap2pointmap_solvers import _lib
    from fmap2pointmap_solvers.batched import thresholds
    L = _lib.load()
    M = 1500
    g = torch.Generator().manual_seed(4242)
    ncad, npc = 900, 400
    CAD = torch.rand(ncad, 3, generator=g) * 2 - 1
    PC = torch.rand(npc, 3, generator=g) * 2 - 1
    p0 = torch.randint(1, ncad, (M,), generator=g)
    p1 = torch.sort(torch.randint(0, npc, (M,), generator=g)).values
    Rm = torch.linalg.qr(torch.randn(3, 3, generator=g)).Q
    good = torch.rand(M, generator=g) < 0.5
    PC[p1[good]] = CAD[p0[good]] @ Rm.t() + 0.01 * torch.randn(int(good.sum()), 3, generator=g)
    CAD[0] = torch.tensor([0.25, bad, -0.5])                   # vertex 0 is used by chunk 3 only
    p0[800:830] = 0
    p = torch.stack([p0, p1], 0).contiguous()
    diam = 3.0
    thr = (ctypes.c_float * 4)(*thresholds(diam))
    out = torch.zeros(2, M, dtype=torch.int64, device="cuda")
    meta = torch.zeros(2, dtype=torch.int32, device="cuda")
    scores = torch.full((3, M), float("nan"), device="cuda")
    ws = _ws(L.f2p_filter_workspace_bytes(M))
    cd, pd, pp = CAD.cuda(), PC.cuda(), p.cuda()
    rc = L.f2p_spatial_filter(cd.data_ptr(), ncad, pd.data_ptr(), npc, pp.data_ptr(), M, thr, out.data_ptr(),
                              meta.data_ptr(), meta.data_ptr() + 4, None, scores.data_ptr(), ws.data_ptr(),
                              ws.numel(), None)
    _lib.check(rc, "f2p_spatial_filter")
    torch.cuda.synchronize()
    s0 = oracle.vote_scores(CAD[p0], PC[p1])
    np.testing.assert_array_equal(scores[0].cpu().numpy().view(np.int32), s0.view(np.int32))
    want = oracle.spatial_filter(CAD, PC, p, diam)
    m = int(meta[0])
    assert m == want.shape[1] and int(meta[1]) & 2 == 0
    np.testing.assert_array_equal(out.flatten()[:2 * m].view(2, m).cpu().numpy(), want)


@pytest.mark.parametrize("where", [3, 700, 1499])
def test_kernel3_out_of_range_candidate_sets_status(solvers, where):
    """A candidate index outside CAD / PC raises status bit 1 whichever chunk holds it (per-chunk flags folded by the
    combine kernel; no zero-initialised word is involved, so the order the init CTAs run in cannot lose it)."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    M, ncad, npc = 1500, 300, 400
    g = torch.Generator().manual_seed(9)
    CAD, PC = torch.rand(ncad, 3, generator=g), torch.rand(npc, 3, generator=g)
    p0 = torch.randint(0, ncad, (M,), generator=g)
    p1 = torch.sort(torch.randint(0, npc, (M,), generator=g)).values
    p0[where] = ncad + 5
    p = torch.stack([p0, p1], 0).contiguous().cuda()
    thr = (ctypes.c_float * 4)(0.3, 0.15, 0.055, 0.065)
    out = torch.zeros(2, M, dtype=torch.int64, device="cuda")
    meta = torch.full((2,), 77, dtype=torch.int32, device="cuda")          # stale words must be overwritten
    ws = _ws(L.f2p_filter_workspace_bytes(M))
    cd, pd = CAD.cuda(), PC.cuda()
    for expect in (2, 0):
        rc = L.f2p_spatial_filter(cd.data_ptr(), ncad, pd.data_ptr(), npc, p.data_ptr(), M, thr, out.data_ptr(),
                                  meta.data_ptr(), meta.data_ptr() + 4, None, None, ws.data_ptr(), ws.numel(), None)
        _lib.check(rc, "f2p_spatial_filter")
        torch.cuda.synchronize()
        assert int(meta[1]) & 2 == expect
        p[0, where] = 1                                                        # second pass: all indices valid


def _adversarial_pair(n1, n2, k, seed):
    """Rows of feat_x sorted by DECREASING distance to every scene point (each new row beats all earlier ones), with
    runs of exactly duplicated rows: the gated kernel's candidate lists overflow and the exact fallback must take over."""
    g = torch.Generator().manual_seed(seed)
    u = torch.randn(k, generator=g)
    u /= u.norm()
    scale = torch.linspace(3.0, 1.0, n1).unsqueeze(1)
    ex = scale * u + 1e-3 * torch.randn(n1, k, generator=g)
    ex[n1 // 2: n1 // 2 + 40] = ex[n1 // 2]                      # exact ties: lowest index must win
    ey = 0.05 * torch.randn(n2, k, generator=g)
    C = torch.eye(k)
    return C, ex.contiguous(), ey.contiguous()


@pytest.mark.parametrize("shape", [(1500, 300, 16), (4000, 130, 50)])
def test_naive_overflowing_lists_fall_back_exactly(solvers, oracle, shape):
    C, ex, ey = _adversarial_pair(*shape, seed=5)
    got = solvers.naive_fmap2pointmap(C.cuda(), ex.cuda(), ey.cuda()).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.naive(C, ex, ey, int16_compat=False))


@pytest.mark.parametrize("shape", [(1500, 300, 16), (4000, 130, 50)])
def test_top5_overflowing_lists_fall_back_exactly(solvers, oracle, shape):
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    n1, n2, k = shape
    C, ex, ey = _adversarial_pair(n1, n2, k, seed=6)
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    xT = torch.empty(k, ld1, device="cuda"); nx = torch.empty(ld1, device="cuda")
    yT = torch.empty(k, ld2, device="cuda"); ny = torch.empty(ld2, device="cuda")
    ws = _ws(L.f2p_project_workspace_bytes())
    exd, eyd, Cd = ex.cuda(), ey.cuda(), C.cuda()
    _lib.check(L.f2p_project(exd.data_ptr(), n1, k, k, Cd.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, ws.data_ptr(), ws.numel(), None), "f2p_project")
    _lib.check(L.f2p_project(eyd.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, ws.data_ptr(), ws.numel(), None), "f2p_project")
    wsn = _ws(L.f2p_nn_workspace_bytes(n1, n2, k, 1))
    out = torch.empty(5 * n2, dtype=torch.int64, device="cuda")
    _lib.check(L.f2p_nn_top5(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k, out.data_ptr(),
                             wsn.data_ptr(), wsn.numel(), None), "f2p_nn_top5")
    torch.cuda.synchronize()
    want = oracle.nn_top5(oracle.project(ex, C), ey)
    np.testing.assert_array_equal(out.cpu().numpy().reshape(n2, 5), want)


# ---------------------------------------------------------------------------------------------- stacked batches
def test_stacked_batch_matches_oracle_and_list_api(solvers, oracle, capsys):
    """Uniform batch as stacked [P,...] tensors (the no-per-pair-host-work path) vs the oracle and the list API."""
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.synth import make_pair
    P = [make_pair(700, 260, 30, seed=900 + i) for i in range(5)]
    st = lambda name: torch.stack([getattr(p, name) for p in P], 0)   # noqa: E731
    C, ex, ey, cad, pc = st("C12"), st("evecs_x"), st("evecs_y"), st("CAD"), st("PC")
    diam = [p.diam_cad * (1.0 + 0.3 * i) for i, p in enumerate(P)]
    got = naive_batched(C.cuda(), ex.cuda(), ey.cuda())
    assert tuple(got.shape) == (5, 2, 260) and got.is_cuda
    for i, p in enumerate(P):
        np.testing.assert_array_equal(got[i].cpu().numpy(), oracle.naive(p.C12, p.evecs_x, p.evecs_y, int16_compat=False))
    res = spatial_batched(C.cuda(), ex.cuda(), ey.cuda(), cad.cuda(), pc.cuda(), diam, verbose=False)
    lst = spatial_batched([p.C12.cuda() for p in P], [p.evecs_x.cuda() for p in P], [p.evecs_y.cuda() for p in P],
                          [p.CAD.cuda() for p in P], [p.PC.cuda() for p in P], diam, verbose=False)
    for i, p in enumerate(P):
        want = oracle.spatial(p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC, diam[i])
        np.testing.assert_array_equal(res[i].cpu().numpy(), want)
        np.testing.assert_array_equal(lst[i].cpu().numpy(), want)
        assert res[i].is_contiguous() and res[i].dtype == torch.int64
    # CPU stacked inputs come back on the CPU; a shared C and a shared CAD are broadcast
    res_cpu = spatial_batched(C[0], ex[:1].expand(5, -1, -1), ey, cad[0], pc, diam, verbose=False)
    assert not res_cpu[0].is_cuda
    for i, p in enumerate(P):
        want = oracle.spatial(P[0].C12, P[0].evecs_x, p.evecs_y, P[0].CAD, p.PC, diam[i])
        np.testing.assert_array_equal(res_cpu[i].numpy(), want)


# ---------------------------------------------------------------------------------------------- host-buffer C-ABI
def test_host_buffer_entry_points(solvers, oracle):
    """f2p_naive_host / f2p_spatial_host: what a non-torch caller binds (numpy buffers in and out)."""
    from fmap2pointmap_solvers import _lib
    from fmap2pointmap_solvers.batched import thresholds
    from fmap2pointmap_solvers.synth import make_pair
    L = _lib.load()
    p = make_pair(900, 333, 50, seed=31)
    C, ex, ey, cad, pc = (np.ascontiguousarray(t.numpy()) for t in (p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC))
    ptr = lambda a: a.ctypes.data_as(ctypes.c_void_p)   # noqa: E731
    out = np.zeros((2, 333), np.int64)
    _lib.check(L.f2p_naive_host(ptr(C), ptr(ex), 900, ptr(ey), 333, 50, 0, ptr(out)), "f2p_naive_host")
    np.testing.assert_array_equal(out, oracle.naive(p.C12, p.evecs_x, p.evecs_y, int16_compat=False))
    thr = (ctypes.c_float * 4)(*thresholds(p.diam_cad))
    buf = np.zeros(10 * 333, np.int64)
    cnt, st = ctypes.c_int32(0), ctypes.c_int32(0)
    _lib.check(L.f2p_spatial_host(ptr(C), ptr(ex), 900, ptr(ey), 333, 50, ptr(cad), 900, ptr(pc), 333, thr,
                                  _lib.F2P_INT16_COMPAT, ptr(buf), ctypes.byref(cnt), ctypes.byref(st)), "f2p_spatial_host")
    want = oracle.spatial(p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC, p.diam_cad)
    assert cnt.value == want.shape[1]
    np.testing.assert_array_equal(buf[:2 * cnt.value].reshape(2, -1), want)


# ---------------------------------------------------------------------------------------------- BASELINE cfg5 sweep
@pytest.mark.parametrize("k", [30, 64, 128])
@pytest.mark.parametrize("n1", [2000, 10000, 50000])
def test_cfg5_sweep_both_solvers_vs_oracle(solvers, oracle, n1, k):
    """BASELINE.json configs[4]: n1 in {2k,10k,50k} x k in {30,64,128}, n2 = 2000, both solvers, bit-exact vs the oracle."""
    from fmap2pointmap_solvers.synth import make_pair
    p = make_pair(n1, 2000, k, seed=5000 + n1 + k)
    C, ex, ey, cad, pc = (t.cuda() for t in (p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC))
    got = solvers.naive_fmap2pointmap(C, ex, ey).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.naive(p.C12, p.evecs_x, p.evecs_y, int16_compat=False))
    for scale in (1.0, 2.5):       # 2.5: looser thresholds, so every round keeps a few thousand candidates
        got = solvers.spacial_filtering_fmap2pointmap(C, ex, ey, CAD=cad, PC=pc, diam_cad=p.diam_cad * scale).cpu().numpy()
        np.testing.assert_array_equal(got, oracle.spatial(p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC, p.diam_cad * scale))


# ---------------------------------------------------------------------------------------------- inlier ratio
@pytest.mark.parametrize("name", list(CASES))
def test_inlier_ratio_matches_reference_golden(solvers, golden, name):
    """compute_inlier_ratio (drop-in for utils/utils.py:81-105) on the reference's own maps vs the reference's ratio."""
    from fmap2pointmap_solvers import compute_inlier_ratio
    from golden.cases import aligned_pc, ir_threshold
    c, inp = load_case(golden, name)
    p = torch.from_numpy(golden[name + "/p_pred"])
    got = compute_inlier_ratio(p.t().cuda(), inp["CAD"].cuda(), aligned_pc(name, inp).cuda(), ir_threshold(inp))
    want = np.float32(golden[name + "/ir"])
    if p.shape[1] == 0:
        assert got == 0 and isinstance(got, int)
    else:
        assert got.is_cuda and got.dtype == torch.float32 and got.dim() == 0
        assert np.float32(got.item()).view(np.int32) == want.view(np.int32)
    # CPU tensors in -> CPU scalar out
    got_cpu = compute_inlier_ratio(p.t(), inp["CAD"], aligned_pc(name, inp), ir_threshold(inp))
    assert float(got_cpu) == float(want)


def test_inlier_ratio_batched_on_solver_output(solvers, oracle):
    """Solver output -> inlier ratios without leaving the device: naive [P,2,V2] maps and the packed spatial buffer."""
    from fmap2pointmap_solvers import inlier_ratio_batched
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.synth import make_pair
    P = [make_pair(900, 300, 30, seed=700 + i) for i in range(4)]
    st = lambda name: torch.stack([getattr(p, name) for p in P], 0).cuda()   # noqa: E731
    C, ex, ey, cad, pc = st("C12"), st("evecs_x"), st("evecs_y"), st("CAD"), st("PC")
    g = torch.Generator().manual_seed(1)
    aligned = torch.stack([p.CAD[p.perm.clamp_min(0)] + 1.0 * torch.randn(300, 3, generator=g) for p in P], 0)
    thr = [0.1 * p.diam_cad for p in P]
    maps = naive_batched(C, ex, ey)
    got = inlier_ratio_batched(maps, cad, aligned.cuda(), thr).cpu().numpy()
    for i, p in enumerate(P):
        want, _ = oracle.inlier_ratio(maps[i].t().cpu().numpy(), p.CAD, aligned[i], thr[i])
        assert got[i].view(np.int32) == np.float32(want).view(np.int32)
    diam = [p.diam_cad * 2.0 for p in P]
    flat, counts, status = spatial_batched(C, ex, ey, cad, pc, diam, verbose=False, packed=True)
    got = inlier_ratio_batched(flat, cad, aligned.cuda(), thr, counts=counts).cpu().numpy()
    cnt = counts.cpu().numpy()
    assert cnt.max() > 0
    for i, p in enumerate(P):
        m = flat[i, :2 * cnt[i]].view(2, -1)
        want, _ = oracle.inlier_ratio(m.t().cpu().numpy(), p.CAD, aligned[i], thr[i])
        assert got[i].view(np.int32) == np.float32(want).view(np.int32)
    with pytest.raises(IndexError):
        from fmap2pointmap_solvers import compute_inlier_ratio
        compute_inlier_ratio(torch.tensor([[0, 5000]]).cuda(), cad[0], aligned[0].cuda(), 1.0)


# ---------------------------------------------------------------------------------------------- find_positives
from golden.positives_cases import POS_CASES  # noqa: E402


@pytest.mark.parametrize("name", list(POS_CASES))
def test_find_positives_matches_reference_golden(solvers, oracle, name):
    """find_positives / get_overlap drop-ins (dataset/object.py:281-288, :311-317) vs the reference's own output."""
    import os
    from fmap2pointmap_solvers.positives import find_positives, find_positives_with_overlap, get_overlap
    from golden.positives_cases import build_positives_inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "positives_golden.npz"))
    a, b, r = build_positives_inputs(name)
    if float(g[name + "/in_sum"]) != float(a.astype(np.float64).sum() + b.astype(np.float64).sum()):
        pytest.skip("NumPy RNG stream changed")
    want = g[name + "/P"]
    got = find_positives(a, b, r=r)
    assert isinstance(got, np.ndarray) and got.dtype == np.int64
    np.testing.assert_array_equal(got, want)
    P, o12, o21 = find_positives_with_overlap(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), r)
    assert P.is_cuda and o12.dtype == torch.int8
    np.testing.assert_array_equal(P.cpu().numpy(), want)
    w12, w21 = get_overlap(a.shape[0], b.shape[0], want)
    np.testing.assert_array_equal(o12.cpu().numpy(), w12)
    np.testing.assert_array_equal(o21.cpu().numpy(), w21)


def test_find_positives_edge_shapes(solvers, oracle):
    from fmap2pointmap_solvers.positives import find_positives
    rng = np.random.default_rng(0)
    for n1, n2 in [(0, 5), (5, 0), (1, 1), (33, 31), (257, 1000)]:
        a, b = rng.random((n1, 3)), rng.random((n2, 3))
        got = find_positives(a, b, r=0.4)
        want = oracle.find_positives(a, b, 0.4) if n1 and n2 else np.zeros((0, 2), np.int64)
        np.testing.assert_array_equal(got, want)


# ---------------------------------------------------------------------------------------------- C_from_sparse_P
from golden.lstsq_cases import LSQ_CASES  # noqa: E402

# Tolerances (floating-point kernel): relative Frobenius distance.  The device result is the FP64 normal-equation
# solution rounded once to FP32, so it sits within FP32 rounding of the exact least-squares solution (the FP64 oracle);
# the reference's FP32 gelsy result is itself ~cond(A) * 2^-24 away from it (<= 6e-6 in the golden cases).
LSQ_TOL_VS_FP64 = 5e-7
LSQ_TOL_VS_REFERENCE = 2e-5


def _lsq_golden(name):
    import os
    from golden.lstsq_cases import build_lstsq_inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lstsq_golden.npz"))
    P, e1, e2, k = build_lstsq_inputs(name)
    if float(g[name + "/in_sum"]) != float(e1.astype(np.float64).sum() + e2.astype(np.float64).sum() + P.sum()):
        pytest.skip("NumPy RNG stream changed")
    return P, e1, e2, k, g[name + "/C"].astype(np.float64)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(LSQ_CASES))
def test_c_from_sparse_p_matches_reference_golden(solvers, oracle, name):
    """C_from_sparse_P drop-in (utils/utils.py:67-80), called as scripts/eval.py:74 calls it: int32 P, column-sliced
    bases (read in place through their row stride)."""
    from fmap2pointmap_solvers import C_from_sparse_P
    P, e1, e2, k, want = _lsq_golden(name)
    exact = oracle.c_from_sparse_p(P, e1[:, :k], e2[:, :k])
    t1, t2 = torch.from_numpy(e1).cuda(), torch.from_numpy(e2).cuda()
    C = C_from_sparse_P(torch.from_numpy(P).type(torch.int).cuda(), t1[:, :k], t2[:, :k])
    assert C.is_cuda and C.dtype == torch.float32 and tuple(C.shape) == (k, k)
    got = C.cpu().numpy().astype(np.float64)
    assert np.linalg.norm(got - exact) / np.linalg.norm(exact) < LSQ_TOL_VS_FP64
    assert np.linalg.norm(got - want) / np.linalg.norm(want) < LSQ_TOL_VS_REFERENCE
    # CPU tensors in -> CPU tensor out, same bits (deterministic: fixed-order slab partials)
    C2 = C_from_sparse_P(torch.from_numpy(P), torch.from_numpy(e1)[:, :k], torch.from_numpy(e2)[:, :k])
    assert not C2.is_cuda
    np.testing.assert_array_equal(C2.numpy().view(np.int32), C.cpu().numpy().view(np.int32))


@pytest.mark.gpu
def test_c_from_sparse_p_batched_ragged_status_and_errors(solvers, oracle):
    from fmap2pointmap_solvers import C_from_sparse_P, C_from_sparse_P_batched
    names = ["lsq_small_300x200_k30", "lsq_under_n_lt_k", "lsq_lm_5002x1369_k30", "lsq_dup_rows_rank_def"]
    cases = [_lsq_golden(n) for n in names]
    Ps = [torch.from_numpy(c[0]).cuda() for c in cases]
    e1 = [torch.from_numpy(c[1]).cuda()[:, :30] for c in cases]
    e2 = [torch.from_numpy(c[2]).cuda()[:, :30] for c in cases]
    C, status = C_from_sparse_P_batched(Ps, e1, e2, return_status=True)
    assert tuple(C.shape) == (4, 30, 30)
    assert status.cpu().tolist() == [0, 1, 0, 1]               # bit0: minimum-norm branch (rank deficient)
    for i, c in enumerate(cases):
        exact = oracle.c_from_sparse_p(c[0], c[1][:, :30], c[2][:, :30])
        assert np.linalg.norm(C[i].cpu().numpy() - exact) / np.linalg.norm(exact) < LSQ_TOL_VS_FP64
        single = C_from_sparse_P(Ps[i], e1[i], e2[i])
        np.testing.assert_array_equal(single.cpu().numpy().view(np.int32), C[i].cpu().numpy().view(np.int32))
    # no positives at all: the minimum-norm solution is 0 (torch returns zeros as well)
    Z = C_from_sparse_P(torch.zeros((0, 2), dtype=torch.int64).cuda(), e1[0], e2[0])
    assert float(Z.abs().max()) == 0.0
    # negative indices wrap like torch indexing; out-of-range raises IndexError
    Pn = Ps[0].clone()
    Pn[:, 1] -= e2[0].shape[0]
    np.testing.assert_array_equal(C_from_sparse_P(Pn, e1[0], e2[0]).cpu().numpy().view(np.int32),
                                  C[0].cpu().numpy().view(np.int32))
    Pb = Ps[0].clone()
    Pb[5, 0] = e1[0].shape[0]
    with pytest.raises(IndexError):
        C_from_sparse_P(Pb, e1[0], e2[0])
    with pytest.raises(ValueError):
        C_from_sparse_P(Ps[0], torch.zeros(300, 65).cuda(), torch.zeros(200, 65).cuda())      # k > 64


# ---------------------------------------------------------------------------------------------- TEASER++ front half
@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 2, 33, 257, 1500])
def test_teaser_front_tims_and_scale_inlier_graph(solvers, oracle, n):
    """computeTIMs and the scale-consistency inlier graph (registration.cc:481-519, :427-445, :582-590) against the
    NumPy restatement, bit for bit (FP64 differences, boolean mask, ordered edge list).  The restatement is not pinned
    to the compiled reference (no Eigen in the image)."""
    from fmap2pointmap_solvers.teaser_front import compute_tims, scale_inlier_graph
    rng = np.random.default_rng(100 + n)
    src = rng.random((3, n)) * 0.3
    Rm = np.linalg.qr(rng.normal(size=(3, 3)))[0]
    dst = Rm @ src + rng.normal(0, 0.004, (3, n)) + np.array([[0.1], [-0.2], [0.7]])
    out = rng.random(n) < 0.4                                   # 40 % outlier correspondences
    dst[:, out] = rng.random((3, int(out.sum()))) * 0.3
    vt, mp = compute_tims(src)
    wvt, wmp = oracle.compute_tims(src)
    assert vt.shape == (3, n * (n - 1) // 2) and mp.dtype == np.int32
    np.testing.assert_array_equal(vt.view(np.int64), wvt.view(np.int64))
    np.testing.assert_array_equal(mp, wmp)
    edges, degree, mask = scale_inlier_graph(src, dst, noise_bound=0.01, return_mask=True)
    we, wd, wm = oracle.scale_inlier_graph(src, dst, 0.01)
    np.testing.assert_array_equal(mask, wm)
    np.testing.assert_array_equal(edges, we)
    np.testing.assert_array_equal(degree, wd)
    if n == 257:                                                # tensors in -> tensors out, on the inputs' device
        e2, d2 = scale_inlier_graph(torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda(), 0.01)
        assert e2.is_cuda and e2.dtype == torch.int32
        np.testing.assert_array_equal(e2.cpu().numpy(), we)


# ---------------------------------------------------------------------------------------------- pipelined host batches
@pytest.mark.gpu
@pytest.mark.parametrize("chunks,P", [(1, 7), (3, 7), (16, 7), (2, 70)])
def test_host_pipelined_batches_match_one_shot_calls(solvers, oracle, chunks, P):
    """Stacked HOST batches cut into pair groups (H2D / kernels / D2H on three streams) give the very same maps as the
    one-shot device call, for both solvers, with pinned and pageable inputs and a C shared by the batch."""
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.pipeline import naive_host_pipelined, spatial_host_pipelined
    g = torch.Generator().manual_seed(31)
    n1, n2, k = 300, 130, 20                 # P = 70 in two chunks: the first group is split again (short first upload)
    ex = torch.randn(P, n1, k, generator=g)
    ey = torch.randn(P, n2, k, generator=g)
    C = torch.eye(k).repeat(P, 1, 1) + 0.05 * torch.randn(P, k, k, generator=g)
    cad = torch.rand(P, n1, 3, generator=g)
    pc = torch.rand(P, n2, 3, generator=g)
    diam = [1.7 + 0.1 * i for i in range(P)]
    want = naive_batched(C.cuda(), ex.cuda(), ey.cuda()).cpu()
    got = naive_host_pipelined(C.pin_memory(), ex.pin_memory(), ey.pin_memory(), chunks=chunks)
    assert not got.is_cuda and got.is_pinned()
    assert torch.equal(got, want)
    for i in (0, P - 1):
        np.testing.assert_array_equal(got[i].numpy(), oracle.naive(C[i], ex[i], ey[i], int16_compat=False))
    wflat, wcnt, wst = spatial_batched(C.cuda(), ex.cuda(), ey.cuda(), cad.cuda(), pc.cuda(), diam, verbose=False, packed=True)
    flat, cnt, st = spatial_host_pipelined(C, ex, ey, cad, pc, diam, chunks=chunks)          # pageable inputs
    assert torch.equal(cnt, wcnt.cpu()) and torch.equal(st, wst.cpu())
    for i in range(P):
        m = int(cnt[i])
        assert torch.equal(flat[i, :2 * m], wflat[i, :2 * m].cpu())
    # one C shared by all pairs
    want1 = naive_batched(C[:1].cuda(), ex.cuda(), ey.cuda()).cpu()
    assert torch.equal(naive_host_pipelined(C[0], ex, ey, chunks=chunks), want1)



@pytest.mark.gpu
@pytest.mark.parametrize("k,npos,kind", [(1, 5, "full"), (7, 3, "under"), (9, 40, "dup"), (33, 500, "full"), (64, 2000, "full"),
                                         (64, 20, "under"), (31, 31, "square")])
def test_c_from_sparse_p_shapes_vs_fp64_oracle(solvers, oracle, k, npos, kind):
    """Solver branches the golden cases do not reach: odd k in the Jacobi pseudo-inverse (padded round-robin ordering),
    the 64-column Gram kernel, k = 1, exactly square systems.  Reference = the FP64 oracle (same rank rule)."""
    from fmap2pointmap_solvers import C_from_sparse_P_batched
    rng = np.random.default_rng(1000 * k + npos)
    n1, n2 = 150, 120
    e1 = rng.normal(size=(n1, k)).astype(np.float32)
    e2 = rng.normal(size=(n2, k)).astype(np.float32)
    cols = rng.integers(0, 4 if kind == "dup" else n2, npos)
    if kind == "square":
        cols = rng.permutation(n2)[:npos]                      # distinct rows: full rank almost surely
    P = np.stack([rng.integers(0, n1, npos), cols], 1).astype(np.int64)
    exact = oracle.c_from_sparse_p(P, e1, e2)
    C, status = C_from_sparse_P_batched([torch.from_numpy(P).cuda()], [torch.from_numpy(e1).cuda()],
                                        [torch.from_numpy(e2).cuda()], return_status=True)
    got = C[0].cpu().numpy().astype(np.float64)
    rank_def = kind in ("under", "dup")
    assert int(status[0]) == (1 if rank_def else 0)
    tol = 5e-6 if kind == "square" else LSQ_TOL_VS_FP64        # a square random system can be ill-conditioned
    assert np.linalg.norm(got - exact) <= tol * max(np.linalg.norm(exact), 1e-30)


@pytest.mark.gpu
def test_teaser_front_matches_the_reference_own_fixture(solvers):
    """The device kernel against the reference's own fixed-scale inlier fixture (see oracle/make_golden_teaser.py)."""
    import os
    from fmap2pointmap_solvers.teaser_front import scale_inlier_graph
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "teaser_golden.npz"))
    src, dst, full = g["src"], g["dst"], g["inlier_matrix"]
    n = src.shape[1]
    iu, ju = np.triu_indices(n, 1)
    edges, degree, mask = scale_inlier_graph(src, dst, float(g["noise_bound"]), float(g["cbar2"]), return_mask=True)
    np.testing.assert_array_equal(mask, full[iu, ju])
    np.testing.assert_array_equal(edges, np.stack([iu, ju], 1)[full[iu, ju]].astype(np.int32))
    np.testing.assert_array_equal(degree, full.sum(1).astype(np.int32))


# ---------------------------------------------------------------------------------------------- stage timing aid
@pytest.mark.gpu
def test_stage_timing_reports_stages_and_leaves_results_unchanged(solvers, oracle):
    """f2p_stage_timing / f2p_stage_times (bench.py's per-kernel breakdown): every stage of a spatial batch gets a
    duration, filter_evals = sum_r M_r^2 is bracketed by the candidate and survivor counts, results are untouched."""
    from fmap2pointmap_solvers import _lib
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.synth import make_pair
    L = _lib.load()
    P = [make_pair(700, 260, 30, seed=950 + i) for i in range(3)]
    st = lambda name: torch.stack([getattr(p, name) for p in P], 0).cuda()   # noqa: E731
    C, ex, ey, cad, pc = st("C12"), st("evecs_x"), st("evecs_y"), st("CAD"), st("PC")
    diam = [p.diam_cad for p in P]
    ms = (ctypes.c_float * 9)()
    ev = ctypes.c_double(-1.0)
    assert L.f2p_stage_times(ms, ctypes.byref(ev), None) != 0          # nothing recorded yet: an error, not stale numbers
    _lib.check(L.f2p_stage_timing(1), "f2p_stage_timing")
    try:
        res = spatial_batched(C, ex, ey, cad, pc, diam, verbose=False)
        _lib.check(L.f2p_stage_times(ms, ctypes.byref(ev), None), "f2p_stage_times")
        t = list(ms)
        assert all(x >= 0.0 for x in t) and t[0] > 0 and t[1] > 0 and t[3] > 0 and t[4] > 0
        M0 = 5 * 260
        surv = sum(int(r.shape[1]) ** 2 for r in res)
        assert 3 * len(P) * M0 * M0 >= ev.value >= len(P) * M0 * M0 + 2 * surv      # M0 >= M1 >= M2 >= M'
        got = naive_batched(C, ex, ey)
        _lib.check(L.f2p_stage_times(ms, ctypes.byref(ev), None), "f2p_stage_times")
        assert ms[0] > 0 and ms[1] > 0 and all(ms[i] == 0 for i in range(2, 9)) and ev.value == 0
    finally:
        _lib.check(L.f2p_stage_timing(0), "f2p_stage_timing")
    for i, p in enumerate(P):
        np.testing.assert_array_equal(res[i].cpu().numpy(), oracle.spatial(p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC, diam[i]))
        np.testing.assert_array_equal(got[i].cpu().numpy(), oracle.naive(p.C12, p.evecs_x, p.evecs_y, int16_compat=False))


# ---------------------------------------------------------------------------------------------- screened vote
def _filter_call(L, _lib, CAD, PC, p, thr4, want_scores=False):
    """f2p_spatial_filter on one candidate list with caller-chosen thresholds -> (survivor columns [2,M'], status)."""
    M = p.shape[1]
    thr = (ctypes.c_float * 4)(*[float(t) for t in thr4])
    out = torch.zeros(2, max(M, 1), dtype=torch.int64, device="cuda")
    meta = torch.zeros(2, dtype=torch.int32, device="cuda")
    scores = torch.full((3, max(M, 1)), float("nan"), device="cuda") if want_scores else None
    ws = _ws(L.f2p_filter_workspace_bytes(M))
    cd, pd, pp = CAD.cuda(), PC.cuda(), p.cuda()
    rc = L.f2p_spatial_filter(cd.data_ptr(), CAD.shape[0], pd.data_ptr(), PC.shape[0], pp.data_ptr(), M, thr, out.data_ptr(),
                              meta.data_ptr(), meta.data_ptr() + 4, None, scores.data_ptr() if want_scores else None,
                              ws.data_ptr(), ws.numel(), None)
    _lib.check(rc, "f2p_spatial_filter")
    torch.cuda.synchronize()
    m = int(meta[0])
    return out.flatten()[:2 * m].view(2, m).cpu().numpy(), int(meta[1])


def _oracle_filter_with_thresholds(oracle, X, Y, pick):
    """The reference's three rounds (spacial_filtering.py:51-75) on gathered points, with every round's threshold chosen
    by `pick(scores)` from that round's exact scores.  Returns (kept original columns, the four thresholds)."""
    idx = np.arange(X.shape[0])
    thr = []
    for r in range(2):
        s = oracle.vote_scores(X[idx], Y[idx])
        t = pick(s, r)
        thr.append(t)
        idx = idx[s < t]
    s = oracle.vote_scores(X[idx], Y[idx])
    t2 = pick(s, 2)
    t3 = np.nextafter(t2, np.float32(np.inf)) if len(s) else t2
    keep = s < t2
    if keep.sum() == 0:
        keep = s < t3
    return idx[keep], thr + [t2, t3]


@pytest.mark.gpu
@pytest.mark.parametrize("M", [700, 2309, 4099])
@pytest.mark.parametrize("mode", ["at_score", "next_up", "lowest"])
def test_screened_vote_thresholds_on_exact_scores(solvers, oracle, M, mode):
    """Thresholds placed exactly ON a column's reference score (s < thr false for it, true one ulp higher) or on the
    lowest score (empty -> 0.065 fallback): the screened vote must hand these columns to the ordered arithmetic."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    g = torch.Generator().manual_seed(4000 + M)
    ncad, npc = 3 * M + 7, M + 3
    CAD = torch.rand(ncad, 3, generator=g) * 10
    p0 = torch.randint(0, ncad, (M,), generator=g)
    p1 = torch.sort(torch.randint(0, npc, (M,), generator=g)).values
    PC = torch.rand(npc, 3, generator=g) * 10
    good = torch.rand(M, generator=g) < 0.6
    Rm = torch.linalg.qr(torch.randn(3, 3, generator=g)).Q
    PC[p1[good]] = CAD[p0[good]] @ Rm.t() + 0.05 * torch.randn(int(good.sum()), 3, generator=g)
    p = torch.stack([p0, p1], 0).contiguous()
    X, Y = CAD[p0].numpy(), PC[p1].numpy()

    def pick(s, r):
        if mode == "lowest" and r == 2:
            return np.float32(s.min())                       # nothing below: the fallback threshold decides
        t = np.float32(np.sort(s)[(2 * len(s)) // 3])
        return np.nextafter(t, np.float32(np.inf)) if mode == "next_up" else t

    kept, thr = _oracle_filter_with_thresholds(oracle, X, Y, pick)
    got, status = _filter_call(L, _lib, CAD, PC, p, thr)
    np.testing.assert_array_equal(got, p.numpy()[:, kept])
    exact, _ = _filter_call(L, _lib, CAD, PC, p, thr, want_scores=True)      # the ordered path agrees as well
    np.testing.assert_array_equal(exact, got)


@pytest.mark.gpu
@pytest.mark.parametrize("M", [640, 1500])
def test_screened_vote_all_columns_inside_the_band(solvers, oracle, M):
    """Regular polygons: every column has the same real score, the FP32 scores differ by rounding only, so with the
    threshold on the median score EVERY column is ambiguous and the whole round is decided by the verify pass."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    ang = torch.arange(M, dtype=torch.float64) * (2 * np.pi / M)
    # centred away from the origin: a coordinate like 5*cos(pi/2) ~ 3e-16 would send the pair to the ordered kernel
    CAD = torch.stack([20 + 5 * torch.cos(ang), 30 + 5 * torch.sin(ang), torch.zeros(M, dtype=torch.float64)], 1).float()
    PC = torch.stack([20 + 5.3 * torch.cos(ang), 30 + 5.3 * torch.sin(ang), torch.ones(M, dtype=torch.float64)], 1).float()
    p = torch.stack([torch.arange(M), torch.arange(M)], 0).contiguous()
    X, Y = CAD.numpy(), PC.numpy()
    kept, thr = _oracle_filter_with_thresholds(oracle, X, Y, lambda s, r: np.float32(np.median(s)))
    assert 0 < len(kept) < M
    got, _ = _filter_call(L, _lib, CAD, PC, p, thr)
    np.testing.assert_array_equal(got, p.numpy()[:, kept])


@pytest.mark.gpu
def test_screened_vote_sqrt_approx_error_bound():
    """The error band of the screened vote (csrc/filter.cu) allows MUFU.SQRT a relative error of 2^-21: measured
    exhaustively over every float a squared distance can be, [2^-100, FLT_MAX]."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    first, last = 0x0D800000, 0x7F7FFFFF
    err = ctypes.c_double(-1.0)
    _lib.check(L.f2p_test_sqrt_approx_error(first, last - first + 1, 1, ctypes.byref(err), None), "f2p_test_sqrt_approx_error")
    print("sqrt.approx.ftz.f32 max relative error: 2^%.2f" % np.log2(err.value))
    assert 0.0 < err.value <= 2.0 ** -21
    zero = ctypes.c_double(-1.0)
    _lib.check(L.f2p_test_sqrt_approx_error(0, 1, 1, ctypes.byref(zero), None), "f2p_test_sqrt_approx_error")
    assert not (zero.value > 0)                              # sqrt(0) = 0 exactly (0/0 -> NaN compares false)


@pytest.mark.gpu
def test_screened_vote_band_is_what_protects_the_result(solvers, oracle, monkeypatch):
    """Negative control: with the error band forced to zero the screened scores decide alone, and on the all-ambiguous
    polygon case they pick a different survivor set than the reference; with the proven band they match it."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    M = 1500
    ang = torch.arange(M, dtype=torch.float64) * (2 * np.pi / M)
    CAD = torch.stack([20 + 5 * torch.cos(ang), 30 + 5 * torch.sin(ang), torch.zeros(M, dtype=torch.float64)], 1).float()
    PC = torch.stack([20 + 5.3 * torch.cos(ang), 30 + 5.3 * torch.sin(ang), torch.ones(M, dtype=torch.float64)], 1).float()
    p = torch.stack([torch.arange(M), torch.arange(M)], 0).contiguous()
    kept, thr = _oracle_filter_with_thresholds(oracle, CAD.numpy(), PC.numpy(), lambda s, r: np.float32(np.median(s)))
    _lib.check(L.f2p_test_vote_band(0.0), "f2p_test_vote_band")      # the only way to change the band: an explicit call
    try:
        unprotected, _ = _filter_call(L, _lib, CAD, PC, p, thr)
    finally:
        _lib.check(L.f2p_test_vote_band(-1.0), "f2p_test_vote_band")
    protected, _ = _filter_call(L, _lib, CAD, PC, p, thr)
    np.testing.assert_array_equal(protected, p.numpy()[:, kept])
    assert unprotected.shape != protected.shape or (unprotected != protected).any()


# ---------------------------------------------------------------------------------------------- near-ties / tensor-pipe screen
HARD = [n for n, c in CASES.items() if c.get("gen")]


@pytest.mark.parametrize("name", HARD)
def test_at_risk_columns_agree_with_the_reference(solvers, golden, name, argmin_path):
    """Census: the columns whose two smallest REFERENCE distances are within one ulp after the
    reference's own (MKL) square root are the only ones where this build's IEEE sqrt_rn could pick another row.
    Every one of them must carry the reference's index, on every arg-min path."""
    c, inp = load_case(golden, name)
    risk = golden[name + "/at_risk"]
    d = _cuda(inp)
    got = solvers.naive_fmap2pointmap(d["C12"], d["evecs_x"], d["evecs_y"], int16_compat=True).cpu().numpy()
    want = golden[name + "/p_pred"]
    np.testing.assert_array_equal(got[0][risk], want[0][risk])
    print(f"{name}: {len(risk)} of {c['n2']} columns at risk, all equal to the reference")


@pytest.mark.parametrize("gen,k", [("uniform", 2), ("grid", 8), ("grid", 32), ("offset", 8), ("offset", 32)])
def test_top5_dense_near_ties_vs_oracle(solvers, oracle, gen, k, top5_path):
    """5 nearest rows on near-tie-dense inputs through both gated paths and the tensor-pipe path: ordered by
    (distance, index) exactly like the oracle (the reference's own order of exact ties is undefined: unstable sort)."""
    from golden.cases import hard_inputs
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    n1, n2 = 6000, 700
    inp = hard_inputs(dict(n1=n1, n2=n2, k=k, seed=900 + k, gen=gen))
    ex, ey, C = inp["evecs_x"], inp["evecs_y"], inp["C12"]
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    xT = torch.empty(k, ld1, device="cuda"); nx = torch.empty(ld1, device="cuda")
    yT = torch.empty(k, ld2, device="cuda"); ny = torch.empty(ld2, device="cuda")
    ws = _ws(L.f2p_project_workspace_bytes())
    exd, eyd, Cd = ex.cuda(), ey.cuda(), C.cuda()
    _lib.check(L.f2p_project(exd.data_ptr(), n1, k, k, Cd.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, ws.data_ptr(), ws.numel(), None), "f2p_project")
    _lib.check(L.f2p_project(eyd.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, ws.data_ptr(), ws.numel(), None), "f2p_project")
    wsn = _ws(L.f2p_nn_workspace_bytes(n1, n2, k, 1))
    out = torch.empty(5 * n2, dtype=torch.int64, device="cuda")
    _lib.check(L.f2p_nn_top5(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k, out.data_ptr(),
                             wsn.data_ptr(), wsn.numel(), None), "f2p_nn_top5")
    torch.cuda.synchronize()
    np.testing.assert_array_equal(out.cpu().numpy().reshape(n2, 5), oracle.nn_top5(oracle.project(ex, C), ey))


def _tc_stats():
    from fmap2pointmap_solvers import _lib
    a = (ctypes.c_ulonglong * 8)()
    _lib.check(_lib.load().f2p_tc_last_stats(a), "f2p_tc_last_stats")
    return dict(recorded=a[0], overflowed=a[1], verified=a[2], violations=a[3])


def test_tensor_pipe_screen_counters_and_lower_bound(solvers, oracle, monkeypatch):
    """The screen's run-time self-check: every re-evaluated candidate's tensor-pipe lower bound must be <= the exact FP32
    value (a violation would mean the error-band assumptions of csrc/tc.cu do not hold on this device), and on separated
    data about one candidate per column survives to the exact evaluation."""
    from fmap2pointmap_solvers.synth import make_pair
    monkeypatch.setenv("F2P_NN_PATH", "tc")
    for (n1, n2, k) in [(3000, 1000, 64), (2500, 700, 128), (4000, 333, 30)]:
        P = make_pair(n1, n2, k, seed=n1 + k)
        got = solvers.naive_fmap2pointmap(P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()).cpu().numpy()
        np.testing.assert_array_equal(got, oracle.naive(P.C12, P.evecs_x, P.evecs_y, int16_compat=False))
        s = _tc_stats()
        assert s["violations"] == 0 and s["overflowed"] == 0
        assert n2 <= s["verified"] <= 3 * n2, s


def test_tensor_pipe_screen_falls_back_when_everything_is_in_band(solvers, golden, monkeypatch):
    """offset inputs (100 + 1e-2*rand): every squared distance lies inside the screen's error band, the candidate lists
    overflow, the column tiles are flagged and the FP32 kernels produce the (reference-identical) result."""
    monkeypatch.setenv("F2P_NN_PATH", "tc")
    name = "naive_offset_k32_20000x1000"
    c, inp = load_case(golden, name)
    d = _cuda(inp)
    got = solvers.naive_fmap2pointmap(d["C12"], d["evecs_x"], d["evecs_y"], int16_compat=True).cpu().numpy()
    np.testing.assert_array_equal(got, golden[name + "/p_pred"])
    s = _tc_stats()
    assert s["overflowed"] > 0 and s["violations"] == 0, s


def test_tensor_pipe_screen_rejects_operands_outside_the_proven_range(solvers, oracle, monkeypatch):
    """Squared norms above 2^100: the packing kernel marks the pair, the screen skips it and the FP32 kernels run."""
    from fmap2pointmap_solvers.synth import make_pair
    monkeypatch.setenv("F2P_NN_PATH", "tc")
    P = make_pair(2000, 500, 32, seed=4242)
    ex, ey = P.evecs_x * 3e16, P.evecs_y * 3e16
    got = solvers.naive_fmap2pointmap(P.C12.cuda(), ex.cuda(), ey.cuda()).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.naive(P.C12, ex, ey, int16_compat=False))
    s = _tc_stats()
    assert s["recorded"] == 0 and s["verified"] == 0, s


def test_top5_bootstrap_ties_keep_the_lowest_indices(solvers, oracle, monkeypatch):
    """duplicated CAD rows INSIDE the first 128-row tile, at positions the gated kernel's interleaved
    bootstrap visits out of index order (0,4,8,... before 1,5,9,...): the 5 nearest must be the (distance, index)-least
    — e.g. rows {0,1,4,8,12}, not {0,4,8,12,16} — on the gated path as on every other."""
    from fmap2pointmap_solvers import _lib
    L = _lib.load()
    n1, n2, k = 640, 130, 24
    g = torch.Generator().manual_seed(321)
    ex = torch.randn(n1, k, generator=g)
    ey = 0.01 * torch.randn(n2, k, generator=g)
    dup = [0, 4, 8, 12, 16, 1, 5, 21, 33]
    ex[dup] = 0.001 * torch.randn(1, k, generator=g)          # nine identical rows, by far the nearest to every scene point
    C = torch.eye(k)
    want = oracle.nn_top5(oracle.project(ex, C), ey)
    assert (np.sort(want, 1) == np.array([0, 1, 4, 5, 8])).all()
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    for path in ("gated", "exact", "tc"):
        monkeypatch.setenv("F2P_NN_PATH", path)
        xT = torch.empty(k, ld1, device="cuda"); nx = torch.empty(ld1, device="cuda")
        yT = torch.empty(k, ld2, device="cuda"); ny = torch.empty(ld2, device="cuda")
        ws = _ws(L.f2p_project_workspace_bytes())
        exd, eyd, Cd = ex.cuda(), ey.cuda(), C.cuda()
        _lib.check(L.f2p_project(exd.data_ptr(), n1, k, k, Cd.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, ws.data_ptr(), ws.numel(), None), "f2p_project")
        _lib.check(L.f2p_project(eyd.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, ws.data_ptr(), ws.numel(), None), "f2p_project")
        wsn = _ws(L.f2p_nn_workspace_bytes(n1, n2, k, 1))
        out = torch.empty(5 * n2, dtype=torch.int64, device="cuda")
        _lib.check(L.f2p_nn_top5(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k, out.data_ptr(),
                                 wsn.data_ptr(), wsn.numel(), None), "f2p_nn_top5")
        torch.cuda.synchronize()
        np.testing.assert_array_equal(out.cpu().numpy().reshape(n2, 5), want, err_msg=path)


def test_saved_results_hold_only_their_own_map(solvers, tmp_path):
    """p_pred views of the batch-wide result buffer must not drag that buffer into every .pt file."""
    from fmap2pointmap_solvers.batched import spatial_batched
    from fmap2pointmap_solvers.results_io import attach_results, load_results, save_results
    from fmap2pointmap_solvers.synth import make_pair
    Ps = [make_pair(600, 300, 30, seed=800 + i) for i in range(3)]
    maps = spatial_batched([p.C12.cuda() for p in Ps], [p.evecs_x.cuda() for p in Ps], [p.evecs_y.cuda() for p in Ps],
                           [p.CAD.cuda() for p in Ps], [p.PC.cuda() for p in Ps], [p.diam_cad * 3 for p in Ps], verbose=False)
    samples = [[{"xyz": p.CAD}, {"xyz": p.PC}, {"obj_id": i}] for i, p in enumerate(Ps)]
    attach_results(samples, maps, [p.C12 for p in Ps], [0.5] * 3)
    paths = save_results(samples, str(tmp_path))
    for path, m in zip(paths, maps):
        assert os.path.getsize(path) < 200_000 + m.numel() * 8
    for r, m in zip(load_results(str(tmp_path), map_location="cpu"), maps):
        pp = r[-1]["p_pred"]
        assert pp.untyped_storage().nbytes() == pp.numel() * 8
        assert torch.equal(pp, m.cpu())


def test_shared_cad_side_operands_match_per_pair_operands(solvers, oracle):
    """Runs of consecutive pairs sharing evecs_x / CAD ([G,n1,k] for P pairs, SURVEY 8(d): one CAD per 16 scenes): same
    maps as the batch with every pair's operands spelled out, on the device and through the host pipeline."""
    from fmap2pointmap_solvers.batched import naive_batched, spatial_batched
    from fmap2pointmap_solvers.pipeline import naive_host_pipelined, spatial_host_pipelined
    from fmap2pointmap_solvers.synth import make_pair
    G, rep, n1, n2, k = 3, 8, 700, 260, 30
    objs = [make_pair(n1, n2, k, seed=600 + g) for g in range(G)]
    gen = torch.Generator().manual_seed(7)
    ex = torch.stack([o.evecs_x for o in objs]); cad = torch.stack([o.CAD for o in objs])
    P = G * rep
    C = torch.stack([torch.linalg.qr(torch.randn(k, k, generator=gen)).Q for _ in range(P)])
    ey = torch.stack([objs[i // rep].evecs_y + 0.002 * torch.randn(n2, k, generator=gen) for i in range(P)])
    pc = torch.stack([objs[i // rep].PC + 0.05 * torch.randn(n2, 3, generator=gen) for i in range(P)])
    diam = [objs[i // rep].diam_cad * 3 for i in range(P)]
    full_ex, full_cad = ex.repeat_interleave(rep, 0), cad.repeat_interleave(rep, 0)
    want_n = naive_batched(C.cuda(), full_ex.cuda(), ey.cuda()).cpu()
    got_n = naive_batched(C.cuda(), ex.cuda(), ey.cuda()).cpu()
    assert torch.equal(got_n, want_n)
    np.testing.assert_array_equal(got_n[5].numpy(), oracle.naive(C[5], full_ex[5], ey[5], int16_compat=False))
    assert torch.equal(naive_host_pipelined(C, ex, ey, chunks=3), want_n)
    want_s = spatial_batched(C.cuda(), full_ex.cuda(), ey.cuda(), full_cad.cuda(), pc.cuda(), diam, verbose=False)
    got_s = spatial_batched(C.cuda(), ex.cuda(), ey.cuda(), cad.cuda(), pc.cuda(), diam, verbose=False)
    assert len(got_s) == P and all(torch.equal(a, b) for a, b in zip(got_s, want_s))
    flat, counts, status = spatial_host_pipelined(C, ex, ey, cad, pc, diam, chunks=3)
    for i in range(P):
        m = int(counts[i])
        assert torch.equal(flat[i, :2 * m].view(2, m), want_s[i].cpu())


@pytest.mark.parametrize("path", ["gated", "tc"])
def test_blocked_host_pipeline_matches_the_one_shot_call(solvers, oracle, path, monkeypatch):
    """One pair cut into row x column blocks with persistent (distance, global row) keys: same map as the one-shot
    solver, including exact ties between rows that land in DIFFERENT row blocks (lowest global row wins) and ragged blocks."""
    from fmap2pointmap_solvers.pipeline import naive_host_blocked
    from fmap2pointmap_solvers.synth import make_pair
    monkeypatch.setenv("F2P_NN_PATH", path)
    P = make_pair(3000, 1400, 64, seed=91)
    ex = P.evecs_x.clone()
    ex[2500:2600] = ex[100:200]          # exact duplicates across row blocks (block size 1024 / 1536)
    ex[1700] = ex[5]
    want = solvers.naive_fmap2pointmap(P.C12.cuda(), ex.cuda(), P.evecs_y.cuda()).cpu()
    np.testing.assert_array_equal(want.numpy(), oracle.naive(P.C12, ex, P.evecs_y, int16_compat=False))
    for xb, yb in ((2, 2), (3, 1), (1, 3), (4, 5)):
        got = naive_host_blocked(P.C12, ex.pin_memory(), P.evecs_y, xb=xb, yb=yb)
        assert not got.is_cuda and torch.equal(got, want), (xb, yb)
    got16 = naive_host_blocked(P.C12.unsqueeze(0), ex, P.evecs_y, int16_compat=True)
    assert torch.equal(got16, solvers.naive_fmap2pointmap(P.C12.cuda(), ex.cuda(), P.evecs_y.cuda(), int16_compat=True).cpu())


def test_pair_stream_matches_single_calls_on_ragged_pairs(solvers):
    """The eval loop as a stream (3 pairs in flight, staging slots reused and regrown): every result equals the single
    call's, in order — pairs of different sizes and K-compatible shapes, a 3-D C12, an empty scene, FP64 inputs."""
    from fmap2pointmap_solvers.pipeline import naive_pair_stream
    from fmap2pointmap_solvers.synth import make_pair
    shapes = [(700, 300, 20), (2500, 1100, 20), (90, 40, 20), (1500, 0, 20), (3000, 2000, 20), (700, 300, 20), (64, 1, 20)]
    ps = [make_pair(n1, max(n2, 1), k, seed=300 + i) for i, (n1, n2, k) in enumerate(shapes)]
    feed = []
    for i, (P, (n1, n2, k)) in enumerate(zip(ps, shapes)):
        C = P.C12.unsqueeze(0) if i % 2 else P.C12
        ex = P.evecs_x.double() if i == 2 else (P.evecs_x.pin_memory() if i % 3 == 0 else P.evecs_x)
        feed.append((C, ex, P.evecs_y[:n2]))
    for depth in (1, 3):
        got = list(naive_pair_stream(iter(feed), depth=depth))
        assert len(got) == len(feed)
        for (C, ex, ey), g in zip(feed, got):
            want = solvers.naive_fmap2pointmap(C.cuda(), ex.float().cuda(), ey.cuda()).cpu()
            assert not g.is_cuda and g.dtype == torch.int64 and torch.equal(g, want)


def test_spatial_pair_stream_matches_single_calls(solvers):
    """Spatial solver over the native pair stream: ragged samples, more samples than slots, every map equal to the
    single call's (same survivors in the same order), including a sample whose thresholds leave nothing."""
    from fmap2pointmap_solvers.pipeline import spatial_pair_stream
    from fmap2pointmap_solvers.synth import make_pair
    shapes = [(900, 333, 30), (2000, 800, 30), (400, 150, 30), (1200, 500, 30), (900, 333, 30)]
    ps = [make_pair(n1, n2, k, seed=400 + i) for i, (n1, n2, k) in enumerate(shapes)]
    diams = [p.diam_cad for p in ps]
    diams[2] = diams[2] * 1e-6           # nothing survives
    feed = [(p.C12, p.evecs_x, p.evecs_y, p.CAD, p.PC, d) for p, d in zip(ps, diams)]
    for depth in (1, 2, 4):
        got = list(spatial_pair_stream(iter(feed), depth=depth))
        assert len(got) == len(feed)
        for (C, ex, ey, cad, pc, d), g in zip(feed, got):
            want = solvers.spacial_filtering_fmap2pointmap(C.cuda(), ex.cuda(), ey.cuda(), CAD=cad.cuda(), PC=pc.cuda(),
                                                           diam_cad=d).cpu()
            assert g.dtype == torch.int64 and g.shape == want.shape and torch.equal(g, want)
    assert got[2].shape == (2, 0)


def test_pair_streams_fail_cleanly_and_recover(solvers, capsys):
    """A bad sample in the middle of a stream raises (after the earlier results were delivered), an abandoned generator
    leaves nothing behind, and the next stream on the same thread works; the spatial stream prints "skipped" like the
    single call and raises IndexError for a CAD that is shorter than the eigenbasis."""
    from fmap2pointmap_solvers.pipeline import naive_pair_stream, spatial_pair_stream
    from fmap2pointmap_solvers.synth import make_pair
    P = make_pair(600, 200, 20, seed=77)
    good = (P.C12, P.evecs_x, P.evecs_y)
    bad = (P.C12, P.evecs_x[:, :19], P.evecs_y)
    want = solvers.naive_fmap2pointmap(P.C12.cuda(), P.evecs_x.cuda(), P.evecs_y.cuda()).cpu()
    got = []
    with pytest.raises(RuntimeError):
        for r in naive_pair_stream(iter([good, good, bad, good]), depth=2):
            got.append(r)
    assert all(torch.equal(r, want) for r in got)
    it = naive_pair_stream(iter([good] * 5), depth=3)
    assert torch.equal(next(it), want)
    it.close()                                             # abandoned with pairs in flight
    assert all(torch.equal(r, want) for r in naive_pair_stream(iter([good] * 4), depth=3))
    capsys.readouterr()
    res = list(spatial_pair_stream(iter([(P.C12, P.evecs_x, P.evecs_y, P.CAD, P.PC, P.diam_cad * 1e-6)])))
    assert res[0].shape == (2, 0) and "skipped" in capsys.readouterr().out
    with pytest.raises(IndexError):
        list(spatial_pair_stream(iter([(P.C12, P.evecs_x, P.evecs_y, P.CAD[:300], P.PC, P.diam_cad)])))


@pytest.mark.parametrize("path", ["gated", "tc"])
@pytest.mark.parametrize("gen,k", [("grid", 32), ("offset", 24), ("uniform", 20)])
def test_blocked_and_sharded_argmin_on_tie_dense_inputs(solvers, oracle, path, gen, k, monkeypatch):
    """Exact multi-way ties and cancellation-heavy inputs cut into row x column blocks: every later row block starts from
    thresholds seeded by the keys its scene block already holds (T0 = up((d + 1 ulp)^2)), so a tie with a held best — in
    the same sqrt_rn class, lower or higher row — must still be recorded and resolved by (distance, global row).  The same
    keys path serves the multi-GPU row split (here on a world of one)."""
    from fmap2pointmap_solvers.distributed import naive_row_sharded
    from fmap2pointmap_solvers.pipeline import naive_host_blocked
    from golden.cases import hard_inputs
    monkeypatch.setenv("F2P_NN_PATH", path)
    inp = hard_inputs(dict(n1=7000, n2=1500, k=k, seed=4242 + k, gen=gen))
    C, ex, ey = inp["C12"], inp["evecs_x"], inp["evecs_y"]
    want = oracle.naive(C, ex, ey, int16_compat=False)
    one = solvers.naive_fmap2pointmap(C.cuda(), ex.cuda(), ey.cuda()).cpu().numpy()
    np.testing.assert_array_equal(one, want)
    for xb, yb in ((2, 2), (5, 1), (3, 2)):
        np.testing.assert_array_equal(naive_host_blocked(C, ex, ey, xb=xb, yb=yb).numpy(), want, err_msg=f"{xb}x{yb}")
    # row blocks arriving in DESCENDING order: the lower rows come last and must still win their ties
    from fmap2pointmap_solvers import _lib
    L = _lib.load(); dev = torch.device("cuda", torch.cuda.current_device()); st = _lib.stream_ptr(dev)
    n1, n2 = ex.shape[0], ey.shape[0]
    Cd, exd, eyd = C.cuda(), ex.cuda(), ey.cuda()
    ld2 = (n2 + 127) // 128 * 128
    yT = torch.empty(k, ld2, device=dev); ny = torch.empty(ld2, device=dev)
    pb = (L.f2p_project_workspace_bytes() + 255) // 256 * 256
    cuts = [(0, 2304), (2304, 4608), (4608, n1)]
    ws = torch.empty(pb + max(L.f2p_nn_workspace_bytes(b - a, n2, k, 0) for a, b in cuts) + 1024, dtype=torch.uint8, device=dev)
    _lib.check(L.f2p_project(eyd.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, ws.data_ptr(), pb, st), "f2p_project")
    keys = torch.full((ld2,), -1, dtype=torch.int64, device=dev)
    for a, b in reversed(cuts):
        ld1 = (b - a + 127) // 128 * 128
        xT = torch.empty(k, ld1, device=dev); nx = torch.empty(ld1, device=dev)
        _lib.check(L.f2p_project(exd[a:b].data_ptr(), b - a, k, k, Cd.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, ws.data_ptr(), pb, st), "f2p_project")
        wn = ws[pb:]
        _lib.check(L.f2p_nn_argmin_keys(xT.data_ptr(), nx.data_ptr(), b - a, ld1, a, yT.data_ptr(), ny.data_ptr(), n2, ld2, k,
                                        keys.data_ptr(), wn.data_ptr(), wn.numel(), st), "f2p_nn_argmin_keys")
    np.testing.assert_array_equal((keys[:n2] & 0xFFFFFFFF).cpu().numpy(), want[0])
    got = naive_row_sharded(Cd, exd, eyd)                                 # not under torch.distributed: a world of one
    np.testing.assert_array_equal(got.cpu().numpy(), want)


def test_batch_mixing_screened_and_ordered_pairs(solvers, oracle):
    """One batch, one launch sequence: pairs 0, 2, 3 vote through the screened kernel, pair 1 (every CAD vertex carries a
    tiny nonzero coordinate, outside the packed square root's proven range) is left to the ordered kernel through its tile
    list.  Every pair's survivors equal the oracle's; the ragged shapes make the pairs' tile counts differ."""
    from fmap2pointmap_solvers.batched import spatial_batched
    from fmap2pointmap_solvers.synth import make_pair
    shapes = [(900, 333), (1200, 410), (700, 150), (1500, 600)]
    ps = [make_pair(n1, n2, 30, seed=610 + i) for i, (n1, n2) in enumerate(shapes)]
    cads = [p.CAD.clone() for p in ps]
    cads[1][:, 1] = 1e-30 * (1.0 + torch.arange(cads[1].shape[0], dtype=torch.float32))
    diams = [p.diam_cad * 2.5 for p in ps]                 # loose thresholds: all three rounds keep candidates
    got = spatial_batched([p.C12.cuda() for p in ps], [p.evecs_x.cuda() for p in ps], [p.evecs_y.cuda() for p in ps],
                          [c.cuda() for c in cads], [p.PC.cuda() for p in ps], diams, verbose=False)
    for p, c, d, g in zip(ps, cads, diams, got):
        want = oracle.spatial(p.C12, p.evecs_x, p.evecs_y, c, p.PC, d)
        assert g.shape[1] == want.shape[1]
        np.testing.assert_array_equal(g.cpu().numpy(), want)
    assert sum(g.shape[1] for g in got) > 0
