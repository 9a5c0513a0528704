"""CPU tests of the host side: the C-ABI library loads and exports every symbol include/f2p.h declares, thresholds are
formed as torch forms them, pair sharding + all-gather works under gloo with world_size 2, and the product path
refuses to run without CUDA (no fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from fmap2pointmap_solvers import _lib
    hdr = open(os.path.join(ROOT, "include", "f2p.h")).read()
    declared = set(re.findall(r"\b(f2p_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.exported_symbols()), declared ^ set(_lib.exported_symbols())
    L = _lib.load()   # raises AttributeError if a declared symbol is not exported
    assert L.f2p_version() == 100
    assert ctypes.sizeof(_lib.F2PPairDesc) == 6 * 8 + 6 * 4 + 4 * 4
    assert L.f2p_project_workspace_bytes() >= 256 and L.f2p_filter_workspace_bytes(1000) > 1000 * 40


def test_no_cpu_fallback():
    import fmap2pointmap_solvers as S
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        S.naive_fmap2pointmap(torch.eye(3), torch.randn(5, 3), torch.randn(4, 3))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        S.spacial_filtering_fmap2pointmap(torch.eye(3), torch.randn(5, 3), torch.randn(4, 3), torch.rand(5, 3),
                                          torch.rand(4, 3), 1.0)


def test_public_api_surface():
    import inspect
    import fmap2pointmap_solvers as S
    assert list(inspect.signature(S.naive_fmap2pointmap).parameters) == ["C12", "evecs_x", "evecs_y", "kwargs"]
    assert list(inspect.signature(S.spacial_filtering_fmap2pointmap).parameters) == ["C12", "evecs_x", "evecs_y", "CAD", "PC", "diam_cad"]
    f = lambda: None  # noqa: E731
    assert S.choose_fmap2pointmap_solver(f) is f
    ns = {}
    exec("from fmap2pointmap_solvers import *", ns)
    assert {"naive_fmap2pointmap", "spacial_filtering_fmap2pointmap", "choose_fmap2pointmap_solver"} <= set(ns)


@pytest.mark.parametrize("diam", [17.3, np.float64(0.2345), np.float32(11.0), torch.tensor(9.7), torch.tensor(3.3, dtype=torch.float64)])
def test_thresholds_match_torch_comparison_semantics(diam):
    from fmap2pointmap_solvers.batched import thresholds
    from oracle import f2p_oracle as O
    thr = thresholds(diam)
    np.testing.assert_array_equal(np.array(thr, np.float32), O.thresholds(diam))
    # torch semantics: (x < c*diam) on an FP32 tensor is decided exactly at fl32 thresholds
    for c, t in zip((0.3, 0.15, 0.055, 0.065), thr):
        t32 = np.float32(t)
        probe = torch.tensor([np.nextafter(t32, np.float32(-np.inf)), t32, np.nextafter(t32, np.float32(np.inf))])
        want = (probe < c * diam).tolist()
        assert want == [True, False, False], (c, diam, want)


def test_shard_range_covers_everything():
    from fmap2pointmap_solvers.distributed import shard_range
    for n in (0, 1, 7, 8, 9, 4096):
        for w in (1, 2, 3, 8):
            got = []
            for r in range(w):
                lo, hi = shard_range(n, w, r)
                got += list(range(lo, hi))
            assert got == list(range(n))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from fmap2pointmap_solvers.distributed import solve_sharded
    from fmap2pointmap_solvers.synth import make_pair
    from oracle import f2p_oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shapes = [(120, 40), (300, 77), (64, 64), (200, 5), (90, 130)]
    Ps = [make_pair(a, b, 16, seed=300 + i) for i, (a, b) in enumerate(shapes)]

    def compute(kind, C, ex, ey, cad, pc, diam):   # the oracle stands in for the CUDA path: this test is about sharding
        if kind == "naive":
            return [torch.from_numpy(O.naive(c, x, y, int16_compat=False)) for c, x, y in zip(C, ex, ey)]
        return [torch.from_numpy(O.spatial(c, x, y, a, b, d)) for c, x, y, a, b, d in zip(C, ex, ey, cad, pc, diam)]

    ok = True
    try:
      from fmap2pointmap_solvers.distributed import shard_range
      lo, hi = shard_range(len(Ps), world, rank)

      def own(vals):      # only this rank's slice is populated: the other entries are placeholders
          return [v if lo <= i < hi else None for i, v in enumerate(vals)]
      for kind in ("naive", "spatial"):
        res = solve_sharded(kind, own([p.C12 for p in Ps]), own([p.evecs_x for p in Ps]), own([p.evecs_y for p in Ps]),
                            own([p.CAD for p in Ps]), own([p.PC for p in Ps]), own([p.diam_cad * 3 for p in Ps]),
                            compute=compute)
        want = compute(kind, [p.C12 for p in Ps], [p.evecs_x for p in Ps], [p.evecs_y for p in Ps],
                       [p.CAD for p in Ps], [p.PC for p in Ps], [p.diam_cad * 3 for p in Ps])
        ok = ok and len(res) == len(want) and all(torch.equal(a, b) for a, b in zip(res, want))
    except Exception as e:  # report instead of hanging the parent
        import traceback
        traceback.print_exc()
        ok = False
    q.put((rank, ok))
    dist.destroy_process_group()


def _worker_packed(rank, world, port, q):
    import torch.distributed as dist
    from fmap2pointmap_solvers.distributed import allgather_packed_maps
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ok = True
    try:
        def rank_data(r):      # what rank r holds: 3 pairs, worst-case width 40, dense [2,M'] prefixes, junk behind them
            g = torch.Generator().manual_seed(50 + r)
            counts = torch.tensor([[3, 0, 7], [11, 2, 0]][r], dtype=torch.int32)
            flat = torch.randint(-40000, 2**31 - 1, (3, 40), generator=g, dtype=torch.int64)
            return flat, counts
        flat, counts = rank_data(rank)
        maps, allc = allgather_packed_maps(flat, counts)
        ok = ok and tuple(maps.shape) == (2, 3, 22) and maps.dtype == torch.int64 and allc.dtype == torch.int32
        for r in range(world):
            f, c = rank_data(r)
            ok = ok and torch.equal(allc[r], c)
            for i in range(3):
                m = int(c[i])
                ok = ok and torch.equal(maps[r, i, :2 * m].view(2, m), f[i, :2 * m].view(2, m))
    except Exception:
        import traceback
        traceback.print_exc()
        ok = False
    q.put((rank, ok))
    dist.destroy_process_group()


def _worker_rows(rank, world, port, q):
    import numpy as np
    import torch.distributed as dist
    from fmap2pointmap_solvers.distributed import naive_row_sharded, row_cuts
    from fmap2pointmap_solvers.synth import make_pair
    from oracle import f2p_oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def local_keys(C, ex_rows, ey, row0):
        # stand-in for the CUDA kernels (this test is about the exchange): the oracle's arg-min over this rank's rows and
        # the distance it belongs to, as the library's key  bits(distance) << 32 | global row
        n2 = ey.shape[0]
        if ex_rows.shape[0] == 0:
            return torch.full((n2,), -1, dtype=torch.int64)
        idx = torch.from_numpy(O.naive(C, ex_rows, ey, int16_compat=False)[0])
        d = torch.cdist(ex_rows @ C.t(), ey)[idx, torch.arange(n2)]
        return (d.view(torch.int32).to(torch.int64) << 32) | (idx + row0)

    ok = True
    try:
        for n1, n2, k, seed in ((300, 77, 16, 1), (1000, 40, 8, 2), (100, 9, 8, 3)):      # the last: rank 1 gets no rows
            P = make_pair(n1, n2, k, seed=seed)
            ex = P.evecs_x.clone()
            if n1 >= 300:
                ex[n1 - 5] = ex[3]          # an exact duplicate in the other rank's range: the lowest global row must win
            want = torch.from_numpy(O.naive(P.C12, ex, P.evecs_y, int16_compat=False))
            for grid in (None, (2, 1), (1, 2)):          # rows split, scene points split
                got = naive_row_sharded(P.C12.unsqueeze(0), ex, P.evecs_y, local_keys=local_keys, grid=grid)
                ok = ok and got.dtype == torch.int64 and torch.equal(got, want)
        cuts = row_cuts(100, world)
        ok = ok and cuts[0] == (0, 100) and cuts[1] == (100, 100)
    except Exception:
        import traceback
        traceback.print_exc()
        ok = False
    q.put((rank, ok))
    dist.destroy_process_group()


def test_row_sharded_single_pair_gloo_world2():
    """One pair, CAD rows split over two ranks, one all-reduce(MIN) of (distance, global row) keys (the
    single-huge-pair extension): same map as the unsplit solver, ties across the split to the lowest row."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 33500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker_rows, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, True), (1, True)]


def test_row_cuts_cover_all_rows_on_tile_boundaries():
    from fmap2pointmap_solvers.distributed import block_grid, row_cuts
    for w in (1, 2, 4, 8):
        r, c = block_grid(100000, 100000, w)
        assert r * c == w
    assert block_grid(100000, 100, 8) == (8, 1) and block_grid(100, 100000, 8) == (1, 8)
    for n in (1, 127, 128, 129, 5000, 100000):
        for w in (1, 2, 3, 8):
            cuts = row_cuts(n, w)
            assert len(cuts) == w and cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:])) and all(lo % 128 == 0 or lo == n for lo, _ in cuts)


def test_packed_map_allgather_gloo_world2():
    """Counts first, then only the used prefix of every pair's result slots, int32 on the wire."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker_packed, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, True), (1, True)]


def test_pair_sharding_allgather_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, True), (1, True)]


def test_result_files_roundtrip_in_reference_layout(tmp_path):
    """`.pt` hand-off (scripts/eval.py:104-119 writer, scripts/test_RANSAC.py:331-376 reader)."""
    from fmap2pointmap_solvers import results_io as R
    samples = []
    for i in range(3):
        samples.append([{"xyz": torch.rand(10, 3)}, {"xyz": torch.rand(7, 3)},
                        {"obj_id": 5 + i, "P": torch.zeros(4, 2, dtype=torch.int64), "diam_cad": 17.0}])
    p = [torch.randint(0, 7, (2, 4 + i)) for i in range(3)]
    C = torch.rand(3, 30, 30)
    irs = [torch.tensor(0.5), 0, torch.tensor(1.0)]
    R.attach_results(samples, p, C, irs)
    paths = R.save_results(samples, str(tmp_path), start_idx=10)
    assert [os.path.basename(q) for q in paths] == ["10_obj_5.pt", "11_obj_6.pt", "12_obj_7.pt"]
    back = R.load_results(str(tmp_path))
    assert len(back) == 3
    for i, r in enumerate(back):
        P_pred, cad_xyz, ir, obj_id = R.correspondences(r)
        assert P_pred.shape == (4 + i, 2) and obj_id == 5 + i
        np.testing.assert_array_equal(P_pred, p[i].T.numpy())
        assert torch.equal(r[-1]["C_pred"], C[i]) and torch.equal(cad_xyz, samples[i][0]["xyz"])
    with pytest.raises(ValueError):
        R.attach_results(samples, [t.float() for t in p], C, irs)
