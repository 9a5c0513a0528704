"""f2p_nn_argmin_static — the allocation-free variant of kernel 2 (include/f2p.h; no counterpart in the reference)
against the existing entry point f2p_nn_argmin, through the C-ABI.

Collected LAST on purpose (file name): a failure here must not hide other tests.
Index output, so the tolerance is the one every index test here uses: bit-exact.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fmap2pointmap_solvers import _lib
    return _lib, _lib.load()


def _operands(_lib, L, n1, n2, k, seed):
    """k-major operands of one pair, made by the library's own projection kernel."""
    from fmap2pointmap_solvers.synth import make_pair
    P = make_pair(n1, n2, k, seed=seed)
    ex, ey, C = P.evecs_x.cuda(), P.evecs_y.cuda(), P.C12.cuda()
    ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
    xT = torch.empty(k, ld1, device="cuda"); nx = torch.empty(ld1, device="cuda")
    yT = torch.empty(k, ld2, device="cuda"); ny = torch.empty(ld2, device="cuda")
    ws = torch.empty(L.f2p_project_workspace_bytes() + 256, dtype=torch.uint8, device="cuda")
    _lib.check(L.f2p_project(ex.data_ptr(), n1, k, k, C.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1, ws.data_ptr(), ws.numel(), None), "f2p_project")
    _lib.check(L.f2p_project(ey.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, ws.data_ptr(), ws.numel(), None), "f2p_project")
    torch.cuda.synchronize()
    return P, xT, nx, yT, ny, ld1, ld2


def _existing(_lib, L, xT, nx, n1, ld1, yT, ny, n2, ld2, k):
    ws = torch.empty(L.f2p_nn_workspace_bytes(n1, n2, k, 0) + 256, dtype=torch.uint8, device="cuda")
    out = torch.full((n2,), -1, dtype=torch.int64, device="cuda")
    _lib.check(L.f2p_nn_argmin(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k,
                               out.data_ptr(), ws.data_ptr(), ws.numel(), None), "f2p_nn_argmin")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _static(_lib, L, xT, nx, n1, ld1, yT, ny, n2, ld2, k):
    out = torch.full((n2,), -1, dtype=torch.int64, device="cuda")     # the caller still owns inputs and outputs
    _lib.check(L.f2p_nn_argmin_static(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, k,
                                      out.data_ptr(), None), "f2p_nn_argmin_static")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_static_scratch_argmin_matches_the_existing_entry_point_twice_in_a_row(lib):
    """Two calls in a row on different pairs — the second one SMALLER, so the shared array still holds the first call's
    keys beyond its extent — then the first pair again: every result equals f2p_nn_argmin's, bit for bit."""
    _lib, L = lib
    from oracle import f2p_oracle as O
    cases = [(3000, 1500, 64, 71), (900, 333, 30, 72), (3000, 1500, 64, 71), (20000, 4096, 128, 73)]
    for n1, n2, k, seed in cases:
        P, xT, nx, yT, ny, ld1, ld2 = _operands(_lib, L, n1, n2, k, seed)
        want = _existing(_lib, L, xT, nx, n1, ld1, yT, ny, n2, ld2, k)
        got = _static(_lib, L, xT, nx, n1, ld1, yT, ny, n2, ld2, k)
        np.testing.assert_array_equal(got, want)
        assert got.min() >= 0 and got.max() < n1
        if n1 <= 3000:      # and both equal the oracle's arg-min
            np.testing.assert_array_equal(got, O.naive(P.C12, P.evecs_x, P.evecs_y, int16_compat=False)[0])


def test_static_scratch_argmin_rejects_shapes_beyond_its_limit(lib):
    _lib, L = lib
    n1, k = 512, 16
    for n2, kk in ((4097, k), (256, 129)):
        ld1, ld2 = (n1 + 127) // 128 * 128, (n2 + 127) // 128 * 128
        xT = torch.zeros(kk, ld1, device="cuda"); nx = torch.zeros(ld1, device="cuda")
        yT = torch.zeros(kk, ld2, device="cuda"); ny = torch.zeros(ld2, device="cuda")
        out = torch.empty(n2, dtype=torch.int64, device="cuda")
        rc = L.f2p_nn_argmin_static(xT.data_ptr(), nx.data_ptr(), n1, ld1, yT.data_ptr(), ny.data_ptr(), n2, ld2, kk,
                                    out.data_ptr(), None)
        assert rc == -3                                              # F2P_ENOTSUP
        msg = L.f2p_last_error().decode()
        assert "static scratch" in msg and "4096" in msg and "128" in msg
