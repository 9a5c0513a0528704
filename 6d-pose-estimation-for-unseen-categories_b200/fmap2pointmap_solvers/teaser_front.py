"""Pairwise-consistency front half of the TEASER++ solve on the device: the step right after the
point map in `scripts/test_teaser.py` (`solver.solve(src, dst)` at :431, configured at :363-370).

`compute_tims(v)` mirrors `RobustRegistrationSolver::computeTIMs` (TEASER-plusplus/teaser/src/registration.cc:481-519):
`v` is `[3,N]` as TEASER takes it, the result `(vtilde [3, N(N-1)/2] float64, map [2, N(N-1)/2] int32)`.
`scale_inlier_graph(src, dst, noise_bound, cbar2)` fuses computeTIMs for both clouds with
`ScaleInliersSelector::solveForScale` (:427-445) and the inlier-graph edge loop (:582-590) and returns the edge list in
the reference's `addEdge` order plus the vertex degrees, without materialising the TIMs.  NumPy in -> NumPy out,
tensors in -> tensors out.  The arithmetic runs in csrc/tims.cu (FP64); there is no CPU path.
Parity: the inlier mask is pinned to the reference's own fixed-scale fixture (tests/golden/teaser_golden.npz, from
TEASER-plusplus/test/teaser/data/registration_test); the 3-term summation order is an assumption (csrc/tims.cu).
"""
import math

import numpy as np
import torch

from . import _lib


def _points(v, device):
    """[3,N] (TEASER's layout) -> float64 [N,3] contiguous on the device."""
    t = torch.from_numpy(np.asarray(v)) if not isinstance(v, torch.Tensor) else v.detach()
    if t.dim() != 2 or t.shape[0] != 3:
        raise ValueError(f"expected a [3,N] matrix, got {tuple(t.shape)}")
    return t.to(device=device, dtype=torch.float64).t().contiguous()


def compute_tims(v):
    lib = _lib.load()
    as_numpy = not isinstance(v, torch.Tensor)
    device = _lib.cuda_device_for(v)
    with torch.cuda.device(device):
        p = _points(v, device)
        n = int(p.shape[0])
        T = n * (n - 1) // 2
        vt = torch.empty((T, 3), dtype=torch.float64, device=device)
        mp = torch.empty((T, 2), dtype=torch.int32, device=device)
        rc = lib.f2p_compute_tims(p.data_ptr(), n, vt.data_ptr(), mp.data_ptr(), _lib.stream_ptr(device))
        _lib.check(rc, "f2p_compute_tims")
        vt, mp = vt.t(), mp.t()                      # [3,T] / [2,T] views: Eigen's column-major matrices
        if as_numpy:
            return vt.cpu().numpy(), mp.cpu().numpy()
        return vt.to(v.device), mp.to(v.device)


def scale_inlier_graph(src, dst, noise_bound, cbar2=1.0, return_mask=False):
    """(edges int32 [E,2], degree int32 [N]) and, with `return_mask`, scale_inliers_mask_ (bool [N(N-1)/2], TIM order)."""
    lib = _lib.load()
    as_numpy = not isinstance(src, torch.Tensor)
    device = _lib.cuda_device_for(src, dst)
    beta = 2.0 * float(noise_bound) * math.sqrt(float(cbar2))          # registration.cc:440
    with torch.cuda.device(device):
        s, d = _points(src, device), _points(dst, device)
        if s.shape != d.shape:
            raise ValueError("src and dst must have the same number of points")
        n = int(s.shape[0])
        total = torch.zeros(1, dtype=torch.int64, device=device)
        degree = torch.zeros(n, dtype=torch.int32, device=device)
        mask = torch.empty(n * (n - 1) // 2, dtype=torch.uint8, device=device) if return_mask else None
        ws = torch.empty(lib.f2p_tim_workspace_bytes(n) + 256, dtype=torch.uint8, device=device)
        st = _lib.stream_ptr(device)
        rc = lib.f2p_tim_graph_count(s.data_ptr(), d.data_ptr(), n, beta, total.data_ptr(), degree.data_ptr(),
                                     mask.data_ptr() if return_mask else None, ws.data_ptr(), ws.numel(), st)
        _lib.check(rc, "f2p_tim_graph_count")
        e = int(total.item())                        # the one host sync: E sizes the edge list
        edges = torch.empty((e, 2), dtype=torch.int32, device=device)
        rc = lib.f2p_tim_graph_write(s.data_ptr(), d.data_ptr(), n, beta, edges.data_ptr(), e, ws.data_ptr(),
                                     ws.numel(), st)
        _lib.check(rc, "f2p_tim_graph_write")
        out = (edges, degree) + ((mask.bool(),) if return_mask else ())
        if as_numpy:
            return tuple(t.cpu().numpy() for t in out)
        return tuple(t.to(src.device) for t in out)
