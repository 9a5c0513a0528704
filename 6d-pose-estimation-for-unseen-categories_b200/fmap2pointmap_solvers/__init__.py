"""B200-native drop-in for the reference package of the same name (`/root/reference/fmap2pointmap_solvers`).

    from fmap2pointmap_solvers import *      # scripts/eval.py:21, scripts/train.py:19

exposes the same three callables with the same signatures and return contract; underneath, Python/PyTorch host code
calls hand-written sm_100a CUDA kernels through the C-ABI of libf2p_sm100.so (include/f2p.h).  There is no CPU
fallback.  Batched and multi-GPU entry points are additions (`naive_batched`, `spatial_batched`, `distributed`), as is the
batched form of the inlier ratio that `scripts/eval.py:88` computes right after the solver (`inlier`).
"""
from ._gin import configurable as _configurable
from .spacial_filtering import spacial_filtering_fmap2pointmap
from .naive import naive_fmap2pointmap
from .batched import naive_batched, spatial_batched
from .inlier import compute_inlier_ratio, inlier_ratio_batched
from .gt_map import C_from_sparse_P, C_from_sparse_P_batched


@_configurable
def choose_fmap2pointmap_solver(solver):
    """Returns the (gin-bound) solver callable (reference: fmap2pointmap_solvers/__init__.py:6-8)."""
    return solver


__all__ = ["naive_fmap2pointmap", "spacial_filtering_fmap2pointmap", "choose_fmap2pointmap_solver",
           "naive_batched", "spatial_batched", "compute_inlier_ratio", "inlier_ratio_batched",
           "C_from_sparse_P", "C_from_sparse_P_batched"]
