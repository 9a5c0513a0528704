"""Drop-in `spacial_filtering_fmap2pointmap` backed by hand-written sm_100a kernels
(reference: fmap2pointmap_solvers/spacial_filtering.py:4-75)."""
from ._gin import configurable
from .batched import spatial_batched


@configurable
def spacial_filtering_fmap2pointmap(C12, evecs_x, evecs_y, CAD, PC, diam_cad):
    """
    Functional map -> 5 nearest CAD vertices per scene point -> three rounds of spatial-consistency voting.

    Args:
        C12: functional map. Shape [K, K] or [1, K, K]
        evecs_x: eigenvectors of the CAD model. Shape [V1, K]
        evecs_y: eigenvectors of the partial scene cloud. Shape [V2, K]
        CAD: CAD vertex positions [>=V1, 3];  PC: scene point positions [>=V2, 3] (zero-padded rows allowed)
        diam_cad: CAD diameter (Python/NumPy float or 0-d tensor)
    Returns:
        int64 [2, M'] (0 <= M' <= 5*V2) on the device of `evecs_x`: surviving (CAD index, PC index) pairs in their
        original order.  Prints "skipped" and returns [2, 0] when nothing survives, as the reference does.
    """
    return spatial_batched([C12], [evecs_x], [evecs_y], [CAD], [PC], [diam_cad], int16_compat=True)[0]
