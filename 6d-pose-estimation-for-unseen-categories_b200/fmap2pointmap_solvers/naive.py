"""Drop-in `naive_fmap2pointmap` backed by hand-written sm_100a kernels (reference: fmap2pointmap_solvers/naive.py:5-35)."""
import torch

from ._gin import configurable
from .batched import naive_batched, naive_single


@configurable
def naive_fmap2pointmap(C12, evecs_x, evecs_y, **kwargs):
    """
    Convert a functional map to a point-to-point map on the GPU (no [V1,V2] matrix is formed).

    Args:
        C12: functional map (shape x -> shape y). Shape [K, K] or [1, K, K]
        evecs_x: eigenvectors of shape x. Shape [V1, K]
        evecs_y: eigenvectors of shape y. Shape [V2, K]
        **kwargs: ignored, as in the reference (scripts/eval.py:87 passes CAD=, PC=, diam_cad=), except
            `int16_compat=True`, which reproduces the reference's int16 wrap of row 1 for V2 > 32767.
    Returns:
        int64 [2, V2] on the device of `evecs_x`: row 0 = nearest shape-x vertex of every shape-y point in the
        spectral embedding (lowest index on exact distance ties), row 1 = the shape-y point index.
    """
    return naive_single(C12, evecs_x, evecs_y, int16_compat=bool(kwargs.get("int16_compat", False)))


def nn_query(feat_x, feat_y, dim=-2):
    """Nearest shape-x row for every shape-y row (reference: naive.py:23-35). [V2] int64."""
    if dim not in (-2, 0):
        raise ValueError("nn_query reduces over the shape-x axis (dim=-2)")
    k = feat_x.shape[1]
    eye = torch.eye(k, dtype=torch.float32, device=feat_x.device)
    # identity projection is exact: fma(e, 1, 0...) reproduces e bit-for-bit
    return naive_batched([eye], [feat_x], [feat_y])[0][0]
