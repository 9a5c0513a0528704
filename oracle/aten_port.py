"""ATen-op port of the reference solvers: the CPU baseline that bench.py times.  TEST/BENCH INFRASTRUCTURE ONLY.

/root/reference does not exist on the GPU box, so the reference's own Python cannot be imported there.  Its
path is nothing but a sequence of ATen calls, so this module re-issues the same calls in the same order on CPU
tensors: it has the reference's cost profile (materialised [V1,V2] distance matrix, full sort, [M,M,3]
broadcasts) and, on the same torch/MKL build, the reference's bits.  tests/test_oracle_golden.py pins it
against the golden vectors made from the real reference.

Call-site map (reference file:line -> here):
  naive.py:17-21        -> naive()                 squeeze, project, nearest row, pack with int16 row
  naive.py:33-34        -> _nearest()              cdist + argmin(-2)
  spacial_filtering.py:32-38 -> _five_nearest()    cdist + sort(-2) + first 5 + pack
  spacial_filtering.py:44-47 -> _pairwise()        broadcast difference, square, sum, sqrt
  spacial_filtering.py:52-73 -> _vote_rounds()     three rounds, 0.055 / 0.065 branch, "skipped"
"""
import torch

_ROUND_FACTORS = (0.3, 0.15)
_FINAL_FACTORS = (0.055, 0.065)
_TOPK = 5


def _as2d(C12):
    return C12.squeeze(0) if C12.dim() == 3 else C12


def _pc_index_row(count, device):
    # linspace -> int16 is what wraps for more than 32767 scene points on the CPU path
    return torch.linspace(0, count - 1, count, device=device).type(torch.int16)


def _nearest(spectral_x, spectral_y):
    return torch.cdist(spectral_x, spectral_y).argmin(dim=-2)


def naive(C12, evecs_x, evecs_y, **_ignored):
    cad_of_pc = _nearest(torch.matmul(evecs_x, _as2d(C12).t()), evecs_y)
    return torch.stack([cad_of_pc, _pc_index_row(cad_of_pc.shape[0], cad_of_pc.device)], 0)


def _five_nearest(spectral_x, spectral_y):
    order = torch.cdist(spectral_x, spectral_y).sort(dim=-2)[1]
    cad_idx = order.t()[:, :_TOPK]
    pc_idx = _pc_index_row(cad_idx.shape[0], cad_idx.device).unsqueeze(1).repeat(1, _TOPK)
    return torch.stack([cad_idx, pc_idx], 0).reshape(2, -1)


def _pairwise(points):
    return torch.sqrt(torch.sum((points[:, None] - points) ** 2, dim=2))


def _scores(CAD, PC, pairs):
    return torch.absolute(_pairwise(CAD[pairs[0]]) - _pairwise(PC[pairs[1]])).mean(0)


def _vote_rounds(CAD, PC, pairs, diam_cad, info=None):
    counts = []
    for factor in _ROUND_FACTORS:
        counts.append(pairs.shape[1])
        pairs = pairs[:, _scores(CAD, PC, pairs) < factor * diam_cad]
    counts.append(pairs.shape[1])
    s = _scores(CAD, PC, pairs)
    strict = s < _FINAL_FACTORS[0] * diam_cad
    skipped = False
    if strict.sum() == 0:
        loose = s < _FINAL_FACTORS[1] * diam_cad
        pairs = pairs[:, loose]
        if loose.sum() == 0:
            skipped = True
            print("skipped")
    else:
        pairs = pairs[:, strict]
    if info is not None:
        info["round_counts"] = counts
        info["skipped"] = skipped
    return pairs


def spatial(C12, evecs_x, evecs_y, CAD, PC, diam_cad, info=None):
    pairs = _five_nearest(torch.matmul(evecs_x, _as2d(C12).t()), evecs_y)
    if info is not None:
        info["candidates"] = pairs
    return _vote_rounds(CAD, PC, pairs, diam_cad, info)
