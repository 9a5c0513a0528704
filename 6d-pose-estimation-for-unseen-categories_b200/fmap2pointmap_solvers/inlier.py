"""Inlier ratio of predicted correspondences (the step right after the solver in `scripts/eval.py:88`).

`compute_inlier_ratio` keeps the reference's signature and return contract (`utils/utils.py:81-105`): `pred_corr` is
the `[M,2]` index tensor `p_pred.t()`, the result a 0-d float32 tensor on the input device (Python `0` for an empty
map).  `inlier_ratio_batched` scores a whole batch of maps with one launch sequence; together with
`spatial_batched(..., packed=True)` the solver output never leaves the device between the two steps.
All arithmetic runs in `inlier_count_kernel` (csrc/inlier.cu) through the C-ABI; there is no CPU path.
"""
import ctypes

import numpy as np
import torch

from . import _lib


def _thr32(threshold):
    """`sq_dist < threshold` compares in FP32: a Python / NumPy number is rounded to FP32, a tensor is cast."""
    if isinstance(threshold, torch.Tensor):
        return float(threshold.detach().to(torch.float32).cpu().item())
    return float(np.float32(float(threshold)))


def _descs(n, ncad, npc, widths, cad_step, pc_step, out_offs):
    d = (_lib.F2PPairDesc * max(n, 1))()
    for i in range(n):
        d[i].ncad, d[i].npc, d[i].n2 = ncad, npc, int(widths[i])
        d[i].cad_off, d[i].pc_off, d[i].out_off = i * cad_step, i * pc_step, int(out_offs[i])
    return d


def compute_inlier_ratio(pred_corr, CAD, PC_aligned, threshold):
    """Drop-in for `utils.compute_inlier_ratio` (utils/utils.py:81-105)."""
    total_corr = len(pred_corr)
    if total_corr == 0:
        return 0
    if pred_corr.dim() != 2 or pred_corr.shape[1] != 2:
        raise RuntimeError(f"pred_corr must be [M,2], got {tuple(pred_corr.shape)}")
    lib = _lib.load()
    device = _lib.cuda_device_for(pred_corr, CAD, PC_aligned)
    with torch.cuda.device(device):
        maps = pred_corr.detach().to(device=device, dtype=torch.int64).t().contiguous()      # dense [2,M]
        cad = _lib.dev_f32(CAD, device).contiguous()
        pc = _lib.dev_f32(PC_aligned, device).contiguous()
        if cad.dim() != 2 or cad.shape[1] != 3 or pc.dim() != 2 or pc.shape[1] != 3:
            raise RuntimeError("CAD / PC_aligned must be [N,3]")
        d = _descs(1, int(cad.shape[0]), int(pc.shape[0]), [total_corr], 0, 0, [0])
        thr = (ctypes.c_float * 1)(_thr32(threshold))
        res = torch.empty(2, dtype=torch.int32, device=device)     # [inliers, ratio bits]
        ws = _lib.workspace(device, lib.f2p_inlier_workspace_bytes(1))
        rc = lib.f2p_inlier_ratio_batched(cad.data_ptr(), pc.data_ptr(), maps.data_ptr(), d, 1, 0, None, thr,
                                          res.data_ptr(), res.data_ptr() + 4, ws.data_ptr(), ws.numel(),
                                          _lib.stream_ptr(device))
        _lib.check(rc, "f2p_inlier_ratio_batched")
        out = res[1:].view(torch.float32)[0]
        if torch.isnan(out):
            raise IndexError("pred_corr holds an index out of range of CAD / PC_aligned")
        return out if pred_corr.is_cuda else out.cpu()


def inlier_ratio_batched(maps, CAD, PC_aligned, thresholds, counts=None):
    """Inlier ratios of a uniform batch, float32 [P] on the device (no host sync; NaN marks an out-of-range index).

    maps        naive: int64 [P,2,V2] (what `naive_batched` returns for stacked inputs);
                spatial: the packed int64 [P, 2*5*V2] buffer of `spatial_batched(..., packed=True)` with `counts`
    CAD         [P,ncad,3] or one shared [ncad,3];  PC_aligned [P,npc,3] (scene cloud under the ground-truth pose)
    thresholds  one number / 0-d tensor per pair (eval.py uses 0.1 * diam_cad)
    """
    lib = _lib.load()
    device = _lib.cuda_device_for(maps, CAD, PC_aligned)
    with torch.cuda.device(device):
        m = maps.detach().to(device=device, dtype=torch.int64).contiguous()
        n = int(m.shape[0])
        cad = _lib.dev_f32(CAD, device).contiguous()
        pc = _lib.dev_f32(PC_aligned, device).contiguous()
        if cad.dim() == 2:
            cad = cad.unsqueeze(0)
        if pc.dim() != 3 or pc.shape[0] != n or pc.shape[2] != 3 or cad.shape[2] != 3 or cad.shape[0] not in (1, n):
            raise RuntimeError("CAD must be [P,ncad,3] or [ncad,3] and PC_aligned [P,npc,3]")
        spatial = counts is not None
        per_pair = int(m[0].numel()) if n else 0
        width = per_pair // 10 if spatial else per_pair // 2          # V2
        if (spatial and m.dim() != 2) or (not spatial and (m.dim() != 3 or m.shape[1] != 2)):
            raise RuntimeError("maps must be [P,2,V2] (naive) or the packed [P,2*5*V2] buffer with counts (spatial)")
        ratio = torch.empty(n, dtype=torch.float32, device=device)
        if n == 0:
            return ratio
        if len(thresholds) != n:
            raise ValueError("thresholds must have one entry per pair")
        thr = (ctypes.c_float * n)(*[_thr32(t) for t in thresholds])
        d = _descs(n, int(cad.shape[1]), int(pc.shape[1]), [width] * n, 0 if cad.shape[0] == 1 else int(cad.shape[1]) * 3,
                   int(pc.shape[1]) * 3, [i * per_pair for i in range(n)])
        inl = torch.empty(n, dtype=torch.int32, device=device)
        cnt = counts.detach().to(device=device, dtype=torch.int32).contiguous() if spatial else None
        ws = _lib.workspace(device, lib.f2p_inlier_workspace_bytes(n))
        rc = lib.f2p_inlier_ratio_batched(cad.data_ptr(), pc.data_ptr(), m.data_ptr(), d, n, 1 if spatial else 0,
                                          cnt.data_ptr() if spatial else None, thr, inl.data_ptr(), ratio.data_ptr(),
                                          ws.data_ptr(), ws.numel(), _lib.stream_ptr(device))
        _lib.check(rc, "f2p_inlier_ratio_batched")
        return ratio
