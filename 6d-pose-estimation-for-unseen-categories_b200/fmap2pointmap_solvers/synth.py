"""Seeded synthetic object/scene pairs of the shapes BASELINE.json names.

There is no network for the BOP datasets and the reference's eigenbasis cache needs packages that are not
installed, so every benchmark / parity case is generated here: eigenbases, a near-isometric functional map,
a CAD cloud and a partial scene cloud that is a noisy rigid transform of a CAD subset plus outliers (so the
spatial-consistency filter of `spacial_filtering.py:51-75` actually prunes).
All tensors are FP32 on the CPU; generation is deterministic in (seed, shapes).
"""
import math
from dataclasses import dataclass

import torch


@dataclass
class Pair:
    C12: torch.Tensor      # [k,k]
    evecs_x: torch.Tensor  # [n1,k]  CAD eigenbasis
    evecs_y: torch.Tensor  # [n2,k]  partial-cloud eigenbasis
    CAD: torch.Tensor      # [n1,3]
    PC: torch.Tensor       # [n2,3]
    diam_cad: float        # Python float, as scripts/eval.py:87 passes it
    perm: torch.Tensor     # [n2] ground-truth CAD index of every PC point (-1 for outliers)


def _rotation(g: torch.Generator) -> torch.Tensor:
    q, r = torch.linalg.qr(torch.randn(3, 3, generator=g))
    q = q * torch.sign(torch.diagonal(r))
    if torch.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q


def make_pair(n1: int, n2: int, k: int, seed: int = 1234, sigma: float = 0.3, outlier_frac: float = 0.3,
              xyz_noise: float = 0.05) -> Pair:
    g = torch.Generator().manual_seed(int(seed))
    s = 1.0 / math.sqrt(max(n1, 1))
    evecs_x = torch.randn(n1, k, generator=g) * s
    if n2 <= n1:
        perm = torch.randperm(n1, generator=g)[:n2]
    else:
        perm = torch.randint(0, max(n1, 1), (n2,), generator=g)
    q, r = torch.linalg.qr(torch.randn(k, k, generator=g))
    C = (q * torch.sign(torch.diagonal(r))).contiguous()
    evecs_y = ((evecs_x[perm] + sigma * s * torch.randn(n2, k, generator=g)) @ C.t()).contiguous()
    CAD = torch.rand(n1, 3, generator=g) * 10.0
    R = _rotation(g)
    t = torch.randn(3, generator=g) * 5.0
    PC = CAD[perm] @ R.t() + t + xyz_noise * torch.randn(n2, 3, generator=g)
    gt = perm.clone()
    n_out = int(outlier_frac * n2)
    if n_out > 0:
        out_idx = torch.randperm(n2, generator=g)[:n_out]
        PC[out_idx] = torch.rand(n_out, 3, generator=g) * 10.0 @ R.t() + t
        evecs_y[out_idx] = (torch.randn(n_out, k, generator=g) * s) @ C.t()
        gt[out_idx] = -1
    diam = float(torch.linalg.norm(CAD.max(0).values - CAD.min(0).values)) if n1 > 0 else 1.0
    return Pair(C, evecs_x.contiguous(), evecs_y, CAD.contiguous(), PC.contiguous(), diam, gt)
