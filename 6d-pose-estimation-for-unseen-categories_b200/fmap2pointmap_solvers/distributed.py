"""Pair-sharded multi-GPU driver: one process per GPU, contiguous pair slices, one all-gather to return index maps.

The reference has no distributed code; pairs are independent (`scripts/eval.py:83`), so the path shards
with no data-path collective.  The only exchange is returning the int64 index maps to every rank: NCCL all-gather over
NVLink on a B200 box, gloo in the CPU tests (which inject a compute function; the product compute is CUDA only).
"""
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of ceil(n/world) items per rank."""
    per = (n_items + world_size - 1) // world_size if world_size > 0 else n_items
    lo = min(rank * per, n_items)
    return lo, min(lo + per, n_items)


def pack_maps(maps: Sequence[torch.Tensor], widths: Sequence[int], per_rank: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """Packs this rank's [2,M_i] maps into a fixed [per_rank, 2, Wmax] int64 payload (−1 padded) plus counts."""
    wmax = max(list(widths) + [1])
    payload = torch.full((per_rank, 2, wmax), -1, dtype=torch.int64, device=device)
    counts = torch.full((per_rank,), -1, dtype=torch.int64, device=device)
    for i, m in enumerate(maps):
        payload[i, :, :m.shape[1]] = m.to(device)
        counts[i] = m.shape[1]
    return payload, counts


def solve_sharded(solver: str, C12: Sequence[torch.Tensor], evecs_x: Sequence[torch.Tensor],
                  evecs_y: Sequence[torch.Tensor], CAD: Optional[Sequence[torch.Tensor]] = None,
                  PC: Optional[Sequence[torch.Tensor]] = None, diam_cad: Optional[Sequence] = None,
                  group=None, compute: Optional[Callable] = None, device=None) -> List[torch.Tensor]:
    """Every rank passes a pair list of the SAME length; only its own slice [lo, hi) has to be populated (other entries
    may be None).  Each rank converts its slice and all ranks return the full list of index maps in global pair order.
    The exchange is counts first, then a payload as wide as the widest map any rank actually produced (int32 on the
    wire) — not the 5*V2 worst case.

    solver: "naive" | "spatial".  `compute(solver, C, ex, ey, cad, pc, diam) -> list of [2,M]` defaults to the CUDA path.
    """
    if solver not in ("naive", "spatial"):
        raise ValueError("solver must be 'naive' or 'spatial'")
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n = len(evecs_x)
    lo, hi = shard_range(n, world, rank)
    per = (n + world - 1) // world if world > 0 else n
    if compute is None:
        from .batched import naive_batched, spatial_batched

        def compute(kind, C, ex, ey, cad, pc, diam):
            if kind == "naive":
                return naive_batched(C, ex, ey)
            return spatial_batched(C, ex, ey, cad, pc, diam, verbose=False)
    sl = slice(lo, hi)
    mine = compute(solver, list(C12[sl]), list(evecs_x[sl]), list(evecs_y[sl]),
                   list(CAD[sl]) if CAD is not None else None, list(PC[sl]) if PC is not None else None,
                   list(diam_cad[sl]) if diam_cad is not None else None) if hi > lo else []
    if world == 1:
        return list(mine)
    if device is None:
        device = mine[0].device if mine else (torch.device("cuda", torch.cuda.current_device())
                                              if torch.cuda.is_available() and dist.get_backend(group) == "nccl"
                                              else torch.device("cpu"))
    # step 1: counts (-1 = slot not used by that rank); concatenated-output form: accepted by NCCL and gloo alike
    counts = torch.full((per,), -1, dtype=torch.int64, device=device)
    for i, m in enumerate(mine):
        counts[i] = m.shape[1]
    all_counts = torch.empty((world * per,), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(all_counts, counts, group=group)
    counts_host = all_counts.view(world, per).cpu()
    # step 2: payload as wide as the widest map of ANY rank (every rank reads the same gathered counts, so they agree),
    # int32 on the wire (indices < 2^31; row 1 may hold the reference's negative int16 wrap, which int32 keeps)
    wmax = max(int(counts_host.max()), 1)
    payload = torch.full((per, 2, wmax), -1, dtype=torch.int32, device=device)
    for i, m in enumerate(mine):
        payload[i, :, :m.shape[1]] = m.to(device=device, dtype=torch.int32)
    all_payload = torch.empty((world * per, 2, wmax), dtype=torch.int32, device=device)
    dist.all_gather_into_tensor(all_payload, payload, group=group)
    all_payload = all_payload.view(world, per, 2, wmax)
    out = []
    for r in range(world):
        for i in range(per):
            c = int(counts_host[r, i])
            if c >= 0:
                out.append(all_payload[r, i, :, :c].to(torch.int64).contiguous())
    return out


def allgather_packed_maps(flat: torch.Tensor, counts: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns every rank's spatial maps from the packed form of `spatial_batched(..., packed=True)`.

    flat: int64 [P, W] — pair i's dense [2, M'_i] map is `flat[i, :2*M'_i]`; counts: int32 [P] = M'_i.
    The result buffer is sized for the worst case (W = 2*5*V2 slots per pair) and mostly unused behind the dense prefix,
    so the exchange is two steps: the counts first (tiny), then only the first 2*max(M') slots of every
    pair, as int32 on the wire (indices are < 2^31; row 1 may hold the reference's negative int16 wrap, which int32
    keeps).  Reading the maximum is the one host sync.  Returns (maps int64 [world, P, 2*Mmax], counts int32 [world, P]):
    pair i of rank r is `maps[r, i, :2*counts[r, i]].view(2, -1)`.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    P = int(flat.shape[0])
    counts = counts.to(torch.int32).contiguous()
    all_counts = torch.empty((world * P,), dtype=torch.int32, device=counts.device)
    if world > 1:
        dist.all_gather_into_tensor(all_counts, counts, group=group)
    else:
        all_counts.copy_(counts)
    mmax = int(all_counts.max().item()) if all_counts.numel() else 0
    width = min(2 * mmax, int(flat.shape[1]))
    wire = flat[:, :width].to(torch.int32).contiguous()
    all_wire = torch.empty((world * P, width), dtype=torch.int32, device=flat.device)
    if world > 1 and width > 0:
        dist.all_gather_into_tensor(all_wire, wire, group=group)
    else:
        all_wire.copy_(wire)
    return all_wire.to(torch.int64).view(world, P, width), all_counts.view(world, P)
