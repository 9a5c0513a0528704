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


# ---- one huge pair over several GPUs ("single huge pair": shard V1, lexicographic (d, idx) minimum) ----------
_KEY_NONE = torch.iinfo(torch.int64).max


def row_cuts(n1: int, world_size: int) -> List[Tuple[int, int]]:
    """Contiguous CAD-row ranges [lo, hi) per rank, boundaries on multiples of 128 (the kernels' row tile); high ranks
    of a small problem may get an empty range."""
    step = ((n1 + world_size - 1) // world_size + 127) // 128 * 128 if world_size > 0 else n1
    return [(min(r * step, n1), min((r + 1) * step, n1)) for r in range(world_size)]


def reduce_keys(keys: torch.Tensor, group=None) -> torch.Tensor:
    """In-place minimum over the ranks of per-scene-point keys `bits(distance) << 32 | global row` (int64; distances are
    non-negative, so the signed order is the (distance, row) order; `_KEY_NONE` = no candidate).  One all-reduce: the
    result on every rank is the key of the nearest row over ALL ranks' rows, ties to the lowest global row."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(keys, op=dist.ReduceOp.MIN, group=group)
    return keys


def _local_keys_cuda(C12, ex_rows, evecs_y, row0, device):
    """Exact keys of this rank's CAD rows against every scene point (csrc: f2p_project + f2p_nn_argmin_keys)."""
    from . import _lib
    lib = _lib.load()
    with torch.cuda.device(device):
        C = _lib.dev_f32(C12, device).contiguous()
        ex = _lib.rowmajor_2d(_lib.dev_f32(ex_rows, device)).contiguous()
        ey = _lib.rowmajor_2d(_lib.dev_f32(evecs_y, device)).contiguous()
        n1, k, n2 = int(ex.shape[0]), int(C.shape[0]), int(ey.shape[0])
        ld1, ld2 = max((n1 + 127) // 128 * 128, 128), (n2 + 127) // 128 * 128
        keys = torch.full((ld2,), -1, dtype=torch.int64, device=device)                 # every byte 0xFF: no candidate
        if n1 > 0 and n2 > 0:                                                           # an empty block has nothing to add
            st = _lib.stream_ptr(device)
            xT = torch.empty((k, ld1), dtype=torch.float32, device=device); nx = torch.empty((ld1,), dtype=torch.float32, device=device)
            yT = torch.empty((k, ld2), dtype=torch.float32, device=device); ny = torch.empty((ld2,), dtype=torch.float32, device=device)
            pb = (lib.f2p_project_workspace_bytes() + 255) // 256 * 256
            ws = _lib.workspace(device, pb + lib.f2p_nn_workspace_bytes(n1, n2, k, 0) + 1024)
            _lib.check(lib.f2p_project(ex.data_ptr(), n1, k, k, C.data_ptr(), xT.data_ptr(), nx.data_ptr(), ld1,
                                       ws.data_ptr(), pb, st), "f2p_project")
            _lib.check(lib.f2p_project(ey.data_ptr(), n2, k, k, None, yT.data_ptr(), ny.data_ptr(), ld2, ws.data_ptr(), pb, st),
                       "f2p_project")
            wn = ws[pb:]
            _lib.check(lib.f2p_nn_argmin_keys(xT.data_ptr(), nx.data_ptr(), n1, ld1, int(row0), yT.data_ptr(), ny.data_ptr(), n2,
                                              ld2, k, keys.data_ptr(), wn.data_ptr(), wn.numel(), st), "f2p_nn_argmin_keys")
        return keys[:n2]


def block_grid(n1: int, n2: int, world_size: int, col_weight: float = 2.0) -> Tuple[int, int]:
    """(row shards R, column shards Cg), R * Cg = world: every rank gets the same n1*n2/world distance evaluations whatever
    the grid, so the choice minimises what does NOT shrink with the other dimension — per-row work (projection, packing)
    ~ n1/R plus per-scene-point work (transposition, packing, thresholds, exact verification) ~ col_weight * n2/Cg."""
    best, best_cost = (world_size, 1), None
    for r in range(1, world_size + 1):
        if world_size % r:
            continue
        c = world_size // r
        cost = n1 / r + col_weight * n2 / c
        if best_cost is None or cost < best_cost:
            best, best_cost = (r, c), cost
    return best


def naive_row_sharded(C12, evecs_x, evecs_y, group=None, int16_compat: bool = False, local_keys: Optional[Callable] = None,
                      device=None, grid: Optional[Tuple[int, int]] = None) -> torch.Tensor:
    """`naive_fmap2pointmap` for ONE pair split over the ranks of `group` (strong scaling of a single huge pair, SURVEY
    8(e)).  Every rank passes the same C12 [K,K], evecs_x [V1,K] and evecs_y [V2,K].  The ranks form an R x Cg grid
    (`grid`, default `block_grid`): rank r = (r // Cg, r % Cg) evaluates the CAD rows `row_cuts(V1, R)[r // Cg]` against
    the scene points `row_cuts(V2, Cg)[r % Cg]` and keeps, per scene point, the exact key (distance bits, GLOBAL row);
    scene points outside its block stay "no candidate".  ONE all-reduce(MIN) of the V2 keys then both merges the row
    shards and assembles the column shards: every rank holds the pair's arg-min.  Per-element distances do not depend on
    the split, so the int64 [2,V2] result is bit-identical to the single-GPU call, ties to the lowest
    row.  `local_keys(C, ex_rows, ey_cols, row0) -> int64 [cols]` replaces the CUDA kernels in the CPU tests."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    C = C12.squeeze(0) if C12.dim() == 3 else C12
    n1, n2 = int(evecs_x.shape[0]), int(evecs_y.shape[0])
    if n1 < 1:
        raise RuntimeError("evecs_x has no rows (argmin of an empty dimension)")
    R, Cg = grid if grid is not None else block_grid(n1, n2, world)
    if R * Cg != world:
        raise ValueError(f"grid {R} x {Cg} does not match the {world} ranks of the group")
    lo, hi = row_cuts(n1, R)[rank // Cg]
    c0, c1 = row_cuts(n2, Cg)[rank % Cg]
    if local_keys is None:
        from . import _lib
        if device is None:
            device = _lib.cuda_device_for(evecs_x, evecs_y, C)
        part = _local_keys_cuda(C, evecs_x[lo:hi], evecs_y[c0:c1], lo, device)
    else:
        part = local_keys(C, evecs_x[lo:hi], evecs_y[c0:c1], lo)
    keys = torch.full((n2,), _KEY_NONE, dtype=torch.int64, device=part.device)
    keys[c0:c1] = torch.where(part < 0, torch.full_like(part, _KEY_NONE), part)   # "no candidate" must lose the signed minimum
    reduce_keys(keys, group)
    if local_keys is None:
        from . import _lib
        lib = _lib.load()
        out = torch.empty((2, n2), dtype=torch.int64, device=keys.device)
        if n2 > 0:
            with torch.cuda.device(keys.device):
                _lib.check(lib.f2p_keys_to_map(keys.data_ptr(), n2, _lib.F2P_INT16_COMPAT if int16_compat else 0, out.data_ptr(),
                                               _lib.stream_ptr(keys.device)), "f2p_keys_to_map")
        return out
    idx = torch.arange(n2, dtype=torch.int64, device=keys.device)
    if int16_compat:
        idx = ((idx + 32768) % 65536) - 32768
    return torch.stack([keys & 0xFFFFFFFF, idx], 0)
