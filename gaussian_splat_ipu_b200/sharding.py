"""Multi-GPU sharding of the splat path — host-side integer logic plus the one
collective the path has (gathering framebuffer bands / views to the streaming rank).

The reference is single-device (`numIpus = 1`, src/main/splat.cpp:41-46); its unit of parallelism
is the framebuffer tile (one `GSplat` vertex per tile, src/splat/ipu_rasteriser.cpp:334-385) and it
already exports the per-tile workload (`getIPUHistogram`, :104-106).  Both shardings below reuse
exactly that: tiles are independent after binning, and the previous frame's per-tile counts predict
the next frame's cost.

* tile-row bands of ONE frame: `balance_bands` + `B200Splatter.set_band` + `gather_bands`
* whole views of a multi-view batch: `assign_views` (no data-path collective)

One process per GPU; `torch.distributed` is only plumbing (NCCL over NVLink on GPUs, gloo on CPU).
"""
from __future__ import annotations

import numpy as np


def row_costs(tile_counts: np.ndarray, tiles_across: int, tile_pixels: int = 640, fixed: float = 64.0,
              cap: float = 2048.0) -> np.ndarray:
    """Predicted cost of every tile row from per-tile list lengths (a11): blend work is
    pixels x list entries walked per tile; `fixed` stands for the per-tile constant (launch,
    framebuffer write) so that empty rows are not free.  `cap`: pixels stop at T(1-alpha) < 1e-4
    (codelets.cpp:405-408), so the entries a tile walks saturate (measured: 1000-1900 per pixel on
    the most expensive tiles, whatever the list length).  Mirror of the library's balancer
    (gsplat_balance_bands, csrc/gsplat_group.inc)."""
    c = np.minimum(np.asarray(tile_counts, dtype=np.float64), cap).reshape(-1, tiles_across)
    return (c * tile_pixels + fixed * tile_pixels).sum(axis=1)


def balance_bands(cost_per_row: np.ndarray, n_bands: int) -> list[int]:
    """Cut `len(cost_per_row)` tile rows into `n_bands` contiguous bands of (nearly) equal cost.
    Returns n_bands+1 non-decreasing row indices, first 0, last = number of rows.  Every band gets
    at least one row while rows remain (n_bands <= rows), so no rank idles on a tiny image."""
    cost = np.asarray(cost_per_row, dtype=np.float64)
    rows = cost.size
    if n_bands <= 0:
        raise ValueError("n_bands must be positive")
    if n_bands >= rows:                       # one row per band, the rest empty
        return [min(i, rows) for i in range(n_bands + 1)]
    prefix = np.concatenate([[0.0], np.cumsum(cost)])
    total = prefix[-1]
    cuts = [0]
    for b in range(1, n_bands):
        target = total * b / n_bands
        r = int(np.searchsorted(prefix, target, side="left"))
        # pick the nearer of the two candidate boundaries
        if r > 0 and abs(prefix[r - 1] - target) <= abs(prefix[min(r, rows)] - target):
            r -= 1
        r = max(r, cuts[-1] + 1)              # at least one row per band
        r = min(r, rows - (n_bands - b))      # leave a row for every later band
        cuts.append(r)
    cuts.append(rows)
    return cuts


def band_imbalance(cost_per_row: np.ndarray, cuts: list[int]) -> float:
    """max band cost / mean band cost (1.0 = perfect)."""
    cost = np.asarray(cost_per_row, dtype=np.float64)
    per = np.array([cost[a:b].sum() for a, b in zip(cuts[:-1], cuts[1:])])
    return float(per.max() / max(per.mean(), 1e-30))


def assign_views(view_costs, n_ranks: int) -> list[list[int]]:
    """Greedy longest-first assignment of whole views to ranks (multi-view batches): returns, per
    rank, the view indices it renders, in ascending order.  Deterministic (ties by index) so every
    rank computes the same assignment without communicating."""
    costs = np.asarray(view_costs, dtype=np.float64)
    order = sorted(range(costs.size), key=lambda i: (-costs[i], i))
    load = [0.0] * n_ranks
    out: list[list[int]] = [[] for _ in range(n_ranks)]
    for i in order:
        r = min(range(n_ranks), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += float(costs[i])
    return [sorted(v) for v in out]


def round_robin_views(n_views: int, rank: int, world: int) -> list[int]:
    """Views of rank `rank` when no cost estimate exists yet (first pass over a batch)."""
    return list(range(rank, n_views, world))


class DeviceArray:
    """Zero-copy view of device memory owned by a gsplat context, for torch / NCCL
    (`torch.as_tensor(DeviceArray(...), device="cuda")`)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


def band_slices(cuts: list[int], tile_h: int, width: int, channels: int = 3):
    """(offset, length) in elements of every band inside a row-major H x W x channels image."""
    row = width * channels
    return [(a * tile_h * row, (b - a) * tile_h * row) for a, b in zip(cuts[:-1], cuts[1:])]


def gather_bands(local_image, cuts: list[int], tile_h: int, width: int, channels: int = 3, dst: int = 0, group=None):
    """Gathers the row bands of one frame to rank `dst`.

    `local_image`: this rank's full-size (H*W*channels,) tensor (uint8 BGR or float), of which only
    rows [cuts[rank]*tile_h, cuts[rank+1]*tile_h) are valid.  On `dst` the other ranks' bands are
    received straight into the right slices of `local_image` (variable sizes, no padding, no
    staging copy); other ranks send theirs.  Empty bands are skipped on both sides.  Returns
    `local_image` on `dst`, None elsewhere.  Works on CUDA tensors (NCCL) and CPU tensors (gloo).
    """
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if len(cuts) != world + 1:
        raise ValueError("cuts must have world_size + 1 entries")
    sl = band_slices(cuts, tile_h, width, channels)
    flat = local_image.reshape(-1)
    ops = []
    if rank == dst:
        for r in range(world):
            off, ln = sl[r]
            if r != dst and ln:
                ops.append(dist.P2POp(dist.irecv, flat[off:off + ln], r, group))
    else:
        off, ln = sl[rank]
        if ln:
            ops.append(dist.P2POp(dist.isend, flat[off:off + ln], dst, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return local_image if rank == dst else None


def allgather_tile_counts(counts, group=None):
    """Sum of the per-tile counts over ranks (each rank only has its band's tiles non-zero) so every
    rank can run `balance_bands` for the next frame on identical data."""
    import torch.distributed as dist
    dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts
