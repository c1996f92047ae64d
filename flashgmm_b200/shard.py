"""Work partitioning over the GPUs of one box: every image -- and every
768x512 tile of a larger image -- is an independent unit with its own streams; nothing is reduced
or exchanged, so there is no data-path collective. One process per GPU takes its share of the work
items; the only cross-rank traffic is bookkeeping (per-item byte counts, the max-over-ranks step
time), which runs over whatever backend ``torch.distributed`` was initialised with (NCCL on the GPU
box, gloo in the CPU tests).

The reference has no tiling convention (it codes whole images); the tile grid here is this
library's: tiles of ``tile_h x tile_w`` pixels, the image padded at the bottom / right up to a
whole number of tiles.
"""
from __future__ import annotations

from typing import List, NamedTuple, Sequence, Tuple


class Tile(NamedTuple):
    """Pixel box [y0, y1) x [x0, x1) of the padded image; `pad_h`, `pad_w` = rows / columns of it
    that lie outside the real image."""
    y0: int
    y1: int
    x0: int
    x1: int
    pad_h: int
    pad_w: int


def tile_image(height: int, width: int, tile_h: int = 512, tile_w: int = 768) -> List[Tile]:
    """Row-major tile grid covering a `height x width` image. 2160 x 3840 -> 5 x 5 tiles of
    512 x 768 (bottom padded to 2560)."""
    if height <= 0 or width <= 0 or tile_h <= 0 or tile_w <= 0:
        raise ValueError("image and tile sizes must be positive")
    tiles = []
    for y0 in range(0, height, tile_h):
        for x0 in range(0, width, tile_w):
            y1, x1 = y0 + tile_h, x0 + tile_w
            tiles.append(Tile(y0, y1, x0, x1, max(0, y1 - height), max(0, x1 - width)))
    return tiles


def assign(costs: Sequence[float], world_size: int) -> List[List[int]]:
    """Work items (by index) per rank. Longest-processing-time first: items sorted by cost,
    each handed to the currently least loaded rank (ties: lowest rank) -- round-robin when all
    costs are equal. Deterministic, so every rank computes the same table without talking."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = sorted(range(len(costs)), key=lambda i: (-float(costs[i]), i))
    load = [0.0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += float(costs[i])
    for lst in out:
        lst.sort()
    return out


def my_items(costs: Sequence[float], world_size: int, rank: int) -> List[int]:
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside [0, {world_size})")
    return assign(costs, world_size)[rank]


def max_over_ranks(seconds: float, device=None) -> float:
    """Step time of the job = the slowest rank's (bench.py's timing rule)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(seconds)
    t = torch.tensor([float(seconds)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_item_sizes(n_items: int, mine: Sequence[Tuple[int, int]], device=None) -> List[int]:
    """Every rank contributes (item index, coded bytes) for its items; all ranks get the full
    per-item table (e.g. for a container index or the job's bits-per-pixel). Items nobody coded
    stay at -1."""
    import torch
    import torch.distributed as dist

    table = torch.full((n_items,), -1, dtype=torch.int64, device=device)
    for idx, nbytes in mine:
        if not 0 <= idx < n_items:
            raise ValueError(f"item index {idx} outside [0, {n_items})")
        table[idx] = int(nbytes)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(table, op=dist.ReduceOp.MAX)
    return [int(v) for v in table.tolist()]
