"""Timepoint sharding across the GPUs of one box (one process per GPU).

The reference parallelises `segment`/`link` with a spawn pool over timepoints
(ultrack/utils/multiprocessing.py:26-56) and SLURM-style `batch_index` windows
(:82-105).  Here every rank of a `torch.distributed` group owns one contiguous block of
timepoints; `segment` needs no communication, `link(t)` needs the node tables of t and
t+1, so the only exchange is the first frame of the next rank's block (frames.exchange).
"""
from __future__ import annotations

import os
from typing import Optional, Sequence, Tuple


def batch_index_range(total_size: int, n_workers: int, batch_index: Optional[int]) -> Sequence[int]:
    """ultrack/utils/multiprocessing.py:82-105."""
    if batch_index is None:
        return range(total_size)
    start = list(range(0, total_size, n_workers))[batch_index]
    return range(start, min(start + n_workers, total_size))


def dist_info() -> Tuple[int, int]:
    """(rank, world_size) of the default process group, (0, 1) outside of one."""
    try:
        import torch.distributed as dist
    except Exception:  # noqa: BLE001
        return 0, 1
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", "0"))


def block_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of `n` items owned by `rank`; sizes differ by at most one."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard(points: Sequence[int], world: int, rank: int) -> Sequence[int]:
    lo, hi = block_bounds(len(points), world, rank)
    return points[lo:hi]


def owner_of(index: int, n: int, world: int) -> int:
    """Rank whose block holds position `index` of an `n`-item range."""
    for r in range(world):
        lo, hi = block_bounds(n, world, r)
        if lo <= index < hi:
            return r
    raise IndexError(index)


def owned(points: Sequence[int], total: int, world: int, rank: int) -> list:
    """The members of `points` (timepoints of one call, possibly a `batch_index` window) that
    fall into `rank`'s block of the `total` frames of the timelapse.  `segment` and `link` both
    use this rule, so the rank that links the pair (t, t+1) is the one that segmented frame t."""
    lo, hi = block_bounds(total, world, rank)
    return [t for t in points if lo <= t < hi]


def barrier() -> None:
    rank, world = dist_info()
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
