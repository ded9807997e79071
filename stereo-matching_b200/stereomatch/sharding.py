"""Frame sharding for multi-GPU runs (new; the reference is single-GPU).

Stereo pairs are independent units (reference stereomatch/pipeline.py:69-94 is a pure function
of one pair), so an N-GPU run is N replicas, one process per GPU, each taking a contiguous
slice of the frame batch.  There is NO collective on the data path; torch.distributed is used
only for the barrier and the max-over-ranks of the device time.  A single frame is never
split: SGM's vertical and diagonal recurrences cross every row boundary.
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[begin, end) of rank's contiguous slice; slices are disjoint, cover range(n_items) and
    differ in length by at most one (the first n_items % world_size ranks get the extra item)."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if n_items < 0:
        raise ValueError("n_items must be >= 0")
    base, extra = divmod(n_items, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def frame_seeds(first_seed: int, frames_per_rank: int, rank: int) -> range:
    """Weak-scaling batch: rank r owns frames [r*F, (r+1)*F) of the global synthetic batch;
    frame f uses seed first_seed + f."""
    return range(first_seed + rank * frames_per_rank, first_seed + (rank + 1) * frames_per_rank)


def max_over_ranks(value: float, device=None) -> float:
    """MAX all-reduce of a scalar (device time in ms); identity when not distributed."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, device=None) -> float:
    """SUM all-reduce of a scalar (units processed); identity when not distributed."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
