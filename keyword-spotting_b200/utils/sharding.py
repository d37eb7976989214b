"""Batch sharding across ranks: clips are independent, so rank r of W takes the
contiguous slice [r*B/W, (r+1)*B/W) and there is no collective on the data path.  The only
collectives are bookkeeping: a barrier and a MAX over ranks of the elapsed time."""
from __future__ import annotations

from typing import Tuple


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced (sizes differ by at most one), covering, non-overlapping."""
    if world < 1 or not 0 <= rank < world or total < 0:
        raise ValueError("bad shard arguments")
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def job_throughput(units_per_rank, elapsed_s_per_rank) -> float:
    """Whole-job rate: all units processed / slowest rank's time (max over ranks)."""
    return float(sum(units_per_rank)) / max(elapsed_s_per_rank)


def max_over_ranks(value: float, device=None) -> float:
    """MAX all-reduce of a scalar when torch.distributed is initialised, identity otherwise."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
