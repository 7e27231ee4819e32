"""Batch sharding across GPUs: images are independent, so each rank takes a
contiguous slice of the batch and nothing is exchanged on the data path.  The only
cross-rank traffic is bookkeeping — per-image detection counts gathered to rank 0 so the
host can index the per-GPU results (bytes << 1 MB), and the max-over-ranks of a timing.
"""

from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(total: int, world: int) -> list[tuple[int, int]]:
    """Contiguous, balanced [lo, hi) slices: the first ``total % world`` ranks get one more."""
    base, extra = divmod(total, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def my_shard(total: int, rank: int | None = None, world: int | None = None) -> tuple[int, int]:
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    return shard_bounds(total, world)[rank]


def gather_counts(local_count: torch.Tensor, total: int) -> torch.Tensor | None:
    """All ranks pass their slice's per-image counts (int32, any device the backend
    supports); rank 0 gets the int32 [total] vector in batch order, others None."""
    world, rank = dist.get_world_size(), dist.get_rank()
    bounds = shard_bounds(total, world)
    width = max(hi - lo for lo, hi in bounds)
    padded = torch.zeros(width, dtype=torch.int32, device=local_count.device)
    padded[: local_count.numel()] = local_count.to(torch.int32)
    bucket = [torch.zeros_like(padded) for _ in range(world)] if rank == 0 else None
    dist.gather(padded, bucket, dst=0)
    if rank != 0:
        return None
    return torch.cat([b[: hi - lo] for b, (lo, hi) in zip(bucket, bounds)])


def max_over_ranks(value: float, device: torch.device | str = "cpu") -> float:
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
