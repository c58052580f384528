"""Batch / ensemble sharding across the GPUs of one box.

Every sample's rollout is independent, so rank ``r`` of ``G`` runs the full model on a
contiguous slice of the members and there is no data-path collective; the only exchange is
the all-reduce of the ``[T, C, 4]`` fp64 metric partials in ``metrics.reduce_sums``.
"""
from __future__ import annotations

from typing import Tuple


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split of ``total`` members: the first ``total % world`` ranks
    take one extra member.  Returns ``(start, stop)``."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside [0, {world})")
    base, extra = divmod(total, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)
