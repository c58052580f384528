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


def bind_to_device_cpus(device_index: int) -> bool:
    """Pin the calling process to the CPU cores next to GPU ``device_index`` (NVML's ideal affinity), so
    that the pinned host buffers a rank allocates afterwards -- the ones ``dlwp_rollout_host`` copies
    from and into -- live on that GPU's NUMA node.  With one process per GPU on a two-socket box this
    keeps eight ranks' host copies off the inter-socket link.  Returns False (and changes nothing) when
    NVML or the affinity call is unavailable."""
    try:
        import os

        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return False
        os.sched_setaffinity(0, cpus)
        return True
    except Exception:
        return False
