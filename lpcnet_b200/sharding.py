"""Stream -> GPU partitioning for multi-GPU runs.

Streams share only read-only weights (src/lpcnet.c:146-167 touches one LPCNetState), so the
path shards with no data-path collective: rank r of `world` owns a contiguous range of
streams, runs them on its own GPU, and nothing crosses NVLink.  torch.distributed is used
only for the timing barrier and the max-over-ranks reduction of the elapsed time.
"""
from __future__ import annotations


def shard_range(n_streams: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [lo, hi) of the global stream ids owned by `rank` (sizes differ by <= 1)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(n_streams, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def owner_of(stream: int, n_streams: int, world: int) -> int:
    """Inverse of shard_range."""
    base, rem = divmod(n_streams, world)
    edge = rem * (base + 1)
    if stream < edge:
        return stream // (base + 1)
    return rem + (stream - edge) // base


def aggregate_rate(samples_per_rank: list[int], seconds_per_rank: list[float]) -> float:
    """Whole-job samples/s: all ranks' samples over the slowest rank's time."""
    return sum(samples_per_rank) / max(seconds_per_rank)
