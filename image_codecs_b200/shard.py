"""Sharding of independent image batches across GPUs (one process per GPU, no collective).

The path shards by image: nothing is shared between images (the reference allocates fresh state per
call, jpeg.c:2445-2450) and stbi__jpeg_reset (jpeg.c:1142) clears all carried state at restart
markers, so there is no exchange step and no data-path collective.  Entropy decode
time is proportional to compressed bytes, so shards are balanced by bytes, not by image count.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def contiguous_shard(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of rank's contiguous slice when n_items are split as evenly as possible."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def balanced_shards(sizes: Sequence[int], world: int) -> List[List[int]]:
    """Greedy longest-processing-time partition of item indices by size (compressed bytes).

    Returns `world` index lists; every index appears exactly once; within a shard indices are
    ascending so outputs can be scattered back in order.  Deterministic (ties broken by index).
    """
    if world <= 0:
        raise ValueError("bad world")
    order = sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))
    load = [0] * world
    shards: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        shards[r].append(i)
        load[r] += int(sizes[i])
    for s in shards:
        s.sort()
    return shards


def shard_for_rank(sizes: Sequence[int], rank: int, world: int) -> List[int]:
    return balanced_shards(sizes, world)[rank]
