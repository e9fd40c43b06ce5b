"""Multi-GPU partitioning: requests / streams are independent, so ranks take contiguous,
byte-balanced ranges of units and never exchange data; the host concatenates the per-rank outputs."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def partition_by_bytes(unit_bytes: Sequence[int], world_size: int) -> List[Tuple[int, int]]:
    """Contiguous [lo, hi) unit ranges per rank, balanced by byte count.  A unit (one request with all its backend
    responses, or one stream with all its chunks) is never split across ranks."""
    n = len(unit_bytes)
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    csum = np.concatenate([[0], np.cumsum(np.asarray(unit_bytes, dtype=np.int64))])
    total = int(csum[-1])
    cuts = [0]
    for r in range(1, world_size):
        target = total * r // world_size
        k = int(np.searchsorted(csum, target, side="left"))
        k = min(max(k, cuts[-1]), n)
        cuts.append(k)
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world_size)]


def shard_offsets(offsets: np.ndarray, group_off: np.ndarray, lo: int, hi: int):
    """Slice a two-level ragged layout (groups of items over a flat byte buffer) to groups [lo, hi).
    Returns (byte_lo, byte_hi, item_off_local, group_off_local)."""
    i0, i1 = int(group_off[lo]), int(group_off[hi])
    b0, b1 = int(offsets[i0]), int(offsets[i1])
    return b0, b1, offsets[i0:i1 + 1] - b0, group_off[lo:hi + 1] - i0
