"""Contiguous sharding of the schedule path across GPUs / ranks: every spec and every Cron's
history segment is independent, so there is no data-path collective -- ranks only meet for the benchmark's barrier and
max-over-ranks timing.  Mirrors the C++ sharding inside libb200cron.so's multi-device context (csrc/b200cron.cu)."""
from __future__ import annotations

import numpy as np


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """[lo, hi) of rank's specs; boundaries are even so 128-bit loads stay aligned on every shard"""
    lo = (n * rank // world) & ~1
    hi = n * (rank + 1) // world
    if rank + 1 < world:
        hi &= ~1
    return lo, hi


def history_cuts(seg_off: np.ndarray, world: int) -> np.ndarray:
    """segment index cuts (world + 1 entries) balancing records per rank; no segment straddles two ranks"""
    seg_off = np.asarray(seg_off, dtype=np.uint64)
    nseg = len(seg_off) - 1
    nrec = int(seg_off[-1] - seg_off[0])
    cuts = np.empty(world + 1, dtype=np.int64)
    cuts[0], cuts[world] = 0, nseg
    for g in range(1, world):
        target = int(seg_off[0]) + nrec * g // world
        c = int(np.searchsorted(seg_off, np.uint64(target), side="left"))
        cuts[g] = min(max(c, cuts[g - 1]), nseg)
    return cuts


def max_over_ranks(values, device=None):
    """element-wise max of a small list of floats over all ranks (identity when torch.distributed is not initialised)"""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()
