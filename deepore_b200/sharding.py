"""Multi-GPU partitioning of the hot path: windows are independent units, so the genome is cut into
contiguous runs of windows, one run per rank, balanced by window count.  Each rank builds its own interval shards
(reads overlapping its windows, i.e. its own +-flank halo of the BAM) and runs encode -> model -> vote locally; there is
NO data-path collective.  Results (int8 call per image row + row->position map) are gathered on the host in window
order.  The reference's counterpart is the per-chromosome process fan-out of src/utils/pileupProcessor.sh:19-25.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def rank_range(n_windows: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [w0, w1) of rank `rank`: sizes differ by at most one, earlier ranks get the remainder."""
    base, rem = divmod(int(n_windows), int(world))
    w0 = rank * base + min(rank, rem)
    return w0, w0 + base + (1 if rank < rem else 0)


def all_ranges(n_windows: int, world: int) -> List[Tuple[int, int]]:
    return [rank_range(n_windows, world, r) for r in range(world)]


def gather_on_host(local: np.ndarray, group=None, dst: int = 0):
    """Concatenate per-rank result arrays (first axis = windows, rank order = window order) on rank `dst` through the
    process group's host path (gloo or the CPU side of nccl's object gather).  Returns the array on dst, None elsewhere."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    parts = [None] * dist.get_world_size(group) if dist.get_rank(group) == dst else None
    dist.gather_object(np.ascontiguousarray(local), parts, dst=dst, group=group)
    if parts is None:
        return None
    return np.concatenate(parts, axis=0)
