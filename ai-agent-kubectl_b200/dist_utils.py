"""Host-side helpers for the data-parallel (replica) layout: every rank owns a
contiguous shard of the requests / strings, there is no collective on the data path;
only the timing is reduced (MAX over ranks)."""
from __future__ import annotations

from typing import Tuple


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [begin, end) of `total` units owned by `rank`; shards differ by at most one unit
    and cover every unit exactly once."""
    base, extra = divmod(total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def max_over_ranks(values, device=None):
    """MAX-reduce a list of floats over the default process group (no-op when not initialised)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]
