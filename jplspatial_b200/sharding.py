"""Multi-GPU plumbing of the renderer: sources are independent units, so they are sharded across ranks in contiguous
ranges (each rank owns its sources' rings and filter state for the whole session: no migration, no data-path collective)
and the ONLY exchange is the sum of the per-rank partial speaker buses ([frames][L] floats, 24.6 KB for 7.1.4) once per block.

One process per GPU; `torch.distributed` (NCCL on the GPU box, gloo in the CPU tests) carries the reduce.
Pure Python + torch.distributed (host-side logic only).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple


def shard_range(num_units: int, world: int, rank: int, weights: Optional[Sequence[float]] = None) -> Tuple[int, int]:
    """Contiguous [begin, end) of the units (sources) rank `rank` owns.

    Without weights: the first `num_units % world` ranks get one extra unit. With weights (cost per unit, e.g. source
    channels x virtual sources for mixed layouts): boundaries are placed where the running cost crosses k/world of the total,
    so every unit belongs to exactly one rank and ranges stay contiguous and ordered."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    if num_units < 0:
        raise ValueError("num_units must be >= 0")
    if weights is None:
        base, extra = divmod(num_units, world)
        begin = rank * base + min(rank, extra)
        return begin, begin + base + (1 if rank < extra else 0)
    if len(weights) != num_units:
        raise ValueError("weights must have one entry per unit")
    total = float(sum(weights))
    if total <= 0.0:
        return shard_range(num_units, world, rank)
    bounds = [0]
    acc, k = 0.0, 1
    for i, w in enumerate(weights):
        if w < 0:
            raise ValueError("weights must be >= 0")
        acc += float(w)
        while k < world and acc >= total * k / world:
            bounds.append(i + 1)
            k += 1
    while len(bounds) < world:
        bounds.append(num_units)
    bounds.append(num_units)
    return bounds[rank], bounds[rank + 1]


def all_shards(num_units: int, world: int, weights: Optional[Sequence[float]] = None):
    return [shard_range(num_units, world, r, weights) for r in range(world)]


def reduce_bus(bus, dst: int = 0, group=None, all_ranks: bool = False):
    """Sum the per-rank partial speaker buses. `bus` is a torch tensor [frames, L] (CUDA tensor with NCCL, CPU tensor with gloo);
    after the call rank `dst` (or every rank if all_ranks) holds the full bus. No-op when torch.distributed is not initialised."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return bus
    if all_ranks:
        dist.all_reduce(bus, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.reduce(bus, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return bus
