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


class PipelinedBusReduce:
    """Overlaps the reduce of block k's bus with the render of block k+1: `depth` bus buffers are used in turn;
    `begin(k)` hands out a zeroed buffer once the reduce that last used it has completed (stream-ordered, no host sync),
    `submit(k)` starts the asynchronous reduce of that buffer (NCCL runs it on its own stream after the work enqueued so far).
    The per-block latency still contains the reduce; the throughput does not. With one rank it degenerates to plain buffers."""

    def __init__(self, frames: int, channels: int, device, depth: int = 2, dst: int = 0, group=None, all_ranks: bool = False):
        import torch
        self.buses = [torch.zeros((frames, channels), dtype=torch.float32, device=device) for _ in range(max(1, depth))]
        self.pending = [None] * len(self.buses)
        self.dst, self.group, self.all_ranks = dst, group, all_ranks

    def _active(self):
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1

    def begin(self, k: int):
        j = k % len(self.buses)
        if self.pending[j] is not None:
            self.pending[j].wait()          # the current stream waits for that reduce; the host does not block (NCCL) / blocks briefly (gloo)
            self.pending[j] = None
        self.buses[j].zero_()
        return self.buses[j]

    def submit(self, k: int):
        j = k % len(self.buses)
        if not self._active():
            return None
        import torch.distributed as dist
        if self.all_ranks:
            self.pending[j] = dist.all_reduce(self.buses[j], op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        else:
            self.pending[j] = dist.reduce(self.buses[j], dst=self.dst, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return self.pending[j]

    def result(self, k: int):
        """the reduced bus of block k (valid on `dst`, or everywhere with all_ranks) once its reduce has completed"""
        j = k % len(self.buses)
        if self.pending[j] is not None:
            self.pending[j].wait()
            self.pending[j] = None
        return self.buses[j]

    def drain(self):
        for j, w in enumerate(self.pending):
            if w is not None:
                w.wait()
                self.pending[j] = None


def make_peer_bus_exchange(ctx, max_floats: int, group=None):
    """Sets up the peer-memory bus combine (include/jplg.h "Multi-GPU") across the ranks of `group`: rank 0 allocates the gather buffer, its 64-byte
    CUDA IPC handle is broadcast with torch.distributed (a CUDA uint8 tensor with NCCL, a CPU tensor with gloo), every other rank maps it.
    Returns a jplspatial_b200.BusExchange; the per-block data path then never touches torch.distributed / NCCL."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from . import BusExchange
    from ._capi import IPC_HANDLE_BYTES
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", ctx.device) if backend == "nccl" else torch.device("cpu")
    if rank == 0:
        x, handle = BusExchange.create_root(ctx, world, max_floats)
        t = torch.from_numpy(handle.copy()).to(dev)
    else:
        x = None
        t = torch.zeros(IPC_HANDLE_BYTES, dtype=torch.uint8, device=dev)
    dist.broadcast(t, src=0, group=group)
    if rank != 0:
        x = BusExchange.open_peer(ctx, world, rank, max_floats, t.cpu().numpy())
    dist.barrier(group)
    return x
