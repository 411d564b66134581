"""CPU tests of the multi-GPU host logic (world_size 2, gloo): contiguous source shards, per-rank partial buses summed by ONE
reduce per block. The per-rank render here is the oracle port (test infrastructure, CPU) standing in for the GPU kernels, so the
test exercises exactly what bench.py does at N > 1: shard -> render partial bus -> reduce to rank 0 -> compare with the unsharded bus."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_partition_the_sources():
    from jplspatial_b200.sharding import shard_range, all_shards
    for n in (0, 1, 7, 16384, 262144, 100003):
        for world in (1, 2, 3, 4, 8):
            sh = all_shards(n, world)
            assert sh[0][0] == 0 and sh[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
            sizes = [e - b for b, e in sh]
            assert max(sizes) - min(sizes) <= 1
    assert shard_range(262144, 8, 3) == (98304, 131072)
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_weighted_shards_balance_cost():
    """Mixed layouts (BASELINE config 4: channel counts from {1,2,4,5}): balance by source channels x virtual sources, not by source count."""
    from jplspatial_b200.sharding import all_shards
    rng = np.random.default_rng(7)
    ch = rng.choice([1, 2, 4, 5], size=4096)
    cost = ch * np.select([ch == 1, ch == 2, ch == 4, ch == 5], [288, 144, 60, 48])
    sh = all_shards(len(cost), 8, cost.tolist())
    assert sh[0][0] == 0 and sh[-1][1] == len(cost) and all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
    loads = np.array([cost[b:e].sum() for b, e in sh], np.float64)
    assert loads.max() / loads.mean() < 1.02
    assert all_shards(5, 2, [0, 0, 0, 0, 0]) == all_shards(5, 2)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from jplspatial_b200.sharding import shard_range, reduce_bus
    from oracle import portpy
    from tests import util
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        N, L, frames, blocks = 48, 12, 512, 2
        M = __import__("jplspatial_b200").ChannelMask
        port_ = portpy.Port()
        params = util.pan_params(N * (blocks + 1), util.SEED + 40, edge=False)
        mix = port_.mdap(M.Surround_7_1_4, M.Mono, params).reshape(blocks + 1, N, L)
        band = (0.1 + 0.9 * util.uniform01((blocks + 1) * N * 4, util.SEED + 41)).reshape(blocks + 1, N, 4)
        delay = ((0.1 + 2.9 * util.uniform01((blocks + 1) * N, util.SEED + 42)) / 343.2).reshape(blocks + 1, N).astype(np.float32)
        audio = util.white_noise((blocks, N, frames, 1), util.SEED + 43)
        b, e = shard_range(N, world, rank)
        fx = [port_.dse(48000.0, 1, L, band[0, i], float(delay[0, i]), mix[0, i]) for i in range(b, e)]
        buses = []
        for blk in range(blocks):
            partial = np.zeros((frames, L), np.float32)
            for k, i in enumerate(range(b, e)):
                fx[k].update(band[blk + 1, i], float(delay[blk + 1, i]), mix[blk + 1, i])
                fx[k].process(audio[blk, i], partial, frames)
            t = torch.from_numpy(partial)
            reduce_bus(t, dst=0)
            buses.append(t.numpy().copy())
        if rank == 0:
            np.save(out_path, np.stack(buses))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_reduce_equals_unsharded_bus(tmp_path):
    import torch.multiprocessing as mp
    from oracle import portpy
    if not portpy.available():
        pytest.fail("oracle/libjploracle.so missing: run __graft_entry__.build()")
    port = _free_port()
    out2 = str(tmp_path / "bus_w2.npy"); out1 = str(tmp_path / "bus_w1.npy")
    mp.spawn(_worker, args=(2, port, out2), nprocs=2, join=True)
    mp.spawn(_worker, args=(1, _free_port(), out1), nprocs=1, join=True)
    a, b = np.load(out2), np.load(out1)
    assert a.shape == b.shape == (2, 512, 12)
    assert np.abs(b).max() > 0.1
    # the two-rank sum re-associates the fp32 additions: (s0 + ... + s23) + (s24 + ... + s47) vs sequential
    np.testing.assert_allclose(a, b, rtol=0, atol=48 * np.finfo(np.float32).eps * np.abs(b).max())


def _worker_pipelined(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from jplspatial_b200.sharding import PipelinedBusReduce
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames, L, blocks = 64, 12, 7
        red = PipelinedBusReduce(frames, L, torch.device("cpu"), depth=2, dst=0)
        got = []
        for k in range(blocks):
            bus = red.begin(k)
            assert not bus.any(), "a recycled bus buffer must come back zeroed"
            bus += float((rank + 1) * (k + 1))          # this rank's partial bus of block k
            red.submit(k)
            if k >= 1:
                got.append(red.result(k - 1).clone())   # block k-1's reduce was in flight while block k was produced
        got.append(red.result(blocks - 1).clone())
        red.drain()
        if rank == 0:
            np.save(out_path, torch.stack(got).numpy())
    finally:
        dist.destroy_process_group()


def test_pipelined_bus_reduce_two_ranks(tmp_path):
    """double-buffered asynchronous reduce: every block's bus arrives complete and in order on the destination rank"""
    import torch.multiprocessing as mp
    out = str(tmp_path / "pipe.npy")
    mp.spawn(_worker_pipelined, args=(2, _free_port(), out), nprocs=2, join=True)
    a = np.load(out)
    assert a.shape == (7, 64, 12)
    for k in range(7):
        assert np.all(a[k] == 3.0 * (k + 1))             # (1 + 2) * (k + 1)
