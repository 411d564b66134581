"""Worker of tests/test_gpu_exchange.py, launched with torchrun (one rank per GPU): the peer-memory bus combine (jplg_bus_exchange_*) against a
rank-ordered sum of the same partial buses gathered with torch.distributed. Exits non-zero on mismatch.

Phases: (1) lock step, every second block ACCUMULATED into a non-zero output; (2) six blocks in flight without any per-block synchronisation (the
schedule bench.py uses); (3) one rank submits late; (4) a rank skips a block: the root's gather must time out, raise the error, leave its output
untouched, and the exchange must work again after the collective jplg_bus_exchange_clear_error."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import jplspatial_b200 as J                      # noqa: E402
from jplspatial_b200.sharding import make_peer_bus_exchange  # noqa: E402
from tests import util                           # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
ctx = J.Context(local)
stream = ctx.torch_stream()
N, L, F, FS = 300 + 17 * rank, 12, 512, 48000.0
LOCK, PIPE, LATE = 5, 6, 2
BLOCKS = LOCK + PIPE + LATE
seed = util.SEED + 1000 * rank
band = (0.1 + 0.9 * util.uniform01(N * 4, seed + 1)).reshape(N, 4)
delay = ((1.0 + 99.0 * util.uniform01(N, seed + 2)) / 343.2).astype(np.float32)
mix = util.uniform01(N * L, seed + 3).reshape(N, L)
rend = J.DirectSoundBatch(ctx, FS, N, 1, L, band, delay, mix, ring_size=16384, max_block_frames=F)
x = make_peer_bus_exchange(ctx, F * L)
ok = True
base = (util.uniform01(F * L, util.SEED + 77).reshape(F, L) - 0.5).astype(np.float32)     # what an accumulated output holds before the block
accumulate = [b % 2 == 1 for b in range(BLOCKS)]
with torch.cuda.stream(stream):
    scratch = [torch.zeros((F, L), dtype=torch.float32, device=dev) for _ in range(BLOCKS)]
    outs = [(torch.from_numpy(base).to(dev) if accumulate[b] else torch.zeros((F, L), dtype=torch.float32, device=dev)) for b in range(BLOCKS)]
    d_audio = [torch.from_numpy(util.white_noise((N, F, 1), seed + 10 + b)).to(dev) for b in range(BLOCKS)]
ctx.synchronize()


def block(b):
    # scratch[b] was zero, so after the block it equals this rank's partial bus exactly (the renderer's device bus is overwritten by the next block)
    rend.ProcessInterleaved(d_audio[b], scratch[b], F)
    x.submit(rend.device_bus_ptr(), F * L, b, out=outs[b] if rank == 0 else None, accumulate=accumulate[b])


for b in range(LOCK):                       # (1) lock step
    block(b)
    x.wait()
    ctx.synchronize()
for b in range(LOCK, LOCK + PIPE):          # (2) no synchronisation of any kind between the blocks
    block(b)
x.wait()
ctx.synchronize()
for b in range(LOCK + PIPE, BLOCKS):        # (3) the last rank is late (well inside the time-out)
    if rank == world - 1:
        time.sleep(0.3)
    block(b)
x.wait()
ctx.synchronize()
torch.cuda.synchronize()
parts = np.stack([s.cpu().numpy() for s in scratch])
gathered = [None] * world
dist.gather_object(parts, gathered if rank == 0 else None, dst=0)
err = x.error()
if rank == 0:
    for b in range(BLOCKS):
        want = gathered[0][b].copy()
        for r in range(1, world):
            want = want + gathered[r][b]            # rank order, float32: the order the gather kernel uses
        if accumulate[b]:
            want = base + want
        got = outs[b].cpu().numpy()
        if not util.bits_equal(got, want):
            ok = False
            print(f"block {b}: exchange differs from the rank-ordered sum, max abs {np.abs(got - want).max()}", flush=True)
        if not np.abs(want).max() > 0:
            ok = False
    print("exchange worker: world", world, "blocks", BLOCKS, "ok" if ok else "FAILED", "error flag", err, flush=True)

# (4) a missing rank: fail closed, then recover
dist.barrier()
x.set_timeout_ms(100)
sentinel = torch.full((F, L), 123.0, dtype=torch.float32, device=dev)
if rank != world - 1:
    x.submit(rend.device_bus_ptr(), F * L, BLOCKS, out=sentinel if rank == 0 else None, accumulate=False)
    x.wait()
ctx.synchronize()
torch.cuda.synchronize()
dist.barrier()
if rank == 0:
    e = x.poll_error()
    untouched = bool((sentinel == 123.0).all().item())
    print("time-out phase: root error", e, "output untouched", untouched, flush=True)
    if e == 0 or not untouched:
        ok = False
x.clear_error()
dist.barrier()
x.set_timeout_ms(2000)
with torch.cuda.stream(stream):
    fresh = torch.zeros((F, L), dtype=torch.float32, device=dev)
    part = torch.zeros((F, L), dtype=torch.float32, device=dev)
rend.ProcessInterleaved(d_audio[0], part, F)
x.submit(rend.device_bus_ptr(), F * L, 0, out=fresh if rank == 0 else None, accumulate=False)
x.wait()
ctx.synchronize()
gathered = [None] * world
dist.gather_object(part.cpu().numpy(), gathered if rank == 0 else None, dst=0)
err2 = x.error()
if rank == 0:
    want = gathered[0].copy()
    for r in range(1, world):
        want = want + gathered[r]
    good = util.bits_equal(fresh.cpu().numpy(), want)
    print("after clear_error: bit-equal", good, "error flag", err2, flush=True)
    ok = ok and good
flag = torch.tensor([0 if (ok and err == 0 and err2 == 0) else 1], device=dev)
dist.all_reduce(flag)
x.close(); rend.close(); ctx.close()
dist.destroy_process_group()
sys.exit(1 if int(flag.item()) else 0)
