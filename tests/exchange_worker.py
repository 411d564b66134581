"""Worker of tests/test_gpu_exchange.py, launched with torchrun (one rank per GPU): the peer-memory bus combine (jplg_bus_exchange_*) against a
rank-ordered sum of the same partial buses gathered with torch.distributed. Exits non-zero on mismatch."""
import os
import sys

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
N, L, F, FS, BLOCKS = 300 + 17 * rank, 12, 512, 48000.0, 9
seed = util.SEED + 1000 * rank
band = (0.1 + 0.9 * util.uniform01(N * 4, seed + 1)).reshape(N, 4)
delay = ((1.0 + 99.0 * util.uniform01(N, seed + 2)) / 343.2).astype(np.float32)
mix = util.uniform01(N * L, seed + 3).reshape(N, L)
rend = J.DirectSoundBatch(ctx, FS, N, 1, L, band, delay, mix, ring_size=16384, max_block_frames=F)
x = make_peer_bus_exchange(ctx, F * L)
ok = True
with torch.cuda.stream(stream):
    scratch = torch.zeros((F, L), dtype=torch.float32, device=dev)
    outs = [torch.zeros((F, L), dtype=torch.float32, device=dev) for _ in range(BLOCKS)]
    parts = []
for b in range(BLOCKS):
    audio = util.white_noise((N, F, 1), seed + 10 + b)
    with torch.cuda.stream(stream):
        d_audio = torch.from_numpy(audio).to(dev)
        scratch.zero_()
        rend.ProcessInterleaved(d_audio, scratch, F)
        # keep a copy of this rank's partial bus for the check (the renderer's device bus is overwritten by the next block)
        x.submit(rend.device_bus_ptr(), F * L, b, out=outs[b] if rank == 0 else None, accumulate=(b % 2 == 1))
    x.wait()
    ctx.synchronize()
    parts.append(scratch.cpu().numpy().copy())      # scratch was zero before the block, so it now equals this rank's partial bus exactly
torch.cuda.synchronize()
gathered = [None] * world
dist.gather_object(np.stack(parts), gathered if rank == 0 else None, dst=0)
err = x.error()
if rank == 0:
    for b in range(BLOCKS):
        want = gathered[0][b].copy()
        for r in range(1, world):
            want = want + gathered[r][b]            # rank order, float32: the order the gather kernel uses
        got = outs[b].cpu().numpy()
        if not util.bits_equal(got, want):
            ok = False
            print(f"block {b}: exchange differs from the rank-ordered sum, max abs {np.abs(got - want).max()}", flush=True)
        if not np.abs(want).max() > 0:
            ok = False
    print("exchange worker: world", world, "blocks", BLOCKS, "ok" if ok else "FAILED", "error flag", err, flush=True)
flag = torch.tensor([0 if (ok and err == 0) else 1], device=dev)
dist.all_reduce(flag)
x.close(); rend.close(); ctx.close()
dist.destroy_process_group()
sys.exit(1 if int(flag.item()) else 0)
