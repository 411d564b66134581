"""BASELINE config 4 (per-GPU share): 262,144 multichannel sources -> 9.1.6 over 8 GPUs = 32,768 sources per GPU, channel counts drawn from
{1, 2, 4, 6 (5.1: the LFE channel is rendered but not panned, PannerBase.h:937)} in equal parts; one renderer per source layout, all adding
into the same [512][16] bus, NCCL reduce of the bus across ranks when launched under torchrun. Device-resident timing with CUDA events.

    python scratch/bench_c4.py [--per-layout 8192] [--mode parity|throughput]
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scratch/bench_c4.py"""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util

ap = argparse.ArgumentParser()
ap.add_argument("--per-layout", type=int, default=8192)
ap.add_argument("--mode", default="parity")
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--ring", type=int, default=0)
ap.add_argument("--streams", type=int, default=4, help="4: one context (= one CUDA stream) per source layout so the four renderers' launches overlap "
                "(their CTA counts 64 / 128 / 256 / 384 leave tails on 148 SMs when run back to back); 1: everything on one stream")
args = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
M = J.ChannelMask
L, F, FS = 16, 512, 48000.0
mode = J.MODE_THROUGHPUT if args.mode == "throughput" else J.MODE_PARITY
ctx = J.Context(local); stream = ctx.torch_stream(); dev = torch.device("cuda", local)
ctxs = [ctx] + [J.Context(local) for _ in range(3)] if args.streams == 4 else [ctx] * 4
pans = []
for c in ctxs:
    if pans and c is ctx:
        pans.append(pans[0]); continue
    pn = J.VBAPanner(c); assert pn.Initialize(M.Surround_9_1_6); pans.append(pn)
LAYOUTS = [("Mono", M.Mono, 1), ("Stereo", M.Stereo, 2), ("Quad", M.Quad, 4), ("Surround_5_1", M.Surround_5_1, 6)]
n = args.per_layout
units = []
for li, (name, mask, C) in enumerate(LAYOUTS):
    pan, uctx = pans[li], ctxs[li]
    ustream = uctx.torch_stream()
    lay = pan.InitializeSourceLayout(mask)
    seed = 0xC4 + 97 * rank + li
    params = np.stack([util.pan_params(n, seed + 10 * b, edge=False) for b in range(2)])
    band = (0.1 + 0.9 * util.uniform01(2 * n * 4, seed + 1)).reshape(2, n, 4).astype(np.float32)
    delay = ((1.0 + 299.0 * util.uniform01(2 * n, seed + 2)) / 343.2).reshape(2, n).astype(np.float32)
    audio = util.white_noise((n, F, C), seed + 3)
    # rows of the renderer's mix map = source channels in ascending EChannel order; the layout's panned channels (LFE skipped) scatter into them
    _, ids = lay.groups()
    chan_bits = [b for b in range(22) if (int(mask) >> b) & 1]
    rows = [chan_bits.index(int(np.log2(int(i)))) if int(i) in [1 << b for b in chan_bits] else int(i) for i in ids]
    rend = J.DirectSoundBatch(uctx, FS, n, C, L, band[0], delay[0], np.zeros((n, C * L), np.float32), ring_size=args.ring, max_block_frames=F, mode=mode)
    with torch.cuda.stream(ustream):
        u = dict(lay=lay, rend=rend, C=C, pan=pan, ctx=uctx, stream=ustream, bus=torch.zeros((F, L), dtype=torch.float32, device=dev),
                 ready=torch.cuda.Event(), done=torch.cuda.Event(), rows=torch.tensor(rows, device=dev, dtype=torch.long),
                 params=torch.from_numpy(params).to(dev), band=torch.from_numpy(band).to(dev), delay=torch.from_numpy(delay).to(dev),
                 audio=torch.from_numpy(audio).to(dev), panned=torch.empty((n, lay.num_channels, L), dtype=torch.float32, device=dev),
                 mix=torch.zeros((n, C, L), dtype=torch.float32, device=dev))
    units.append(u)
with torch.cuda.stream(stream):
    bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
ctx.synchronize()

def step(k):
    j = k & 1
    with torch.cuda.stream(stream):
        start = torch.cuda.Event(); start.record(stream)
    for u in units:
        st = u["stream"]
        with torch.cuda.stream(st):
            if st is not stream:
                st.wait_event(start)        # the previous block's final sum has read this unit's bus
            u["bus"].zero_()
            u["pan"].ProcessVBAPData(u["lay"], u["params"][j], out=u["panned"])
            if u["panned"].shape[1] == u["C"]:
                mix = u["panned"]
            else:
                u["mix"].index_copy_(1, u["rows"], u["panned"]); mix = u["mix"]
            u["rend"].UpdateParameters(u["band"][j], u["delay"][j], mix)
            u["rend"].ProcessInterleaved(u["audio"], u["bus"], F)
            u["done"].record(st)
    with torch.cuda.stream(stream):
        for u in units:
            if u["stream"] is not stream:
                stream.wait_event(u["done"])
        torch.add(units[0]["bus"], units[1]["bus"], out=bus); bus.add_(units[2]["bus"]); bus.add_(units[3]["bus"])   # fixed order
        if dist is not None:
            dist.reduce(bus, dst=0)

for k in range(args.warmup): step(k)
torch.cuda.synchronize()
if dist is not None: dist.barrier()
for c in set(ctxs): c.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(stream): e0.record(stream)
for k in range(args.steps): step(args.warmup + k)
with torch.cuda.stream(stream): e1.record(stream)
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
if dist is not None: dist.all_reduce(t, op=dist.ReduceOp.MAX)
ms = float(t.item())
if rank == 0:
    total = 4 * n * world
    print(json.dumps({"config": "BASELINE config 4 share: per GPU %d sources (%d each of mono / stereo / quad / 5.1 = %d source-channels) -> 9.1.6, 48 kHz / 512 frames" % (4 * n, n, 13 * n),
                      "n_gpus": world, "mode": args.mode, "ms_per_block": ms, "total_sources": total, "realtime_sources": total * (F / FS) / (ms * 1e-3),
                      "bus_nonzero": bool(bus.abs().max().item() > 0),
                      "streams": args.streams}))
