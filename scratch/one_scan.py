"""A few throughput-mode blocks through the time-parallel path, for ncu: python scratch/one_scan.py [blocks] [sources]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
import bench
blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 6
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
L, F = 12, 512
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
audio, params, band, delay = bench.make_inputs(N, 2, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix = [pan.ProcessVBAPData(lay, params[k]).reshape(N, L) for k in range(2)]
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix[0], ring_size=65536, mode=J.MODE_THROUGHPUT)
rend.SetTimeParallel(1)
with torch.cuda.stream(stream):
    d_audio = torch.from_numpy(audio).to(dev)
    bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
ctx.synchronize()
for k in range(blocks):
    rend.UpdateParameters(band[k % 2], delay[k % 2], mix[k % 2])
    rend.ProcessInterleaved(d_audio[k % 2], bus, F)
ctx.synchronize()
print("done", rend.time_parallel_blocks, float(bus.abs().max().item()))
