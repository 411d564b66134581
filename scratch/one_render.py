"""A few blocks of BASELINE config 3 (device resident, serial order) for ncu: python scratch/one_render.py [blocks] [sources] [mode]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
import bench
blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 6
N = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
mode = J.MODE_THROUGHPUT if (len(sys.argv) > 3 and sys.argv[3] == "throughput") else J.MODE_PARITY
L, F = 12, 512
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
audio, params, band, delay = bench.make_inputs(N, 2, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix0 = pan.ProcessVBAPData(lay, params[0]).reshape(N, L)
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix0, ring_size=16384, mode=mode)
with torch.cuda.stream(stream):
    d_audio = torch.from_numpy(audio).to(dev); d_params = torch.from_numpy(params).to(dev)
    d_band = torch.from_numpy(band).to(dev); d_delay = torch.from_numpy(delay).to(dev)
    bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
ctx.synchronize()
for k in range(blocks):
    rend.UpdatePanned(pan, lay, d_params[k % 2], d_band[k % 2], d_delay[k % 2])
    rend.ProcessInterleaved(d_audio[k % 2], bus, F)
ctx.synchronize()
print("done", float(bus.abs().max().item()))
