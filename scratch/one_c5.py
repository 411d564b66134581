"""Scratch: a few blocks of config 5's two banks (4,096 x FDN16 reverb, 4,096 x 16-tap ER bus) for ncu captures: python scratch/one_c5.py [blocks]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util
N, L, F, FS, TAPS = 4096, 12, 512, 48000.0, 16
blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 6
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
x = util.white_noise((4, N, F), 5) * np.float32(0.5)
rt = (0.3 + 2.7 * util.uniform01(N * 4, 6)).reshape(N, 4).astype(np.float32)
rev = J.ReverbBusBatch(ctx, N, FS, L); rev.SetRT60(rt)
er = J.ERBusBatch(ctx, N, TAPS, FS, L)
u = util.uniform01(N * TAPS * (6 + L), 7).reshape(N, TAPS, 6 + L)
for i in range(N):
    er.SetTaps(i, 0.1 + 0.9 * u[i, :, :4], 0.002 + 0.078 * u[i, :, 4], np.arange(1, TAPS + 1), u[i, :, 6:])
with torch.cuda.stream(stream):
    d_x = torch.from_numpy(x).to(dev)
    d_rev = torch.zeros((N, F, L), dtype=torch.float32, device=dev); d_er = torch.zeros((N, F, L), dtype=torch.float32, device=dev)
ctx.synchronize()
for k in range(blocks):
    rev.ProcessInterleaved(d_x[k % 4], d_rev, F)
    er.ProcessInterleaved(d_x[k % 4].reshape(N, F, 1), d_er, F)
ctx.synchronize()
print("ok", float(d_rev.abs().max()), float(d_er.abs().max()))
