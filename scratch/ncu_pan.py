"""ncu target: a few launches of k_mdap_3d<16> (BASELINE config 2 shape: mono -> 7.1.4) and k_scene_frontend at 65,536 sources."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util
from tests.frontend_inputs import make_positions
M = J.ChannelMask
n = 65536
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
pan = J.VBAPanner(ctx); assert pan.Initialize(M.Surround_7_1_4); lay = pan.InitializeSourceLayout(M.Mono)
fe = J.DirectPathBatch(ctx, functions=[("model", 1, 1.0, 200.0, 1.0)])
src = make_positions(n, util.SEED + 1, 150.0); lis = make_positions(1, util.SEED + 2, 1.0)
with torch.cuda.stream(stream):
    p = torch.from_numpy(util.pan_params(n, util.SEED + 3)).to(dev)
    s = torch.from_numpy(src).to(dev); l = torch.from_numpy(lis).to(dev)
    for _ in range(4):
        mix = pan.ProcessVBAPData(lay, p)
        o = fe.Process(n, ("direct", "pan_params", "band_gains", "delay_seconds", "distance_attenuation"), sources=s, listeners=l, focus=0.5)
ctx.synchronize()
print("ok", float(mix.sum()))
