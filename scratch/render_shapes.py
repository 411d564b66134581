"""k_render launch time (CUDA events around the kernel, context stream) for a few (channels per source, output channels, sources) shapes."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util
F, FS = 512, 48000.0
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
res = []
for C, L, N in [(1, 12, 16384), (1, 12, 18944), (1, 12, 32768), (1, 12, 37888), (1, 16, 16384), (1, 16, 18944), (2, 16, 8192), (4, 16, 4096), (6, 16, 2688), (6, 16, 8192)]:
    band = (0.1 + 0.9 * util.uniform01(N * 4, 1)).reshape(N, 4); delay = ((1.0 + 299.0 * util.uniform01(N, 2)) / 343.2).astype(np.float32)
    mix = util.uniform01(N * C * L, 3).reshape(N, C * L)
    r = J.DirectSoundBatch(ctx, FS, N, C, L, band, delay, mix, ring_size=16384, max_block_frames=F)
    with torch.cuda.stream(stream):
        audio = torch.from_numpy(util.white_noise((N, F, C), 4)).to(dev); bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
        for _ in range(3): r.ProcessInterleaved(audio, bus, F)
    ctx.synchronize(); ctx.set_kernel_timing(True); ctx.reset_kernel_timing()
    with torch.cuda.stream(stream):
        for _ in range(20): r.ProcessInterleaved(audio, bus, F)
    ctx.synchronize()
    kt = {k: ctx.kernel_time(k) for k in ("push", "render", "bus_reduce")}
    ctx.set_kernel_timing(False)
    sc = N * C; ctas = -(-N // (128 // C))
    res.append({"C": C, "L": L, "sources": N, "source_channels": sc, "ctas": ctas, "render_us": 1e3 * kt["render"][0] / kt["render"][1],
                "push_us": 1e3 * kt["push"][0] / kt["push"][1], "ns_per_source_channel": 1e6 * kt["render"][0] / kt["render"][1] / sc})
    print(res[-1], flush=True)
    r.close(); del audio, bus
json.dump(res, open("gpurun_out/render_shapes.json", "w"), indent=1)
