"""config 4 pieces one at a time (nothing concurrent): per layout, kernel durations of the fused panning update, push, render, bus reduce"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import jplspatial_b200 as J
from tests import util
M = J.ChannelMask
FS, F, L = 48000.0, 512, 16
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
out = {}
for li, (name, C) in enumerate((("Mono", 1), ("Stereo", 2), ("Quad", 4), ("Surround_5_1", 6))):
    ctx = J.Context(0)
    dev = torch.device("cuda", 0)
    pan = J.VBAPanner(ctx); assert pan.Initialize(M.Surround_9_1_6)
    lay = pan.InitializeSourceLayout(getattr(M, name))
    params = np.stack([util.pan_params(n, 5 + 10 * b, edge=False) for b in range(2)])
    band = (0.1 + 0.9 * util.uniform01(2 * n * 4, 7)).reshape(2, n, 4).astype(np.float32)
    dist0 = 1.0 + 299.0 * util.uniform01(n, 8); speed = (util.uniform01(n, 11) * 2.0 - 1.0) * 30.0
    tt = (np.arange(2, dtype=np.float32) * np.float32(F / FS))[:, None]
    delay = (np.maximum(dist0[None] + speed[None] * tt, 0.5) / np.float32(343.2)).astype(np.float32)
    audio = util.white_noise((2, n, F, C), 9)
    rend = J.DirectSoundBatch(ctx, FS, n, C, L, band[0], delay[0], np.zeros((n, C * L), np.float32), ring_size=0, max_block_frames=F)
    st = ctx.torch_stream()
    with torch.cuda.stream(st):
        bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
        dp, db, dd, da = (torch.from_numpy(x).to(dev) for x in (params, band, delay, audio))
    ctx.synchronize()
    for k in range(4):
        rend.UpdatePanned(pan, lay, dp[k & 1], db[k & 1], dd[k & 1]); rend.ProcessInterleaved(da[k & 1], bus, F)
    ctx.synchronize()
    ctx.set_kernel_timing(True); ctx.reset_kernel_timing()
    steps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for k in range(steps):
        rend.UpdatePanned(pan, lay, dp[k & 1], db[k & 1], dd[k & 1]); rend.ProcessInterleaved(da[k & 1], bus, F)
    e1.record(st)
    ctx.synchronize()
    out[name] = {k: round(ctx.kernel_time(k)[0] / steps, 4) for k in ("mdap", "push", "render", "bus_reduce")}
    out[name]["step_ms"] = round(e0.elapsed_time(e1) / steps, 4)
    out[name]["source_channels"] = n * C
    ctx.set_kernel_timing(False)
    rend.close()
    print(name, out[name], flush=True)
print(json.dumps(out))
