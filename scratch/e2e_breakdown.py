"""Scratch: wall-time breakdown of the host-buffer (e2e) step, plus raw pinned H2D bandwidth."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
import bench

N, L, F = 16384, 12, 512
ctx = J.Context(0)
audio, params, band, delay = bench.make_inputs(N, 4, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix0 = pan.ProcessVBAPData(lay, params[0]).reshape(N, L)
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix0)
def pinned(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory(); return t, t.numpy()
keep = []
def P(a):
    t, n = pinned(a); keep.append(t); return n
h_audio = [P(audio[j]) for j in range(4)]; h_params = [P(params[j]) for j in range(4)]; h_band = [P(band[j]) for j in range(4)]; h_delay = [P(delay[j]) for j in range(4)]
h_mix = P(np.zeros((N, 1, L), np.float32)); h_bus = P(np.zeros((F, L), np.float32))
T = {"mdap": 0.0, "update": 0.0, "process": 0.0}
reps = 40
for k in range(reps + 5):
    j = k % 4
    t0 = time.perf_counter(); pan.ProcessVBAPData(lay, h_params[j], out=h_mix)
    t1 = time.perf_counter(); rend.UpdateParameters(h_band[j], h_delay[j], h_mix)
    t2 = time.perf_counter(); h_bus[:] = 0; rend.ProcessInterleaved(h_audio[j], h_bus, F)
    t3 = time.perf_counter()
    if k >= 5:
        T["mdap"] += t1 - t0; T["update"] += t2 - t1; T["process"] += t3 - t2
print("host-call wall times (us):", {k: round(v / reps * 1e6, 1) for k, v in T.items()}, "pipeline mode", os.environ.get("JPLG_HOST_PIPELINE", "default"))
# raw H2D
d = torch.empty(N * F, dtype=torch.float32, device="cuda")
src = torch.from_numpy(h_audio[0].reshape(-1))
for _ in range(3): d.copy_(src, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): d.copy_(src, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print(f"pinned H2D 33.5 MB: {dt * 1e6:.0f} us = {N * F * 4 / dt / 1e9:.1f} GB/s")
small = torch.empty(F * L, dtype=torch.float32, device="cuda"); hs = torch.from_numpy(h_bus.reshape(-1))
t0 = time.perf_counter()
for _ in range(200): hs.copy_(small, non_blocking=True); torch.cuda.synchronize()
print(f"small D2H + sync round trip: {(time.perf_counter() - t0) / 200 * 1e6:.1f} us")
