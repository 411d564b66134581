"""Scratch: which transfers set the e2e step? A = all host (bench e2e), B = only the audio from the host, C = only the small arrays from the host."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
import bench

N, L, F, POOL = 16384, 12, 512, 4
ctx = J.Context(0); dev = torch.device("cuda", 0); stream = ctx.torch_stream()
audio, params, band, delay = bench.make_inputs(N, POOL, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix0 = pan.ProcessVBAPData(lay, params[0]).reshape(N, L)
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix0, max_block_frames=F)
keep = []
def P(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory(); keep.append(t); return t.numpy()
h_audio = [P(audio[j]) for j in range(POOL)]; h_params = [P(params[j]) for j in range(POOL)]; h_band = [P(band[j]) for j in range(POOL)]; h_delay = [P(delay[j]) for j in range(POOL)]
h_mix = P(np.zeros((N, 1, L), np.float32)); h_buses = (P(np.zeros((F, L), np.float32)), P(np.zeros((F, L), np.float32)))
with torch.cuda.stream(stream):
    d_audio = torch.from_numpy(audio).to(dev); d_params = torch.from_numpy(params).to(dev); d_band = torch.from_numpy(band).to(dev); d_delay = torch.from_numpy(delay).to(dev)
    d_mix = torch.empty((N, 1, L), dtype=torch.float32, device=dev); d_bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
ctx.synchronize()

def run(variant, steps=60):
    def submit(k):
        j = k % POOL; hb = h_buses[k & 1]
        small_host = variant in ("A", "C"); audio_host = variant in ("A", "B")
        with torch.cuda.stream(stream):
            if small_host:
                pan.ProcessVBAPData(lay, h_params[j], out=h_mix, async_host=True)
                rend.UpdateParameters(h_band[j], h_delay[j], h_mix, async_host=True)
            else:
                pan.ProcessVBAPData(lay, d_params[j], out=d_mix)
                rend.UpdateParameters(d_band[j], d_delay[j], d_mix)
            if audio_host:
                hb[:] = 0.0
                rend.ProcessInterleaved(h_audio[j], hb, F, async_host=True)
            else:
                rend.ProcessInterleaved(d_audio[j], d_bus, F)
    def collect(k):
        if variant in ("A", "B"): rend.WaitBlock()
        else: ctx.synchronize()
    W = 40
    for k in range(W):
        submit(k)
        if k: collect(k - 1)
    collect(W - 1); ctx.synchronize()
    t0 = time.perf_counter()
    for k in range(steps):
        submit(W + k)
        if k: collect(W + k - 1)
    collect(W + steps - 1); ctx.synchronize()
    return (time.perf_counter() - t0) / steps * 1e3

for v in ("A", "B", "C", "A", "B", "A"):
    print(v, f"{run(v):.4f} ms/step")
