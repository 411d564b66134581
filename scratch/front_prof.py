"""Scratch (PROF=1 build): clock64 stamps of k_scan_front's phases for CTA 0 + the kernel's event-timed duration."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, jplspatial_b200 as J, bench
N, L, F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 12, 512
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
audio, params, band, delay = bench.make_inputs(N, 2, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4); lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix = [pan.ProcessVBAPData(lay, params[k]).reshape(N, L) for k in range(2)]
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix[0], ring_size=65536, mode=J.MODE_THROUGHPUT); rend.SetTimeParallel(1)
with torch.cuda.stream(stream):
    d_audio = torch.from_numpy(audio).to(dev); bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
for k in range(8):
    rend.UpdateParameters(band[k % 2], delay[k % 2], mix[k % 2]); rend.ProcessInterleaved(d_audio[k % 2], bus, F)
ctx.synchronize()
out = (ctypes.c_ulonglong * 16)(); ctypes.CDLL(J.lib_path).jplg_debug_front_prof(out)
names = {0: "entry", 1: "P0 push + P1 delay", 2: "P2 taps", 3: "P3 Thiran", 4: "P4 zero-state", 5: "P5 scan", 6: "P6 final", 15: "end"}
prev = out[0]
for i in (1, 2, 3, 4, 5, 6, 15):
    print(f"{names[i]:22s} {out[i] - prev:7d} cycles  (cumulative {out[i] - out[0]:7d} = {(out[i] - out[0]) / 1.96e3:6.2f} us)"); prev = out[i]

mx = (ctypes.c_ulonglong * 16)(); ctypes.CDLL(J.lib_path).jplg_debug_mix_prof(mx)
print("k_scan_mix CTA (0,0):", [int(mx[i] - mx[0]) for i in (1, 2, 3, 4, 5, 15) if mx[i]], "cycles after entry at its barriers / end")
