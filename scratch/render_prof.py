"""Scratch (PROF=1 build): per-role cycle accounting of k_render for CTA 0."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
import bench
N, L, F = 16384, 12, 512
ctx = J.Context(0)
audio, params, band, delay = bench.make_inputs(N, 2, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix0 = pan.ProcessVBAPData(lay, params[0]).reshape(N, L)
mode = J.MODE_THROUGHPUT if "throughput" in sys.argv else J.MODE_PARITY
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix0, mode=mode)
bus = np.zeros((F, L), np.float32)
for k in range(4):
    rend.UpdateParameters(band[k % 2], delay[k % 2], mix0)
    rend.ProcessInterleaved(audio[k % 2], bus, F)
ctx.synchronize()
lib = ctypes.CDLL(J.lib_path)
out = (ctypes.c_ulonglong * 15)()
lib.jplg_debug_render_prof(out)
for slot, name in enumerate(["D delay", "A front", "M mix", "Ba back lanes 0,1", "Bb back lanes 2,3"]):
    tot, wait, spin = out[3 * slot], out[3 * slot + 1], out[3 * slot + 2]
    print(f"{name}: lifetime {tot} cycles, in bar.sync {wait} ({100.0 * wait / max(tot, 1):.1f} %), in mbarrier / counter spin waits {spin} ({100.0 * spin / max(tot, 1):.1f} %), busy {tot - wait - spin}")
