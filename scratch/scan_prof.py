"""Scratch (PROF=1 build): time stamps of k_scan_front_pipe's stages for CTA 0."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import subprocess
subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "one_scan.py"), "6", sys.argv[1] if len(sys.argv) > 1 else "1024"], check=True) if False else None
import torch, jplspatial_b200 as J, bench
N, L, F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 12, 512
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
audio, params, band, delay = bench.make_inputs(N, 2, 0xB200)
pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4); lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
mix = [pan.ProcessVBAPData(lay, params[k]).reshape(N, L) for k in range(2)]
rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix[0], ring_size=65536, mode=J.MODE_THROUGHPUT); rend.SetTimeParallel(1)
with torch.cuda.stream(stream):
    d_audio = torch.from_numpy(audio).to(dev); bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
for k in range(6):
    rend.UpdateParameters(band[k % 2], delay[k % 2], mix[k % 2]); rend.ProcessInterleaved(d_audio[k % 2], bus, F)
ctx.synchronize()
out = (ctypes.c_ulonglong * 16)(); ctypes.CDLL(J.lib_path).jplg_debug_scan_prof(out)
t0 = out[0]
names = ["start", "push done", "P2 done (warp 0)", "P2 done (warp 3)", "P4 done (warp 0)", "P4 done (warp 3)", "P1 done", "P3 done", "P5 done (sc 0)", "P6 done (warp 0)", "P6 done (warp 3)", "after final barrier"]
for i, n in enumerate(names): print(f"{n:24s} {out[i] - t0:8d} cycles = {(out[i] - t0) / 1.965e3:6.2f} us")

cta = (ctypes.c_ulonglong * 3072)(); ctypes.CDLL(J.lib_path).jplg_debug_scan_cta(cta)
n = (N + 3) // 4
base = cta[0]; st = np.array([cta[3 * i] - base for i in range(n)], np.float64); en = np.array([cta[3 * i + 1] - base for i in range(n)], np.float64); sm = np.array([cta[3 * i + 2] for i in range(n)])
t0 = st.min()
print("CTAs", n, "start us: min %.2f max %.2f" % (0.0, (st.max() - t0) / 1e3), "end us: min %.2f max %.2f" % ((en.min() - t0) / 1e3, (en.max() - t0) / 1e3), "duration us: min %.2f median %.2f max %.2f" % ((en - st).min() / 1e3, np.median(en - st) / 1e3, (en - st).max() / 1e3))
import collections
per = collections.Counter(sm.tolist()); print("CTAs per SM:", collections.Counter(per.values()))
late = np.argsort(st)[-8:]; print("latest starts:", [(int(i), round((st[i] - t0) / 1e3, 2), round((en[i] - st[i]) / 1e3, 2), int(sm[i])) for i in late])

print("CTA 0: %d cycles in %.2f us -> SM clock %.0f MHz" % (out[13] - out[12], (en[0] - st[0]) / 1e3, (out[13] - out[12]) / max(en[0] - st[0], 1) * 1e3))
