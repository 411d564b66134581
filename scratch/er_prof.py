"""Scratch (PROF=1 build): per-role cycle accounting of k_render<ER> for CTA 0 (4,096 ER buses x 16 taps) + event-timed kernels."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util
N, L, F, FS, TAPS = 4096, 12, 512, 48000.0, 16
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
x = util.white_noise((4, N, F), 5) * np.float32(0.5)
er = J.ERBusBatch(ctx, N, TAPS, FS, L)
u = util.uniform01(N * TAPS * (6 + L), 7).reshape(N, TAPS, 6 + L)
for i in range(N):
    er.SetTaps(i, 0.1 + 0.9 * u[i, :, :4], 0.002 + 0.078 * u[i, :, 4], np.arange(1, TAPS + 1), u[i, :, 6:])
with torch.cuda.stream(stream):
    d_x = torch.from_numpy(x).to(dev); d_er = torch.zeros((N, F, L), dtype=torch.float32, device=dev)
ctx.synchronize()
for k in range(6):
    er.ProcessInterleaved(d_x[k % 4].reshape(N, F, 1), d_er, F)
ctx.synchronize()
ctx.set_kernel_timing(True); ctx.reset_kernel_timing()
for k in range(20):
    er.ProcessInterleaved(d_x[k % 4].reshape(N, F, 1), d_er, F)
ctx.synchronize()
print({n: ctx.kernel_time(n) for n in ("er_push", "render", "other")})
lib = ctypes.CDLL(J.lib_path)
out = (ctypes.c_ulonglong * 15)()
lib.jplg_debug_render_prof(out)
for slot, name in enumerate(["D delay", "A front", "M mix", "Ba back lanes 0,1", "Bb back lanes 2,3"]):
    tot, wait, spin = out[3 * slot], out[3 * slot + 1], out[3 * slot + 2]
    print(f"{name}: lifetime {tot} cycles, in bar.sync {wait} ({100.0 * wait / max(tot, 1):.1f} %), in mbarrier / counter spin waits {spin} ({100.0 * spin / max(tot, 1):.1f} %), busy {tot - wait - spin}")
