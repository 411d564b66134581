"""BASELINE config 5: batched 16th-order FDN late reverb + early-reflection tap rendering, 4,096 instances (mono in, 7.1.4 out,
16 taps per ER bus). Device-resident timing with CUDA events + the reference's CPU code on a bounded sample."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util

N, L, F, FS, TAPS = 4096, 12, 512, 48000.0, 16
mode = J.MODE_THROUGHPUT if "throughput" in sys.argv else J.MODE_PARITY
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
x = util.white_noise((4, N, F), 5) * np.float32(0.5)
rt = (0.3 + 2.7 * util.uniform01(N * 4, 6)).reshape(N, 4).astype(np.float32)
rev = J.ReverbBusBatch(ctx, N, FS, L, mode=mode); rev.SetRT60(rt)
er = J.ERBusBatch(ctx, N, TAPS, FS, L, mode=mode)
u = util.uniform01(N * TAPS * (6 + L), 7).reshape(N, TAPS, 6 + L)
for i in range(N):
    er.SetTaps(i, 0.1 + 0.9 * u[i, :, :4], 0.002 + 0.078 * u[i, :, 4], np.arange(1, TAPS + 1), u[i, :, 6:])
with torch.cuda.stream(stream):
    d_x = torch.from_numpy(x).to(dev)
    d_rev = torch.zeros((N, F, L), dtype=torch.float32, device=dev)
    d_er = torch.zeros((N, F, L), dtype=torch.float32, device=dev)
ctx.synchronize()
def step(k):
    with torch.cuda.stream(stream):
        rev.ProcessInterleaved(d_x[k % 4], d_rev, F)
        er.ProcessInterleaved(d_x[k % 4].reshape(N, F, 1), d_er, F)
for k in range(10): step(k)
ctx.synchronize()
ctx.set_kernel_timing(True); ctx.reset_kernel_timing()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 50
with torch.cuda.stream(stream): e0.record(stream)
for k in range(steps): step(10 + k)
with torch.cuda.stream(stream): e1.record(stream)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
kt = {n: ctx.kernel_time(n) for n in ("fdn", "er_push", "render", "other")}
ctx.set_kernel_timing(False)
res = {"config": "BASELINE config 5: 4,096 x (ReverbBus FDN16 + ERBus with 16 taps), mono in -> 7.1.4, 48 kHz / 512 frames", "mode": "throughput" if mode else "parity",
       "ms_per_block": ms, "realtime_instances": N * (F / FS) / (ms * 1e-3),
       "kernel_ms": {k: v[0] / max(v[1], 1) for k, v in kt.items()}}
# CPU reference sample
from oracle import refpy
if refpy.available("sse2"):
    ref = refpy.Ref("sse2"); n = 64; cores = os.cpu_count() or 1
    rb = [ref.reverb(FS) for _ in range(n)]; eb = [ref.erbus(FS, L) for _ in range(n)]
    for i in range(n):
        rb[i].set_rt60(rt[i]); eb[i].set_taps(0.1 + 0.9 * u[i, :, :4], 0.002 + 0.078 * u[i, :, 4], np.arange(1, TAPS + 1), u[i, :, 6:])
    o1 = np.zeros((n, F, L), np.float32); o2 = np.zeros((n, F, L), np.float32)
    for k in range(2):
        ref.reverb_process_batch(rb, x[k % 4, :n], o1, F, threads=cores); ref.erbus_process_batch(eb, x[k % 4, :n], o2, F, threads=cores)
    t0 = time.perf_counter(); reps = 6; tr = te = 0.0
    for k in range(reps):
        tr += ref.reverb_process_batch(rb, np.ascontiguousarray(x[k % 4, :n]), o1, F, threads=cores)
        te += ref.erbus_process_batch(eb, np.ascontiguousarray(x[k % 4, :n]), o2, F, threads=cores)
    res["cpu_reference"] = {"instances": n, "threads": cores, "reverb_ms_per_block": tr / reps * 1e3, "erbus_ms_per_block": te / reps * 1e3,
                            "realtime_instances": n * (F / FS) / ((tr + te) / reps)}
print(json.dumps(res))
