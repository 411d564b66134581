"""Scene front-end: k_scene_frontend on device-resident arrays, CUDA events on the context stream, with the reference's
own per-source sequence (oracle/_ref: ref_frontend_timed) timed beside it. Also times front-end -> MDAP chained on the device.
Algorithmic bytes per source (all outputs of the fused tick): in 36 (position) + 8 (cone) + 4 (size) = 48 B; out direct 16 + pan 20 + band 16 + delay 4 + att 4 = 60 B."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
from tests import util
from tests.frontend_inputs import make_positions

M = J.ChannelMask
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
peak = 6538.0
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
res = {"kernel": "k_scene_frontend", "bytes_per_source": 108, "peak_gbs": peak, "runs": []}
fe = J.DirectPathBatch(ctx, functions=[("model", J.MODEL_INVERSE, 1.0, 200.0, 1.0)])
pan = J.VBAPanner(ctx); assert pan.Initialize(M.Surround_7_1_4)
lay = pan.InitializeSourceLayout(M.Mono)
want = ("direct", "pan_params", "band_gains", "delay_seconds", "distance_attenuation")
for n in (16384, 262144, 2097152):
    src = make_positions(n, util.SEED + 1, 150.0); lis = make_positions(1, util.SEED + 2, 1.0)
    cones = np.sort((util.uniform01(2 * n, util.SEED + 3) * np.float32(6.28)).reshape(n, 2), axis=1).astype(np.float32)
    sizes = (util.uniform01(n, util.SEED + 4) * np.float32(20.0)).astype(np.float32)
    with torch.cuda.stream(stream):
        t = {k: torch.from_numpy(v).to(dev) for k, v in dict(sources=src, listeners=lis, cones=cones, object_sizes=sizes).items()}
        out = {k: torch.empty((n, w), dtype=torch.float32, device=dev) for k, w in (("direct", 4), ("pan_params", 5), ("band_gains", 4), ("delay_seconds", 1), ("distance_attenuation", 1))}
        mix = torch.empty((n, 1, 12), dtype=torch.float32, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    def tick(chain):
        fe.Process(n, want, focus=0.75, spread=0.0, out=out, **t)
        if chain:
            pan.ProcessVBAPData(lay, out["pan_params"], out=mix)
    run = {"sources": n}
    for chain in (False, True):
        with torch.cuda.stream(stream):
            for _ in range(5): tick(chain)
        ctx.synchronize()
        times = []
        for _ in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                sink = flush.sum()                 # L2 flush between timed iterations: READ 256 MB (> 126 MB L2) so no dirty lines are written back under the timed kernel
                e0.record(stream); tick(chain); e1.record(stream)
            torch.cuda.synchronize(); times.append(e0.elapsed_time(e1))
        ms = float(np.median(times))
        key = "frontend_plus_mdap_ms" if chain else "frontend_ms"
        run[key] = ms
        if not chain:
            run["achieved_gbs"] = n * 108 / (ms * 1e-3) / 1e9
            run["frac_of_hbm_peak"] = run["achieved_gbs"] / peak
            run["sources_per_second"] = n / (ms * 1e-3)
    res["runs"].append(run)
    del t, out, mix, flush
# the reference's CPU code on a bounded sample
from oracle import refpy
if refpy.available("sse2"):
    ref = refpy.Ref("sse2"); n = 262144; cores = os.cpu_count() or 1
    src = make_positions(n, util.SEED + 1, 150.0); lis = make_positions(1, util.SEED + 2, 1.0)
    cones = np.sort((util.uniform01(2 * n, util.SEED + 3) * np.float32(6.28)).reshape(n, 2), axis=1).astype(np.float32)
    sizes = (util.uniform01(n, util.SEED + 4) * np.float32(20.0)).astype(np.float32)
    model = np.array([1, 1.0, 200.0, 1.0], np.float32)
    for th in (1, cores):
        best = min(ref.frontend_timed(src, lis, cones, sizes, model, 0.75, threads=th)[0] for _ in range(5))
        res[f"cpu_reference_{th}_threads"] = {"sources": n, "ms": best * 1e3, "sources_per_second": n / best}
print(json.dumps(res))
