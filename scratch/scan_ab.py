"""A/B of the time-parallel throughput path (jplg_scan_kernels.cu) against the sequential k_render, device-resident blocks:
python scratch/scan_ab.py -> one line per (sources, variant): ms per block (ProcessInterleaved only: push + render + bus reduce)"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import jplspatial_b200 as J
import bench
L, F = 12, 512
ctx = J.Context(0); stream = ctx.torch_stream(); dev = torch.device("cuda", 0)
out = []
for N in [int(a) for a in sys.argv[1:]] or [256, 1024, 2048, 4096, 8192, 16384]:
    audio, params, band, delay = bench.make_inputs(N, 2, 0xB200)
    pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
    lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
    mix = [pan.ProcessVBAPData(lay, params[k]).reshape(N, L) for k in range(2)]
    with torch.cuda.stream(stream):
        d_audio = torch.from_numpy(audio).to(dev)
        bus = torch.zeros((F, L), dtype=torch.float32, device=dev)
    row = {"sources": N}
    for name, mode, policy in [("parity", J.MODE_PARITY, 0), ("throughput_sequential", J.MODE_THROUGHPUT, 0), ("throughput_scan", J.MODE_THROUGHPUT, 1)]:
        rend = J.DirectSoundBatch(ctx, 48000.0, N, 1, L, band[0], delay[0], mix[0], ring_size=65536, mode=mode)
        rend.SetTimeParallel(policy)
        steps, warm = 60, 10
        with torch.cuda.stream(stream):
            for k in range(warm):
                rend.UpdateParameters(band[k % 2], delay[k % 2], mix[k % 2]); rend.ProcessInterleaved(d_audio[k % 2], bus, F)
            ctx.synchronize()
            tot = 0.0
            for k in range(steps):
                rend.UpdateParameters(band[k % 2], delay[k % 2], mix[k % 2])
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record(stream); rend.ProcessInterleaved(d_audio[k % 2], bus, F); e1.record(stream)
                ctx.synchronize(); tot += e0.elapsed_time(e1)
        row[name] = round(tot / steps, 5)
        row[name + "_scan_blocks"] = rend.time_parallel_blocks
        del rend
    row["scan_speedup_vs_sequential"] = round(row["throughput_sequential"] / row["throughput_scan"], 2)
    row["scan_speedup_vs_parity"] = round(row["parity"] / row["throughput_scan"], 2)
    print(json.dumps(row), flush=True); out.append(row)
