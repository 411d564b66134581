import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import jplspatial_b200 as J
from tests import util
from tests.test_gpu_render import FS, make_scene
from oracle import refpy
ref = refpy.Ref("sse2")
ctx = J.Context(0)
N, C, L, blocks, frames = 256, 1, 12, 24, 512
sc = make_scene(N, C, L, blocks, frames, util.SEED + 401)
sc["delay"][:, : N // 8] = (np.round(sc["delay"][:, : N // 8] * FS) / FS).astype(np.float32)
fx = [ref.dse(FS, C, L, sc["band"][0, i], float(sc["delay"][0, i]), sc["mix"][0, i]) for i in range(N)]
rs = {p: J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0], mode=J.MODE_THROUGHPUT) for p in (0, 1)}
for p in rs: rs[p].SetTimeParallel(p)
rper = np.zeros((N, frames, L), np.float32); g = {p: np.zeros((N, frames, L), np.float32) for p in rs}
for b in range(blocks):
    for i in range(N):
        fx[i].update(sc["band"][b + 1, i], float(sc["delay"][b + 1, i]), sc["mix"][b + 1, i]); rper[i] = 0; fx[i].process(sc["audio"][b, i], rper[i], frames)
    rms = np.sqrt((rper.astype(np.float64) ** 2).mean(axis=(1, 2), keepdims=True))
    tol = 1e-5 * np.maximum(np.maximum(np.abs(rper), rms), 1e-3)
    for p in rs:
        rs[p].UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], sc["mix"][b + 1])
        bus = np.zeros((frames, L), np.float32)
        rs[p].ProcessInterleaved(sc["audio"][b], bus, frames, per_source=g[p])
        ratio = np.abs(g[p].astype(np.float64) - rper) / tol
        per_src = ratio.max(axis=(1, 2))
        worst = np.argsort(-per_src)[:4]
        fr = [int(np.argmax(ratio[i].max(axis=1))) for i in worst]
        print(f"block {b} policy {p}: worst ratio {ratio.max():.3f}; sources {list(map(int, worst))} ratios {[round(float(per_src[i]),2) for i in worst]} frames {fr}; per-tile max {[round(float(ratio[:, t*32:(t+1)*32].max()),2) for t in range(16)]}")
    st0, _, _ = rs[0].state_export(); st1, _, _ = rs[1].state_export()
    d = np.abs(st0 - st1)
    if b % 6 == 5: print("   state diff max per float:", [f"{x:.1e}" for x in d.max(axis=0)[:32]], "integer-delay sources are 0..", N // 8 - 1, "static are 0..", N // 4 - 1)
