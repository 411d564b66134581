"""GPU parity tests at the shapes bench.py times:

* BASELINE config 3 at its FULL size — 16,384 mono sources -> 7.1.4, 128 full CTAs, the vectorised ring push at full grid — two blocks through the
  production (non-debug) kernel and the fused panning update: per-source audio bit-exact against the unmodified reference on a strided sample of
  512 sources, the bus of the production kernel bit-equal to the debug kernel's and within the fp32 summation bound of an fp64 sum of ALL sources.
* BASELINE config 4's shape — mono / stereo / quad / 5.1 sources -> 9.1.6 (the 5.1 sources rendered with six channels, the LFE row unpanned), MDAP ->
  fused update -> render per layout, all layouts adding into one bus: per-source bit-exact, bus within the bound.
"""
import numpy as np
import pytest
import torch

import jplspatial_b200 as J
from tests import util

pytestmark = pytest.mark.gpu
FS, F = 48000.0, 512
M = J.ChannelMask


def _scene(n, blocks, seed):
    params = np.stack([util.pan_params(n, seed + 10 * b, edge=False) for b in range(blocks + 1)])
    band = (0.1 + 0.9 * util.uniform01((blocks + 1) * n * 4, seed + 1)).reshape(blocks + 1, n, 4).astype(np.float32)
    dist0 = 1.0 + 299.0 * util.uniform01(n, seed + 2)
    speed = (util.uniform01(n, seed + 3) * 2.0 - 1.0) * 30.0
    speed[: n // 4] = 0.0
    t = (np.arange(blocks + 1, dtype=np.float32) * np.float32(F / FS))[:, None]
    delay = (np.maximum(dist0[None] + speed[None] * t, 0.5) / np.float32(343.2)).astype(np.float32)
    return params, band, delay


def _rows_of(mask):
    """renderer row (channel index in ascending EChannel order) of every panned channel group: LFE and top channels are not panned"""
    rows, idx = [], 0
    for b in range(22):
        if (int(mask) >> b) & 1:
            if b != 3 and b < 13:
                rows.append(idx)
            idx += 1
    return rows


def _bus_bound(per):
    return 4.0 * np.finfo(np.float32).eps * np.abs(per.astype(np.float64)).sum(0) + 1e-12


def test_config3_full_size_two_blocks(gpu_ctx, ref):
    N, L, blocks = 16384, 12, 2
    dev = torch.device("cuda", gpu_ctx.device)
    stream = gpu_ctx.torch_stream()
    params, band, delay = _scene(N, blocks, util.SEED + 300)
    audio = util.white_noise((blocks, N, F, 1), util.SEED + 301)
    pan = J.VBAPanner(gpu_ctx)
    assert pan.Initialize(M.Surround_7_1_4)
    lay = pan.InitializeSourceLayout(M.Mono)
    mix0 = pan.ProcessVBAPData(lay, params[0]).reshape(N, L)

    def run(debug):
        r = J.DirectSoundBatch(gpu_ctx, FS, N, 1, L, band[0], delay[0], mix0, ring_size=0, max_block_frames=F)
        with torch.cuda.stream(stream):
            d_par = torch.from_numpy(params).to(dev); d_band = torch.from_numpy(band).to(dev); d_del = torch.from_numpy(delay).to(dev)
            d_audio = torch.from_numpy(audio).to(dev)
            bus = torch.zeros((blocks, F, L), dtype=torch.float32, device=dev)
            per = torch.zeros((blocks, N, F, L), dtype=torch.float32, device=dev) if debug else None
        gpu_ctx.synchronize()
        for b in range(blocks):
            r.UpdatePanned(pan, lay, d_par[b + 1], d_band[b + 1], d_del[b + 1])            # the fused path bench.py uses
            r.ProcessInterleaved(d_audio[b], bus[b], F, per_source=per[b] if debug else None)
        gpu_ctx.synchronize()
        out = bus.cpu().numpy(), (per.cpu().numpy() if debug else None)
        r.close()
        return out

    bus_prod, _ = run(False)
    bus_dbg, per = run(True)
    assert util.bits_equal(bus_prod, bus_dbg), "production kernel's bus differs from the debug kernel's"
    # every source contributes: bus against the fp64 sum of all 16,384 per-source outputs
    for b in range(blocks):
        exact = per[b].astype(np.float64).sum(0)
        assert np.all(np.abs(bus_prod[b] - exact) <= _bus_bound(per[b])), f"block {b}: bus outside the fp32 summation bound"
        assert np.abs(exact).max() > 1.0
    # strided sample of 512 sources against the unmodified reference (MDAP gains and rendered audio, bit for bit)
    sample = np.arange(0, N, N // 512)
    rp = ref.panner(M.Surround_7_1_4)
    rl = rp.layout(M.Mono)
    rmix = np.stack([rp.mdap(rl, params[b][sample]).reshape(len(sample), L) for b in range(blocks + 1)])
    assert util.bits_equal(rmix[0], mix0[sample])
    fx = [ref.dse(FS, 1, L, band[0, s], float(delay[0, s]), rmix[0, i]) for i, s in enumerate(sample)]
    for b in range(blocks):
        for i, s in enumerate(sample):
            fx[i].update(band[b + 1, s], float(delay[b + 1, s]), rmix[b + 1, i])
            out = np.zeros((F, L), np.float32)
            fx[i].process(audio[b, s], out, F)
            assert util.bits_equal(out, per[b, s]), f"block {b} source {s}: rendered audio differs from the reference, max {np.abs(out - per[b, s]).max()}"


@pytest.mark.parametrize("n", [512])
def test_config4_shape_mixed_layouts_to_9_1_6(gpu_ctx, ref, n):
    L, blocks = 16, 2
    dev = torch.device("cuda", gpu_ctx.device)
    stream = gpu_ctx.torch_stream()
    pan = J.VBAPanner(gpu_ctx)
    assert pan.Initialize(M.Surround_9_1_6)
    rp = ref.panner(M.Surround_9_1_6)
    with torch.cuda.stream(stream):
        bus = torch.zeros((blocks, F, L), dtype=torch.float32, device=dev)
    all_per = [[] for _ in range(blocks)]
    for li, (name, C) in enumerate((("Mono", 1), ("Stereo", 2), ("Quad", 4), ("Surround_5_1", 6))):
        mask = getattr(M, name)
        lay = pan.InitializeSourceLayout(mask)
        rl = rp.layout(mask)
        rows = _rows_of(mask)
        assert len(rows) == lay.num_channels == rl.num_channels and (C == 6) == (len(rows) != C)
        params, band, delay = _scene(n, blocks, util.SEED + 400 + li)
        audio = util.white_noise((blocks, n, F, C), util.SEED + 450 + li)

        def full_map(panned):           # [n][panned channels][L] -> [n][C][L], unpanned rows (LFE) zero
            m = np.zeros((n, C, L), np.float32)
            m[:, rows, :] = panned.reshape(n, len(rows), L)
            return m
        gmix0 = pan.ProcessVBAPData(lay, params[0])
        rmix = [rp.mdap(rl, params[b]) for b in range(blocks + 1)]
        assert util.bits_equal(gmix0, rmix[0].reshape(gmix0.shape)), f"{name}: MDAP gains differ from the reference"
        r = J.DirectSoundBatch(gpu_ctx, FS, n, C, L, band[0], delay[0], full_map(gmix0).reshape(n, C * L), ring_size=0, max_block_frames=F)
        with torch.cuda.stream(stream):
            d_par = torch.from_numpy(params).to(dev); d_band = torch.from_numpy(band).to(dev); d_del = torch.from_numpy(delay).to(dev)
            d_audio = torch.from_numpy(audio).to(dev)
            per = torch.zeros((blocks, n, F, L), dtype=torch.float32, device=dev)
        gpu_ctx.synchronize()
        for b in range(blocks):
            r.UpdatePanned(pan, lay, d_par[b + 1], d_band[b + 1], d_del[b + 1])            # panned rows scattered, LFE row zero
            r.ProcessInterleaved(d_audio[b], bus[b], F, per_source=per[b])                 # every layout adds into the same bus
        gpu_ctx.synchronize()
        per_h = per.cpu().numpy()
        r.close()
        fx = [ref.dse(FS, C, L, band[0, i], float(delay[0, i]), full_map(rmix[0])[i].reshape(-1)) for i in range(n)]
        for b in range(blocks):
            fm = full_map(rmix[b + 1])
            for i in range(n):
                fx[i].update(band[b + 1, i], float(delay[b + 1, i]), fm[i].reshape(-1))
                out = np.zeros((F, L), np.float32)
                fx[i].process(audio[b, i], out, F)
                assert util.bits_equal(out, per_h[b, i]), f"{name} block {b} source {i}: differs from the reference, max {np.abs(out - per_h[b, i]).max()}"
            all_per[b].append(per_h[b])
    bus_h = bus.cpu().numpy()
    for b in range(blocks):
        per = np.concatenate(all_per[b], 0)
        exact = per.astype(np.float64).sum(0)
        # four accumulating launches: the bound of one re-associated sum per launch plus the three float additions between them
        assert np.all(np.abs(bus_h[b] - exact) <= 2.0 * _bus_bound(per)), f"block {b}: bus outside the fp32 summation bound"
        assert np.abs(exact).max() > 1.0
