"""GPU parity tests for stages 2-4 (delay/Doppler read, LR4 four-band filter, gain-ramped mix) through the C-ABI
against the unmodified reference's DirectSoundEffect (oracle/_ref, SSE2 build).

Bars: per-source rendered audio BIT-EXACT in parity mode; <= 1e-5 relative per sample in throughput mode
(tolerance |d| <= 1e-5 * max(|ref|, rms_block(ref), 1e-3)); the N-source bus (re-associated sum) against the
reference's sequential float sum and an fp64 sum with a tolerance scaled by the sum of |terms|."""
import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util

pytestmark = pytest.mark.gpu
FS = 48000.0


def make_scene(N, C, L, blocks, frames, seed, integer_delays=False):
    rng_u = lambda n, k: util.uniform01(n, seed + k)
    sc = {}
    sc["audio"] = util.white_noise((blocks, N, frames, C), seed + 100)
    dist0 = 1.0 + 299.0 * rng_u(N, 1)                      # metres
    speed = (rng_u(N, 2) * 2.0 - 1.0) * 30.0               # m/s radial
    speed[: N // 4] = 0.0                                   # static sources
    t = (np.arange(blocks + 1, dtype=np.float32) * np.float32(frames / FS))[:, None]
    dist = np.maximum(dist0[None, :] + speed[None, :] * t, 0.5).astype(np.float32)
    delay = (dist / np.float32(343.2)).astype(np.float32)  # seconds (inc/SpeedOfSound.h:32)
    if integer_delays:
        delay = (np.round(delay * FS) / FS).astype(np.float32)
    sc["delay"] = delay
    sc["band"] = (0.1 + 0.9 * rng_u((blocks + 1) * N * 4, 3)).reshape(blocks + 1, N, 4).astype(np.float32)
    mix = rng_u((blocks + 1) * N * C * L, 4).reshape(blocks + 1, N, C * L).astype(np.float32)
    mix[:, :, 3::7] = 0.0                                   # some silent speakers (LFE-like)
    mix[:, N // 2:, :] = mix[0:1, N // 2:, :]               # half the sources keep constant pan
    sc["mix"] = mix
    return sc


def run_reference(ref, sc, N, C, L, blocks, frames):
    fx = [ref.dse(FS, C, L, sc["band"][0, i], float(sc["delay"][0, i]), sc["mix"][0, i]) for i in range(N)]
    per = np.zeros((blocks, N, frames, L), np.float32)
    bus = np.zeros((blocks, frames, L), np.float32)
    for b in range(blocks):
        for i in range(N):
            fx[i].update(sc["band"][b + 1, i], float(sc["delay"][b + 1, i]), sc["mix"][b + 1, i])
            out = np.zeros((frames, L), np.float32)
            fx[i].process(sc["audio"][b, i], out, frames)
            per[b, i] = out
            bus[b] += out  # the host loop of the reference: every effect adds into the same bus, in source order
    return per, bus


def run_gpu(ctx, sc, N, C, L, blocks, frames, mode, debug=True, ring_size=0):
    r = J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0], ring_size=ring_size,
                           max_block_frames=max(frames, 512), mode=mode)
    per = np.zeros((blocks, N, frames, L), np.float32)
    bus = np.zeros((blocks, frames, L), np.float32)
    for b in range(blocks):
        r.UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], sc["mix"][b + 1])
        r.ProcessInterleaved(sc["audio"][b], bus[b], frames, per_source=per[b] if debug else None)
    return per, bus, r


@pytest.mark.parametrize("N,C,L,blocks,frames", [
    (96, 1, 12, 6, 512),     # mono -> 7.1.4, BASELINE config 3 shape (small N)
    (40, 2, 12, 4, 512),     # stereo sources
    (33, 1, 16, 3, 512),     # -> 9.1.6, ragged CTA
    (20, 6, 16, 3, 512),     # 5.1-style 6-channel sources (C does not divide the CTA width)
    (17, 1, 4, 3, 100),      # short blocks (single chunk, ragged tiles)
    (9, 1, 12, 2, 1000),     # long blocks: 480 + 480 + 40
])
def test_per_source_bit_exact(gpu_ctx, ref, N, C, L, blocks, frames):
    sc = make_scene(N, C, L, blocks, frames, util.SEED + N)
    rper, rbus = run_reference(ref, sc, N, C, L, blocks, frames)
    gper, gbus, _ = run_gpu(gpu_ctx, sc, N, C, L, blocks, frames, J.MODE_PARITY)
    nbad = util.count_bit_mismatch(gper, rper)
    assert nbad == 0, f"{nbad} samples differ, max abs {np.abs(gper - rper).max()}"
    # bus: same terms, different summation order
    exact = rper.astype(np.float64).sum(1)
    tol = 4.0 * np.finfo(np.float32).eps * np.abs(rper.astype(np.float64)).sum(1) + 1e-12
    assert np.all(np.abs(gbus - exact) <= tol)
    assert np.all(np.abs(gbus - rbus) <= 2 * tol)


def test_integer_delays_bit_exact(gpu_ctx, ref):
    """Static sources at exactly-integer delays ring at Nyquist without decay in the reference (Thiran a = 1,
    SURVEY §7.3-4): only a bit-exact recurrence reproduces that."""
    N, C, L, blocks, frames = 32, 1, 12, 5, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 77, integer_delays=True)
    rper, _ = run_reference(ref, sc, N, C, L, blocks, frames)
    gper, _, _ = run_gpu(gpu_ctx, sc, N, C, L, blocks, frames, J.MODE_PARITY)
    assert util.count_bit_mismatch(gper, rper) == 0


def test_throughput_mode_tolerance(gpu_ctx, ref):
    N, C, L, blocks, frames = 64, 1, 12, 8, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 5)
    rper, rbus = run_reference(ref, sc, N, C, L, blocks, frames)
    gper, gbus, _ = run_gpu(gpu_ctx, sc, N, C, L, blocks, frames, J.MODE_THROUGHPUT)
    rms = np.sqrt((rper.astype(np.float64) ** 2).mean(axis=(2, 3), keepdims=True))
    tol = 1e-5 * np.maximum(np.maximum(np.abs(rper), rms), 1e-3)
    err = np.abs(gper.astype(np.float64) - rper)
    assert np.all(err <= tol), f"worst ratio {np.max(err / tol)}"
    resid_db = 10 * np.log10((err ** 2).mean() + 1e-300)
    assert resid_db < -100.0, resid_db


def test_bus_without_debug_matches_debug(gpu_ctx):
    N, C, L, blocks, frames = 300, 1, 12, 3, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 9)
    _, bus_a, _ = run_gpu(gpu_ctx, sc, N, C, L, blocks, frames, J.MODE_PARITY, debug=True)
    _, bus_b, _ = run_gpu(gpu_ctx, sc, N, C, L, blocks, frames, J.MODE_PARITY, debug=False)
    assert util.bits_equal(bus_a, bus_b)   # deterministic reduction tree, identical in both kernels
    _, bus_c, _ = run_gpu(gpu_ctx, sc, N, C, L, blocks, frames, J.MODE_PARITY, debug=False)
    assert util.bits_equal(bus_b, bus_c)   # run-to-run reproducible


def test_bus_accumulates_into_output(gpu_ctx):
    N, C, L, frames = 8, 1, 12, 512
    sc = make_scene(N, C, L, 1, frames, util.SEED + 13)
    r = J.DirectSoundBatch(gpu_ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0])
    bus0 = np.zeros((frames, L), np.float32)
    r.ProcessInterleaved(sc["audio"][0], bus0, frames)
    r2 = J.DirectSoundBatch(gpu_ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0])
    bus1 = np.full((frames, L), 0.25, np.float32)
    r2.ProcessInterleaved(sc["audio"][0], bus1, frames)
    assert util.bits_equal(bus1, (np.float32(0.25) + bus0).astype(np.float32))


def test_state_export_import_roundtrip(gpu_ctx):
    N, C, L, frames = 24, 2, 12, 512
    sc = make_scene(N, C, L, 4, frames, util.SEED + 17)
    _, busA, rA = run_gpu(gpu_ctx, sc, N, C, L, 2, frames, J.MODE_PARITY, debug=False, ring_size=4096)
    st, rings, pos = rA.state_export(with_rings=True)
    assert pos == 2 * 544 and st.shape  # 512-frame block = 512 + 32 ring pushes (reference chunk quirk) == (N * C, J.STATE_FLOATS)
    rB = J.DirectSoundBatch(gpu_ctx, FS, N, C, L, sc["band"][2], sc["delay"][2], sc["mix"][2], ring_size=4096)
    rB.state_import(st, rings, pos)
    outA = np.zeros((frames, L), np.float32)
    outB = np.zeros((frames, L), np.float32)
    for r, o in ((rA, outA), (rB, outB)):
        r.UpdateParameters(sc["band"][3], sc["delay"][3], sc["mix"][3])
        r.ProcessInterleaved(sc["audio"][2], o, frames)
    assert util.bits_equal(outA, outB)


def test_crossover_bank_vs_reference(gpu_ctx, ref):
    n, frames = 37, 700
    x = util.white_noise((n, frames), util.SEED + 41)
    g0 = util.uniform01(n * 4, util.SEED + 42).reshape(n, 4)
    g1 = util.uniform01(n * 4, util.SEED + 43).reshape(n, 4)
    bank = J.CrossoverBank(gpu_ctx, n, FS)
    y1 = bank.ProcessBlock(x, g0, g1)
    y2 = bank.ProcessBlock(x, g1, g0)  # state carried across calls
    for i in range(n):
        rc = ref.crossover(FS)
        assert util.bits_equal(y1[i], rc.process_ramp(x[i], g0[i], g1[i]))
        assert util.bits_equal(y2[i], rc.process_ramp(x[i], g1[i], g0[i]))
    # CrossoverFilterTest.cpp:74-123: bands of an impulse sum back with preserved energy; split == fused path
    bank.Reset()
    imp = np.zeros((n, 4096), np.float32); imp[:, 0] = 1.0
    bands = bank.SplitBlock(imp)
    rc = ref.crossover(FS)
    assert util.bits_equal(bands[0], rc.split(imp[0]))
    summed = bands.sum(-1)
    assert abs(np.sqrt((summed[0].astype(np.float64) ** 2).sum()) - 1.0) < 1e-4


def test_delay_bank_vs_reference(gpu_ctx, ref):
    """DelayLineTest.cpp:58-215 known answers + Thiran/linear taps against the reference DelayLine."""
    n = 5
    bank = J.DelayBank(gpu_ctx, n, 1000)
    assert bank.GetSize() == 1024
    x = util.white_noise((n, 300), util.SEED + 51)
    bank.Push(x)
    rd = [ref.delay(1000) for _ in range(n)]
    for i in range(n):
        rd[i].push(x[i])
    # zero delay = last written sample; fixed integer delays
    d = np.tile(np.array([0, 1, 7, 299], np.float32), (n, 1))
    out = bank.Read(d, J.DelayBank.NEAREST)
    assert np.array_equal(out[:, 0], x[:, -1]) and np.array_equal(out[:, 1], x[:, -2]) and np.array_equal(out[:, 3], x[:, 0])
    # linear tap 0.25 between two samples (DelayLineTest linear tap case)
    lin = bank.Read(np.full((n, 1), 3.25, np.float32), J.DelayBank.LINEAR)
    for i in range(n):
        assert np.float32(lin[i, 0]) == np.float32(rd[i].tap_linear(3.25))
    delays = (util.uniform01(n * 64, util.SEED + 52).reshape(n, 64) * 250.0).astype(np.float32)
    th = bank.Read(delays, J.DelayBank.THIRAN1)
    for i in range(n):
        r = np.array([rd[i].tap_thiran(float(v)) for v in delays[i]], np.float32)
        assert util.bits_equal(th[i], r)
    # DelayTapXFade (DelayLine.h:73-122): cosine crossfade between the previous and the new integer delay, state carried across taps and calls.
    # The gains go through libm's cosf in the reference: 1e-6 absolute on samples in [-1, 1), bit-exact where alpha is 0 or 1.
    xd = (util.uniform01(n * 32, util.SEED + 54).reshape(n, 32) * 250.0).astype(np.float32)
    xa = util.uniform01(n * 32, util.SEED + 55).reshape(n, 32).astype(np.float32)
    xa[:, :4] = [0.0, 1.0, 0.5, 0.25]
    for rep in range(2):
        xf = bank.ReadXFade(xd, xa)
        for i in range(n):
            r = np.array([rd[i].tap_xfade(float(v), float(a)) for v, a in zip(xd[i], xa[i])], np.float32)
            assert np.abs(xf[i] - r).max() <= 1e-6
            assert util.bits_equal(xf[i, :2], r[:2])
    # wrap: push past the ring size and read across the seam
    big = util.white_noise((n, 1500), util.SEED + 53)
    bank.Push(big[:, :1000]); bank.Push(big[:, 1000:])
    for i in range(n):
        rd[i].push(big[i])
    d2 = np.tile(np.arange(0, 1000, 37, dtype=np.float32), (n, 1))
    o2 = bank.Read(d2, J.DelayBank.NEAREST)
    for i in range(n):
        assert np.array_equal(o2[i], np.array([rd[i].read_int(int(v)) for v in d2[i]], np.float32))


def test_channel_mixing_vs_reference(gpu_ctx, ref):
    n, frames, nd, ns = 6, 333, 12, 2
    dest = util.white_noise((n, frames, nd), util.SEED + 61)
    src = util.white_noise((n, frames, ns), util.SEED + 62)
    g0 = util.uniform01(n, util.SEED + 63); g1 = util.uniform01(n, util.SEED + 64)
    out = J.add_and_apply_gain_ramp(gpu_ctx, dest, src, 5, 1, frames, g0, g1)
    for i in range(n):
        r = ref.add_gain_ramp(dest[i], src[i], 5, 1, nd, ns, frames, float(g0[i]), float(g1[i]))
        assert util.bits_equal(out[i], r)
    flat = util.white_noise((n, frames), util.SEED + 65)
    assert util.bits_equal(J.apply_gain(gpu_ctx, flat, g0), (flat * g0[:, None]).astype(np.float32))
    # Add, both overloads, and DownmixToMono (ChannelMixing.h:32-48, 129-136)
    added = J.add(gpu_ctx, dest, src, 7, 1)
    added_mono = J.add(gpu_ctx, dest, flat, 3)
    down = J.downmix_to_mono(gpu_ctx, flat, dest)
    for i in range(n):
        assert util.bits_equal(added[i], ref.mix_add(dest[i], src[i], 7, 1, nd, ns, frames))
        assert util.bits_equal(added_mono[i], ref.mix_add(dest[i], flat[i], 3, 0, nd, 0, frames))
        assert util.bits_equal(down[i], ref.downmix_to_mono(flat[i], dest[i], nd))


def test_empty_and_error_paths(gpu_ctx):
    p = J.VBAPanner(gpu_ctx)
    assert not p.Initialize(J.ChannelMask.Mono)          # single target channel (VBAPanning2D.h:133-134)
    assert not p.Initialize(0)                            # invalid map
    assert p.Initialize(J.ChannelMask.Quad)
    assert p.GetSpeakerGains(np.zeros((0, 3), np.float32)).shape == (0, 4)
    with pytest.raises(J.JplgError):
        p.InitializeSourceLayout(J.ChannelMask.Surround_5_0_2)   # source map with top channels (PannerBase.h:196-199)
    r = J.DirectSoundBatch(gpu_ctx, FS, 4, 1, 4)
    bus = np.zeros((0, 4), np.float32)
    r.ProcessInterleaved(np.zeros((4, 0, 1), np.float32), bus, 0)  # zero frames is a no-op
    with pytest.raises(J.JplgError):
        r.ProcessInterleaved(np.zeros((4, 4096, 1), np.float32), np.zeros((4096, 4), np.float32), 4096)  # > max_block_frames


@pytest.mark.parametrize("C,pipeline", [(1, 1), (2, 1), (1, 2), (1, 0)])
def test_pipelined_host_path_equals_device_path(C, pipeline):
    """JPLG_HOST_PIPELINE=1/2: JPLG_MEM_HOST blocks with >= 32 CTAs are rendered in slices on auxiliary streams (transfer of slice k+1 overlaps the
    render of slice k; in mode 2 pinned input is read in place by k_push). The result must equal the single-launch device path bit for bit
    (same CTA partials, same reduction tree), for pinned (torch) and pageable (numpy) host buffers."""
    import os
    import torch
    os.environ["JPLG_HOST_PIPELINE"] = str(pipeline)
    try:
        ctx = J.Context(0)
    finally:
        del os.environ["JPLG_HOST_PIPELINE"]
    N, L, blocks, frames = 4096 // C + 37, 12, 3, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 77)
    dev = torch.device("cuda", 0)

    def make():
        return J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0], ring_size=4096, max_block_frames=512, mode=J.MODE_PARITY)

    ra, rb, rc = make(), make(), make()
    stream = ctx.torch_stream()
    for b in range(blocks):
        for r in (ra, rb, rc):
            r.UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], sc["mix"][b + 1])
        with torch.cuda.stream(stream):
            d_in = torch.from_numpy(sc["audio"][b]).to(dev)
            d_bus = torch.zeros((frames, L), dtype=torch.float32, device=dev)
            ra.ProcessInterleaved(d_in, d_bus, frames)
        ctx.synchronize()
        want = d_bus.cpu().numpy()
        pinned = torch.from_numpy(sc["audio"][b].copy()).pin_memory()
        bus_p = torch.zeros((frames, L), dtype=torch.float32).pin_memory()
        rb.ProcessInterleaved(pinned, bus_p, frames)
        bus_n = np.zeros((frames, L), np.float32)
        rc.ProcessInterleaved(sc["audio"][b], bus_n, frames)
        assert np.abs(want).max() > 0
        assert util.bits_equal(bus_p.numpy(), want), f"block {b}: pinned host path differs"
        assert util.bits_equal(bus_n, want), f"block {b}: pageable host path differs"
    for r in (ra, rb, rc):
        r.close()


def test_async_host_path_equals_sync_host_path():
    """JPLG_MEM_HOST_ASYNC: MDAP -> update -> process_block enqueued without host waits, the input transfer of block k+1 on the copy stream
    while block k renders (two staging slots, at most two blocks in flight). Every block's bus must equal the synchronous JPLG_MEM_HOST path
    bit for bit; submitting a third block before waiting must not corrupt the first."""
    import torch
    ctx = J.Context(0)
    N, C, L, blocks, frames = 4133, 1, 12, 7, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 78)
    pan = J.VBAPanner(ctx); assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
    lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
    params = [util.pan_params(N, util.SEED + 500 + b) for b in range(blocks)]

    def make():
        return J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0], ring_size=4096, max_block_frames=512, mode=J.MODE_PARITY)

    def pin(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()

    rs, ra, rp = make(), make(), make()
    want = []
    base_bus = (util.uniform01(frames * L, util.SEED + 501).reshape(frames, L) - 0.5).astype(np.float32)   # the bus is accumulated into, not overwritten
    for b in range(blocks):
        mix = pan.ProcessVBAPData(lay, params[b])
        rs.UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], mix)
        bus = base_bus.copy()
        rs.ProcessInterleaved(sc["audio"][b], bus, frames)
        want.append(bus)
    # page-locked buffers: the small arrays and the bus are read / written by the kernels through their device alias, only the audio is copied
    keep = [pin(x) for b in range(blocks) for x in (params[b], sc["band"][b + 1], sc["delay"][b + 1], sc["audio"][b])]
    h = [k[1] for k in keep]
    kmix, hmix = pin(np.zeros((N, 1, L), np.float32))
    buses = [pin(base_bus.copy()) for _ in range(blocks)]
    for b in range(blocks):
        p, band, delay, audio = h[4 * b:4 * b + 4]
        pan.ProcessVBAPData(lay, p, out=hmix, async_host=True)
        ra.UpdateParameters(band, delay, hmix, async_host=True)
        ra.ProcessInterleaved(audio, buses[b][1], frames, async_host=True)
        if b in (1, 2, 5):          # sometimes two blocks in flight, sometimes the library has to wait itself (third submission)
            ra.WaitBlock()
    for _ in range(blocks):
        ra.WaitBlock()              # waiting with nothing in flight is a no-op
    ctx.synchronize()
    for b in range(blocks):
        assert np.abs(want[b]).max() > 0
        assert util.bits_equal(buses[b][1], want[b]), f"block {b}: asynchronous host path differs from the synchronous one"
    # pageable buffers: same call, staged copies instead of the alias
    pmix = np.zeros((N, 1, L), np.float32)
    pbuses = [base_bus.copy() for _ in range(blocks)]
    for b in range(blocks):
        pan.ProcessVBAPData(lay, params[b], out=pmix, async_host=True)
        rp.UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], pmix, async_host=True)
        rp.ProcessInterleaved(sc["audio"][b], pbuses[b], frames, async_host=True)
        rp.WaitBlock()
    ctx.synchronize()
    for b in range(blocks):
        assert util.bits_equal(pbuses[b], want[b]), f"block {b}: asynchronous path with pageable buffers differs"
    rs.close(); ra.close(); rp.close()
    del pan, lay
    ctx.close()


def test_push_ahead_equals_plain_path():
    """jplg_renderer_push_ahead: block k+1's ring pushes on a side stream while block k renders. Same buses, bit for bit, as pushing inside
    process_block; a process_block with a different block overwrites the stale pushes."""
    import torch
    ctx = J.Context(0)
    N, C, L, blocks, frames = 2085, 1, 12, 6, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 79)
    dev = torch.device("cuda", 0)
    stream = ctx.torch_stream()

    def make():
        return J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0], ring_size=0, max_block_frames=512, mode=J.MODE_PARITY)   # reference ring (131,072): holds the longest delay of the scene (42 k samples) + blocks ahead

    ra, rb = make(), make()
    with torch.cuda.stream(stream):
        d_audio = [torch.from_numpy(sc["audio"][b]).to(dev) for b in range(blocks)]
        bus_a = [torch.zeros((frames, L), dtype=torch.float32, device=dev) for _ in range(blocks)]
        bus_b = [torch.zeros((frames, L), dtype=torch.float32, device=dev) for _ in range(blocks)]
    ctx.synchronize()
    for b in range(blocks):
        ra.UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], sc["mix"][b + 1])
        ra.ProcessInterleaved(d_audio[b], bus_a[b], frames)
    for b in range(blocks):
        rb.UpdateParameters(sc["band"][b + 1], sc["delay"][b + 1], sc["mix"][b + 1])
        rb.ProcessInterleaved(d_audio[b], bus_b[b], frames)          # block 0 pushes itself, the others were pushed ahead
        if b + 1 < blocks:
            rb.PushAhead(d_audio[b + 1], frames)
            if b == 2:
                with pytest.raises(J.JplgError):
                    rb.PushAhead(d_audio[b + 1], frames)            # only one block ahead
    ctx.synchronize()
    for b in range(blocks):
        assert np.abs(bus_a[b].cpu().numpy()).max() > 0
        assert util.bits_equal(bus_b[b].cpu().numpy(), bus_a[b].cpu().numpy()), f"block {b}: push-ahead path differs"
    # the guard: delays that only exist on the device need a declared bound; a ring that cannot hold delay + two blocks of pushes + a chunk is refused
    rc = make()
    with torch.cuda.stream(stream):
        d_delay = torch.from_numpy(sc["delay"][1]).to(dev)
    rc.UpdateParameters(None, d_delay, None)
    with pytest.raises(J.JplgError):
        rc.PushAhead(d_audio[0], frames)
    rc.SetMaxDelaySeconds(float(sc["delay"].max()))
    rc.PushAhead(d_audio[0], frames)
    rc.close()
    rs = J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], sc["mix"][0], ring_size=32768, max_block_frames=512)   # 0.87 s of delay = 42 k samples > the ring
    with pytest.raises(J.JplgError):
        rs.PushAhead(d_audio[0], frames)
    rs.close()
    # a different block than the one pushed ahead: same result as without the stale pushes
    with torch.cuda.stream(stream):
        xa = torch.zeros((frames, L), dtype=torch.float32, device=dev); xb = torch.zeros((frames, L), dtype=torch.float32, device=dev)
    rb.PushAhead(d_audio[0], frames)
    rb.ProcessInterleaved(d_audio[1], xb, frames)
    ra.ProcessInterleaved(d_audio[1], xa, frames)
    ctx.synchronize()
    assert util.bits_equal(xa.cpu().numpy(), xb.cpu().numpy())
    ra.close(); rb.close()
    ctx.close()


@pytest.mark.gpu
def test_background_context_panning_one_block_ahead_equals_serial():
    """jplg_context_create_ex(JPLG_CONTEXT_BACKGROUND): the panning data of block k+1 computed by a second, lower-priority context while block k
    renders (the reference's game thread beside its audio thread), ordered with events, gives the buses of the serial order bit for bit."""
    import torch
    ctx = J.Context(0)
    bg = J.Context(0, background=True)
    N, C, L, blocks, frames = 2085, 1, 12, 6, 512
    sc = make_scene(N, C, L, blocks, frames, util.SEED + 83)
    params = np.stack([util.pan_params(N, util.SEED + 90 + b, edge=True) for b in range(blocks)])
    dev = torch.device("cuda", 0)
    stream, stream_bg = ctx.torch_stream(), bg.torch_stream()
    pans = []
    for c in (ctx, bg):
        p = J.VBAPanner(c)
        assert p.Initialize(J.ChannelMask.Surround_7_1_4)
        pans.append((p, p.InitializeSourceLayout(J.ChannelMask.Mono)))
    mix0 = pans[0][0].ProcessVBAPData(pans[0][1], params[0]).reshape(N, C, L)
    assert util.bits_equal(pans[1][0].ProcessVBAPData(pans[1][1], params[0]).reshape(N, C, L), mix0)

    def make():
        return J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], mix0, ring_size=0, max_block_frames=512, mode=J.MODE_PARITY)

    ra, rb = make(), make()
    with torch.cuda.stream(stream):
        d_audio = [torch.from_numpy(sc["audio"][b]).to(dev) for b in range(blocks)]
        d_params = torch.from_numpy(params).to(dev)
        d_band = torch.from_numpy(sc["band"]).to(dev)
        d_delay = torch.from_numpy(sc["delay"]).to(dev)
        d_mix = torch.empty((N, C, L), dtype=torch.float32, device=dev)
        d_mix2 = [torch.empty((N, C, L), dtype=torch.float32, device=dev) for _ in range(2)]
        bus_a = [torch.zeros((frames, L), dtype=torch.float32, device=dev) for _ in range(blocks)]
        bus_b = [torch.zeros((frames, L), dtype=torch.float32, device=dev) for _ in range(blocks)]
    ctx.synchronize()
    # serial: everything on the rendering context
    for b in range(blocks):
        pans[0][0].ProcessVBAPData(pans[0][1], d_params[b], out=d_mix)
        ra.UpdateParameters(d_band[b + 1], d_delay[b + 1], d_mix)
        ra.ProcessInterleaved(d_audio[b], bus_a[b], frames)
    ctx.synchronize()
    # pipelined: the background context works one block ahead into two mix-map buffers
    ev_mdap = [torch.cuda.Event() for _ in range(2)]
    ev_upd = torch.cuda.Event()

    def issue(b, after_update):
        with torch.cuda.stream(stream_bg):
            if after_update:
                stream_bg.wait_event(ev_upd)
                rb.WaitRenderStarted(bg)
            pans[1][0].ProcessVBAPData(pans[1][1], d_params[b], out=d_mix2[b & 1])
            ev_mdap[b & 1].record(stream_bg)

    rb.SetMaxDelaySeconds(float(sc["delay"].max()))      # device-side updates below: push_ahead needs the bound declared
    issue(0, False)
    for b in range(blocks):
        with torch.cuda.stream(stream):
            stream.wait_event(ev_mdap[b & 1])
            rb.UpdateParameters(d_band[b + 1], d_delay[b + 1], d_mix2[b & 1])
            ev_upd.record(stream)
            rb.ProcessInterleaved(d_audio[b], bus_b[b], frames)
        if b + 1 < blocks:
            issue(b + 1, True)
            rb.PushAhead(d_audio[b + 1], frames)
    ctx.synchronize(); bg.synchronize()
    for b in range(blocks):
        assert np.abs(bus_a[b].cpu().numpy()).max() > 0
        assert util.bits_equal(bus_b[b].cpu().numpy(), bus_a[b].cpu().numpy()), f"block {b}: pipelined panning differs"
    del d_audio, d_params, d_band, d_delay, d_mix, d_mix2, bus_a, bus_b
    ra.close(); rb.close()
    del pans, p
    bg.close(); ctx.close()


@pytest.mark.gpu
def test_update_panned_equals_mdap_then_update():
    """jplg_renderer_update_panned (panning kernel with the UpdateParameters epilogue, double-buffered targets) against jplg_panner_mdap +
    jplg_renderer_update: same buses bit for bit, (a) on the renderer's own context, (b) from a background context one block ahead, (c) with a
    classic partial update refining a pending panned update. Mono and stereo sources."""
    import torch
    for C, mask in ((1, J.ChannelMask.Mono), (2, J.ChannelMask.Stereo)):
        ctx = J.Context(0)
        bg = J.Context(0, background=True)
        N, L, blocks, frames = 1031, 12, 5, 512
        sc = make_scene(N, C, L, blocks, frames, util.SEED + 87 + C)
        params = np.stack([util.pan_params(N, util.SEED + 120 + b, edge=True) for b in range(blocks)])
        dev = torch.device("cuda", 0)
        stream, stream_bg = ctx.torch_stream(), bg.torch_stream()
        pans = []
        for c in (ctx, bg):
            p = J.VBAPanner(c)
            assert p.Initialize(J.ChannelMask.Surround_7_1_4)
            pans.append((p, p.InitializeSourceLayout(mask)))
        mix0 = pans[0][0].ProcessVBAPData(pans[0][1], params[0]).reshape(N, C, L)

        def make():
            return J.DirectSoundBatch(ctx, FS, N, C, L, sc["band"][0], sc["delay"][0], mix0, ring_size=0, max_block_frames=512, mode=J.MODE_PARITY)

        ra, rb, rc_, rd = make(), make(), make(), make()
        with torch.cuda.stream(stream):
            d_audio = [torch.from_numpy(sc["audio"][b]).to(dev) for b in range(blocks)]
            d_params = torch.from_numpy(params).to(dev)
            d_band = torch.from_numpy(sc["band"]).to(dev)
            d_delay = torch.from_numpy(sc["delay"]).to(dev)
            d_mix = torch.empty((N, C, L), dtype=torch.float32, device=dev)
            buses = {k: [torch.zeros((frames, L), dtype=torch.float32, device=dev) for _ in range(blocks)] for k in "abcd"}
        ctx.synchronize()
        for b in range(blocks):                      # reference order: two calls
            pans[0][0].ProcessVBAPData(pans[0][1], d_params[b], out=d_mix)
            ra.UpdateParameters(d_band[b + 1], d_delay[b + 1], d_mix)
            ra.ProcessInterleaved(d_audio[b], buses["a"][b], frames)
        for b in range(blocks):                      # (a) fused, same context
            rb.UpdatePanned(pans[0][0], pans[0][1], d_params[b], d_band[b + 1], d_delay[b + 1])
            rb.ProcessInterleaved(d_audio[b], buses["b"][b], frames)
        ctx.synchronize()
        # (b) fused on the background context, issued one block ahead (the library orders the two streams itself)
        rc_.UpdatePanned(pans[1][0], pans[1][1], d_params[0], d_band[1], d_delay[1])
        for b in range(blocks):
            rc_.ProcessInterleaved(d_audio[b], buses["c"][b], frames)
            if b + 1 < blocks:
                rc_.WaitRenderStarted(bg)
                rc_.UpdatePanned(pans[1][0], pans[1][1], d_params[b + 1], d_band[b + 2], d_delay[b + 2])
        ctx.synchronize(); bg.synchronize()
        # (c) a pending panned update refined by a classic one (other band gains), against the same in two calls
        with torch.cuda.stream(stream):
            xa = torch.zeros((frames, L), dtype=torch.float32, device=dev)
        pans[0][0].ProcessVBAPData(pans[0][1], d_params[1], out=d_mix)
        ra.UpdateParameters(d_band[3], d_delay[2], d_mix)
        ra.ProcessInterleaved(d_audio[0], xa, frames)
        rd_in = buses["d"][0]
        for b in range(blocks):                      # bring rd to ra's state first
            rd.UpdatePanned(pans[0][0], pans[0][1], d_params[b], d_band[b + 1], d_delay[b + 1])
            rd.ProcessInterleaved(d_audio[b], buses["d"][b], frames)
        with torch.cuda.stream(stream):
            xd = torch.zeros((frames, L), dtype=torch.float32, device=dev)
        rd.UpdatePanned(pans[0][0], pans[0][1], d_params[1], d_band[2], d_delay[2])
        rd.UpdateParameters(d_band[3], None, None)
        rd.ProcessInterleaved(d_audio[0], xd, frames)
        ctx.synchronize()
        for b in range(blocks):
            want = buses["a"][b].cpu().numpy()
            assert np.abs(want).max() > 0
            for k in "bcd":
                assert util.bits_equal(buses[k][b].cpu().numpy(), want), f"C={C} block {b}: fused path {k} differs"
        assert util.bits_equal(xd.cpu().numpy(), xa.cpu().numpy()), f"C={C}: classic update on top of a pending panned update differs"
        del d_audio, d_params, d_band, d_delay, d_mix, buses, xa, xd, rd_in
        for r_ in (ra, rb, rc_, rd):
            r_.close()
        del pans, p
        bg.close(); ctx.close()
