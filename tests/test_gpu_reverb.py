"""GPU parity tests for the batched ReverbBus (FDN16 late reverb, BASELINE config 5) through the C-ABI against the
unmodified reference's ReverbBus (oracle/_ref, SSE2 build).

Bars: decay gains (dBToGain = the reference's SIMD exp) and rendered audio BIT-EXACT in parity mode, over enough
blocks that every feedback path (delays 1609..4259 samples) has wrapped several times; throughput mode (fused
multiply-adds at the reference's FMA call sites) within 1e-5 * max(|ref|, rms, 1e-3) per sample."""
import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util

pytestmark = pytest.mark.gpu
FS = 48000.0
FDN_DELAYS = np.array([2927, 2593, 2273, 3697, 1877, 3877, 2477, 3461, 1609, 3779, 3541, 4259, 1669, 3539, 3637, 4013], np.float32)


def make_input(n, blocks, frames, seed):
    x = util.white_noise((blocks, n, frames), seed) * np.float32(0.5)
    x[0, :, 0] += 1.0                     # an impulse on top of the noise
    x[blocks // 2:, n // 2:] = 0.0        # half the buses ring out in silence (pure feedback)
    return np.ascontiguousarray(x, np.float32)


def rt60_for(n, seed, step):
    u = util.uniform01(n * 4, seed + 17 * step).reshape(n, 4)
    rt = (0.3 + 2.7 * u).astype(np.float32)
    rt[0] = [0.01, 20.0, 1.0, 0.03]       # clamped to [0.03, 10] by SetRT60
    return rt


def run_reference(ref, x, rts, L, frames):
    blocks, n = x.shape[:2]
    buses = [ref.reverb(FS) for _ in range(n)]
    out = np.zeros((blocks, n, frames, L), np.float32)
    for b in range(blocks):
        for i in range(n):
            if b in rts:
                buses[i].set_rt60(rts[b][i])
            buses[i].process(x[b, i].reshape(frames, 1), out[b, i], frames)
    return out


def run_gpu(ctx, x, rts, L, frames, mode):
    blocks, n = x.shape[:2]
    bank = J.ReverbBusBatch(ctx, n, FS, L, max_block_frames=max(frames, 512), mode=mode)
    out = np.zeros((blocks, n, frames, L), np.float32)
    for b in range(blocks):
        if b in rts:
            bank.SetRT60(rts[b])
        bank.ProcessInterleaved(x[b], out[b], frames)
    return out, bank


def test_decay_gains_bit_exact(gpu_ctx, ref):
    n = 37
    bank = J.ReverbBusBatch(gpu_ctx, n, FS, 12)
    g0 = bank.DecayGains()
    slope0 = np.full((1, 4), -60.0, np.float32)
    want0 = np.stack([ref.db_to_gain(slope0 * (d * np.float32(1.0 / FS))) for d in FDN_DELAYS])[:, 0]
    assert util.bits_equal(g0, np.broadcast_to(want0, (n, 16, 4)))
    rt = rt60_for(n, 5, 0)
    bank.SetRT60(rt)
    slope = (np.float32(-60.0) / np.clip(rt, np.float32(0.03), np.float32(10.0))).astype(np.float32)
    inv = np.float32(np.float32(1.0) / np.float32(FS))
    want = np.stack([np.stack([ref.db_to_gain(slope[i:i + 1] * (d * inv))[0] for d in FDN_DELAYS]) for i in range(n)])
    assert util.bits_equal(bank.DecayGains(), want)
    sl = (-2500.0 + 2499.0 * util.uniform01(n * 4, 9).reshape(n, 4)).astype(np.float32)   # SetDecaySlope clamps to [-2000, -6]
    bank.SetDecaySlope(sl)
    slc = np.clip(sl, np.float32(-60.0 / 0.03), np.float32(-60.0 / 10.0)).astype(np.float32)
    want = np.stack([np.stack([ref.db_to_gain(slc[i:i + 1] * (d * inv))[0] for d in FDN_DELAYS]) for i in range(n)])
    assert util.bits_equal(bank.DecayGains(), want)


@pytest.mark.parametrize("n,L,frames,blocks", [(21, 12, 512, 14), (9, 2, 512, 12), (5, 16, 100, 40), (3, 1, 512, 10), (8, 5, 333, 12)])
def test_reverb_bit_exact(gpu_ctx, ref, n, L, frames, blocks):
    x = make_input(n, blocks, frames, 0xABCDE + n)
    rts = {0: rt60_for(n, n, 0), blocks // 2: rt60_for(n, n, 1)}
    want = run_reference(ref, x, rts, L, frames)
    got, _ = run_gpu(gpu_ctx, x, rts, L, frames, J.MODE_PARITY)
    assert np.abs(want[-1]).max() > 1e-4, "the tail must still be audible or the test checks nothing"
    bad = util.count_bit_mismatch(got, want)
    assert bad == 0, f"{bad} of {got.size} samples differ; max |d| = {np.abs(got - want).max()}"


def test_reverb_accumulates_into_output(gpu_ctx, ref):
    n, L, frames = 4, 12, 512
    x = make_input(n, 5, frames, 77)
    bank = J.ReverbBusBatch(gpu_ctx, n, FS, L)
    buses = [ref.reverb(FS) for _ in range(n)]
    for b in range(5):
        got = np.full((n, frames, L), 0.25, np.float32)
        want = np.full((n, frames, L), 0.25, np.float32)
        bank.ProcessInterleaved(x[b], got, frames)
        for i in range(n):
            buses[i].process(x[b, i].reshape(frames, 1), want[i], frames)
        assert util.bits_equal(got, want)


def test_reverb_throughput_mode_tolerance(gpu_ctx, ref):
    n, L, frames, blocks = 16, 12, 512, 12
    x = make_input(n, blocks, frames, 4242)
    rts = {0: rt60_for(n, 3, 0)}
    want = run_reference(ref, x, rts, L, frames)
    got, _ = run_gpu(gpu_ctx, x, rts, L, frames, J.MODE_THROUGHPUT)
    rms = np.sqrt((want.astype(np.float64) ** 2).mean(axis=(2, 3), keepdims=True))
    tol = 1e-5 * np.maximum(np.maximum(np.abs(want), rms), 1e-3)
    assert np.all(np.abs(got.astype(np.float64) - want) <= tol), f"max excess {(np.abs(got - want) / tol).max()}"


@pytest.mark.parametrize("n_in,frames", [(2, 480), (6, 256), (4, 100)])
def test_reverb_multichannel_input_downmix_bit_exact(gpu_ctx, ref, n_in, frames):
    """stereo / quad / 5.1 reverb sends at <= 480 frames: DownmixToMono + ApplyGain(1 / numIn) in front of the FDN (LateReverb.cpp:80-92), device and host buffers"""
    n, L, blocks = 6, 12, (14 if frames >= 256 else 48)     # long enough for every feedback delay (1,609 .. 4,259 samples) to come round
    x = (util.white_noise((blocks, n, frames, n_in), 777 + n_in) * np.float32(0.5)).astype(np.float32)
    x[blocks // 2:, n // 2:] = 0.0
    buses = [ref.reverb(FS) for _ in range(n)]
    want = np.zeros((blocks, n, frames, L), np.float32)
    for b in range(blocks):
        for i in range(n):
            buses[i].process(x[b, i], want[b, i], frames, n_in=n_in)
    bank = J.ReverbBusBatch(gpu_ctx, n, FS, L)
    got = np.zeros_like(want)
    for b in range(blocks):
        bank.ProcessInterleaved(x[b], got[b], frames, in_channels=n_in)
    assert util.count_bit_mismatch(got, want) == 0, f"max abs {np.abs(got - want).max()}"
    assert np.abs(want[-1]).max() > 1e-4


def test_reverb_argument_errors(gpu_ctx):
    with pytest.raises(J.JplgError):
        J.ReverbBusBatch(gpu_ctx, 4, FS, 17)            # more outputs than FDN channels (asserted by the reference)
    with pytest.raises(J.JplgError):
        J.ReverbBusBatch(gpu_ctx, 4, FS, 12, max_block_frames=2048)
    bank = J.ReverbBusBatch(gpu_ctx, 2, FS, 12)
    out = np.zeros((2, 512, 12), np.float32)
    with pytest.raises(J.JplgError):     # multichannel input beyond one 480-frame chunk: the reference over-runs its scratch array (LateReverb.cpp:83)
        bank.ProcessInterleaved(np.zeros((2, 512, 2), np.float32), out, 512, in_channels=2)
    with pytest.raises(J.JplgError):
        bank.ProcessInterleaved(np.zeros((2, 600), np.float32), np.zeros((2, 600, 12), np.float32), 600)
