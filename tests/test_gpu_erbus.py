"""GPU parity tests for the batched ERBus (early-reflection taps, BASELINE config 5) through the C-ABI against the
unmodified reference's ERBus (oracle/_ref, SSE2 build).

A tap's contribution depends only on the bus' mono ring and its own parameters, so the reference contribution of
tap `id` is an ERBus that is only ever given that one tap (and an empty list while the tap is absent): per-tap audio
must be BIT-EXACT in parity mode, through the whole add -> update -> fade-out -> re-create life cycle. The bus output
(sum over taps, re-associated on the GPU) is checked against the fp64 sum of the bit-exact per-tap outputs within the
fp32 summation bound, and against the reference bus' own sequential float sum."""
import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util

pytestmark = pytest.mark.gpu
FS = 48000.0


def schedule(bus, blocks, max_taps, seed):
    """-> per block: list of tap ids requested for this bus (ids come and go, some come back)"""
    u = util.uniform01(blocks * 64, seed + 31 * bus).reshape(blocks, 64)
    ids = list(range(1, max_taps // 2 + 1))
    nxt = 100
    out = []
    for b in range(blocks):
        if b and b % 3 == 0 and ids:
            ids = ids[1:]                                   # the oldest tap disappears -> fades out
        if b and b % 2 == 0 and len(ids) < max_taps - 2:
            ids = ids + [nxt]; nxt += 1                     # a new reflection appears -> fades in
        if b == 7 and 1 not in ids:
            ids = ids + [1]                                 # an id that was removed earlier comes back as a NEW tap
        if bus % 5 == 4 and b == 4:
            ids = []                                        # one bus in five loses every tap at once
        out.append(list(ids))
    return out, u


def tap_params(bus, tap_id, block, L, seed):
    u = util.uniform01(8 + L, seed + 7919 * bus + 104729 * tap_id + 13 * block)
    fg = (0.1 + 0.9 * u[:4]).astype(np.float32)
    base = util.uniform01(2, seed + 7919 * bus + 104729 * tap_id)          # delay drifts slowly around a per-tap base
    delay = np.float32(0.002 + 0.078 * base[0] + 0.0004 * (u[4] - 0.5))
    g = (u[8:8 + L] * (u[5] > 0.2)).astype(np.float32)
    g[3 % L] = 0.0
    return fg, delay, g


def run_case(gpu_ctx, ref, B, max_taps, L, in_ch, frames, blocks, seed, mode=J.MODE_PARITY, check_per_tap=True):
    x = util.white_noise((blocks, B, frames, in_ch), seed + 1)
    bank = J.ERBusBatch(gpu_ctx, B, max_taps, FS, L, max_block_frames=max(frames, 512), mode=mode)
    T = bank.slots_per_bus
    ref_bus = [ref.erbus(FS, L) for _ in range(B)]
    scheds = [schedule(i, blocks, max_taps, seed)[0] for i in range(B)]
    # one single-tap reference bus per (bus, id), alive from block 0 so its ring always holds the same history as the real bus
    ref_tap = {(i, t): ref.erbus(FS, L) for i in range(B) for t in sorted({t for ids in scheds[i] for t in ids})}
    worst_bus = 0.0
    for b in range(blocks):
        for i in range(B):
            ids = scheds[i][b]
            fgs, ds, gs = [], [], []
            for t in ids:
                fg, d, g = tap_params(i, t, b, L, seed)
                fgs.append(fg); ds.append(d); gs.append(g)
            bank.SetTaps(i, fgs, ds, ids, gs)
            ref_bus[i].set_taps(fgs, ds, ids, gs)
            assert bank.GetNumTaps(i) == ref_bus[i].num_taps()
            for (bi, t), rb in ref_tap.items():
                if bi != i:
                    continue
                if t in ids:
                    k = ids.index(t)
                    rb.set_taps([fgs[k]], [ds[k]], [t], [gs[k]])
                else:
                    rb.set_taps([], [], [], [])               # absent: fades out over the next chunk, then gone
        out = np.zeros((B, frames, L), np.float32)
        per = np.zeros((B * T, frames, L), np.float32)
        bank.ProcessInterleaved(x[b], out, frames, in_channels=in_ch, per_tap=per)
        want_bus = np.zeros((B, frames, L), np.float32)
        for i in range(B):
            ref_bus[i].process(x[b, i], want_bus[i], frames, n_in=in_ch)
        used = np.zeros(B * T, bool)
        for (i, t), rb in ref_tap.items():
            want = np.zeros((frames, L), np.float32)
            rb.process(x[b, i], want, frames, n_in=in_ch)
            s = bank.tap_slot(i, t)
            if not check_per_tap:
                continue
            if s < 0:
                assert not np.any(want), f"block {b} bus {i} tap {t}: the reference still renders a tap the bank has dropped"
                continue
            used[s] = True
            bad = util.count_bit_mismatch(per[s], want)
            assert bad == 0, f"block {b} bus {i} tap {t}: {bad} samples differ, max |d| {np.abs(per[s] - want).max()}"
        if check_per_tap:
            assert not np.any(per[~used]), "dead slots must contribute exact zeros"
            exact = per.reshape(B, T, frames, L).astype(np.float64).sum(1)
            bound = 4 * np.finfo(np.float32).eps * np.abs(per.reshape(B, T, frames, L)).astype(np.float64).sum(1) + 1e-12
            assert np.all(np.abs(out - exact) <= bound)
        scale = np.maximum(np.abs(want_bus), np.sqrt((want_bus.astype(np.float64) ** 2).mean()) + 1e-3)
        worst_bus = max(worst_bus, float((np.abs(out - want_bus) / scale).max()))
    return worst_bus


@pytest.mark.parametrize("B,max_taps,L,in_ch,frames,blocks", [(5, 16, 12, 1, 512, 10), (3, 8, 16, 2, 512, 9), (2, 30, 4, 1, 100, 24), (9, 6, 12, 3, 480, 9)])
def test_erbus_taps_bit_exact_through_life_cycle(gpu_ctx, ref, B, max_taps, L, in_ch, frames, blocks):
    worst = run_case(gpu_ctx, ref, B, max_taps, L, in_ch, frames, blocks, seed=0xE2 + B)
    assert worst <= 2e-6, f"bus sum differs from the reference's sequential sum by {worst} (relative)"


def test_erbus_throughput_mode_tolerance(gpu_ctx, ref):
    worst = run_case(gpu_ctx, ref, 4, 16, 12, 1, 512, 8, seed=99, mode=J.MODE_THROUGHPUT, check_per_tap=False)
    assert worst <= 1e-5


def test_erbus_argument_errors(gpu_ctx):
    with pytest.raises(J.JplgError):
        J.ERBusBatch(gpu_ctx, 2, 200, FS, 12)
    bank = J.ERBusBatch(gpu_ctx, 1, 8, FS, 12)
    ids = np.arange(1, 10)
    with pytest.raises(J.JplgError):
        bank.SetTaps(0, np.ones((9, 4)), np.full(9, 0.01), ids, np.ones((9, 12)))
