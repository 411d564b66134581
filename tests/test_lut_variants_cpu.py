"""CPU checks of the LUT storage variants (jplg_panner_create_ex): host-side LUT construction only, no device needed."""
import numpy as np

import jplspatial_b200 as J
from jplspatial_b200 import _capi

M = J.ChannelMask


def _lut(ctx, mask, variant):
    p = J.VBAPanner(ctx)
    assert p.Initialize(mask, lut_variant=variant), getattr(p, "last_error", "")
    return p.GetLUT()


def test_kb786_gains_sit_on_the_24_bit_grid(host_ctx):
    g0, s0 = _lut(host_ctx, M.Surround_7_1_4, _capi.LUT_KB_983)
    g2, s2 = _lut(host_ctx, M.Surround_7_1_4, _capi.LUT_KB_786)
    assert np.array_equal(s0, s2)                                    # direct speaker indices either way
    assert not np.array_equal(g0, g2)
    grid = np.arange(1 << 24, dtype=np.uint32).astype(np.float32) * np.float32(1.0 / 16777215.0)     # Gain24Bit::operator float for every code
    idx = np.clip(np.searchsorted(grid, g2.ravel()), 0, grid.size - 1)
    assert np.array_equal(grid[idx], g2.ravel())
    pos = g0 > 0
    assert np.abs(g2[pos] - g0[pos]).max() <= 2.0 ** -23             # the packing error of Gain24Bit(float) for gains in (0, 1]


def test_kb851_without_lfe_is_the_default_table(host_ctx):
    g0, s0 = _lut(host_ctx, M.Surround_7_0_4, _capi.LUT_KB_983)
    g1, s1 = _lut(host_ctx, M.Surround_7_0_4, _capi.LUT_KB_851)
    assert np.array_equal(g0, g1) and np.array_equal(s0, s1)


def test_undefined_variants_are_refused(host_ctx):
    p = J.VBAPanner(host_ctx)
    assert not p.Initialize(M.Surround_7_1_4, lut_variant=_capi.LUT_KB_851) and "LFE" in p.last_error
    assert not p.Initialize(M.Surround_9_1_6, lut_variant=_capi.LUT_KB_655)
    assert not p.Initialize(M.Surround_9_1_6, lut_variant=_capi.LUT_KB_65)
    assert not p.Initialize(M.Surround_5_1, lut_variant=_capi.LUT_KB_786) and "3D" in p.last_error
    assert not p.Initialize(M.Surround_7_1_4, lut_variant=-1)
