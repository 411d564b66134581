"""GPU tests of the LUT storage variants (jplg_panner_create_ex; ELUTSize, inc/Panning/VBAPLUT3D.h:93-104,186-204) against the unmodified
reference's own VBAPanner3D<Traits, ELUTSize> instantiations (oracle/_ref, ref_lut_variant): bit for bit."""
import numpy as np
import pytest

import jplspatial_b200 as J
from jplspatial_b200 import _capi
from tests import util

pytestmark = pytest.mark.gpu
M = J.ChannelMask


def _dirs(n, seed):
    v = util.uniform01(3 * n, seed).reshape(n, 3).astype(np.float32) * 2.0 - 1.0
    v /= np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-3).astype(np.float32)
    return v.astype(np.float32)


@pytest.mark.parametrize("variant,kind,target,source", [
    (_capi.LUT_KB_786, 2, "Surround_7_1_4", "Mono"),
    (_capi.LUT_KB_786, 2, "Surround_9_1_6", "Stereo"),
    (_capi.LUT_KB_786, 2, "Surround_5_0_2", "Quad"),
    (_capi.LUT_KB_851, 1, "Surround_7_0_4", "Mono"),
    (_capi.LUT_KB_851, 1, "Surround_5_0_2", "Stereo"),
    (_capi.LUT_KB_851, 1, "Surround_9_0_6", "Surround_5_0"),
    (_capi.LUT_KB_983, 0, "Surround_7_1_4", "Mono"),
])
def test_variant_matches_the_references_own_variant(gpu_ctx, ref, variant, kind, target, source):
    tm, sm = getattr(M, target), getattr(M, source)
    params = util.pan_params(2000, util.SEED + 600 + kind)
    dirs = _dirs(3000, util.SEED + 610 + kind)
    got = ref.lut_variant(kind, tm, sm, params, dirs)
    assert got is not None
    rmix, rpt = got
    pan = J.VBAPanner(gpu_ctx)
    assert pan.Initialize(tm, lut_variant=variant), getattr(pan, "last_error", "")
    lay = pan.InitializeSourceLayout(sm)
    gpt = pan.GetSpeakerGains(dirs)
    assert util.bits_equal(gpt, rpt), f"point gains differ: {np.abs(gpt - rpt).max()} on {np.count_nonzero((gpt != rpt).any(1))} directions"
    gmix = pan.ProcessVBAPData(lay, params)
    assert util.bits_equal(gmix.reshape(rmix.shape), rmix), f"MDAP gains differ: {np.abs(gmix.reshape(rmix.shape) - rmix).max()}"
    if variant == _capi.LUT_KB_786:
        # the packing is visible: the table is not the default one
        dflt = J.VBAPanner(gpu_ctx); assert dflt.Initialize(tm)
        assert not util.bits_equal(dflt.GetSpeakerGains(dirs), gpt)


def test_undefined_variants_are_refused(gpu_ctx):
    pan = J.VBAPanner(gpu_ctx)
    assert not pan.Initialize(M.Surround_7_1_4, lut_variant=_capi.LUT_KB_851) and "LFE" in pan.last_error      # upstream reads past its triangle table
    assert not pan.Initialize(M.Surround_7_1_4, lut_variant=_capi.LUT_KB_655)
    assert not pan.Initialize(M.Surround_7_1_4, lut_variant=_capi.LUT_KB_65)
    assert not pan.Initialize(M.Quad, lut_variant=_capi.LUT_KB_786)                                               # 2D panners have one LUT type
    assert not pan.Initialize(M.Surround_7_1_4, lut_variant=7)
