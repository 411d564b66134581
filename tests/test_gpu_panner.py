"""GPU parity tests for stage 1 (VBAP / MDAP gains) through the C-ABI against the unmodified reference
(oracle/_ref, SSE2 build). Bar: north_star says 1e-6 absolute; these kernels are bit-exact, so the tests ask for
bit equality and report the absolute error as well."""
import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util

pytestmark = pytest.mark.gpu
M = J.ChannelMask


def _panner(ctx, mask):
    p = J.VBAPanner(ctx)
    assert p.Initialize(mask)
    return p


def test_hello_quad_360_directions(gpu_ctx, ref):
    """BASELINE config 1 / Examples/HelloJPLSpatial/main.cpp:34-60."""
    p = _panner(gpu_ctx, M.Quad)
    rp = ref.panner(M.Quad)
    th = np.deg2rad(np.arange(360, dtype=np.float64))
    dirs = np.stack([np.sin(th), np.zeros_like(th), -np.cos(th)], 1).astype(np.float32)
    dirs = np.concatenate([np.array([[0, 0, -1], [1, 0, 0], [0, 0, 1], [-1, 0, 0]], np.float32), dirs])
    g = p.GetSpeakerGains(dirs)
    assert util.bits_equal(g, rp.point_gains(dirs))
    # expected text of Examples/HelloJPLSpatial/README.md:38-41 (quad order FL FR BL BR)
    exp = np.array([[.7071, .7071, 0, 0], [0, .7071, 0, .7071], [0, 0, .7071, .7071], [.7071, 0, .7071, 0]], np.float32)
    assert np.allclose(g[:4], exp, atol=5e-5)


@pytest.mark.parametrize("name", ["Stereo", "LCR", "Quad", "Surround_5_1", "Surround_7_1", "Octagonal",
                                  "Surround_5_0_2", "Surround_7_1_4", "Surround_9_1_6", "Surround_7_0_6"])
def test_point_gains_bit_exact(gpu_ctx, ref, name):
    mask = getattr(M, name)
    p = _panner(gpu_ctx, mask)
    rp = ref.panner(mask)
    dirs = np.concatenate([util.special_directions(), util.unit_directions(20000, util.SEED + 11)])
    if not p.is_3d():
        dirs = dirs[(np.abs(dirs[:, 0]) + np.abs(dirs[:, 2])) > 0]  # 2D query normalises (x,0,z): exclude the two vertical axes
    g = p.GetSpeakerGains(dirs)
    r = rp.point_gains(dirs)
    assert util.count_bit_mismatch(g, r) == 0, f"max abs err {np.abs(g - r).max()}"


MDAP_CASES = [
    ("Surround_7_1_4", "Mono", 65536),      # BASELINE config 2
    ("Surround_7_1_4", "Stereo", 8192),
    ("Surround_7_1_4", "Surround_5_1", 4096),
    ("Surround_9_1_6", "Mono", 8192),
    ("Surround_9_1_6", "Quad", 4096),
    ("Surround_5_0_2", "Mono", 8192),
    ("Quad", "Mono", 16384),
    ("Surround_7_1", "Stereo", 8192),
    ("Surround_5_1", "Surround_5_1", 4096),
    ("Stereo", "Mono", 4096),
]


@pytest.mark.parametrize("target,source,n", MDAP_CASES)
def test_mdap_bit_exact(gpu_ctx, ref, target, source, n):
    tm, sm = getattr(M, target), getattr(M, source)
    p = _panner(gpu_ctx, tm)
    lay = p.InitializeSourceLayout(sm)
    rp = ref.panner(tm)
    rl = rp.layout(sm)
    assert (lay.num_channels, lay.num_virtual_sources) == (rl.num_channels, rl.num_virtual_sources)
    params = util.pan_params(n, util.SEED + 5)
    if not p.is_3d():
        params = params[(np.abs(params[:, 0]) + np.abs(params[:, 2])) > 1e-3]
    g = p.ProcessVBAPData(lay, params)
    r = rp.mdap(rl, params)
    bad = np.nonzero((g.view(np.uint32) != r.view(np.uint32)).reshape(len(params), -1).any(1))[0]
    err = np.abs(g - r).max()
    assert err <= 1e-6, f"{len(bad)} sources differ, max abs err {err}, first {bad[:5]}"
    assert len(bad) == 0, f"{len(bad)} of {len(params)} sources not bit-exact (max abs err {err})"


def test_mdap_with_reference_lut_bytes(gpu_ctx, ref):
    """Kernel parity decoupled from LUT construction: upload the reference's own LUT bytes."""
    p = _panner(gpu_ctx, M.Surround_7_1_4)
    rp = ref.panner(M.Surround_7_1_4)
    g, s = rp.lut()
    p.SetLUT(g, s)
    lay = p.InitializeSourceLayout(M.Mono)
    params = util.pan_params(4096, util.SEED + 9)
    assert util.bits_equal(p.ProcessVBAPData(lay, params), rp.mdap(rp.layout(M.Mono), params))


def test_mdap_power_preservation(gpu_ctx):
    """VBAPPannedGainsL2Normalized (SpatializationTests/src/Tests/VBAP3DTest.cpp:510-653): sum_out (sum_src g)^2 == 1, tol 2e-6."""
    p = _panner(gpu_ctx, M.Surround_7_1_4)
    for src in ("Mono", "Stereo", "Quad", "Surround_5_1"):
        lay = p.InitializeSourceLayout(getattr(M, src))
        params = util.pan_params(2048, util.SEED + 21)
        g = p.ProcessVBAPData(lay, params).astype(np.float64)
        power = ((g.sum(1)) ** 2).sum(1)
        assert np.abs(power - 1.0).max() <= 3e-6, np.abs(power - 1.0).max()


def test_process_virtual_sources(gpu_ctx, ref):
    for name in ("Quad", "Surround_5_0_2", "Surround_7_1_4"):
        mask = getattr(M, name)
        p = _panner(gpu_ctx, mask)
        rp = ref.panner(mask)
        dirs = util.unit_directions(64 * 7, util.SEED + 31).reshape(64, 7, 3)
        if not p.is_3d():
            dirs[..., 1] = 0
        w = util.uniform01(64 * 7, util.SEED + 32).reshape(64, 7, 1)
        vs = np.concatenate([dirs, w], 2).astype(np.float32)
        for b in range(64):
            vs[b] = J.normalize_weights(vs[b])
        g = p.ProcessVirtualSources(vs)
        r = np.stack([rp.process_virtual_sources(vs[b]) for b in range(64)])
        assert util.count_bit_mismatch(g, r) == 0, np.abs(g - r).max()
    # VBAP2DTest.cpp:445-501: quad, one VS at 0 deg -> {.7071,.7071,0,0}; four VS at the speakers -> {.5,.5,.5,.5}
    p = _panner(gpu_ctx, M.Quad)
    g = p.ProcessVirtualSources(np.array([[[0, 0, -1, 1.0]]], np.float32))
    assert np.allclose(g[0], [0.7071, 0.7071, 0, 0], atol=1e-4)
    s = 0.70710678
    vs = np.array([[[-s, 0, -s, .25], [s, 0, -s, .25], [-s, 0, s, .25], [s, 0, s, .25]]], np.float32)
    assert np.allclose(p.ProcessVirtualSources(vs)[0], [0.5, 0.5, 0.5, 0.5], atol=1e-4)
