"""GPU parity tests for the scene front-end (the per-source body of SpatialManager::AdvanceSimulation and the
parameter producers) and the orientation-aware ProcessVBAPData (8f.4), through the C-ABI against the unmodified reference (oracle/_ref).

Bars, per output:
  * direct path (distance, dots, relative position), sanitised direction, base models Inverse / Linear, curve segments Constant /
    Linear / SCurve, air absorption dB, band gains (the reference's SIMD exp), delay: BIT-EXACT.
  * outputs that go through glibc's libm in the reference (cone factor: cosf of the cone angles; spread: atanf; Exponential model and
    the Sine / Log / Exp curve types: powf / sinf / cosf / logf): the device evaluates in double and rounds once; bar = 1e-6 absolute
    (the tolerance of the reference's own DirectPathServiceTest) and in practice <= 2 ulp.
The reference's own known-answer vectors for this path (DirectPathServiceTest.cpp) are restated here against the GPU as well."""
import math

import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util
from tests.frontend_inputs import make_positions

pytestmark = pytest.mark.gpu
M = J.ChannelMask


def ulp_diff(a, b):
    a = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7fffffff), a)
    b = np.where(b < 0, -(b & 0x7fffffff), b)
    return np.abs(a - b)


def test_direct_path_bit_exact(gpu_ctx, ref):
    n, m = 20000, 7
    src = make_positions(n, util.SEED + 100)
    lis = make_positions(m, util.SEED + 200, 10.0)
    idx = (util.splitmix64(n, util.SEED + 300) % np.uint64(m)).astype(np.uint32)
    # edge cases: on top of the listener, straight above / below (nudge branch), listener frame axes
    src[0, :3] = lis[idx[0], :3]
    for k, s in ((1, 1.0), (2, -1.0)):
        src[k, :3] = lis[idx[k], :3] + np.float32(s * 5.0) * lis[idx[k], 3:6]
    fe = J.DirectPathBatch(gpu_ctx)
    out = fe.Process(n, ("direct", "relative"), sources=src, listeners=lis, listener_index=idx)
    r = ref.direct_path(src, lis, idx)
    assert util.bits_equal(out["direct"][:, :3], r[:, :3])
    assert util.bits_equal(out["relative"], r[:, 3:12])
    # single listener, no index array
    out0 = fe.Process(n, ("direct",), sources=src, listeners=lis[:1])
    assert util.bits_equal(out0["direct"][:, :3], ref.direct_path(src, lis[:1])[:, :3])


def test_direct_path_known_answers(gpu_ctx):
    """DirectPathServiceTest.cpp:467-654 (ProcessDirectPath_ComputesCorrectDotProducts), tolerance 1e-5 as there."""
    s = math.sqrt(0.5)
    up = (0, 1, 0)
    cases = [  # source location, listener location, listener forward, expected dot, expected inverse dot
        ((0, 0, -5), (0, 0, 0), (0, 0, -1), 1.0, -1.0), ((0, 5, 0), (0, 0, 0), (0, 0, -1), 0.0, 0.0),
        ((-5, 0, 0), (0, 0, 0), (0, 0, -1), 0.0, 0.0), ((0, 0, 0), (0, 0, 0), (0, 0, -1), 1.0, -1.0),
        ((0, 0, -5), (0, 0, 0), (0, 0, 1), -1.0, -1.0), ((0, 5, 0), (0, 0, 0), (0, 0, 1), 0.0, 0.0),
        ((-5, 0, 0), (0, 0, 0), (0, 0, 1), 0.0, 0.0), ((0, 0, 0), (0, 0, 0), (0, 0, 1), 1.0, 1.0),
        ((0, 0, -10), (0, 0, -5), (0, 0, -1), 1.0, -1.0), ((10, 0, -10), (0, 0, -10), (0, 0, -1), 0.0, 0.0),
        ((-10, 0, -10), (0, 0, -20), (0, 0, -1), -0.7071, 0.7071),
        ((10, 0, -10), (5, 0, -5), (s, 0, -s), 1.0, -0.7071), ((-10, 0, -10), (0, 0, -20), (-s, 0, s), 1.0, 0.7071),
    ]
    fe = J.DirectPathBatch(gpu_ctx)
    for sl, ll, lf, dot, inv in cases:
        src = np.array([[*sl, *up, 0, 0, -1]], np.float32)      # OrientationData::IdentityForward
        lis = np.array([[*ll, *up, *lf]], np.float32)
        d = fe.Process(1, ("direct",), sources=src, listeners=lis)["direct"][0]
        assert abs(d[1] - dot) <= 1e-4 and abs(d[2] - inv) <= 1e-4, (sl, ll, lf, d)


def test_cone_attenuation(gpu_ctx, ref):
    n = 8192
    dots = (util.white_noise((n,), util.SEED + 400)).astype(np.float32)
    cones = (util.uniform01(2 * n, util.SEED + 401) * np.float32(2 * math.pi)).reshape(n, 2).astype(np.float32)
    cones.sort(axis=1)
    cones[:8] = np.array([[math.radians(60), math.radians(120)]] * 8, np.float32)
    dots[:8] = np.array([1.0, -1.0, math.cos(math.radians(45)), 0.0, math.cos(math.radians(30)), math.cos(math.radians(60)), 0.3, -0.3], np.float32)
    fe = J.DirectPathBatch(gpu_ctx)
    out = fe.Process(n, ("direct",), dot_in=dots, cones=cones)["direct"][:, 3]
    r = ref.angle_attenuation(0, dots, cones)
    assert np.abs(out - r).max() <= 1e-6
    # a 1-ulp difference between glibc's cosf and the correctly rounded cosine of a cone angle is amplified by the subtraction in
    # (cutoffInner - dot) / (cutoffInner - cutoffOuter); most factors are still the same float
    assert (ulp_diff(out, r) == 0).mean() > 0.8
    # known answers of DirectPathServiceTest.cpp:113-170 (InsideInnerCone, OutsideOuterCone) with lerp(1, outerGain, factor)
    assert out[0] == 0.0 and out[1] == 1.0 and 0.0 < out[2] < 1.0


def test_distance_attenuation_models_and_curves(gpu_ctx, ref):
    n = 4096
    d = (util.uniform01(n, util.SEED + 500) * np.float32(120.0) - np.float32(5.0)).astype(np.float32)
    curve_a = [(0.0, 1.0, J.CURVE_LINEAR), (10.0, 0.5, J.CURVE_SCURVE), (20.0, 0.25, J.CURVE_CONSTANT), (40.0, 0.1, J.CURVE_LINEAR), (100.0, 0.0, J.CURVE_LINEAR)]
    curve_b = [(50.0, 0.2, J.CURVE_EXP_3), (0.0, 1.0, J.CURVE_SINE_FADE_IN), (5.0, 0.9, J.CURVE_SINE_FADE_OUT), (12.0, 0.7, J.CURVE_LOG_1_41),
               (25.0, 0.5, J.CURVE_LOG_3), (35.0, 0.3, J.CURVE_EXP_1_41), (80.0, 0.05, J.CURVE_LINEAR)]   # unsorted on purpose: SortPoints
    funcs = [("model", J.MODEL_NONE, 1.0, 50.0, 1.0), ("model", J.MODEL_INVERSE, 1.0, 50.0, 1.0), ("model", J.MODEL_INVERSE, 2.0, 80.0, 0.5),
             ("model", J.MODEL_LINEAR, 1.0, 100.0, 1.0), ("model", J.MODEL_EXPONENTIAL, 1.0, 100.0, 2.0), ("model", J.MODEL_INVERSE, 10.0, 5.0, 1.0),
             ("curve", curve_a), ("curve", curve_b), ("curve", []), ("curve", [(3.0, 0.7, J.CURVE_LINEAR)])]
    fe = J.DirectPathBatch(gpu_ctx, functions=funcs)
    exact = {0, 1, 2, 3, 5, 6, 8, 9}
    for k, f in enumerate(funcs):
        out = fe.Process(n, ("distance_attenuation",), distance_in=d, function_index=np.full(n, k, np.uint32))["distance_attenuation"][:, 0]
        r = ref.attenuation_model(f[1], f[2], f[3], f[4], d) if f[0] == "model" else ref.attenuation_curve(np.array(f[1], np.float32).reshape(-1, 3), d)
        if k in exact:
            assert util.bits_equal(out, r), (k, np.abs(out - r).max())
        else:
            assert np.abs(out - r).max() <= 1e-6 and ulp_diff(out, r).max() <= 4, (k, np.abs(out - r).max(), ulp_diff(out, r).max())
    # mixed function indices in one launch; an index outside the table evaluates to 1 (no curve assigned)
    idx = (util.splitmix64(n, util.SEED + 501) % np.uint64(len(funcs) + 1)).astype(np.uint32)
    out = fe.Process(n, ("distance_attenuation",), distance_in=d, function_index=idx)["distance_attenuation"][:, 0]
    assert np.all(out[idx == len(funcs)] == 1.0)
    sel = idx == 6
    assert util.bits_equal(out[sel], ref.attenuation_curve(np.array(curve_a, np.float32), d[sel]))
    # AttenuationCurveTest.cpp-style known answers (the live subset of that file: linear segment, out-of-range clamps to the end values)
    ka = J.DirectPathBatch(gpu_ctx, functions=[("curve", [(0.0, 1.0, J.CURVE_LINEAR), (10.0, 0.5, J.CURVE_LINEAR)])])
    v = ka.Process(4, ("distance_attenuation",), distance_in=np.array([5.0, 15.0, -5.0, 10.0], np.float32))["distance_attenuation"][:, 0]
    assert np.allclose(v, [0.75, 0.5, 1.0, 0.5], atol=1e-6)


@pytest.mark.parametrize("env", [None, (5.0, 80.0, 95.0), (35.0, 15.0, 101.325)])
def test_air_absorption_band_gains_and_delay(gpu_ctx, ref, env):
    n = 16384
    d = (util.uniform01(n, util.SEED + 600) ** 2 * np.float32(3000.0)).astype(np.float32)
    d[:4] = [0.0, 1.0, 114.60987, 1844.9231]      # at the default knee distances
    if env is None:
        air, kw, c = None, {}, 0.0
    else:
        air = J.AirAbsorptionCache.CacheParameters(None, *env)
        kw = dict(temp_c=env[0], humidity=env[1], pressure=env[2])
        c = J.speed_of_sound_for_temperature(env[0])
        assert c == ref.speed_of_sound(env[0])
    fe = J.DirectPathBatch(gpu_ctx, air=air, speed_of_sound=c)
    out = fe.Process(n, ("air_db", "band_gains", "delay_seconds"), distance_in=d)
    rdb, rg = ref.air_for_distance(d, **kw)
    assert util.bits_equal(out["air_db"], rdb)
    assert util.bits_equal(out["band_gains"], rg)
    inv = np.float32(ref.inv_speed_of_sound()) if env is None else np.float32(1.0) / np.float32(c)
    assert util.bits_equal(out["delay_seconds"][:, 0], d * inv)


def test_full_frontend_feeds_panner_and_matches_reference(gpu_ctx, ref):
    """The whole per-source body against the reference's own sequence (ref_frontend_timed), then the produced rows go straight into
    the MDAP kernel on the device (torch tensors, no host round trip) and must give the reference's mix maps."""
    import torch
    n = 8192
    src = make_positions(n, util.SEED + 700, 150.0)
    lis = make_positions(1, util.SEED + 701, 1.0)
    src[0, :3] = lis[0, :3] + np.float32(7.0) * lis[0, 3:6]        # straight above: sanitised direction
    src[1, :3] = lis[0, :3] - np.float32(3.0) * lis[0, 3:6]
    cones = np.sort((util.uniform01(2 * n, util.SEED + 702) * np.float32(2 * math.pi)).reshape(n, 2), axis=1).astype(np.float32)
    sizes = (util.uniform01(n, util.SEED + 703) * np.float32(20.0)).astype(np.float32)
    model = np.array([J.MODEL_INVERSE, 1.0, 200.0, 1.0], np.float32)
    _, r = ref.frontend_timed(src, lis, cones, sizes, model, focus=0.75)
    fe = J.DirectPathBatch(gpu_ctx, functions=[("model", J.MODEL_INVERSE, 1.0, 200.0, 1.0)])
    want = ("direct", "pan_params", "band_gains", "delay_seconds", "distance_attenuation")
    out = fe.Process(n, want, sources=src, listeners=lis, cones=cones, object_sizes=sizes, focus=0.75, spread=0.0)
    assert util.bits_equal(out["direct"][:, :3], r["direct"][:, :3])
    assert np.abs(out["direct"][:, 3] - r["direct"][:, 3]).max() <= 1e-6
    assert util.bits_equal(out["distance_attenuation"][:, 0], r["distance_attenuation"])
    assert util.bits_equal(out["band_gains"], r["band_gains"])
    assert util.bits_equal(out["delay_seconds"][:, 0], r["delay_seconds"])
    assert util.bits_equal(out["pan_params"][:, :4], r["pan_params"][:, :4])
    assert np.abs(out["pan_params"][:, 4] - r["pan_params"][:, 4]).max() <= 1e-6 and ulp_diff(out["pan_params"][:, 4], r["pan_params"][:, 4]).max() <= 2
    assert util.bits_equal(out["pan_params"][:2, :3], ref.sanitize_source_dir(ref.direct_path(src[:2], lis)[:, 3:6]))
    # sources without an object size keep the spread they were given
    out2 = fe.Process(n, ("pan_params",), sources=src, listeners=lis, object_sizes=np.full(n, -1.0, np.float32), focus=0.5, spread=0.25)
    assert np.all(out2["pan_params"][:, 3] == 0.5) and np.all(out2["pan_params"][:, 4] == 0.25)

    # device-resident chain: front-end -> MDAP, all on the context stream (own context: closed before torch tears CUDA down)
    ctx = J.Context(0)
    fe2 = J.DirectPathBatch(ctx, functions=[("model", J.MODEL_INVERSE, 1.0, 200.0, 1.0)])
    pan = J.VBAPanner(ctx); assert pan.Initialize(M.Surround_7_1_4)
    lay = pan.InitializeSourceLayout(M.Mono)
    dev = torch.device("cuda:0")
    with torch.cuda.stream(ctx.torch_stream()):
        t = {k: torch.from_numpy(v).to(dev) for k, v in dict(sources=src, listeners=lis, cones=cones, object_sizes=sizes).items()}
        o = fe2.Process(n, ("direct", "pan_params"), focus=0.75, spread=0.0, **t)
        mix = pan.ProcessVBAPData(lay, o["pan_params"])
    ctx.synchronize()
    gp = o["pan_params"].cpu().numpy()
    gmix = mix.cpu().numpy()
    torch.cuda.synchronize()
    del t, o, mix, fe2, lay, pan
    ctx.close()
    assert util.bits_equal(gp, out["pan_params"])
    rp = ref.panner(M.Surround_7_1_4)
    rmix = rp.mdap(rp.layout(M.Mono), gp)      # same rows (the spread column may differ from the reference's by an ulp, see above)
    assert util.bits_equal(gmix.reshape(rmix.shape), rmix)


def test_spread_from_source_size_both_flavours(gpu_ctx, ref):
    n = 4096
    src = make_positions(n, util.SEED + 800, 80.0)
    lis = make_positions(1, util.SEED + 801, 1.0)
    sizes = (util.uniform01(n, util.SEED + 802) * np.float32(50.0)).astype(np.float32)
    dist = ref.direct_path(src, lis)[:, 0]
    for flavour in (0, 1):
        fe = J.DirectPathBatch(gpu_ctx, spread_flavour=flavour)
        sp = fe.Process(n, ("pan_params",), sources=src, listeners=lis, object_sizes=sizes)["pan_params"][:, 4]
        r = ref.spread_from_size(flavour, sizes, dist)
        assert np.abs(sp - r).max() <= 1e-6 and ulp_diff(sp, r).max() <= 2


@pytest.mark.parametrize("target,source,n", [("Surround_7_1_4", "Mono", 4096), ("Surround_7_1_4", "Stereo", 2048), ("Surround_9_1_6", "Quad", 1024)])
def test_mdap_oriented_bit_exact(gpu_ctx, ref, target, source, n):
    """ProcessVBAPData(PanUpdateDataWithOrientation) (PannerBase.h:501-515): rows produced by the front-end with pan_stride = 11."""
    src = make_positions(n, util.SEED + 900, 40.0)
    lis = make_positions(3, util.SEED + 901, 5.0)
    idx = (util.splitmix64(n, util.SEED + 902) % np.uint64(3)).astype(np.uint32)
    fs = np.stack([util.uniform01(n, util.SEED + 903), util.uniform01(n, util.SEED + 904)], axis=1).astype(np.float32)
    fs[:9] = [[f, s] for f in (0.0, 0.5, 1.0) for s in (0.0, 0.5, 1.0)]
    fe = J.DirectPathBatch(gpu_ctx)
    rows = fe.Process(n, ("pan_params", "relative"), sources=src, listeners=lis, listener_index=idx, focus_spread=fs, pan_stride=11)
    p = rows["pan_params"]
    assert util.bits_equal(p[:, 5:11], rows["relative"][:, 3:9]) and util.bits_equal(p[:, 3:5], fs)
    pan = J.VBAPanner(gpu_ctx); assert pan.Initialize(getattr(M, target))
    lay = pan.InitializeSourceLayout(getattr(M, source))
    g = pan.ProcessVBAPDataOriented(lay, p)
    rp = ref.panner(getattr(M, target))
    r = rp.mdap_oriented(rp.layout(getattr(M, source)), p)
    nbad = util.count_bit_mismatch(g, r)
    assert nbad == 0, f"{nbad} gains differ, max abs {np.abs(g - r).max()}"


def test_frontend_error_paths(gpu_ctx):
    fe = J.DirectPathBatch(gpu_ctx)
    assert fe.Process(0, ("direct",), distance_in=np.zeros(0, np.float32))["direct"].shape == (0, 4)
    with pytest.raises(J.JplgError):
        fe.Process(4, ("pan_params",), distance_in=np.zeros(4, np.float32), pan_stride=7)
    with pytest.raises(J.JplgError):
        fe.Process(4, ("pan_params",), distance_in=np.zeros(4, np.float32), pan_stride=11)   # oriented rows need positions
    with pytest.raises(J.JplgError):
        fe.Process(4, ("direct",), sources=np.zeros((4, 9), np.float32))                      # sources without listeners
