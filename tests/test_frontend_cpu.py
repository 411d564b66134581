"""CPU tests of the scene front-end: the oracle restatement (oracle/port) and the product's HOST builders
(AirAbsorptionCache, speed of sound, SanitizeSourceDir's fallback) against the committed golden fixture tests/golden/frontend.npz, which
oracle/tools/gen_golden_frontend.py wrote by running the unmodified reference. Everything here is bit-exact: the port runs on the CPU and
calls the same libm the reference does. No device is needed."""
import os

import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util
from tests.frontend_inputs import CURVE_A, CURVE_B, MODELS, ENVS

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "frontend.npz")
M = J.ChannelMask


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def port():
    from oracle import portpy
    if not portpy.available():
        pytest.skip("oracle/libjploracle.so not built")
    return portpy.Port()


def test_port_direct_path_and_cone(port, gold):
    assert util.bits_equal(port.direct_path(gold["src"], gold["lis"], gold["idx"]), gold["direct"])
    assert util.bits_equal(port.angle_attenuation(gold["dots"], gold["cones"]), gold["cone_factor"])


def test_port_distance_attenuation(port, gold):
    for k, m in enumerate(MODELS):
        assert util.bits_equal(port.attenuation_model(*m, gold["dist"]), gold[f"model{k}"]), k
    assert util.bits_equal(port.attenuation_curve(CURVE_A, gold["dist"]), gold["curve_a"])
    assert util.bits_equal(port.attenuation_curve(CURVE_B, gold["dist"]), gold["curve_b"])
    # the live subset of AttenuationCurveTest.cpp (its table is compiled out upstream): linear segment, clamps outside the range, empty curve
    lin = [(0.0, 1.0, 1), (10.0, 0.5, 1)]
    assert np.allclose(port.attenuation_curve(lin, [5.0, 15.0, -5.0]), [0.75, 0.5, 1.0], atol=1e-6)
    assert port.attenuation_curve(np.zeros((0, 3), np.float32), [5.0])[0] == 1.0


def test_port_air_absorption_and_speed_of_sound(port, gold):
    for k, env in enumerate(ENVS):
        cache, knee = port.air_cache(gold["bands"], *env)
        assert util.bits_equal(cache, gold[f"air_cache{k}"]) and np.array_equal(knee, gold[f"air_knee{k}"])
        db, g = port.air_for_distance(cache, knee, gold["far"])
        assert util.bits_equal(db, gold[f"air_db{k}"]) and util.bits_equal(g, gold[f"air_gain{k}"])
        assert np.float32(port.speed_of_sound(env[0])) == gold[f"sos{k}"]


def test_port_sanitize_and_spread(port, gold):
    assert util.bits_equal(port.sanitize_source_dir(gold["san_in"]), gold["san_out"])
    for f in (0, 1):
        assert util.bits_equal(port.spread_from_size(f, gold["sizes"], gold["direct"][:, 0]), gold[f"spread{f}"])


def test_port_mdap_oriented(port, gold):
    assert util.bits_equal(port.mdap_oriented(M.Surround_7_1_4, M.Mono, gold["oriented_params"]), gold["oriented_mix"])


def test_host_builders_match_reference_fixture(gold):
    """jplg_air_absorption_cache_build / jplg_speed_of_sound_for_temperature / jplg_sanitize_fallback_dir: host code of the product,
    the same role as the LUT builders (the reference also computes these once per environment on the CPU)."""
    assert util.bits_equal(J.default_band_centers(), gold["bands"])
    names = ("band_centers", "lower_third_centers", "center_db", "lower_third_db", "weighted_db", "decay_slope", "distance_to_knee", "linearized_loss_factor")
    for k, env in enumerate(ENVS):
        c = J.AirAbsorptionCache.CacheParameters(None, *env).as_dict()
        assert util.bits_equal(np.concatenate([c[n] for n in names]), gold[f"air_cache{k}"])
        assert np.array_equal(c["has_knee"], gold[f"air_knee{k}"])
        assert np.float32(J.speed_of_sound_for_temperature(env[0])) == gold[f"sos{k}"]
    assert util.bits_equal(J.sanitize_fallback_dir(), gold["san_out"][5])
    assert np.float32(1.0) / np.float32(343.2) == gold["inv_sos"]


def test_port_matches_live_reference_on_larger_inputs(port, ref):
    from tests.frontend_inputs import make_positions
    n = 4000
    src = make_positions(n, util.SEED + 1100, 200.0); lis = make_positions(5, util.SEED + 1101, 20.0)
    idx = (util.splitmix64(n, util.SEED + 1102) % np.uint64(5)).astype(np.uint32)
    r = ref.direct_path(src, lis, idx)
    assert util.bits_equal(port.direct_path(src, lis, idx), r)
    assert util.bits_equal(port.sanitize_source_dir(r[:, 3:6]), ref.sanitize_source_dir(r[:, 3:6]))
    rng = np.random.default_rng(7)
    for _ in range(40):
        env = tuple(float(np.float32(v)) for v in (rng.uniform(-20, 45), rng.uniform(5, 100), rng.uniform(80, 110)))
        bands = np.sort(rng.uniform(30, 12000, 4)).astype(np.float32)
        c = ref.air_cache(bands, *env)
        cache, knee = port.air_cache(bands, *env)
        names = ("band_centers", "lower_third_centers", "center_db", "lower_third_db", "weighted_db", "decay_slope", "distance_to_knee", "linearized_loss_factor")
        assert util.bits_equal(cache, np.concatenate([c[k] for k in names]))
        h = J.AirAbsorptionCache.CacheParameters(bands, *env).as_dict()
        assert util.bits_equal(np.concatenate([h[k] for k in names]), cache)


def test_frontend_needs_a_device(host_ctx):
    """No CPU fallback: creating the front-end on a host-only context fails loudly."""
    with pytest.raises(J.JplgError) as e:
        J.DirectPathBatch(host_ctx)
    assert "no CPU fallback" in str(e.value) or e.value.status != 0
