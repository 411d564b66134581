"""CPU tests (no GPU) of the drop-in boundary: the C-ABI library loads and exports every symbol include/jplg.h declares, the
host-side builders it runs at create time (ChannelMap helpers, LUT construction, source layouts, filter coefficients) reproduce the
reference bit for bit (golden fixtures), and every data-path call fails loudly without a device (no CPU fallback)."""
import ctypes as C
import hashlib
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def J():
    import jplspatial_b200 as J
    return J


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "jplg.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"JPLG_API\s+[\w\s\*]+?\b(jplg_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol(J):
    names = _declared_symbols()
    assert len(names) >= 60, names
    lib = C.CDLL(J.lib_path)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/jplg.h but not exported by libjplg.so: {missing}"


def test_header_cites_reference_interfaces():
    text = open(os.path.join(ROOT, "include", "jplg.h")).read()
    for cite in ("PannerBase.h:315", "DirectSound.cpp:95-229", "CrossoverFilter.h:416-535", "DelayLine.h", "ChannelMixing.h:32-136", "ChannelMap.h"):
        assert cite in text, cite


def test_library_is_sm100a_and_has_no_cpu_fallback(J, host_ctx):
    assert "sm_100a" in J.version()
    pan = J.VBAPanner(host_ctx)
    assert pan.Initialize(J.ChannelMask.Quad)
    with pytest.raises(J.JplgError) as e:
        pan.GetSpeakerGains(np.float32([[0, 0, -1]]))
    assert e.value.status == 5  # JPLG_ERR_NO_DEVICE
    with pytest.raises(J.JplgError):
        J.DirectSoundBatch(host_ctx, 48000.0, 4, 1, 4)
    lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
    with pytest.raises(J.JplgError):
        pan.ProcessVBAPData(lay, np.float32([[0, 0, -1, 0.5, 0.5]]))


def test_background_context_flag(J):
    """jplg_context_create_ex: flags = 0 is jplg_context_create, JPLG_CONTEXT_BACKGROUND is accepted (host-only contexts have no stream to
    prioritise), anything else is rejected; jplg_renderer_wait_render_started checks its arguments."""
    lib = C.CDLL(J.lib_path)
    lib.jplg_context_create_ex.restype = C.c_int32
    lib.jplg_context_create_ex.argtypes = [C.c_int32, C.c_uint32, C.POINTER(C.c_void_p)]
    for flags, want in ((0, 0), (1, 0), (2, 1), (0x80000001, 1)):
        h = C.c_void_p()
        st = lib.jplg_context_create_ex(-1, flags, C.byref(h))
        assert st == want, (flags, st)                       # 1 = JPLG_ERR_INVALID_ARGUMENT
        if st == 0:
            lib.jplg_context_destroy.argtypes = [C.c_void_p]
            lib.jplg_context_destroy(h)
        else:
            assert not h.value
    bg = J.Context(-1, background=True)
    pan = J.VBAPanner(bg)
    assert pan.Initialize(J.ChannelMask.Quad)                # host builders work on a background context too
    lib.jplg_renderer_wait_render_started.restype = C.c_int32
    lib.jplg_renderer_wait_render_started.argtypes = [C.c_void_p, C.c_void_p]
    assert lib.jplg_renderer_wait_render_started(None, bg.h) == 1
    lib.jplg_renderer_update_panned.restype = C.c_int32
    lib.jplg_renderer_update_panned.argtypes = [C.c_void_p] * 6 + [C.c_int32]
    assert lib.jplg_renderer_update_panned(None, None, None, None, None, None, 1) == 1      # no renderer / panner: invalid argument, nothing touched
    del pan
    bg.close()


def test_channel_map_helpers(J):
    """inc/ChannelMap.h:153-249"""
    M, E = J.ChannelMask, J.EChannel
    m = J.ChannelMap.FromChannelMask(M.Surround_7_1_4)
    assert m.GetNumChannels() == 12 and m.HasLFE() and m.HasTopChannels()
    assert m.GetChannelIndex(E.FrontLeft) == 0 and m.GetChannelIndex(E.LFE) == 3 and m.GetChannelIndex(E.TopBackRight) == 11
    assert m.GetChannelIndex(E.WideLeft) == 0xFFFFFFFF
    assert J.ChannelMap.FromNumChannels(2).mask == M.Stereo and J.ChannelMap.FromNumChannels(4).mask == M.Quad
    assert J.ChannelMap.FromNumChannels(6).mask == M.Surround_5_1 and J.ChannelMap.FromNumChannels(8).mask == M.Surround_7_1
    assert J.ChannelMap.FromChannelMask(M.Surround_9_1_6).GetNumChannels() == 16
    assert not J.ChannelMap.FromChannelMask(M.Quad).HasTopChannels()


def test_initialize_rejects_what_the_reference_rejects(J, host_ctx):
    """Panning2D/3D::IsValidTargetChannelMap (inc/Panning/VBAPanning2D.h:124-137, VBAPanning3D.h:157-166); invalid source maps (PannerBase.h:188-203)"""
    assert not J.VBAPanner(host_ctx).Initialize(0)
    assert not J.VBAPanner(host_ctx).Initialize(J.ChannelMask.Mono)
    pan = J.VBAPanner(host_ctx)
    assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
    with pytest.raises(J.JplgError):
        pan.InitializeSourceLayout(J.ChannelMask.Surround_5_0_2)   # top channels do not take part in panning
    with pytest.raises(J.JplgError):
        pan.InitializeSourceLayout(0)


def _sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def test_host_builders_reproduce_reference_luts_and_layouts(J, host_ctx):
    """jplg_panner_create builds the LUT on the host (jplg_host.cpp): bytes equal the reference's for the 25 standard layouts;
    source-layout constants equal too (golden: tests/golden/lut_hashes.json, generated from the reference build)."""
    table = json.load(open(os.path.join(GOLD, "lut_hashes.json")))
    for name, e in table.items():
        pan = J.VBAPanner(host_ctx)
        assert pan.Initialize(getattr(J.ChannelMask, name)), name
        assert pan.GetNumChannels() == e["channels"] and pan.is_3d() == e["is3d"]
        assert int(np.float32(pan.GetShortestSpeakerAperture()).view(np.uint32)) == e["aperture_bits"], name
        if pan.is_3d():
            assert pan.triangulation_is_tabulated()
        g, s = pan.GetLUT()
        assert (_sha(g, s) if pan.is_3d() else _sha(g)) == e["sha256"], name
        for sname, want in e["layouts"].items():
            smask = getattr(J.ChannelMask, sname)
            if want is None:
                with pytest.raises(J.JplgError):
                    pan.InitializeSourceLayout(smask)
                continue
            lay = pan.InitializeSourceLayout(smask)
            assert (lay.num_channels, lay.num_virtual_sources, lay.num_rings, lay.samples_per_ring) == (want["channels"], want["vs"], want["rings"], want["per_ring"])
            for v, key in ((lay.phi_term, "phi_bits"), (lay.offset_term, "offset_bits"), (lay.vs_weight, "weight_bits")):
                assert int(np.float32(v).view(np.uint32)) == want[key], (name, sname, key)
            q, ids = lay.groups()
            assert _sha(q) == want["quats_sha256"] and ids.tolist() == want["ids"]


def test_speaker_triangle_table_matches_golden(J, host_ctx):
    tris = json.load(open(os.path.join(GOLD, "speaker_tris.json")))
    for key, flat in tris.items():
        pan = J.VBAPanner(host_ctx)
        assert pan.Initialize(int(key, 16))
        assert pan.triangles().reshape(-1).tolist() == flat


def test_generic_triangulator_for_non_standard_layout(J, host_ctx):
    """A 3D mask outside the reference's 14 standard layouts goes through the generic hull: a closed hull must come out (parity reported, not promised)."""
    E = J.EChannel
    mask = E.FrontLeft | E.FrontRight | E.BackLeft | E.BackRight | E.TopFrontCenter | E.TopBackCenter
    pan = J.VBAPanner(host_ctx)
    assert pan.Initialize(mask)
    assert not pan.triangulation_is_tabulated()
    t = pan.triangles()
    verts = 6 + 1
    assert t.shape[0] == 2 * verts - 4
    g, s = pan.GetLUT()
    code = np.arange(65536)
    valid = ((code & 255) != 255) & ((code >> 8) != 255)          # component 255 is not a valid code (DirectionEncoding.h:260-263)
    assert (s[valid] != 255).all(), "every valid cell must be covered by a triangle"
    np.testing.assert_allclose((g[valid].astype(np.float64) ** 2).sum(1), 1.0, atol=1e-5)


def test_crossover_coefficients_match_golden(J, host_ctx):
    bank = J.CrossoverBank(host_ctx, 1, 48000.0)
    want = np.load(os.path.join(GOLD, "crossover.npz"))["coeffs"]
    assert np.array_equal(bank.coeffs().view(np.uint32), want.view(np.uint32))


def test_small_host_helpers(J):
    """NormalizeWeights (PannerBase.h:879-892), GetSpreadFromSourceSize (:69-74)"""
    vs = np.float32([[0, 0, -1, 2.0], [1, 0, 0, -1.0], [0, 0, 1, 1.0]])
    out = J.normalize_weights(vs)
    np.testing.assert_allclose(out[:, 3], [0.5, -0.25, 0.25])
    z = J.normalize_weights(np.float32([[0, 0, -1, 0.0]]))
    assert z[0, 3] == 0.0
    s = J.get_spread_from_source_size(1.0, 1.0)
    assert 0.0 < s <= 1.0
    assert J.get_spread_from_source_size(1.0, 100.0) < s
