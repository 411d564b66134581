"""TEST INFRASTRUCTURE — ctypes wrapper over oracle/libjploracle.so, the plain-C++ CPU restatement of the hot path
(oracle/port/jpl_oracle.cpp). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may
import this module; the product package never does."""
import ctypes as C
import json
import os
import numpy as np

_here = os.path.dirname(os.path.abspath(__file__))
_vp, _u32, _f32 = C.c_void_p, C.c_uint32, C.c_float


def port_path():
    return os.path.join(_here, "libjploracle.so")


def available():
    return os.path.exists(port_path())


def speaker_triangles(mask):
    """Golden speaker-triangle table of the reference for a 3D layout (tests/golden/speaker_tris.json, dumped from
    SpeakerTriangulation::TriangulateSpeakerLayout by oracle/tools/gen_golden.py). LFE does not take part."""
    path = os.path.join(os.path.dirname(_here), "tests", "golden", "speaker_tris.json")
    table = json.load(open(path))
    key = "0x%06x" % (mask & ~(1 << 3))
    if key not in table:
        return None
    return np.ascontiguousarray(np.array(table[key], np.uint8).reshape(-1, 3))


def _f(a):
    return np.ascontiguousarray(a, np.float32)


class Port:
    def __init__(self):
        self.lib = L = C.CDLL(port_path())
        L.orc_panner_create.restype = _vp
        L.orc_panner_create.argtypes = [_u32, _vp, _u32]
        L.orc_panner_aperture.restype = _f32
        L.orc_smoother_alpha.restype = _f32
        L.orc_dse_create.restype = _vp
        L.orc_dse_create.argtypes = [_f32, _u32, _u32, _vp, _f32, _vp]
        L.orc_dse_update.argtypes = [_vp, _vp, _f32, _vp]
        L.orc_dse_process.argtypes = [_vp, _vp, _vp, _u32]
        L.orc_dse_process_batch.restype = C.c_double
        L.orc_dse_process_batch.argtypes = [_vp, _u32, _vp, _u32, _vp, C.c_int]
        L.orc_dse_destroy.argtypes = [_vp]
        L.orc_slerp_simd.argtypes = [_vp, _vp, _f32, _u32]
        L.orc_xover_coeffs.argtypes = [_f32, _f32, _f32, _f32, _vp]
        L.orc_xover_process_ramp.argtypes = [_vp, _vp, _vp, _u32, _vp, _vp, _vp]
        L.orc_xover_split.argtypes = [_vp, _vp, _vp, _u32, _vp]
        L.orc_add_gain_ramp.argtypes = [_vp, _vp, _u32, _u32, _u32, _u32, _u32, _f32, _f32]
        for name in ("orc_panner_destroy", "orc_panner_is3d", "orc_panner_num_channels", "orc_panner_aperture"):
            getattr(L, name).argtypes = [_vp]
        L.orc_panner_lut.argtypes = [_vp, _vp, _vp]
        L.orc_point_gains.argtypes = [_vp, _vp, _u32, _vp]
        L.orc_layout_info.argtypes = [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp]
        L.orc_mdap.argtypes = [_vp, _u32, _vp, _u32, _vp]
        L.orc_process_virtual_sources.argtypes = [_vp, _vp, _u32, _vp]
        L.orc_reverb_create.restype = _vp
        L.orc_reverb_create.argtypes = [_f32, _f32, _f32, _f32]
        L.orc_reverb_set_rt60.argtypes = [_vp, _vp]
        L.orc_reverb_set_decay_slope.argtypes = [_vp, _vp]
        L.orc_reverb_process.argtypes = [_vp, _vp, _vp, _u32, _u32]
        L.orc_reverb_destroy.argtypes = [_vp]
        L.orc_db_to_gain.argtypes = [_vp, _u32, _vp]
        L.orc_erbus_create.restype = _vp
        L.orc_erbus_create.argtypes = [_f32, _u32]
        L.orc_erbus_destroy.argtypes = [_vp]
        L.orc_erbus_num_taps.restype = _u32
        L.orc_erbus_num_taps.argtypes = [_vp]
        L.orc_erbus_set_taps.argtypes = [_vp, _u32, _vp, _vp, _vp, _vp]
        L.orc_erbus_process.argtypes = [_vp, _vp, _u32, _vp, _u32]
        self._panners = {}

    # ---- late reverb / early reflections
    def reverb(self, fs=48000.0, splits=(0.0, 0.0, 0.0)):
        return PortReverb(self, fs, splits)

    def erbus(self, fs, num_channels):
        return PortERBus(self, fs, num_channels)

    def db_to_gain(self, db):
        d = np.ascontiguousarray(db, np.float32)
        out = np.empty_like(d)
        self.lib.orc_db_to_gain(d.ctypes.data, d.size, out.ctypes.data)
        return out

    # ---- math primitives
    def _map1(self, fn, x):
        x = _f(x); o = np.empty_like(x)
        getattr(self.lib, fn)(_vp(x.ctypes.data), _u32(x.size), _vp(o.ctypes.data)); return o

    def sincos(self, x):
        x = _f(x); s = np.empty_like(x); c = np.empty_like(x)
        self.lib.orc_sincos(_vp(x.ctypes.data), _u32(x.size), _vp(s.ctypes.data), _vp(c.ctypes.data)); return s, c

    def acos(self, x): return self._map1("orc_acos", x)
    def tan(self, x): return self._map1("orc_tan", x)
    def invsqrt_nr1(self, x): return self._map1("orc_invsqrt_nr1", x)
    def invsqrt_nr2(self, x): return self._map1("orc_invsqrt_nr2", x)

    def oct_encode_simd(self, d):
        d = _f(d); o = np.empty(d.shape[0], np.uint32)
        self.lib.orc_oct_encode_simd(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def oct_encode_scalar(self, d):
        d = _f(d); o = np.empty(d.shape[0], np.uint32)
        self.lib.orc_oct_encode_scalar(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def oct_decode_scalar(self, codes):
        c = np.ascontiguousarray(codes, np.uint32); o = np.empty((c.size, 3), np.float32)
        self.lib.orc_oct_decode_scalar(_vp(c.ctypes.data), _u32(c.size), _vp(o.ctypes.data)); return o

    def diamond_scalar(self, xy):
        xy = _f(xy); o = np.empty(xy.shape[0], np.float32)
        self.lib.orc_diamond_scalar(_vp(xy.ctypes.data), _u32(xy.shape[0]), _vp(o.ctypes.data)); return o

    def from_diamond(self, p):
        p = _f(p); o = np.empty((p.size, 2), np.float32)
        self.lib.orc_from_diamond(_vp(p.ctypes.data), _u32(p.size), _vp(o.ctypes.data)); return o

    def slerp_simd(self, v0, v1, t):
        v0 = _f(v0).copy(); v1 = _f(v1)
        self.lib.orc_slerp_simd(_vp(v0.ctypes.data), _vp(v1.ctypes.data), _f32(t), _u32(v0.shape[0])); return v0

    def quat_lookat(self, d):
        d = _f(d); o = np.empty((d.shape[0], 4), np.float32)
        self.lib.orc_quat_lookat(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def smoother_alpha(self, ticks):
        return float(self.lib.orc_smoother_alpha(_u32(ticks)))

    def add_gain_ramp(self, dest, src, dest_ch, src_ch, n_dest, n_src, frames, g0, g1):
        d = _f(dest).copy(); s = _f(src)
        self.lib.orc_add_gain_ramp(d.ctypes.data, s.ctypes.data, dest_ch, src_ch, n_dest, n_src, frames, g0, g1); return d

    # ---- panner
    def panner(self, mask, tris=None):
        return PortPanner(self, mask, tris)

    def mdap(self, target_mask, source_mask, params):
        key = target_mask
        if key not in self._panners:
            self._panners[key] = self.panner(target_mask)
        return self._panners[key].mdap(source_mask, params)

    # ---- scene front-end
    def direct_path(self, sources, listeners, listener_index=None):
        s = _f(sources).reshape(-1, 9); l = _f(listeners).reshape(-1, 9)
        out = np.empty((s.shape[0], 12), np.float32)
        for i in range(s.shape[0]):
            li = 0 if listener_index is None else int(listener_index[i])
            self.lib.orc_direct_path(_vp(s[i].ctypes.data), _vp(l[li].ctypes.data), _vp(out[i].ctypes.data))
        return out

    def angle_attenuation(self, dots, cones):
        self.lib.orc_angle_attenuation.restype = _f32; self.lib.orc_angle_attenuation.argtypes = [_f32, _f32, _f32]
        c = _f(cones).reshape(-1, 2)
        return np.array([self.lib.orc_angle_attenuation(float(d), float(a), float(b)) for d, (a, b) in zip(_f(dots), c)], np.float32)

    def attenuation_model(self, model, mn, mx, rolloff, distance):
        self.lib.orc_attenuation_model.restype = _f32; self.lib.orc_attenuation_model.argtypes = [C.c_int, _f32, _f32, _f32, _f32]
        return np.array([self.lib.orc_attenuation_model(int(model), mn, mx, rolloff, float(d)) for d in _f(distance)], np.float32)

    def attenuation_curve(self, points, distance):
        """points are sorted by distance first (AttenuationCurve::SortPoints; distinct distances assumed here)"""
        self.lib.orc_attenuation_curve.restype = _f32; self.lib.orc_attenuation_curve.argtypes = [_vp, _u32, _f32]
        p = _f(points).reshape(-1, 3)
        p = np.ascontiguousarray(p[np.argsort(p[:, 0], kind="stable")]) if p.shape[0] else p
        return np.array([self.lib.orc_attenuation_curve(p.ctypes.data, p.shape[0], float(d)) for d in _f(distance)], np.float32)

    def speed_of_sound(self, t):
        self.lib.orc_speed_of_sound.restype = _f32; self.lib.orc_speed_of_sound.argtypes = [_f32]
        return float(self.lib.orc_speed_of_sound(t))

    def air_cache(self, bands, temp_c=20.0, humidity=50.0, pressure=101.325, nyquist=22050.0):
        self.lib.orc_air_cache.argtypes = [_vp, _f32, _f32, _f32, _f32, _vp, _vp]
        b = _f(bands); out = np.empty(32, np.float32); knee = np.empty(4, np.uint32)
        self.lib.orc_air_cache(b.ctypes.data, temp_c, humidity, pressure, nyquist, out.ctypes.data, knee.ctypes.data)
        return out, knee

    def air_for_distance(self, cache, knee, distance):
        self.lib.orc_air_for_distance.argtypes = [_vp, _vp, _f32, _vp, _vp]
        d = _f(distance); db = np.empty((d.size, 4), np.float32); g = np.empty((d.size, 4), np.float32)
        for i in range(d.size):
            self.lib.orc_air_for_distance(cache.ctypes.data, knee.ctypes.data, float(d[i]), db[i].ctypes.data, g[i].ctypes.data)
        return db, g

    def sanitize_source_dir(self, dirs):
        d = _f(dirs).reshape(-1, 3); out = np.empty_like(d)
        for i in range(d.shape[0]):
            self.lib.orc_sanitize_source_dir(_vp(d[i].ctypes.data), _vp(out[i].ctypes.data))
        return out

    def spread_from_size(self, flavour, size, distance):
        self.lib.orc_spread_from_size.restype = _f32; self.lib.orc_spread_from_size.argtypes = [C.c_int, _f32, _f32]
        return np.array([self.lib.orc_spread_from_size(int(flavour), float(s), float(d)) for s, d in zip(_f(size), _f(distance))], np.float32)

    def mdap_oriented(self, target_mask, source_mask, params):
        if target_mask not in self._panners:
            self._panners[target_mask] = self.panner(target_mask)
        pn = self._panners[target_mask]
        p = _f(params).reshape(-1, 11)
        info = pn.layout_info(source_mask)
        ch = int(info["channels"]) if isinstance(info, dict) else int(info[0][0])
        out = np.zeros((p.shape[0], ch, pn.L), np.float32)
        self.lib.orc_mdap_oriented.argtypes = [_vp, _u32, _vp, _u32, _vp]
        assert self.lib.orc_mdap_oriented(pn.h, source_mask, p.ctypes.data, p.shape[0], out.ctypes.data)
        return out

    # ---- DSP
    def xover_coeffs(self, fs=48000.0, f1=176.0, f2=775.0, f3=3408.0):
        o = np.zeros(27, np.float32); self.lib.orc_xover_coeffs(fs, f1, f2, f3, o.ctypes.data); return o

    def crossover(self, fs=48000.0):
        return PortCrossover(self, fs)

    def dse(self, fs, n_src, n_out, band_gains, delay_seconds, mix_map):
        return PortDSE(self, fs, n_src, n_out, band_gains, delay_seconds, mix_map)

    def dse_batch_process(self, effects, inp, frames, bus, threads=1):
        arr = (_vp * len(effects))(*[e.h for e in effects])
        x = _f(inp)
        return float(self.lib.orc_dse_process_batch(arr, len(effects), x.ctypes.data, frames, bus.ctypes.data, threads))


class PortPanner:
    def __init__(self, port, mask, tris=None):
        self.lib = port.lib
        if tris is None and mask >= (1 << 11):
            tris = speaker_triangles(mask)
        t = None if tris is None else np.ascontiguousarray(tris, np.uint8)
        self.h = _vp(self.lib.orc_panner_create(mask, None if t is None else t.ctypes.data, 0 if t is None else t.shape[0]))
        self.ok = bool(self.h)
        if self.ok:
            self.is3d = bool(self.lib.orc_panner_is3d(self.h))
            self.L = int(self.lib.orc_panner_num_channels(self.h))
            self.aperture = float(self.lib.orc_panner_aperture(self.h))

    def lut(self):
        if self.is3d:
            g = np.zeros((65536, 3), np.float32); s = np.zeros((65536, 3), np.uint8)
            self.lib.orc_panner_lut(self.h, g.ctypes.data, s.ctypes.data); return g, s
        g = np.zeros((1024, self.L), np.float32)
        self.lib.orc_panner_lut(self.h, g.ctypes.data, None); return g, None

    def point_gains(self, dirs):
        d = _f(dirs); o = np.zeros((d.shape[0], self.L), np.float32)
        self.lib.orc_point_gains(self.h, d.ctypes.data, d.shape[0], o.ctypes.data); return o

    def layout_info(self, source_mask):
        info = (_u32 * 4)(); a, b, c = _f32(), _f32(), _f32()
        q = np.zeros((8, 4), np.float32); ids = np.zeros(8, np.uint32)
        ok = self.lib.orc_layout_info(self.h, source_mask, C.cast(info, _vp), C.cast(C.byref(a), _vp), C.cast(C.byref(b), _vp), C.cast(C.byref(c), _vp),
                                      q.ctypes.data, ids.ctypes.data)
        if not ok:
            return None
        n = int(info[0])
        return {"channels": n, "vs": int(info[1]), "rings": int(info[2]), "per_ring": int(info[3]), "phi": a.value, "offset": b.value,
                "weight": c.value, "quats": q[:n].copy(), "ids": ids[:n].copy()}

    def mdap(self, source_mask, params):
        p = _f(params)
        info = self.layout_info(source_mask)
        assert info is not None, "invalid source layout"
        o = np.zeros((p.shape[0], info["channels"], self.L), np.float32)
        assert self.lib.orc_mdap(self.h, source_mask, p.ctypes.data, p.shape[0], o.ctypes.data)
        return o

    def process_virtual_sources(self, vs):
        v = _f(vs); o = np.zeros(self.L, np.float32)
        self.lib.orc_process_virtual_sources(self.h, v.ctypes.data, v.shape[0], o.ctypes.data); return o

    def __del__(self):
        try:
            if self.h:
                self.lib.orc_panner_destroy(self.h)
        except Exception:
            pass


class PortCrossover:
    def __init__(self, port, fs):
        self.lib = port.lib
        self.coef = port.xover_coeffs(fs)
        self.state = np.zeros(24, np.float32)

    def coeffs(self): return self.coef.copy()
    def reset(self): self.state[:] = 0

    def process_ramp(self, x, g0, g1):
        x = _f(x); g0 = _f(g0); g1 = _f(g1); o = np.empty_like(x)
        self.lib.orc_xover_process_ramp(self.coef.ctypes.data, self.state.ctypes.data, x.ctypes.data, x.size, g0.ctypes.data, g1.ctypes.data, o.ctypes.data); return o

    def split(self, x):
        x = _f(x); o = np.empty((x.size, 4), np.float32)
        self.lib.orc_xover_split(self.coef.ctypes.data, self.state.ctypes.data, x.ctypes.data, x.size, o.ctypes.data); return o


class PortDSE:
    def __init__(self, port, fs, n_src, n_out, band_gains, delay_seconds, mix_map):
        self.lib = port.lib
        b = _f(band_gains); m = _f(mix_map)
        self.h = _vp(self.lib.orc_dse_create(fs, n_src, n_out, b.ctypes.data, float(delay_seconds), m.ctypes.data))

    def update(self, band_gains, delay_seconds, mix_map):
        b = _f(band_gains); m = _f(mix_map)
        self.lib.orc_dse_update(self.h, b.ctypes.data, float(delay_seconds), m.ctypes.data)

    def process(self, inp, out, frames):
        x = _f(inp)
        assert out.dtype == np.float32 and out.flags.c_contiguous
        self.lib.orc_dse_process(self.h, x.ctypes.data, out.ctypes.data, frames)
        return out

    def __del__(self):
        try:
            self.lib.orc_dse_destroy(self.h)
        except Exception:
            pass


class PortReverb:
    """restatement of ReverbBus; same interface as oracle.refpy.RefReverb"""

    def __init__(self, port, fs, splits):
        self.port = port
        self.h = port.lib.orc_reverb_create(_f32(fs), _f32(splits[0]), _f32(splits[1]), _f32(splits[2]))

    def set_rt60(self, rt60):
        a = np.ascontiguousarray(rt60, np.float32)
        self.port.lib.orc_reverb_set_rt60(self.h, a.ctypes.data)

    def set_decay_slope(self, slope):
        a = np.ascontiguousarray(slope, np.float32)
        self.port.lib.orc_reverb_set_decay_slope(self.h, a.ctypes.data)

    def process(self, inp, out, frames, n_in=1):
        assert n_in == 1
        x = np.ascontiguousarray(inp, np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
        self.port.lib.orc_reverb_process(self.h, x.ctypes.data, out.ctypes.data, out.shape[-1], frames)

    def __del__(self):
        try:
            if self.h:
                self.port.lib.orc_reverb_destroy(self.h)
                self.h = None
        except Exception:
            pass


class PortERBus:
    """restatement of ERBus; same interface as oracle.refpy.RefERBus"""

    def __init__(self, port, fs, num_channels):
        self.port, self.L = port, int(num_channels)
        self.h = port.lib.orc_erbus_create(_f32(fs), self.L)

    def set_taps(self, filter_gains, delay_seconds, ids, gains):
        i = np.ascontiguousarray(ids, np.uint32).reshape(-1)
        fg = np.ascontiguousarray(filter_gains, np.float32).reshape(-1, 4) if len(i) else np.zeros((0, 4), np.float32)
        d = np.ascontiguousarray(delay_seconds, np.float32).reshape(-1) if len(i) else np.zeros(0, np.float32)
        g = np.ascontiguousarray(gains, np.float32).reshape(-1, self.L) if len(i) else np.zeros((0, self.L), np.float32)
        self.port.lib.orc_erbus_set_taps(self.h, len(i), fg.ctypes.data, d.ctypes.data, i.ctypes.data, g.ctypes.data)

    def num_taps(self):
        return int(self.port.lib.orc_erbus_num_taps(self.h))

    def process(self, inp, out, frames, n_in=1):
        x = np.ascontiguousarray(inp, np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
        self.port.lib.orc_erbus_process(self.h, x.ctypes.data, n_in, out.ctypes.data, frames)

    def __del__(self):
        try:
            if self.h:
                self.port.lib.orc_erbus_destroy(self.h)
                self.h = None
        except Exception:
            pass
