"""TEST INFRASTRUCTURE — ctypes wrapper over oracle/_ref/libjplref_*.so (the unmodified reference, see
oracle/ref/ref_harness.cpp). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm
may import this module; the product package never does."""
import ctypes as C
import os
import numpy as np

_here = os.path.dirname(os.path.abspath(__file__))
_vp, _u32, _f32 = C.c_void_p, C.c_uint32, C.c_float


def ref_path(variant="sse2"):
    return os.path.join(_here, "_ref", f"libjplref_{variant}.so")


def available(variant="sse2"):
    return os.path.exists(ref_path(variant))


class Ref:
    def __init__(self, variant="sse2"):
        self.lib = L = C.CDLL(ref_path(variant))
        for name in ("ref_panner_create", "ref_layout_create", "ref_xover_create", "ref_delay_create", "ref_dse_create"):
            getattr(L, name).restype = _vp
        L.ref_panner_aperture.restype = _f32
        L.ref_smoother_alpha.restype = _f32
        L.ref_delay_read_int.restype = _f32
        L.ref_delay_tap_thiran.restype = _f32
        L.ref_delay_tap_linear.restype = _f32
        L.ref_dse_process_batch.restype = C.c_double
        L.ref_mdap_timed.restype = C.c_double
        L.ref_panner_create.argtypes = [_u32]
        L.ref_layout_create.argtypes = [_vp, _u32]
        L.ref_xover_create.argtypes = [_f32]
        L.ref_delay_create.argtypes = [_u32]
        L.ref_delay_read_int.argtypes = [_vp, _u32]
        L.ref_delay_tap_thiran.argtypes = [_vp, _f32]
        L.ref_delay_tap_linear.argtypes = [_vp, _f32]
        L.ref_dse_create.argtypes = [_f32, _u32, _u32, _vp, _f32, _vp]
        L.ref_dse_update.argtypes = [_vp, _vp, _f32, _vp]
        L.ref_dse_process.argtypes = [_vp, _vp, _vp, _u32]
        L.ref_slerp_simd.argtypes = [_vp, _vp, _f32, _u32]
        L.ref_smoother_run.argtypes = [_f32, _f32, _u32, _u32, _vp]
        L.ref_add_gain_ramp.argtypes = [_vp, _vp, _u32, _u32, _u32, _u32, _u32, _f32, _f32]
        L.ref_mix_add.argtypes = [_vp, _vp, _u32, _u32, _u32, _u32, _u32]
        L.ref_downmix_to_mono.argtypes = [_vp, _vp, _u32, _u32]
        for name in ("ref_reverb_create", "ref_erbus_create"):
            getattr(L, name).restype = _vp
        L.ref_reverb_create.argtypes = [_f32, _f32, _f32, _f32]
        L.ref_reverb_set_rt60.argtypes = [_vp, _vp]
        L.ref_reverb_set_decay_slope.argtypes = [_vp, _vp]
        L.ref_reverb_process.argtypes = [_vp, _vp, _u32, _vp, _u32, _u32]
        L.ref_reverb_process_batch.restype = C.c_double
        L.ref_reverb_process_batch.argtypes = [_vp, _u32, _vp, _vp, _u32, _u32, C.c_int]
        L.ref_db_to_gain.argtypes = [_vp, _vp]
        L.ref_erbus_create.argtypes = [_f32, _u32]
        L.ref_erbus_num_taps.restype = _u32
        L.ref_erbus_set_taps.argtypes = [_vp, _u32, _vp, _vp, _vp, _vp]
        L.ref_erbus_process.argtypes = [_vp, _vp, _u32, _vp, _u32]
        L.ref_erbus_process_batch.restype = C.c_double
        L.ref_erbus_process_batch.argtypes = [_vp, _u32, _vp, _vp, _u32, C.c_int]
        for name in ("ref_reverb_destroy", "ref_erbus_destroy", "ref_erbus_num_taps"):
            getattr(L, name).argtypes = [_vp]
        for name in ("ref_panner_destroy", "ref_layout_destroy", "ref_xover_destroy", "ref_delay_destroy", "ref_dse_destroy",
                     "ref_panner_ok", "ref_panner_is3d", "ref_panner_num_channels", "ref_panner_aperture", "ref_layout_ok",
                     "ref_xover_reset", "ref_delay_size"):
            getattr(L, name).argtypes = [_vp]

    # ---- late reverb / early reflections (BASELINE config 5)
    def reverb(self, fs=48000.0, splits=(0.0, 0.0, 0.0)):
        return RefReverb(self, fs, splits)

    def erbus(self, fs, num_channels):
        return RefERBus(self, fs, num_channels)

    def db_to_gain(self, db):
        d = np.ascontiguousarray(db, np.float32).reshape(-1, 4)
        out = np.empty_like(d)
        for i in range(d.shape[0]):
            self.lib.ref_db_to_gain(d[i].ctypes.data, out[i].ctypes.data)
        return out

    def reverb_process_batch(self, buses, inp, out, frames, threads=1):
        arr = (_vp * len(buses))(*[b.h for b in buses])
        return self.lib.ref_reverb_process_batch(arr, len(buses), inp.ctypes.data, out.ctypes.data, out.shape[-1], frames, int(threads))

    def erbus_process_batch(self, buses, inp, out, frames, threads=1):
        arr = (_vp * len(buses))(*[b.h for b in buses])
        return self.lib.ref_erbus_process_batch(arr, len(buses), inp.ctypes.data, out.ctypes.data, frames, int(threads))

    # ---- scene front-end; positions are [.,9] = location, up, forward
    def _p(self, a, dtype=np.float32):
        return None if a is None else np.ascontiguousarray(a, dtype)

    def direct_path(self, sources, listeners, listener_index=None):
        s = self._p(sources).reshape(-1, 9); l = self._p(listeners).reshape(-1, 9); li = self._p(listener_index, np.uint32)
        out = np.empty((s.shape[0], 12), np.float32)
        self.lib.ref_direct_path(_vp(s.ctypes.data), _vp(l.ctypes.data), _vp(None if li is None else li.ctypes.data), _u32(s.shape[0]), _vp(out.ctypes.data))
        return out

    def angle_attenuation(self, kind, x, cones, ref_point=None):
        x = self._p(x); c = self._p(cones).reshape(-1, 2); r = self._p(ref_point)
        out = np.empty(c.shape[0], np.float32)
        self.lib.ref_angle_attenuation(C.c_int(kind), _vp(x.ctypes.data), _vp(None if r is None else r.ctypes.data), _vp(c.ctypes.data), _u32(c.shape[0]), _vp(out.ctypes.data))
        return out

    def attenuation_model(self, model, min_d, max_d, rolloff, distance):
        d = self._p(distance); out = np.empty_like(d)
        self.lib.ref_attenuation_model(C.c_int(model), _f32(min_d), _f32(max_d), _f32(rolloff), _vp(d.ctypes.data), _u32(d.size), _vp(out.ctypes.data))
        return out

    def attenuation_curve(self, points, distance, sort_points=True):
        p = self._p(points).reshape(-1, 3); d = self._p(distance); out = np.empty_like(d)
        self.lib.ref_attenuation_curve(_vp(p.ctypes.data), _u32(p.shape[0]), C.c_int(1 if sort_points else 0), _vp(d.ctypes.data), _u32(d.size), _vp(out.ctypes.data))
        return out

    def default_bands(self):
        out = np.empty(4, np.float32); self.lib.ref_default_bands(_vp(out.ctypes.data)); return out

    def air_cache(self, bands=None, temp_c=20.0, humidity=50.0, pressure=101.325, nyquist=22050.0):
        b = self.default_bands() if bands is None else self._p(bands)
        out = np.empty(32, np.float32); knee = np.empty(4, np.uint32)
        self.lib.ref_air_cache(_vp(b.ctypes.data), _f32(temp_c), _f32(humidity), _f32(pressure), _f32(nyquist), _vp(out.ctypes.data), _vp(knee.ctypes.data))
        names = ("band_centers", "lower_third_centers", "center_db", "lower_third_db", "weighted_db", "decay_slope", "distance_to_knee", "linearized_loss_factor")
        d = {k: out[4 * i:4 * i + 4].copy() for i, k in enumerate(names)}
        d["has_knee"] = knee
        return d

    def air_for_distance(self, distance, bands=None, temp_c=20.0, humidity=50.0, pressure=101.325, nyquist=22050.0):
        b = self.default_bands() if bands is None else self._p(bands)
        d = self._p(distance); db = np.empty((d.size, 4), np.float32); g = np.empty((d.size, 4), np.float32)
        self.lib.ref_air_for_distance(_vp(b.ctypes.data), _f32(temp_c), _f32(humidity), _f32(pressure), _f32(nyquist), _vp(d.ctypes.data), _u32(d.size), _vp(db.ctypes.data), _vp(g.ctypes.data))
        return db, g

    def air_for_frequency(self, f, temp_c=20.0, humidity=50.0, pressure=101.325):
        self.lib.ref_air_for_frequency.restype = _f32
        return float(self.lib.ref_air_for_frequency(_f32(f), _f32(temp_c), _f32(humidity), _f32(pressure)))

    def speed_of_sound(self, temp_c=None):
        self.lib.ref_speed_of_sound.restype = _f32; self.lib.ref_speed_of_sound_reference.restype = _f32
        return float(self.lib.ref_speed_of_sound_reference(C.c_int(0))) if temp_c is None else float(self.lib.ref_speed_of_sound(_f32(temp_c)))

    def inv_speed_of_sound(self):
        self.lib.ref_speed_of_sound_reference.restype = _f32
        return float(self.lib.ref_speed_of_sound_reference(C.c_int(1)))

    def sanitize_source_dir(self, dirs):
        d = self._p(dirs).reshape(-1, 3); out = np.empty_like(d)
        self.lib.ref_sanitize_source_dir(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(out.ctypes.data)); return out

    def spread_from_size(self, flavour, size, distance):
        s = self._p(size); d = self._p(distance); out = np.empty_like(s)
        self.lib.ref_spread_from_size(C.c_int(flavour), _vp(s.ctypes.data), _vp(d.ctypes.data), _u32(s.size), _vp(out.ctypes.data)); return out

    def frontend_timed(self, sources, listener, cones, sizes, model, focus=1.0, threads=1):
        """-> (seconds, dict of outputs): the per-source body of AdvanceSimulation + producers on `threads` std::threads"""
        s = self._p(sources).reshape(-1, 9); n = s.shape[0]
        l = self._p(listener).reshape(-1, 9); c = self._p(cones).reshape(-1, 2); z = self._p(sizes); m = self._p(model)
        o = {"direct": np.empty((n, 4), np.float32), "pan_params": np.empty((n, 5), np.float32), "band_gains": np.empty((n, 4), np.float32),
             "delay_seconds": np.empty(n, np.float32), "distance_attenuation": np.empty(n, np.float32)}
        self.lib.ref_frontend_timed.restype = C.c_double
        t = self.lib.ref_frontend_timed(_vp(s.ctypes.data), _vp(l.ctypes.data), _u32(n), _vp(c.ctypes.data), _vp(z.ctypes.data), _vp(m.ctypes.data),
                                        _vp(o["direct"].ctypes.data), _vp(o["pan_params"].ctypes.data), _vp(o["band_gains"].ctypes.data),
                                        _vp(o["delay_seconds"].ctypes.data), _vp(o["distance_attenuation"].ctypes.data), _f32(focus), C.c_int(threads))
        return float(t), o

    # ---- panners
    def panner(self, mask):
        return RefPanner(self, mask)

    def triangulate(self, mask):
        v = np.zeros((32, 3), np.float32); t = np.zeros((64, 3), np.uint8); nv = C.c_int(); nt = C.c_int()
        ok = self.lib.ref_triangulate(_u32(mask), _vp(v.ctypes.data), C.byref(nv), _vp(t.ctypes.data), C.byref(nt))
        if not ok:
            return None, None
        return v[:nv.value].copy(), t[:nt.value].copy()

    # ---- math primitives (n multiple of 4 for the SIMD ones)
    def _f(self, a):
        return np.ascontiguousarray(a, np.float32)

    def sincos(self, x):
        x = self._f(x); s = np.empty_like(x); c = np.empty_like(x)
        self.lib.ref_sincos(_vp(x.ctypes.data), _u32(x.size), _vp(s.ctypes.data), _vp(c.ctypes.data)); return s, c

    def acos(self, x):
        x = self._f(x); o = np.empty_like(x)
        self.lib.ref_acos(_vp(x.ctypes.data), _u32(x.size), _vp(o.ctypes.data)); return o

    def tan(self, x):
        x = self._f(x); o = np.empty_like(x)
        self.lib.ref_tan(_vp(x.ctypes.data), _u32(x.size), _vp(o.ctypes.data)); return o

    def invsqrt(self, x):
        x = self._f(x); a = np.empty_like(x); b = np.empty_like(x); c = np.empty_like(x)
        self.lib.ref_invsqrt(_vp(x.ctypes.data), _u32(x.size), _vp(a.ctypes.data), _vp(b.ctypes.data), _vp(c.ctypes.data)); return a, b, c

    def oct_encode_simd(self, d):
        d = self._f(d); o = np.empty(d.shape[0], np.uint32)
        self.lib.ref_oct_encode_simd(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def oct_encode_scalar(self, d):
        d = self._f(d); o = np.empty(d.shape[0], np.uint32)
        self.lib.ref_oct_encode_scalar(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def oct_decode_scalar(self, codes):
        c = np.ascontiguousarray(codes, np.uint32); o = np.empty((c.size, 3), np.float32)
        self.lib.ref_oct_decode_scalar(_vp(c.ctypes.data), _u32(c.size), _vp(o.ctypes.data)); return o

    def diamond_scalar(self, xy):
        xy = self._f(xy); o = np.empty(xy.shape[0], np.float32)
        self.lib.ref_diamond_scalar(_vp(xy.ctypes.data), _u32(xy.shape[0]), _vp(o.ctypes.data)); return o

    def from_diamond(self, p):
        p = self._f(p); o = np.empty((p.size, 2), np.float32)
        self.lib.ref_from_diamond(_vp(p.ctypes.data), _u32(p.size), _vp(o.ctypes.data)); return o

    def slerp_simd(self, v0, v1, t):
        v0 = self._f(v0).copy(); v1 = self._f(v1)
        self.lib.ref_slerp_simd(_vp(v0.ctypes.data), _vp(v1.ctypes.data), _f32(t), _u32(v0.shape[0])); return v0

    def quat_lookat(self, d):
        d = self._f(d); o = np.empty((d.shape[0], 4), np.float32)
        self.lib.ref_quat_lookat(_vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def smoother_alpha(self, ticks):
        return float(self.lib.ref_smoother_alpha(_u32(ticks)))

    def smoother_run(self, cur, tgt, ticks, n):
        o = np.empty(n, np.float32)
        self.lib.ref_smoother_run(cur, tgt, ticks, n, o.ctypes.data); return o

    def add_gain_ramp(self, dest, src, dest_ch, src_ch, n_dest, n_src, frames, g0, g1):
        d = self._f(dest).copy(); s = self._f(src)
        self.lib.ref_add_gain_ramp(d.ctypes.data, s.ctypes.data, dest_ch, src_ch, n_dest, n_src, frames, g0, g1); return d

    def mix_add(self, dest, src, dest_ch, src_ch, n_dest, n_src, frames):
        """Add (ChannelMixing.h:32-48); n_src == 0 calls the mono-source overload"""
        d = self._f(dest).copy(); s = self._f(src)
        self.lib.ref_mix_add(d.ctypes.data, s.ctypes.data, dest_ch, src_ch, n_dest, n_src, frames); return d

    def downmix_to_mono(self, dest, src, num_channels):
        d = self._f(dest).copy(); s = self._f(src)
        self.lib.ref_downmix_to_mono(d.ctypes.data, s.ctypes.data, num_channels, s.size); return d

    def crossover(self, fs=48000.0):
        return RefCrossover(self, fs)

    def delay(self, min_size):
        return RefDelay(self, min_size)

    def lut_variant(self, kind, target_mask, source_mask, params, dirs):
        """VBAPanner3D<VBAPStandardTraits, ELUTSize kind> (kind 0 KB_983, 1 KB_851, 2 KB_786, 3 KB_655, 4 KB_65), one shot: ProcessVBAPData for
        params [n][5] and GetSpeakerGains for dirs [m][3] -> (mix [n][stride], point [m][L]); None when the panner does not initialise"""
        prm = self._f(params).reshape(-1, 5); d = self._f(dirs).reshape(-1, 3)
        L = bin(int(target_mask)).count("1")
        mix = np.zeros((prm.shape[0], 8 * L), np.float32); pt = np.zeros((d.shape[0], L), np.float32)
        self.lib.ref_lut_variant.restype = C.c_int
        self.lib.ref_lut_variant.argtypes = [C.c_int, _u32, _u32, _vp, _u32, _vp, _vp, _u32, _vp]
        big = np.zeros((prm.shape[0] * 8 * L,), np.float32)
        stride = int(self.lib.ref_lut_variant(int(kind), int(target_mask), int(source_mask), prm.ctypes.data, prm.shape[0], big.ctypes.data,
                                              d.ctypes.data, d.shape[0], pt.ctypes.data))
        if stride == 0:
            return None
        return big[: prm.shape[0] * stride].reshape(prm.shape[0], stride).copy(), pt

    def dse(self, fs, n_src, n_out, band_gains, delay_seconds, mix_map):
        return RefDSE(self, fs, n_src, n_out, band_gains, delay_seconds, mix_map)

    def dse_handles(self, effects):
        """ctypes handle array of a list of RefDSE, built once (timed loops pass it instead of the list)"""
        return (_vp * len(effects))(*[e.h for e in effects])

    def dse_batch_process(self, effects, inp, frames, bus, threads=1):
        """effects: list of RefDSE or dse_handles(list); inp [N][frames*C]; bus [frames*L] accumulated. Returns seconds.
        threads > 1: sources partitioned over a persistent worker pool, per-thread buses summed in thread order."""
        arr = effects if isinstance(effects, C.Array) else self.dse_handles(effects)
        x = self._f(inp)
        return float(self.lib.ref_dse_process_batch(arr, _u32(len(arr)), _vp(x.ctypes.data), _u32(frames), _vp(bus.ctypes.data), C.c_int(threads)))

    def dse_batch_update(self, effects, band, delay, mix, threads=1):
        arr = effects if isinstance(effects, C.Array) else self.dse_handles(effects)
        b = self._f(band); d = self._f(delay); m = self._f(mix)
        if threads > 1 and hasattr(self.lib, "ref_dse_update_batch_mt"):
            self.lib.ref_dse_update_batch_mt(arr, _u32(len(arr)), _vp(b.ctypes.data), _vp(d.ctypes.data), _vp(m.ctypes.data), C.c_int(threads))
        else:
            self.lib.ref_dse_update_batch(arr, _u32(len(arr)), _vp(b.ctypes.data), _vp(d.ctypes.data), _vp(m.ctypes.data))

    def build_info(self):
        if not hasattr(self.lib, "ref_build_info"):
            return "unknown"
        self.lib.ref_build_info.restype = C.c_char_p
        return self.lib.ref_build_info().decode()


class RefPanner:
    def __init__(self, ref, mask):
        self.ref, self.lib = ref, ref.lib
        self.h = _vp(self.lib.ref_panner_create(mask))
        self.ok = bool(self.lib.ref_panner_ok(self.h))
        self.is3d = bool(self.lib.ref_panner_is3d(self.h))
        self.L = int(self.lib.ref_panner_num_channels(self.h)) if self.ok else 0
        self.aperture = float(self.lib.ref_panner_aperture(self.h)) if self.ok else 0.0

    def lut(self):
        if self.is3d:
            g = np.zeros((65536, 3), np.float32); s = np.zeros((65536, 3), np.uint8)
            self.lib.ref_panner_lut3d(self.h, _vp(g.ctypes.data), _vp(s.ctypes.data)); return g, s
        g = np.zeros((1024, self.L), np.float32)
        self.lib.ref_panner_lut2d(self.h, _vp(g.ctypes.data)); return g, None

    def point_gains(self, dirs):
        d = np.ascontiguousarray(dirs, np.float32); o = np.zeros((d.shape[0], self.L), np.float32)
        self.lib.ref_point_gains(self.h, _vp(d.ctypes.data), _u32(d.shape[0]), _vp(o.ctypes.data)); return o

    def layout(self, source_mask):
        return RefLayout(self, source_mask)

    def mdap(self, layout, params, threads=0):
        p = np.ascontiguousarray(params, np.float32); o = np.zeros((p.shape[0], layout.num_channels, self.L), np.float32)
        if threads:
            self.last_seconds = float(self.lib.ref_mdap_timed(self.h, layout.h, _vp(p.ctypes.data), _u32(p.shape[0]), _vp(o.ctypes.data), C.c_int(threads)))
        else:
            self.lib.ref_mdap(self.h, layout.h, _vp(p.ctypes.data), _u32(p.shape[0]), _vp(o.ctypes.data))
        return o

    def mdap_oriented(self, layout, params):
        """params [n,11] = dir, focus, spread, up, forward (3D panners)"""
        p = np.ascontiguousarray(params, np.float32).reshape(-1, 11)
        out = np.zeros((p.shape[0], layout.num_channels, self.L), np.float32)
        self.lib.ref_mdap_oriented(self.h, layout.h, _vp(p.ctypes.data), _u32(p.shape[0]), _vp(out.ctypes.data))
        return out

    def process_virtual_sources(self, vs):
        v = np.ascontiguousarray(vs, np.float32); o = np.zeros(self.L, np.float32)
        self.lib.ref_process_virtual_sources(self.h, _vp(v.ctypes.data), _u32(v.shape[0]), _vp(o.ctypes.data)); return o

    def __del__(self):
        try:
            self.lib.ref_panner_destroy(self.h)
        except Exception:
            pass


class RefLayout:
    def __init__(self, panner, source_mask):
        self.panner, self.lib = panner, panner.lib
        self.h = _vp(self.lib.ref_layout_create(panner.h, source_mask))
        self.ok = bool(self.lib.ref_layout_ok(self.h))
        if self.ok:
            info = (_u32 * 4)(); a, b, c = _f32(), _f32(), _f32()
            self.lib.ref_layout_info(self.h, info, C.byref(a), C.byref(b), C.byref(c))
            self.num_channels, self.num_virtual_sources, self.num_rings, self.samples_per_ring = [int(v) for v in info]
            self.phi_term, self.offset_term, self.vs_weight = a.value, b.value, c.value

    def groups(self):
        q = np.zeros((self.num_channels, 4), np.float32); ids = np.zeros(self.num_channels, np.uint32)
        self.lib.ref_layout_groups(self.h, _vp(q.ctypes.data), _vp(ids.ctypes.data)); return q, ids

    def azimuth(self):
        s = np.zeros(256, np.float32); c = np.zeros(256, np.float32)
        n = self.lib.ref_layout_azimuth(self.h, _vp(s.ctypes.data), _vp(c.ctypes.data)); return s[:4 * n].copy(), c[:4 * n].copy()

    def __del__(self):
        try:
            self.lib.ref_layout_destroy(self.h)
        except Exception:
            pass


class RefCrossover:
    def __init__(self, ref, fs):
        self.lib = ref.lib
        self.h = _vp(self.lib.ref_xover_create(fs))

    def coeffs(self):
        o = np.zeros(27, np.float32); self.lib.ref_xover_coeffs(self.h, _vp(o.ctypes.data)); return o

    def reset(self):
        self.lib.ref_xover_reset(self.h)

    def process_ramp(self, x, g0, g1):
        x = np.ascontiguousarray(x, np.float32); g0 = np.ascontiguousarray(g0, np.float32); g1 = np.ascontiguousarray(g1, np.float32); o = np.empty_like(x)
        self.lib.ref_xover_process_ramp(self.h, _vp(x.ctypes.data), _u32(x.size), _vp(g0.ctypes.data), _vp(g1.ctypes.data), _vp(o.ctypes.data)); return o

    def process_gains(self, x, g):
        x = np.ascontiguousarray(x, np.float32); g = np.ascontiguousarray(g, np.float32); o = np.empty_like(x)
        self.lib.ref_xover_process_gains(self.h, _vp(x.ctypes.data), _u32(x.size), _vp(g.ctypes.data), _vp(o.ctypes.data)); return o

    def split(self, x):
        x = np.ascontiguousarray(x, np.float32); o = np.empty((x.size, 4), np.float32)
        self.lib.ref_xover_split(self.h, _vp(x.ctypes.data), _u32(x.size), _vp(o.ctypes.data)); return o

    def __del__(self):
        try:
            self.lib.ref_xover_destroy(self.h)
        except Exception:
            pass


class RefDelay:
    def __init__(self, ref, min_size):
        self.lib = ref.lib
        self.h = _vp(self.lib.ref_delay_create(min_size))

    def size(self):
        return int(self.lib.ref_delay_size(self.h))

    def push(self, x):
        x = np.ascontiguousarray(x, np.float32); self.lib.ref_delay_push(self.h, _vp(x.ctypes.data), _u32(x.size))

    def read_int(self, d):
        return float(self.lib.ref_delay_read_int(self.h, int(d)))

    def tap_thiran(self, d):
        return float(self.lib.ref_delay_tap_thiran(self.h, float(d)))

    def tap_linear(self, d):
        return float(self.lib.ref_delay_tap_linear(self.h, float(d)))

    def tap_xfade(self, d, alpha):
        self.lib.ref_delay_tap_xfade.restype = _f32; self.lib.ref_delay_tap_xfade.argtypes = [_vp, _f32, _f32]
        return float(self.lib.ref_delay_tap_xfade(self.h, float(d), float(alpha)))

    def __del__(self):
        try:
            self.lib.ref_delay_destroy(self.h)
        except Exception:
            pass


class RefDSE:
    def __init__(self, ref, fs, n_src, n_out, band_gains, delay_seconds, mix_map):
        self.lib = ref.lib
        self.n_src, self.n_out = n_src, n_out
        b = np.ascontiguousarray(band_gains, np.float32); m = np.ascontiguousarray(mix_map, np.float32)
        self.h = _vp(self.lib.ref_dse_create(fs, n_src, n_out, b.ctypes.data, float(delay_seconds), m.ctypes.data))

    def update(self, band_gains, delay_seconds, mix_map):
        b = np.ascontiguousarray(band_gains, np.float32); m = np.ascontiguousarray(mix_map, np.float32)
        self.lib.ref_dse_update(self.h, b.ctypes.data, float(delay_seconds), m.ctypes.data)

    def process(self, inp, out, frames):
        x = np.ascontiguousarray(inp, np.float32)
        assert out.dtype == np.float32 and out.flags.c_contiguous
        self.lib.ref_dse_process(self.h, x.ctypes.data, out.ctypes.data, frames)
        return out

    def __del__(self):
        try:
            self.lib.ref_dse_destroy(self.h)
        except Exception:
            pass


class RefReverb:
    """ReverbBus (inc/Auralization/LateReverb.h:84-108)"""

    def __init__(self, ref, fs, splits):
        self.ref = ref
        self.h = ref.lib.ref_reverb_create(_f32(fs), _f32(splits[0]), _f32(splits[1]), _f32(splits[2]))

    def set_rt60(self, rt60):
        a = np.ascontiguousarray(rt60, np.float32)
        self.ref.lib.ref_reverb_set_rt60(self.h, a.ctypes.data)

    def set_decay_slope(self, slope):
        a = np.ascontiguousarray(slope, np.float32)
        self.ref.lib.ref_reverb_set_decay_slope(self.h, a.ctypes.data)

    def process(self, inp, out, frames, n_in=1):
        """inp [frames][n_in]; out [frames][n_out] accumulated into"""
        x = np.ascontiguousarray(inp, np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
        self.ref.lib.ref_reverb_process(self.h, x.ctypes.data, n_in, out.ctypes.data, out.shape[-1], frames)

    def __del__(self):
        try:
            if self.h:
                self.ref.lib.ref_reverb_destroy(self.h)
                self.h = None
        except Exception:
            pass


class RefERBus:
    """ERBus (inc/Auralization/EarlyReflectionsBus.h:60-88)"""

    def __init__(self, ref, fs, num_channels):
        self.ref, self.L = ref, int(num_channels)
        self.h = ref.lib.ref_erbus_create(_f32(fs), self.L)

    def set_taps(self, filter_gains, delay_seconds, ids, gains):
        fg = np.ascontiguousarray(filter_gains, np.float32).reshape(-1, 4)
        d = np.ascontiguousarray(delay_seconds, np.float32).reshape(-1)
        i = np.ascontiguousarray(ids, np.uint32).reshape(-1)
        g = np.ascontiguousarray(gains, np.float32).reshape(-1, self.L) if len(i) else np.zeros((0, self.L), np.float32)
        self.ref.lib.ref_erbus_set_taps(self.h, len(i), fg.ctypes.data, d.ctypes.data, i.ctypes.data, g.ctypes.data)

    def num_taps(self):
        return int(self.ref.lib.ref_erbus_num_taps(self.h))

    def process(self, inp, out, frames, n_in=1):
        x = np.ascontiguousarray(inp, np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
        self.ref.lib.ref_erbus_process(self.h, x.ctypes.data, n_in, out.ctypes.data, frames)

    def __del__(self):
        try:
            if self.h:
                self.ref.lib.ref_erbus_destroy(self.h)
                self.h = None
        except Exception:
            pass
