"""ctypes binding of include/jplg.h plus a thin object layer mirroring the reference's component API."""
import ctypes as C
import os
import numpy as np

_here = os.path.dirname(os.path.abspath(__file__))
lib_path = os.environ.get("JPLG_LIB_PATH") or os.path.join(_here, "libjplg.so")   # JPLG_LIB_PATH: experiment builds (scratch/), A/B runs only
if not os.path.exists(lib_path):
    raise ImportError(
        f"{lib_path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no CPU/PyTorch fallback for this package)")
lib = C.CDLL(lib_path)

MEM_HOST, MEM_DEVICE, MEM_HOST_ASYNC = 0, 1, 2
MODE_PARITY, MODE_THROUGHPUT = 0, 1
STATE_FLOATS = 64
_vp, _u32, _i32, _f32, _u64 = C.c_void_p, C.c_uint32, C.c_int32, C.c_float, C.c_uint64
_fp = C.c_void_p  # float* as raw address (numpy .ctypes.data or a device pointer)


class RendererConfig(C.Structure):
    _fields_ = [("sample_rate", _f32), ("num_sources", _u32), ("source_channels", _u32), ("output_channels", _u32),
                ("ring_size", _u32), ("max_block_frames", _u32), ("mode", _i32),
                ("split_f1", _f32), ("split_f2", _f32), ("split_f3", _f32)]


def _sig(name, res, args):
    f = getattr(lib, name)
    f.restype = res
    f.argtypes = args


_TRACE = C.CFUNCTYPE(None, C.c_char_p)
_sig("jplg_set_trace_callback", None, [_TRACE])
_sig("jplg_version", C.c_char_p, [])
_sig("jplg_device_count", _i32, [])
_sig("jplg_context_create", _i32, [_i32, C.POINTER(_vp)])
_sig("jplg_context_create_ex", _i32, [_i32, _u32, C.POINTER(_vp)])
_sig("jplg_context_destroy", None, [_vp])
_sig("jplg_last_error", C.c_char_p, [_vp])
_sig("jplg_context_stream", _vp, [_vp])
_sig("jplg_context_synchronize", _i32, [_vp])
_sig("jplg_context_rsqrt_model_verified", _i32, [_vp])
_sig("jplg_context_launch_count", _u64, [_vp])
_sig("jplg_context_set_kernel_timing", None, [_vp, _i32])
_sig("jplg_context_reset_kernel_timing", None, [_vp])
_sig("jplg_context_kernel_time", _i32, [_vp, _i32, C.POINTER(C.c_double), C.POINTER(_u64)])
_sig("jplg_context_probe_fp32", _i32, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)])
_sig("jplg_channel_map_num_channels", _u32, [_u32])
_sig("jplg_channel_map_has_lfe", _i32, [_u32])
_sig("jplg_channel_map_has_top_channels", _i32, [_u32])
_sig("jplg_channel_map_channel_index", _u32, [_u32, _u32])
_sig("jplg_channel_map_from_num_channels", _u32, [_u32])
_sig("jplg_panner_create", _i32, [_vp, _u32, C.POINTER(_vp)])
_sig("jplg_panner_create_ex", _i32, [_vp, _u32, _i32, C.POINTER(_vp)])
_sig("jplg_panner_destroy", None, [_vp])
_sig("jplg_panner_is_3d", _i32, [_vp])
_sig("jplg_panner_num_channels", _u32, [_vp])
_sig("jplg_panner_channel_mask", _u32, [_vp])
_sig("jplg_panner_shortest_aperture", _f32, [_vp])
_sig("jplg_panner_triangulation_is_tabulated", _i32, [_vp])
_sig("jplg_panner_triangles", _u32, [_vp, _vp])
_sig("jplg_panner_get_lut", _i32, [_vp, _fp, _vp])
_sig("jplg_panner_set_lut", _i32, [_vp, _fp, _vp])
_sig("jplg_panner_point_gains", _i32, [_vp, _fp, _u32, _fp, _i32])
_sig("jplg_source_layout_create", _i32, [_vp, _u32, C.POINTER(_vp)])
_sig("jplg_source_layout_destroy", None, [_vp])
_sig("jplg_source_layout_num_channels", _u32, [_vp])
_sig("jplg_source_layout_num_virtual_sources", _u32, [_vp])
_sig("jplg_source_layout_info", None, [_vp, C.POINTER(_u32), C.POINTER(_f32), C.POINTER(_f32), C.POINTER(_f32)])
_sig("jplg_source_layout_groups", None, [_vp, _fp, _vp])
_sig("jplg_panner_mdap", _i32, [_vp, _vp, _fp, _u32, _fp, _i32])
_sig("jplg_panner_mdap_oriented", _i32, [_vp, _vp, _fp, _u32, _fp, _i32])
_sig("jplg_panner_process_virtual_sources", _i32, [_vp, _fp, _u32, _u32, _fp, _i32])
_sig("jplg_normalize_weights", None, [_fp, _u32])
_sig("jplg_spread_from_source_size", _f32, [_f32, _f32])
_sig("jplg_renderer_create", _i32, [_vp, C.POINTER(RendererConfig), _fp, _fp, _fp, C.POINTER(_vp)])
_sig("jplg_renderer_destroy", None, [_vp])
_sig("jplg_renderer_update", _i32, [_vp, _fp, _fp, _fp, _i32])
_sig("jplg_renderer_process_block", _i32, [_vp, _fp, _fp, _u32, _i32])
_sig("jplg_renderer_process_block_debug", _i32, [_vp, _fp, _fp, _fp, _u32, _i32])
_sig("jplg_renderer_wait_block", _i32, [_vp])
_sig("jplg_renderer_update_panned", _i32, [_vp, _vp, _vp, _vp, _vp, _vp, _i32])
_sig("jplg_renderer_push_ahead", _i32, [_vp, _fp, _u32])
_sig("jplg_renderer_wait_render_started", _i32, [_vp, _vp])
_sig("jplg_renderer_device_bus", _vp, [_vp])
_sig("jplg_renderer_set_time_parallel", _i32, [_vp, _i32])
_sig("jplg_renderer_set_max_delay_seconds", _i32, [_vp, _f32])
_sig("jplg_renderer_time_parallel_blocks", _u64, [_vp])
_sig("jplg_renderer_ring_size", _u32, [_vp])
_sig("jplg_renderer_state_export", _i32, [_vp, _fp, _fp, C.POINTER(_u64)])
_sig("jplg_renderer_state_import", _i32, [_vp, _fp, _fp, _u64])
_sig("jplg_crossover_bank_create", _i32, [_vp, _u32, _f32, _f32, _f32, _f32, _i32, C.POINTER(_vp)])
_sig("jplg_crossover_bank_destroy", None, [_vp])
_sig("jplg_crossover_bank_reset", _i32, [_vp])
_sig("jplg_crossover_bank_coeffs", None, [_vp, _fp])
_sig("jplg_crossover_bank_process_ramp", _i32, [_vp, _fp, _fp, _fp, _fp, _u32, _i32])
_sig("jplg_crossover_bank_split", _i32, [_vp, _fp, _fp, _u32, _i32])
_sig("jplg_delay_bank_create", _i32, [_vp, _u32, _u32, C.POINTER(_vp)])
_sig("jplg_delay_bank_destroy", None, [_vp])
_sig("jplg_delay_bank_size", _u32, [_vp])
_sig("jplg_delay_bank_clear", _i32, [_vp])
_sig("jplg_delay_bank_push", _i32, [_vp, _fp, _u32, _i32])
_sig("jplg_delay_bank_read", _i32, [_vp, _fp, _u32, _i32, _fp, _i32])
_sig("jplg_delay_bank_read_xfade", _i32, [_vp, _fp, _fp, _u32, _fp, _i32])
_sig("jplg_mix_add_and_apply_gain_ramp", _i32, [_vp, _fp, _fp, _u32, _u32, _u32, _u32, _u32, _u32, _fp, _fp, _i32])
_sig("jplg_mix_apply_gain", _i32, [_vp, _fp, _u32, _u32, _fp, _i32])
_sig("jplg_mix_add", _i32, [_vp, _fp, _fp, _u32, _u32, _u32, _u32, _u32, _u32, _i32])
_sig("jplg_mix_downmix_to_mono", _i32, [_vp, _fp, _fp, _u32, _u32, _u32, _i32])

DECLARED_SYMBOLS = [n for n in dir(lib) if n.startswith("jplg_")]


class JplgError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"jplg status {status}: {message}")
        self.status = status


def version():
    return lib.jplg_version().decode()


def device_count():
    return int(lib.jplg_device_count())


# ------------------------------------------------------------------------------------------------
from ._masks import EChannel, ChannelMask  # noqa: E402,F401


class ChannelMap:
    """inc/ChannelMap.h:153-249"""

    def __init__(self, mask=0):
        self.mask = int(mask)

    @staticmethod
    def FromChannelMask(mask):
        return ChannelMap(mask)

    @staticmethod
    def FromNumChannels(n):
        return ChannelMap(lib.jplg_channel_map_from_num_channels(n))

    def IsValid(self):
        return self.mask != 0

    def Has(self, ch):
        return (self.mask & ch) == ch

    def HasLFE(self):
        return bool(lib.jplg_channel_map_has_lfe(self.mask))

    def HasTopChannels(self):
        return bool(lib.jplg_channel_map_has_top_channels(self.mask))

    def GetNumChannels(self):
        return int(lib.jplg_channel_map_num_channels(self.mask))

    def GetChannelIndex(self, ch):
        return int(lib.jplg_channel_map_channel_index(self.mask, ch))

    def GetChannelMask(self):
        return self.mask

    def __eq__(self, o):
        return isinstance(o, ChannelMap) and o.mask == self.mask


def _mask(m):
    return m.mask if isinstance(m, ChannelMap) else int(m)


# ------------------------------------------------------------------------------------------------
def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _arg(x, dtype=np.float32):
    """-> (address, mem, keepalive). numpy arrays are host memory, CUDA torch tensors device memory."""
    if x is None:
        return None, None, None
    if _is_torch(x):
        assert x.is_contiguous()
        return x.data_ptr(), (MEM_DEVICE if x.is_cuda else MEM_HOST), x
    a = np.ascontiguousarray(x, dtype=dtype)
    return a.ctypes.data, MEM_HOST, a


def _ctx_alive(obj):
    ctx = getattr(obj, "ctx", None)
    if ctx is None:
        owner = getattr(obj, "panner", None)
        ctx = getattr(owner, "ctx", None)
    return ctx is None or ctx.h is not None


class Context:
    """Device + stream + constant tables. device=-1 gives a host-only context: the builders (LUTs, layouts,
    filter coefficients) work, every data-path call raises JplgError(JPLG_ERR_NO_DEVICE)."""

    def __init__(self, device=0, background=False):
        """background=True: the context's stream has a lower priority than a rendering context's (jplg_context_create_ex)."""
        h = _vp()
        st = lib.jplg_context_create_ex(int(device), 1 if background else 0, C.byref(h))
        if st != 0:
            raise JplgError(st, "jplg_context_create failed (no CUDA device? this library has no CPU fallback)")
        self.h = h
        self.device = device

    def check(self, st):
        if st != 0:
            raise JplgError(st, lib.jplg_last_error(self.h).decode())

    def synchronize(self):
        self.check(lib.jplg_context_synchronize(self.h))

    @property
    def stream(self):
        return lib.jplg_context_stream(self.h)

    def torch_stream(self):
        import torch
        return torch.cuda.ExternalStream(self.stream, device=torch.device("cuda", self.device))

    @property
    def rsqrt_model_verified(self):
        return bool(lib.jplg_context_rsqrt_model_verified(self.h))

    @property
    def launch_count(self):
        return int(lib.jplg_context_launch_count(self.h))

    KERNELS = {"push": 0, "render": 1, "bus_reduce": 2, "mdap": 3, "point_gains": 4, "update": 5, "other": 6, "fdn": 7, "er_push": 8, "frontend": 9, "exchange": 10}

    def set_kernel_timing(self, enabled, only=None):
        """only: iterable of kernel names (see kernel_time) to bracket with events; default every kernel"""
        if enabled and only:
            mask = 0
            for name in only:
                mask |= 1 << self.KERNELS[name]
            lib.jplg_context_set_kernel_timing(self.h, mask << 1)
        else:
            lib.jplg_context_set_kernel_timing(self.h, 1 if enabled else 0)

    def reset_kernel_timing(self):
        lib.jplg_context_reset_kernel_timing(self.h)

    def probe_fp32_lane_rate(self):
        """FP32 ceiling of this device for unfused packed multiplies / adds (jplg_context_probe_fp32)"""
        rate, ms = C.c_double(), C.c_double()
        self.check(lib.jplg_context_probe_fp32(self.h, C.byref(rate), C.byref(ms)))
        return {"lane_ops_per_s": float(rate.value), "probe_ms": float(ms.value)}

    def kernel_time(self, name):
        """-> (total milliseconds, launches) of one of KERNELS since the last reset (CUDA events on the context stream)"""
        ms, cnt = C.c_double(), _u64()
        self.check(lib.jplg_context_kernel_time(self.h, self.KERNELS[name], C.byref(ms), C.byref(cnt)))
        return ms.value, int(cnt.value)

    def close(self):
        if self.h:
            lib.jplg_context_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SourceLayout:
    """VBAPLayoutBase / SourceLayout (inc/Panning/PannerBase.h:268-289)"""

    def __init__(self, panner, source_map):
        h = _vp()
        panner.ctx.check(lib.jplg_source_layout_create(panner.h, _mask(source_map), C.byref(h)))
        self.h = h
        self.panner = panner
        info = (_u32 * 4)()
        a, b, c = _f32(), _f32(), _f32()
        lib.jplg_source_layout_info(h, info, C.byref(a), C.byref(b), C.byref(c))
        self.num_channels, self.num_virtual_sources, self.num_rings, self.samples_per_ring = [int(v) for v in info]
        self.phi_term, self.offset_term, self.vs_weight = a.value, b.value, c.value

    def groups(self):
        q = np.zeros((self.num_channels, 4), np.float32)
        ids = np.zeros(self.num_channels, np.uint32)
        lib.jplg_source_layout_groups(self.h, q.ctypes.data, ids.ctypes.data)
        return q, ids

    def GetNumVirtualSources(self):
        return self.num_virtual_sources

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_source_layout_destroy(self.h)
                self.h = None
        except Exception:
            pass


LUT_KB_983, LUT_KB_851, LUT_KB_786, LUT_KB_655, LUT_KB_65 = 0, 1, 2, 3, 4   # ELUTSize (inc/Panning/VBAPLUT3D.h:93-104)


class VBAPanner:
    """VBAPanner2D / VBAPanner3D (inc/Panning/PannerBase.h:315-389), batched."""

    def __init__(self, ctx):
        self.ctx = ctx
        self.h = None

    def Initialize(self, channel_map, lut_variant=0):
        """lut_variant: the reference's ELUTSize template parameter (LUT_KB_983 default, LUT_KB_851, LUT_KB_786; jplg_panner_create_ex)"""
        h = _vp()
        st = lib.jplg_panner_create_ex(self.ctx.h, _mask(channel_map), int(lut_variant), C.byref(h))
        if st != 0:
            self.last_error = lib.jplg_last_error(self.ctx.h).decode()
            return False  # like the reference: false + trace
        self.h = h
        return True

    def IsInitialized(self):
        return self.h is not None

    def GetNumChannels(self):
        return int(lib.jplg_panner_num_channels(self.h))

    def GetChannelMap(self):
        return ChannelMap(lib.jplg_panner_channel_mask(self.h))

    def GetShortestSpeakerAperture(self):
        return float(lib.jplg_panner_shortest_aperture(self.h))

    def is_3d(self):
        return bool(lib.jplg_panner_is_3d(self.h))

    def triangulation_is_tabulated(self):
        return bool(lib.jplg_panner_triangulation_is_tabulated(self.h))

    def triangles(self):
        n = lib.jplg_panner_triangles(self.h, None)
        t = np.zeros((n, 3), np.uint8)
        if n:
            lib.jplg_panner_triangles(self.h, t.ctypes.data)
        return t

    def GetLUT(self):
        L = self.GetNumChannels()
        if self.is_3d():
            g = np.zeros((65536, 3), np.float32)
            s = np.zeros((65536, 3), np.uint8)
            self.ctx.check(lib.jplg_panner_get_lut(self.h, g.ctypes.data, s.ctypes.data))
            return g, s
        g = np.zeros((1024, L), np.float32)
        self.ctx.check(lib.jplg_panner_get_lut(self.h, g.ctypes.data, None))
        return g, None

    def SetLUT(self, gains, speakers=None):
        g = np.ascontiguousarray(gains, np.float32)
        s = None if speakers is None else np.ascontiguousarray(speakers, np.uint8)
        self.ctx.check(lib.jplg_panner_set_lut(self.h, g.ctypes.data, None if s is None else s.ctypes.data))

    def InitializeSourceLayout(self, source_map):
        return SourceLayout(self, source_map)

    def GetSpeakerGains(self, directions, out=None):
        """directions [n,3] -> gains [n,L]"""
        d, mem, keep = _arg(directions)
        n = int(np.prod(keep.shape[:-1])) if keep is not None else 0
        L = self.GetNumChannels()
        if out is None:
            if mem == MEM_DEVICE:
                import torch
                out = torch.empty((n, L), dtype=torch.float32, device=keep.device)
            else:
                out = np.empty((n, L), np.float32)
        o, omem, okeep = _arg(out)
        assert omem == mem
        self.ctx.check(lib.jplg_panner_point_gains(self.h, d, n, o, mem))
        return out

    def ProcessVBAPData(self, layout, params, out=None, async_host=False):
        """params [n,5] = {dir.xyz, focus, spread} -> mix maps [n, channels, L]. async_host: params / out are PINNED host arrays and the call
        returns without waiting (JPLG_MEM_HOST_ASYNC)."""
        p, mem, keep = _arg(params)
        n = keep.shape[0]
        L = self.GetNumChannels()
        if out is None:
            if mem == MEM_DEVICE:
                import torch
                out = torch.empty((n, layout.num_channels, L), dtype=torch.float32, device=keep.device)
            else:
                out = np.empty((n, layout.num_channels, L), np.float32)
        o, omem, okeep = _arg(out)
        assert omem == mem
        self.ctx.check(lib.jplg_panner_mdap(self.h, layout.h, p, n, o, MEM_HOST_ASYNC if (async_host and mem == MEM_HOST) else mem))
        return out

    def ProcessVBAPDataOriented(self, layout, params, out=None):
        """PanUpdateDataWithOrientation overload (PannerBase.h:375): params [n,11] = {dir.xyz, focus, spread, up.xyz, forward.xyz}"""
        p, mem, keep = _arg(params)
        n = keep.shape[0]
        assert keep.shape[-1] == 11
        L = self.GetNumChannels()
        if out is None:
            if mem == MEM_DEVICE:
                import torch
                out = torch.empty((n, layout.num_channels, L), dtype=torch.float32, device=keep.device)
            else:
                out = np.empty((n, layout.num_channels, L), np.float32)
        o, omem, okeep = _arg(out)
        assert omem == mem
        self.ctx.check(lib.jplg_panner_mdap_oriented(self.h, layout.h, p, n, o, mem))
        return out

    def ProcessVirtualSources(self, virtual_sources, out=None):
        """virtual_sources [batches, n, 4] = {dir.xyz, weight} -> gains [batches, L] (accumulated into `out`, then L2-normalised)"""
        v = np.ascontiguousarray(virtual_sources, np.float32)
        if v.ndim == 2:
            v = v[None]
        b, n = v.shape[0], v.shape[1]
        if out is None:
            out = np.zeros((b, self.GetNumChannels()), np.float32)
        self.ctx.check(lib.jplg_panner_process_virtual_sources(self.h, v.ctypes.data, n, b, out.ctypes.data, MEM_HOST))
        return out

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_panner_destroy(self.h)
                self.h = None
        except Exception:
            pass


def normalize_weights(vs):
    v = np.ascontiguousarray(vs, np.float32)
    lib.jplg_normalize_weights(v.ctypes.data, v.shape[0])
    return v


def get_spread_from_source_size(size, distance):
    return float(lib.jplg_spread_from_source_size(size, distance))


# ------------------------------------------------------------------------------------------------
class DirectSoundBatch:
    """N x DirectSoundEffect (inc/Auralization/DirectSound.h:41-125): ctor, UpdateParameters, ProcessInterleaved."""

    def __init__(self, ctx, sample_rate, num_sources, num_source_channels, num_output_channels,
                 initial_filter_gains=None, initial_delay_seconds=None, initial_mix_maps=None,
                 ring_size=0, max_block_frames=512, mode=MODE_PARITY, splits=(0.0, 0.0, 0.0)):
        self.ctx = ctx
        self.N, self.C, self.L = int(num_sources), int(num_source_channels), int(num_output_channels)
        cfg = RendererConfig(float(sample_rate), self.N, self.C, self.L, int(ring_size), int(max_block_frames), int(mode),
                             float(splits[0]), float(splits[1]), float(splits[2]))
        b, _, kb = _arg(initial_filter_gains)
        d, _, kd = _arg(initial_delay_seconds)
        m, _, km = _arg(initial_mix_maps)
        h = _vp()
        ctx.check(lib.jplg_renderer_create(ctx.h, C.byref(cfg), b, d, m, C.byref(h)))
        self.h = h
        self.ring_size = int(lib.jplg_renderer_ring_size(h))

    def UpdateParameters(self, filter_gains=None, delay_seconds=None, mix_maps=None, async_host=False):
        b, m1, kb = _arg(filter_gains)
        d, m2, kd = _arg(delay_seconds)
        m, m3, km = _arg(mix_maps)
        mems = {x for x in (m1, m2, m3) if x is not None}
        assert len(mems) <= 1, "all update arrays must live in the same memory space"
        mem = mems.pop() if mems else MEM_HOST
        self.ctx.check(lib.jplg_renderer_update(self.h, b, d, m, MEM_HOST_ASYNC if (async_host and mem == MEM_HOST) else mem))

    def UpdatePanned(self, panner, layout, pan_params, band_gains, delay_seconds):
        """ProcessVBAPData + UpdateParameters fused on the device (jplg_renderer_update_panned): CUDA tensors; enqueued on the panner's context stream,
        adopted by the next ProcessInterleaved."""
        pp, m0, _ = _arg(pan_params); bb, m1, _ = _arg(band_gains); dd, m2, _ = _arg(delay_seconds)
        assert m0 == m1 == m2 == MEM_DEVICE
        panner.ctx.check(lib.jplg_renderer_update_panned(self.h, panner.h, layout.h, pp, bb, dd, MEM_DEVICE))

    def WaitRenderStarted(self, waiting_ctx):
        """`waiting_ctx`'s stream waits until the k_render enqueued last is fully resident (jplg_renderer_wait_render_started)."""
        self.ctx.check(lib.jplg_renderer_wait_render_started(self.h, waiting_ctx.h))

    def PushAhead(self, inp, num_frames):
        """issue the NEXT block's ring pushes now, on a side stream (jplg_renderer_push_ahead); `inp` is a CUDA tensor that the next
        ProcessInterleaved call must be given again"""
        i, mem, ki = _arg(inp)
        assert mem == MEM_DEVICE
        self.ctx.check(lib.jplg_renderer_push_ahead(self.h, i, int(num_frames)))

    def WaitBlock(self):
        """host-wait for the oldest asynchronous block (see ProcessInterleaved(async_host=True))"""
        self.ctx.check(lib.jplg_renderer_wait_block(self.h))

    def ProcessInterleaved(self, inp, bus, num_frames, per_source=None, async_host=False):
        """inp [N, frames, C]; bus [frames, L] accumulated into; per_source optional [N, frames, L] (debug).
        async_host: inp / bus are PINNED host arrays; the block is enqueued (its input transfer overlaps the previous block's render) and
        `bus` is valid after WaitBlock()."""
        i, mem, ki = _arg(inp)
        o, omem, ko = _arg(bus)
        assert mem == omem
        if async_host and mem == MEM_HOST:
            mem = MEM_HOST_ASYNC
        if per_source is None:
            self.ctx.check(lib.jplg_renderer_process_block(self.h, i, o, int(num_frames), mem))
        else:
            p, pmem, kp = _arg(per_source)
            assert pmem == mem
            self.ctx.check(lib.jplg_renderer_process_block_debug(self.h, i, o, p, int(num_frames), mem))
        return bus

    def device_bus_ptr(self):
        return lib.jplg_renderer_device_bus(self.h)

    def SetMaxDelaySeconds(self, seconds):
        """upper bound of device-resident target delays (jplg_renderer_set_max_delay_seconds): PushAhead needs it"""
        self.ctx.check(lib.jplg_renderer_set_max_delay_seconds(self.h, float(seconds)))

    def SetTimeParallel(self, policy):
        """throughput mode's time-parallel path (jplg_renderer_set_time_parallel): -1 chosen by size, 0 never, 1 whenever a block qualifies"""
        self.ctx.check(lib.jplg_renderer_set_time_parallel(self.h, int(policy)))

    @property
    def time_parallel_blocks(self):
        return int(lib.jplg_renderer_time_parallel_blocks(self.h))

    def state_export(self, with_rings=False):
        st = np.zeros((self.N * self.C, STATE_FLOATS), np.float32)
        rings = np.zeros((self.N * self.C, self.ring_size), np.float32) if with_rings else None
        pos = _u64()
        self.ctx.check(lib.jplg_renderer_state_export(self.h, st.ctypes.data, None if rings is None else rings.ctypes.data, C.byref(pos)))
        return st, rings, int(pos.value)

    def state_import(self, state, rings, pos):
        st = np.ascontiguousarray(state, np.float32)
        rg = None if rings is None else np.ascontiguousarray(rings, np.float32)
        self.ctx.check(lib.jplg_renderer_state_import(self.h, st.ctypes.data, None if rg is None else rg.ctypes.data, int(pos)))

    def close(self):
        if self.h and _ctx_alive(self):   # a closed context's children are gone with it
            lib.jplg_renderer_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_sig("jplg_reverb_bank_create", _i32, [_vp, _u32, _f32, _f32, _f32, _f32, _u32, _u32, _i32, C.POINTER(_vp)])
_sig("jplg_reverb_bank_destroy", None, [_vp])
_sig("jplg_reverb_bank_set_rt60", _i32, [_vp, _fp, _i32])
_sig("jplg_reverb_bank_set_decay_slope", _i32, [_vp, _fp, _i32])
_sig("jplg_reverb_bank_decay_gains", _i32, [_vp, _fp])
_sig("jplg_reverb_bank_process", _i32, [_vp, _fp, _u32, _fp, _u32, _i32])
_sig("jplg_reverb_bank_process_no_alloc", _i32, [_vp, _fp, _u32, _fp, _u32])
REVERB_NO_ALLOC_MAX_FLOATS = 245760   # JPLG_REVERB_NO_ALLOC_MAX_FLOATS (include/jplg.h)


class ReverbBusBatch:
    """n x ReverbBus (inc/Auralization/LateReverb.h:34-117): 16th-order FDN with four-band decay filters"""

    def __init__(self, ctx, n, sample_rate, output_channels, splits=(0.0, 0.0, 0.0), max_block_frames=512, mode=MODE_PARITY):
        self.ctx, self.n, self.L = ctx, int(n), int(output_channels)
        h = _vp()
        ctx.check(lib.jplg_reverb_bank_create(ctx.h, self.n, float(sample_rate), float(splits[0]), float(splits[1]), float(splits[2]),
                                              self.L, int(max_block_frames), int(mode), C.byref(h)))
        self.h = h

    def SetRT60(self, rt60):
        a, mem, _k = _arg(rt60)
        self.ctx.check(lib.jplg_reverb_bank_set_rt60(self.h, a, mem))

    def SetDecaySlope(self, slope):
        a, mem, _k = _arg(slope)
        self.ctx.check(lib.jplg_reverb_bank_set_decay_slope(self.h, a, mem))

    def DecayGains(self):
        out = np.zeros((self.n, 16, 4), np.float32)
        self.ctx.check(lib.jplg_reverb_bank_decay_gains(self.h, out.ctypes.data))
        return out

    def ProcessInterleaved(self, inp, out, frames, in_channels=1):
        """inp [n][frames] mono, out [n][frames][L] accumulated into (numpy = host buffers, torch CUDA tensors = device buffers)"""
        a, mem_a, _k = _arg(inp)
        if _is_torch(out):
            o, mem_o = out.data_ptr(), (MEM_DEVICE if out.is_cuda else MEM_HOST)
        else:
            assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
            o, mem_o = out.ctypes.data, MEM_HOST
        assert mem_a == mem_o, "input and output must live in the same memory space"
        self.ctx.check(lib.jplg_reverb_bank_process(self.h, a, int(in_channels), o, int(frames), mem_a))

    def ProcessInterleavedNoAlloc(self, inp, out, frames, in_channels=1):
        """jplg_reverb_bank_process_no_alloc: device tensors only, no allocation on the call (the down-mix scratch is a static device array)"""
        assert _is_torch(inp) and _is_torch(out) and inp.is_cuda and out.is_cuda and inp.is_contiguous() and out.is_contiguous()
        self.ctx.check(lib.jplg_reverb_bank_process_no_alloc(self.h, inp.data_ptr(), int(in_channels), out.data_ptr(), int(frames)))

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_reverb_bank_destroy(self.h)
                self.h = None
        except Exception:
            pass


_sig("jplg_erbus_bank_create", _i32, [_vp, _u32, _u32, _f32, _u32, _u32, _u32, _i32, C.POINTER(_vp)])
_sig("jplg_erbus_bank_destroy", None, [_vp])
_sig("jplg_erbus_bank_set_taps", _i32, [_vp, _u32, _u32, _fp, _fp, _vp, _fp])
_sig("jplg_erbus_bank_num_taps", _u32, [_vp, _u32])
_sig("jplg_erbus_bank_slots_per_bus", _u32, [_vp])
_sig("jplg_erbus_bank_tap_slot", _i32, [_vp, _u32, _u32])
_sig("jplg_erbus_bank_process", _i32, [_vp, _fp, _u32, _fp, _u32, _i32])
_sig("jplg_erbus_bank_process_debug", _i32, [_vp, _fp, _u32, _fp, _fp, _u32, _i32])


def _out_arg(out):
    if _is_torch(out):
        assert out.is_contiguous()
        return out.data_ptr(), (MEM_DEVICE if out.is_cuda else MEM_HOST)
    assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
    return out.ctypes.data, MEM_HOST


class ERBusBatch:
    """B x ERBus (inc/Auralization/EarlyReflectionsBus.h:45-249): early-reflection taps on one mono delay line per bus"""

    def __init__(self, ctx, num_buses, max_taps_per_bus, sample_rate, num_channels, ring_size=0, max_block_frames=512, mode=MODE_PARITY):
        self.ctx, self.B, self.L = ctx, int(num_buses), int(num_channels)
        h = _vp()
        ctx.check(lib.jplg_erbus_bank_create(ctx.h, self.B, int(max_taps_per_bus), float(sample_rate), self.L, int(ring_size),
                                             int(max_block_frames), int(mode), C.byref(h)))
        self.h = h
        self.slots_per_bus = int(lib.jplg_erbus_bank_slots_per_bus(h))

    def SetTaps(self, bus, filter_gains, delay_seconds, ids, gains):
        ids = np.ascontiguousarray(ids, np.uint32).reshape(-1)
        n = len(ids)
        fg = np.ascontiguousarray(filter_gains, np.float32).reshape(n, 4) if n else np.zeros((0, 4), np.float32)
        d = np.ascontiguousarray(delay_seconds, np.float32).reshape(n) if n else np.zeros(0, np.float32)
        g = np.ascontiguousarray(gains, np.float32).reshape(n, self.L) if n else np.zeros((0, self.L), np.float32)
        self.ctx.check(lib.jplg_erbus_bank_set_taps(self.h, int(bus), n, fg.ctypes.data, d.ctypes.data, ids.ctypes.data, g.ctypes.data))

    def GetNumTaps(self, bus):
        return int(lib.jplg_erbus_bank_num_taps(self.h, int(bus)))

    def tap_slot(self, bus, tap_id):
        return int(lib.jplg_erbus_bank_tap_slot(self.h, int(bus), int(tap_id)))

    def ProcessInterleaved(self, inp, out, frames, in_channels=1, per_tap=None):
        """inp [B][frames][in_channels], out [B][frames][L] accumulated into; per_tap [B*slots_per_bus][frames][L] (debug)"""
        a, mem_a, _k = _arg(inp)
        o, mem_o = _out_arg(out)
        assert mem_a == mem_o, "input and output must live in the same memory space"
        if per_tap is None:
            self.ctx.check(lib.jplg_erbus_bank_process(self.h, a, int(in_channels), o, int(frames), mem_a))
        else:
            p, mem_p = _out_arg(per_tap)
            assert mem_p == mem_a
            self.ctx.check(lib.jplg_erbus_bank_process_debug(self.h, a, int(in_channels), o, p, int(frames), mem_a))

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_erbus_bank_destroy(self.h)
                self.h = None
        except Exception:
            pass


class CrossoverBank:
    """n x FourBandCrossover (inc/Auralization/CrossoverFilter.h:416-535)"""

    def __init__(self, ctx, n, sample_rate, splits=(0.0, 0.0, 0.0), mode=MODE_PARITY):
        self.ctx, self.n = ctx, int(n)
        h = _vp()
        ctx.check(lib.jplg_crossover_bank_create(ctx.h, self.n, float(sample_rate), float(splits[0]), float(splits[1]), float(splits[2]), int(mode), C.byref(h)))
        self.h = h

    def coeffs(self):
        c = np.zeros(27, np.float32)
        lib.jplg_crossover_bank_coeffs(self.h, c.ctypes.data)
        return c

    def Reset(self):
        self.ctx.check(lib.jplg_crossover_bank_reset(self.h))

    def ProcessBlock(self, inp, start_gains, end_gains):
        x = np.ascontiguousarray(inp, np.float32)
        g0 = np.ascontiguousarray(start_gains, np.float32)
        g1 = np.ascontiguousarray(end_gains, np.float32)
        out = np.empty_like(x)
        self.ctx.check(lib.jplg_crossover_bank_process_ramp(self.h, x.ctypes.data, g0.ctypes.data, g1.ctypes.data, out.ctypes.data, x.shape[1], MEM_HOST))
        return out

    def SplitBlock(self, inp):
        x = np.ascontiguousarray(inp, np.float32)
        out = np.empty(x.shape + (4,), np.float32)
        self.ctx.check(lib.jplg_crossover_bank_split(self.h, x.ctypes.data, out.ctypes.data, x.shape[1], MEM_HOST))
        return out

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_crossover_bank_destroy(self.h)
                self.h = None
        except Exception:
            pass


class DelayBank:
    """n x (DelayLine + DelayTap) (inc/Auralization/DelayLine.h:43-68,129-230)"""
    NEAREST, LINEAR, THIRAN1 = 0, 1, 2

    def __init__(self, ctx, n, min_size):
        self.ctx, self.n = ctx, int(n)
        h = _vp()
        ctx.check(lib.jplg_delay_bank_create(ctx.h, self.n, int(min_size), C.byref(h)))
        self.h = h

    def GetSize(self):
        return int(lib.jplg_delay_bank_size(self.h))

    def Clear(self):
        self.ctx.check(lib.jplg_delay_bank_clear(self.h))

    def Push(self, samples):
        x = np.ascontiguousarray(samples, np.float32).reshape(self.n, -1)
        self.ctx.check(lib.jplg_delay_bank_push(self.h, x.ctypes.data, x.shape[1], MEM_HOST))

    def Read(self, delays, interp=2):
        d = np.ascontiguousarray(delays, np.float32).reshape(self.n, -1)
        out = np.empty_like(d)
        self.ctx.check(lib.jplg_delay_bank_read(self.h, d.ctypes.data, d.shape[1], int(interp), out.ctypes.data, MEM_HOST))
        return out

    def ReadXFade(self, delays, alphas):
        """DelayTapXFade (DelayLine.h:73-122): per tap SetCrossfade(alpha), SetDelay(delay), Process"""
        d = np.ascontiguousarray(delays, np.float32).reshape(self.n, -1)
        a = np.ascontiguousarray(alphas, np.float32).reshape(self.n, -1)
        out = np.empty_like(d)
        self.ctx.check(lib.jplg_delay_bank_read_xfade(self.h, d.ctypes.data, a.ctypes.data, d.shape[1], out.ctypes.data, MEM_HOST))
        return out

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_delay_bank_destroy(self.h)
                self.h = None
        except Exception:
            pass


def add_and_apply_gain_ramp(ctx, dest, source, dest_channel, source_channel, frames, gain_start, gain_end):
    """AddAndApplyGainRamp (inc/Auralization/ChannelMixing.h:89-107) for n (dest, source) pairs: dest [n, frames, nDest], source [n, frames, nSrc]"""
    d = np.array(dest, np.float32, copy=True, order="C")  # returns the mixed copy; the caller's array is left alone
    s = np.ascontiguousarray(source, np.float32)
    g0 = np.ascontiguousarray(gain_start, np.float32)
    g1 = np.ascontiguousarray(gain_end, np.float32)
    ctx.check(lib.jplg_mix_add_and_apply_gain_ramp(ctx.h, d.ctypes.data, s.ctypes.data, d.shape[0], dest_channel, source_channel,
                                                   d.shape[2], s.shape[2], frames, g0.ctypes.data, g1.ctypes.data, MEM_HOST))
    return d


def add(ctx, dest, source, dest_channel, source_channel=0):
    """Add (inc/Auralization/ChannelMixing.h:32-48) for n pairs: dest [n, frames, nDest] += source [n, frames, nSrc] (or [n, frames]: the mono overload)"""
    d = np.array(dest, np.float32, copy=True, order="C")
    s = np.ascontiguousarray(source, np.float32)
    if s.ndim == 2:
        s = s[:, :, None]
    ctx.check(lib.jplg_mix_add(ctx.h, d.ctypes.data, s.ctypes.data, d.shape[0], dest_channel, source_channel, d.shape[2], s.shape[2], d.shape[1], MEM_HOST))
    return d


def downmix_to_mono(ctx, dest, source):
    """DownmixToMono (ChannelMixing.h:129-136) for n pairs: dest [n, frames] += sum_ch source [n, frames, C]"""
    d = np.array(dest, np.float32, copy=True, order="C")
    s = np.ascontiguousarray(source, np.float32)
    ctx.check(lib.jplg_mix_downmix_to_mono(ctx.h, d.ctypes.data, s.ctypes.data, d.shape[0], s.shape[2], d.shape[1], MEM_HOST))
    return d


def apply_gain(ctx, dest, gains):
    d = np.array(dest, np.float32, copy=True, order="C")
    g = np.ascontiguousarray(gains, np.float32)
    ctx.check(lib.jplg_mix_apply_gain(ctx.h, d.ctypes.data, d.shape[0], d.shape[1], g.ctypes.data, MEM_HOST))
    return d


# ------------------------------------------------------------------------------------------------
# Scene front-end (inc/SpatialManager.h:374-420 batched)
class AttenuationFunction(C.Structure):
    _fields_ = [("kind", _i32), ("model", _i32), ("min_distance", _f32), ("max_distance", _f32), ("rolloff", _f32),
                ("first_point", _u32), ("num_points", _u32)]


class AirAbsorptionCache(C.Structure):
    """inc/AirAbsorption.h:51-87"""
    _fields_ = [(k, _f32 * 4) for k in ("band_centers", "lower_third_centers", "center_db", "lower_third_db", "weighted_db",
                                        "decay_slope", "distance_to_knee", "linearized_loss_factor")] + [("has_knee", _u32 * 4)]

    @staticmethod
    def CacheParameters(bands=None, temperature_c=20.0, humidity_percent=50.0, pressure_kpa=101.325, nyquist=22050.0):
        c = AirAbsorptionCache()
        b = None if bands is None else np.ascontiguousarray(bands, np.float32)
        lib.jplg_air_absorption_cache_build(None if b is None else b.ctypes.data, temperature_c, humidity_percent, pressure_kpa, nyquist, C.byref(c))
        return c

    def as_dict(self):
        return {k: np.array(getattr(self, k)) for k, _ in self._fields_}


class FrontendConfig(C.Structure):
    _fields_ = [("functions", C.POINTER(AttenuationFunction)), ("num_functions", _u32), ("points", _fp), ("num_points", _u32),
                ("sort_points", _i32), ("air", C.POINTER(AirAbsorptionCache)), ("speed_of_sound", _f32), ("spread_flavour", _i32)]


class FrontendIO(C.Structure):
    _fields_ = [("n", _u32), ("num_listeners", _u32), ("sources", _fp), ("listeners", _fp), ("listener_index", _fp),
                ("distance_in", _fp), ("dot_in", _fp), ("cones", _fp), ("object_sizes", _fp), ("focus_spread", _fp),
                ("focus", _f32), ("spread", _f32), ("function_index", _fp), ("pan_stride", _u32),
                ("direct", _fp), ("relative", _fp), ("pan_params", _fp), ("distance_attenuation", _fp), ("air_db", _fp),
                ("band_gains", _fp), ("delay_seconds", _fp)]


_sig("jplg_air_absorption_cache_build", None, [_fp, _f32, _f32, _f32, _f32, C.POINTER(AirAbsorptionCache)])
_sig("jplg_default_band_centers", None, [_fp])
_sig("jplg_speed_of_sound_for_temperature", _f32, [_f32])
_sig("jplg_sanitize_fallback_dir", None, [_fp])
_sig("jplg_frontend_create", _i32, [_vp, C.POINTER(FrontendConfig), C.POINTER(_vp)])
_sig("jplg_frontend_destroy", None, [_vp])
_sig("jplg_frontend_process", _i32, [_vp, C.POINTER(FrontendIO), _i32])
DECLARED_SYMBOLS = [n for n in dir(lib) if n.startswith("jplg_")]

MODEL_NONE, MODEL_INVERSE, MODEL_LINEAR, MODEL_EXPONENTIAL = 0, 1, 2, 3
(CURVE_CONSTANT, CURVE_LINEAR, CURVE_SINE_FADE_IN, CURVE_SINE_FADE_OUT, CURVE_SCURVE, CURVE_LOG_1_41, CURVE_LOG_3,
 CURVE_EXP_1_41, CURVE_EXP_3) = range(9)


def default_band_centers():
    out = np.zeros(4, np.float32)
    lib.jplg_default_band_centers(out.ctypes.data)
    return out


def speed_of_sound_for_temperature(t):
    return float(lib.jplg_speed_of_sound_for_temperature(t))


def sanitize_fallback_dir():
    out = np.zeros(3, np.float32)
    lib.jplg_sanitize_fallback_dir(out.ctypes.data)
    return out


class DirectPathBatch:
    """Batched DirectPathService + PanningService::EvaluateDirection glue + parameter producers (the per-source body of
    SpatialManager::AdvanceSimulation). `functions`: list of ("model", model, min, max, rolloff) or ("curve", [(distance, value, type), ...])."""

    _OUT = {"direct": 4, "relative": 9, "pan_params": None, "distance_attenuation": 1, "air_db": 4, "band_gains": 4, "delay_seconds": 1}

    def __init__(self, ctx, functions=(), sort_points=True, air=None, speed_of_sound=0.0, spread_flavour=0):
        self.ctx = ctx
        funcs = (AttenuationFunction * max(len(functions), 1))()
        pts = []
        for i, f in enumerate(functions):
            if f[0] == "model":
                funcs[i] = AttenuationFunction(0, int(f[1]), float(f[2]), float(f[3]), float(f[4]), 0, 0)
            else:
                funcs[i] = AttenuationFunction(1, 0, 0.0, 0.0, 0.0, len(pts), len(f[1]))
                pts.extend(f[1])
        p = np.ascontiguousarray(np.array(pts, np.float32).reshape(-1, 3))
        cfg = FrontendConfig(funcs, len(functions), p.ctypes.data if len(pts) else None, len(pts), 1 if sort_points else 0,
                             C.pointer(air) if air is not None else None, float(speed_of_sound), int(spread_flavour))
        self.h = _vp()
        ctx.check(lib.jplg_frontend_create(ctx.h, C.byref(cfg), C.byref(self.h)))

    def Process(self, n, want, sources=None, listeners=None, listener_index=None, distance_in=None, dot_in=None, cones=None,
                object_sizes=None, focus_spread=None, focus=1.0, spread=0.0, function_index=None, pan_stride=5, out=None):
        """`want`: names of the outputs (see jplg_frontend_io). numpy inputs -> numpy outputs (staged); CUDA torch inputs -> torch outputs
        on the context stream. Returns a dict."""
        io = FrontendIO()
        io.n = n
        io.focus, io.spread, io.pan_stride = focus, spread, pan_stride
        keep, mems = [], set()

        def put(name, x, dtype=np.float32):
            if x is None:
                return
            a, mem, k = _arg(x, dtype)
            setattr(io, name, a)
            keep.append(k)
            mems.add(mem)
        put("sources", sources); put("listeners", listeners); put("listener_index", listener_index, np.uint32)
        put("distance_in", distance_in); put("dot_in", dot_in); put("cones", cones); put("object_sizes", object_sizes)
        put("focus_spread", focus_spread); put("function_index", function_index, np.uint32)
        io.num_listeners = 0 if listeners is None else int(np.prod(listeners.shape) // 9)
        assert len(mems) == 1, "all front-end arrays must live in the same memory space"
        mem = mems.pop()
        res = {} if out is None else dict(out)
        for name in want:
            w = self._OUT[name]
            w = pan_stride if w is None else w
            if name not in res:
                if mem == MEM_DEVICE:
                    import torch
                    res[name] = torch.empty((n, w), dtype=torch.float32, device=next(k for k in keep if _is_torch(k)).device)
                else:
                    res[name] = np.empty((n, w), np.float32)
            a, omem, k = _arg(res[name])
            assert omem == mem
            if not _is_torch(res[name]):
                res[name] = k
            setattr(io, name, a)
        self.ctx.check(lib.jplg_frontend_process(self.h, C.byref(io), mem))
        return res

    def __del__(self):
        try:
            if self.h and _ctx_alive(self):   # a closed context's children are gone with it
                lib.jplg_frontend_destroy(self.h)
                self.h = None
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------
# Multi-GPU: peer-memory combine of the per-GPU partial buses (include/jplg.h "Multi-GPU")
_sig("jplg_bus_exchange_create_root", _i32, [_vp, _u32, _u32, C.POINTER(_vp), _vp])
_sig("jplg_bus_exchange_open_peer", _i32, [_vp, _u32, _u32, _u32, _vp, C.POINTER(_vp)])
_sig("jplg_bus_exchange_destroy", None, [_vp])
_sig("jplg_bus_exchange_submit", _i32, [_vp, _fp, _u32, _u64, _fp, _i32])
_sig("jplg_bus_exchange_error", _i32, [_vp])
_sig("jplg_bus_exchange_wait", _i32, [_vp])
_sig("jplg_bus_exchange_poll_error", _i32, [_vp])
_sig("jplg_bus_exchange_set_timeout_ms", _i32, [_vp, _u32])
_sig("jplg_bus_exchange_clear_error", _i32, [_vp])
DECLARED_SYMBOLS = [n for n in dir(lib) if n.startswith("jplg_")]
IPC_HANDLE_BYTES = 64


class BusExchange:
    """One per rank. Rank 0: BusExchange.create_root(ctx, world, max_floats) -> (exchange, handle bytes to distribute);
    other ranks: BusExchange.open_peer(ctx, world, rank, max_floats, handle)."""

    def __init__(self, ctx, h, world, rank):
        self.ctx, self.h, self.world, self.rank = ctx, h, world, rank

    @staticmethod
    def create_root(ctx, world, max_floats):
        h = _vp()
        handle = np.zeros(IPC_HANDLE_BYTES, np.uint8)
        ctx.check(lib.jplg_bus_exchange_create_root(ctx.h, int(world), int(max_floats), C.byref(h), handle.ctypes.data))
        return BusExchange(ctx, h, int(world), 0), handle

    @staticmethod
    def open_peer(ctx, world, rank, max_floats, handle):
        h = _vp()
        hb = np.ascontiguousarray(handle, np.uint8)
        ctx.check(lib.jplg_bus_exchange_open_peer(ctx.h, int(world), int(rank), int(max_floats), hb.ctypes.data, C.byref(h)))
        return BusExchange(ctx, h, int(world), int(rank))

    def submit(self, device_bus_ptr, n, seq, out=None, accumulate=False):
        """device_bus_ptr: raw device address of this rank's partial bus (e.g. DirectSoundBatch.device_bus_ptr()); out: CUDA tensor on rank 0"""
        o = None
        if out is not None:
            o, mem, _ = _arg(out)
            assert mem == MEM_DEVICE
        self.ctx.check(lib.jplg_bus_exchange_submit(self.h, device_bus_ptr, int(n), int(seq), o, 1 if accumulate else 0))

    def wait(self):
        """the context stream waits (on the device) for every block submitted so far; on rank 0 `out` is then complete"""
        self.ctx.check(lib.jplg_bus_exchange_wait(self.h))

    def error(self):
        return int(lib.jplg_bus_exchange_error(self.h))

    def poll_error(self):
        """this rank's own host-mapped error word: no device access, cheap enough to call after every block"""
        return int(lib.jplg_bus_exchange_poll_error(self.h))

    def set_timeout_ms(self, ms):
        self.ctx.check(lib.jplg_bus_exchange_set_timeout_ms(self.h, int(ms)))

    def clear_error(self):
        """collective: call on every rank between two host-side barriers, then restart seq at 0"""
        self.ctx.check(lib.jplg_bus_exchange_clear_error(self.h))

    def close(self):
        if self.h and _ctx_alive(self):   # a closed context's children are gone with it
            lib.jplg_bus_exchange_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
