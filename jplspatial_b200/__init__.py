"""jplspatial_b200 — B200-native many-source spatial renderer (JPL Spatial hot path).

Python is only plumbing here (tests, bench, torch.distributed launch): the product is the C-ABI shared
library `libjplg.so` (include/jplg.h) built from hand-written CUDA for sm_100a. This module mirrors the
reference's component interface (same names / argument meaning, batched) on top of that ABI through ctypes:

    ChannelMask / ChannelMap          inc/ChannelMap.h
    VBAPanner (2D/3D), SourceLayout   inc/Panning/PannerBase.h:315-389
    DirectSoundBatch                  inc/Auralization/DirectSound.h:41-125 (N effects at once)
    CrossoverBank, DelayBank          inc/Auralization/CrossoverFilter.h, DelayLine.h
    ReverbBusBatch                    inc/Auralization/LateReverb.h:34-117 (n FDN16 late-reverb buses)
    ERBusBatch                        inc/Auralization/EarlyReflectionsBus.h:45-249 (B early-reflection buses x taps)
    DirectPathBatch                   inc/SpatialManager.h:374-420 (DirectPathService + attenuation + air absorption + pan glue, batched)

There is NO CPU fallback: without the built library the import fails, without a GPU every data-path call raises.
"""
from ._capi import (  # noqa: F401
    JplgError, lib, lib_path, device_count, version,
    MEM_HOST, MEM_DEVICE, MEM_HOST_ASYNC, MODE_PARITY, MODE_THROUGHPUT,
    EChannel, ChannelMask, ChannelMap,
    Context, VBAPanner, SourceLayout, DirectSoundBatch, CrossoverBank, DelayBank, ReverbBusBatch, ERBusBatch,
    add_and_apply_gain_ramp, apply_gain, add, downmix_to_mono, normalize_weights, get_spread_from_source_size,
    STATE_FLOATS, REVERB_NO_ALLOC_MAX_FLOATS,
    BusExchange, DirectPathBatch, AirAbsorptionCache, default_band_centers, speed_of_sound_for_temperature, sanitize_fallback_dir,
    MODEL_NONE, MODEL_INVERSE, MODEL_LINEAR, MODEL_EXPONENTIAL,
    CURVE_CONSTANT, CURVE_LINEAR, CURVE_SINE_FADE_IN, CURVE_SINE_FADE_OUT, CURVE_SCURVE, CURVE_LOG_1_41, CURVE_LOG_3, CURVE_EXP_1_41, CURVE_EXP_3,
)
