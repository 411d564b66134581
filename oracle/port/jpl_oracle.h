/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * jpl_oracle: a plain CPU restatement of JPL Spatial's hot path (panner gains -> fractional delay ->
 * LR4 four-band filter -> gain-ramped mix), written from the reference's algorithm with every function
 * citing the reference file:line it follows (paths relative to /root/reference/Spatialization/include/JPLSpatial/
 * = "inc/", .../src/Spatialization/ = "src/"). It includes no reference header and no product code.
 *
 * Parity status: PINNED. The restatement is checked (tests/test_oracle_cpu.py) against
 *   (a) outputs of the reference itself compiled here (oracle/_ref, built from /root/reference in place),
 *   (b) the committed golden fixtures generated from that build (tests/golden/, script oracle/tools/gen_golden.py),
 *   (c) the known-answer vectors of the reference's own tests restated as asserts.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may load this library.
 */
#ifndef JPL_ORACLE_H
#define JPL_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- math primitives (lane-wise restatements of the SSE2 code paths) */
void orc_sincos(const float* x, uint32_t n, float* s, float* c);          /* inc/Math/SIMDMath.h:488-534 */
void orc_acos(const float* x, uint32_t n, float* out);                     /* :667-714 */
void orc_tan(const float* x, uint32_t n, float* out);                      /* :617-665 */
void orc_invsqrt_nr1(const float* x, uint32_t n, float* out);              /* inc/Math/SIMD.h:1700-1790 (RSQRTPS + 1 Newton step) */
void orc_invsqrt_nr2(const float* x, uint32_t n, float* out);              /* same, 2 Newton steps (Math::InvSqrt simd) */
void orc_oct_encode_simd(const float* dirs, uint32_t n, uint32_t* out);    /* inc/Math/DirectionEncoding.h:346-386 */
void orc_oct_encode_scalar(const float* dirs, uint32_t n, uint32_t* out);  /* :292-343 */
void orc_oct_decode_scalar(const uint32_t* codes, uint32_t n, float* out); /* :389-436 */
void orc_diamond_scalar(const float* xy, uint32_t n, float* out);          /* :197-209 */
void orc_from_diamond(const float* p, uint32_t n, float* outXY);           /* :212-227 */
void orc_slerp_simd(float* v0, const float* v1, float t, uint32_t n);      /* inc/Math/Vec3Math.h:206-307; v0 [n][3] in/out, v1 ONE target [3] */
void orc_quat_lookat(const float* dirs, uint32_t n, float* out);           /* inc/Math/MinimalQuat.h:84-144 */

/* ---- panner (inc/Panning/*) */
typedef struct orc_panner orc_panner;
/* tris: speaker triangle list for 3D masks ([ntris][3], vertex numbering = non-LFE channels then dummies);
 * ignored for 2D masks. Returns NULL when the reference's Initialize would return false. */
orc_panner* orc_panner_create(uint32_t target_mask, const uint8_t* tris, uint32_t ntris);
void     orc_panner_destroy(orc_panner*);
int      orc_panner_is3d(const orc_panner*);
uint32_t orc_panner_num_channels(const orc_panner*);
float    orc_panner_aperture(const orc_panner*);
void     orc_panner_lut(const orc_panner*, float* gains, uint8_t* speakers);
void     orc_point_gains(const orc_panner*, const float* dirs, uint32_t n, float* out);               /* PannerBase.h:441-445 */
/* info[0..3] = channels, VS per channel, rings, samples per ring; returns 0 if the layout is invalid */
int      orc_layout_info(const orc_panner*, uint32_t source_mask, uint32_t info[4], float* phi, float* offset, float* weight, float* quats, uint32_t* ids);
int      orc_mdap(const orc_panner*, uint32_t source_mask, const float* params, uint32_t n, float* out); /* PannerBase.h:486-728 */
void     orc_process_virtual_sources(const orc_panner*, const float* vs, uint32_t n, float* out);      /* :731-804 */

/* ---- DSP components */
void  orc_xover_coeffs(float fs, float f1, float f2, float f3, float out[27]);                          /* inc/Auralization/CrossoverFilter.h:50-61,139-150,253-258,343-362 */
void  orc_xover_process_ramp(const float coef[27], float state[24], const float* in, uint32_t n, const float g0[4], const float g1[4], float* out); /* :505-535 */
void  orc_xover_split(const float coef[27], float state[24], const float* in, uint32_t n, float* out_bands); /* :462-470 */
float orc_smoother_alpha(uint32_t ticks);                                                                /* inc/Utilities/SmoothedValue.h:36-41 */
void  orc_add_gain_ramp(float* dest, const float* src, uint32_t dest_ch, uint32_t src_ch, uint32_t n_dest, uint32_t n_src, uint32_t frames, float g0, float g1); /* inc/Auralization/ChannelMixing.h:89-107 */

/* ---- DirectSoundEffect (src/Auralization/DirectSound.cpp:34-239) */
typedef struct orc_dse orc_dse;
orc_dse* orc_dse_create(float fs, uint32_t n_src, uint32_t n_out, const float band_gains[4], float delay_seconds, const float* mix_map);
void     orc_dse_destroy(orc_dse*);
void     orc_dse_update(orc_dse*, const float band_gains[4], float delay_seconds, const float* mix_map);
void     orc_dse_process(orc_dse*, const float* in, float* out, uint32_t frames); /* out accumulated into */
/* n effects, one block each, sources partitioned over `threads`; returns seconds (CPU baseline "port") */
double   orc_dse_process_batch(orc_dse** fx, uint32_t n, const float* in, uint32_t frames, float* bus, int threads);


/* ---- ReverbBus (src/Auralization/LateReverb.cpp:31-134, inc/Auralization/ReverbDesign.h:311-349,388-492,573-704) */
typedef struct orc_reverb orc_reverb;
void        orc_db_to_gain(const float* db, uint32_t n, float* out);                                   /* inc/Math/DecibelsAndGain.h:40-56 (SIMD exp, SIMDMath.h:296-316) */
orc_reverb* orc_reverb_create(float fs, float f1, float f2, float f3);
void        orc_reverb_destroy(orc_reverb*);
void        orc_reverb_set_rt60(orc_reverb*, const float rt60[4]);
void        orc_reverb_set_decay_slope(orc_reverb*, const float slope[4]);
void        orc_reverb_process(orc_reverb*, const float* in_mono, float* out, uint32_t n_out, uint32_t frames); /* out accumulated into */

/* ---- ERBus (src/Auralization/EarlyReflectionsBus.cpp:83-417) */
typedef struct orc_erbus orc_erbus;
orc_erbus*  orc_erbus_create(float fs, uint32_t num_channels);
void        orc_erbus_destroy(orc_erbus*);
uint32_t    orc_erbus_num_taps(const orc_erbus*);
void        orc_erbus_set_taps(orc_erbus*, uint32_t n, const float* filter_gains, const float* delay_seconds, const uint32_t* ids, const float* gains);
void        orc_erbus_process(orc_erbus*, const float* in, uint32_t n_in, float* out, uint32_t frames);          /* out accumulated into */

/* ---- scene front-end (inc/SpatialManager.h:374-420 and the services / producers it calls); positions = location, up, forward */
int   orc_mdap_oriented(const orc_panner*, uint32_t source_mask, const float* params /* [n][11] */, uint32_t n, float* out); /* PannerBase.h:501-515 */
void  orc_direct_path(const float* source, const float* listener, float out[12]);        /* inc/Services/DirectPathService.h:280-338 */
float orc_angle_attenuation(float azimuth_cos, float inner, float outer);                 /* :340-364 */
float orc_attenuation_model(int model, float min_d, float max_d, float rolloff, float d); /* inc/DistanceAttenuation.h:66-119 */
float orc_attenuation_curve(const float* points, uint32_t k, float d);                    /* :172-215 (points sorted) */
float orc_speed_of_sound(float temperature_c);                                            /* inc/SpeedOfSound.h:38-47 */
void  orc_air_cache(const float* bands, float t_c, float rh, float p_kpa, float nyquist, float out[32], uint32_t knee[4]); /* inc/AirAbsorption.h:179-210 */
void  orc_air_for_distance(const float* cache, const uint32_t* knee, float d, float* db, float* gain); /* :212-234 + DecibelsAndGain.h:40-50 */
void  orc_sanitize_source_dir(const float d[3], float out[3]);                            /* inc/Services/PanningService.h:480-495 */
float orc_spread_from_size(int flavour, float size, float distance);                      /* inc/SpatialManager.h:617-623, inc/Panning/PannerBase.h:69-74 */

#ifdef __cplusplus
}
#endif
#endif
