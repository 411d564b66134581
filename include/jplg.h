/*
 * jplg.h — C ABI of the B200-native many-source spatial renderer.
 *
 * Drop-in boundary for ONE hot path of JPL Spatial (reference paths are relative to
 * Spatialization/include/JPLSpatial/ = "inc/", Spatialization/src/Spatialization/ = "src/"):
 *
 *   batched VBAP/MDAP panner gains  ->  fractional-delay/Doppler ring reads  ->
 *   LR4 four-band propagation filter with ramped band gains  ->  gain-ramped mix of N sources into
 *   an L-channel speaker bus.
 *
 * The reference has no C ABI (it is a C++ header-template library); its boundary is the component API
 * in namespace JPL. Every entry point below names the reference interface it replaces. The calls are
 * BATCHED (a GPU call per source is meaningless), otherwise argument meaning, channel order
 * (ascending EChannel bit order, LFE included with gain 0), coordinate system (right handed, +Y up,
 * -Z forward) and error behaviour (no exceptions; a status code + last-error string + optional trace
 * callback instead of `false`/JPL_ERROR_TAG) follow the reference.
 *
 * All data buffers are caller-owned. A `mem` argument says whether the data pointers of that call are
 * host memory (JPLG_MEM_HOST: the call copies in/out on the context stream and synchronises before
 * returning) or device memory of the context's GPU (JPLG_MEM_DEVICE: the call only enqueues work on
 * the context stream and returns; no synchronisation).
 *
 * Thread model: same as the reference (inc/Auralization/DirectSound.h:52-65): one "audio" caller per
 * renderer (process_block), one "non-audio" caller (update); panner queries are const and may be
 * called concurrently after creation *from one thread per context stream*.
 */
#ifndef JPLG_H
#define JPLG_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define JPLG_API __declspec(dllexport)
#else
#define JPLG_API __attribute__((visibility("default")))
#endif

/* ---------------------------------------------------------------------------------------------- */
/* status / errors  (replaces: bool returns + JPL_ERROR_TAG, inc/ErrorReporting.h:96-129)          */

typedef int32_t jplg_status;
enum {
    JPLG_OK = 0,
    JPLG_ERR_INVALID_ARGUMENT = 1,
    JPLG_ERR_INVALID_CHANNEL_MAP = 2, /* Initialize() returned false in the reference */
    JPLG_ERR_NOT_INITIALIZED = 3,
    JPLG_ERR_CUDA = 4,                /* sticky: the context is unusable afterwards */
    JPLG_ERR_NO_DEVICE = 5,           /* no CUDA device: the library has NO CPU fallback */
    JPLG_ERR_OUT_OF_MEMORY = 6,
    JPLG_ERR_UNSUPPORTED = 7,
    JPLG_ERR_TRIANGULATION = 8
};

enum { JPLG_MEM_HOST = 0, JPLG_MEM_DEVICE = 1,
       /* PINNED host memory, asynchronous: the call enqueues its copies and kernels and returns without waiting; the buffers must stay
        * untouched until jplg_renderer_wait_block / jplg_context_synchronize. Accepted by jplg_panner_mdap(_oriented), jplg_renderer_update
        * and jplg_renderer_process_block (whose input transfer then runs on a copy stream into one of two staging buffers, so the PCIe
        * transfer of block k+1 overlaps the render of block k). Page-locked buffers (cudaHostAlloc / cudaHostRegister) have a device alias
        * under unified addressing: the small per-block arrays (pan parameters, band gains, delays, mix maps) and the bus are then read /
        * accumulated by the kernels directly over PCIe, so they never queue behind the audio transfer on the copy engine. Pageable
        * buffers are accepted and take staged copies (the driver then blocks the caller for their duration). */
       JPLG_MEM_HOST_ASYNC = 2 };

/* Arithmetic mode of the render kernels.
 *  JPLG_MODE_PARITY     every reference FMA(a,b,c) is a separately rounded mul + add, exactly like the
 *                       reference's SSE2 baseline build (inc/Math/SIMD.h:1843-1850 without JPL_USE_FMADD):
 *                       per-source rendered audio is BIT-EXACT against that build.
 *  JPLG_MODE_THROUGHPUT fused multiply-adds at the reference's FMA call sites (the semantic of its
 *                       JPL_ENABLE_AVX2_FMA build); meets 1e-5 relative against the SSE2 build. */
enum { JPLG_MODE_PARITY = 0, JPLG_MODE_THROUGHPUT = 1 };

typedef struct jplg_context_t*       jplg_context;
typedef struct jplg_panner_t*        jplg_panner;
typedef struct jplg_source_layout_t* jplg_source_layout;
typedef struct jplg_renderer_t*      jplg_renderer;
typedef struct jplg_crossover_bank_t* jplg_crossover_bank;
typedef struct jplg_delay_bank_t*    jplg_delay_bank;
typedef struct jplg_reverb_bank_t*   jplg_reverb_bank;
typedef struct jplg_erbus_bank_t*    jplg_erbus_bank;
typedef struct jplg_frontend_t*      jplg_frontend;
typedef struct jplg_bus_exchange_t*  jplg_bus_exchange;

/* replaces: JPL::SpatialTrace (inc/ErrorReporting.h:44-47) */
typedef void (*jplg_trace_fn)(const char* message);
JPLG_API void jplg_set_trace_callback(jplg_trace_fn fn);

JPLG_API const char* jplg_version(void);
/* 0 if no usable CUDA device is visible. The library never computes on the CPU. */
JPLG_API int32_t jplg_device_count(void);

/* ---------------------------------------------------------------------------------------------- */
/* context: device, stream, constant tables                                                        */

JPLG_API jplg_status jplg_context_create(int32_t device, jplg_context* out);
/* The reference runs SpatialManager::AdvanceSimulation (panning data, scene front-end) on the game thread beside the audio thread that renders
 * (SURVEY.md 3); the device analogue is a second context whose stream has a LOWER priority than the rendering context's: with both eligible, the
 * renderer's CTAs are dispatched first and the background context's kernels fill the SMs the renderer leaves free. Order work between the two
 * contexts with events on jplg_context_stream(). flags = 0 is jplg_context_create. */
enum { JPLG_CONTEXT_BACKGROUND = 1 };
JPLG_API jplg_status jplg_context_create_ex(int32_t device, uint32_t flags, jplg_context* out);
JPLG_API void        jplg_context_destroy(jplg_context ctx);
JPLG_API const char* jplg_last_error(jplg_context ctx);
/* cudaStream_t of the context, as void*; all work of this context is enqueued on it. */
JPLG_API void*       jplg_context_stream(jplg_context ctx);
JPLG_API jplg_status jplg_context_synchronize(jplg_context ctx);
/* 1 when the host's RSQRTPS matched the "2048-entry table + exact exponent scaling" model the device
 * uses to reproduce _mm_rsqrt_ps (inc/Math/SIMD.h:1709) on every probed input, else 0. */
JPLG_API int32_t     jplg_context_rsqrt_model_verified(jplg_context ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
JPLG_API uint64_t    jplg_context_launch_count(jplg_context ctx);
/* Per-kernel device timing with CUDA events recorded on the context stream around each launch (measurement aid,
 * off by default; `enabled`: 0 off, 1 every kernel, or (mask << 1) with mask bits (1 << JPLG_K_*) to bracket only those kernels
 * so the events do not perturb the others). `which`: one of JPLG_K_*. kernel_time synchronises the stream and returns the accumulated
 * milliseconds and launch count since the last reset. */
enum { JPLG_K_PUSH = 0, JPLG_K_RENDER = 1, JPLG_K_BUS_REDUCE = 2, JPLG_K_MDAP = 3, JPLG_K_POINT_GAINS = 4,
       JPLG_K_UPDATE = 5, JPLG_K_OTHER = 6, JPLG_K_FDN = 7, JPLG_K_ER_PUSH = 8, JPLG_K_FRONTEND = 9, JPLG_K_EXCHANGE = 10, JPLG_K_COUNT = 11 };
JPLG_API void        jplg_context_set_kernel_timing(jplg_context ctx, int32_t enabled);
JPLG_API void        jplg_context_reset_kernel_timing(jplg_context ctx);
JPLG_API jplg_status jplg_context_kernel_time(jplg_context ctx, int32_t which, double* total_ms, uint64_t* launches);
/* Measurement aid: the FP32 CUDA-core ceiling of this device for the instruction mix the parity arithmetic is made of (packed FP32x2 multiplies
 * and adds, one rounded operation per lane and issue slot, no fused operations), in lane operations per second; a short blocking probe kernel on
 * every SM (about 2 ms). bench.py uses it as the denominator of the fused render's roofline fraction. */
JPLG_API jplg_status jplg_context_probe_fp32(jplg_context ctx, double* lane_ops_per_second, double* kernel_ms);

/* ---------------------------------------------------------------------------------------------- */
/* ChannelMap helpers  (replaces: inc/ChannelMap.h:153-249; masks are EChannel bit masks :38-79)    */

JPLG_API uint32_t jplg_channel_map_num_channels(uint32_t mask);            /* GetNumChannels :165 */
JPLG_API int32_t  jplg_channel_map_has_lfe(uint32_t mask);                 /* HasLFE :161 */
JPLG_API int32_t  jplg_channel_map_has_top_channels(uint32_t mask);        /* HasTopChannels :162 */
JPLG_API uint32_t jplg_channel_map_channel_index(uint32_t mask, uint32_t channel_bit); /* GetChannelIndex :167-183; 0xFFFFFFFF if absent */
JPLG_API uint32_t jplg_channel_map_from_num_channels(uint32_t n);          /* FromNumChannels :203-219 */

/* ---------------------------------------------------------------------------------------------- */
/* Panners: VBAPanner2D / VBAPanner3D  (replaces: inc/Panning/PannerBase.h:315-389)                 */

/* Initialize(ChannelMap) (:315,455-482). 2D panner for masks without top channels
 * (inc/Panning/VBAPanning2D.h:124-137), 3D panner (KB_983 LUT) otherwise (VBAPanning3D.h:157-166).
 * Builds the LUT on the host exactly as LUTBuilder2D / LUTBuilder do and uploads it. */
JPLG_API jplg_status jplg_panner_create(jplg_context ctx, uint32_t target_mask, jplg_panner* out);
/* The same with one of the reference's LUT storage variants (template parameter ELUTSize of VBAPanner3D, inc/Panning/VBAPLUT3D.h:93-104,186-204;
 * 3D layouts only). The device always queries one repacked 16-byte cell table, so a variant changes what the table HOLDS: exactly what the
 * reference's query of that variant returns.
 *   JPLG_LUT_KB_983  float gains, direct speaker indices: the default, what jplg_panner_create builds
 *   JPLG_LUT_KB_851  float gains, speakers through the per-triangle table (ExtractDynamicData :768-800). Layouts WITHOUT an LFE only: with one the
 *                    reference shifts its triplet indices by the LFE's channel position (:618-620 -> :516-519) and reads past the table's end,
 *                    so there is no defined output to match (JPLG_ERR_UNSUPPORTED)
 *   JPLG_LUT_KB_786  gains packed to 24 bits and back (Gain24Bit, inc/Utilities/GainEncoding.h:30-62), direct speaker indices
 *   JPLG_LUT_KB_655, JPLG_LUT_KB_65  JPLG_ERR_UNSUPPORTED: KB_655 inherits KB_851's indexing, KB_65 returns NaN rows upstream */
#define JPLG_LUT_KB_983 0
#define JPLG_LUT_KB_851 1
#define JPLG_LUT_KB_786 2
#define JPLG_LUT_KB_655 3
#define JPLG_LUT_KB_65  4
JPLG_API jplg_status jplg_panner_create_ex(jplg_context ctx, uint32_t target_mask, int32_t lut_variant, jplg_panner* out);
JPLG_API void        jplg_panner_destroy(jplg_panner p);
JPLG_API int32_t     jplg_panner_is_3d(jplg_panner p);
JPLG_API uint32_t    jplg_panner_num_channels(jplg_panner p);              /* GetNumChannels :346 */
JPLG_API uint32_t    jplg_panner_channel_mask(jplg_panner p);              /* GetChannelMap :349 */
JPLG_API float       jplg_panner_shortest_aperture(jplg_panner p);         /* GetShortestSpeakerAperture :353 */
/* 1 when the 3D speaker triangulation came from the built-in table of the reference's standard layouts,
 * 0 when the generic convex-hull triangulator was used (LUT parity then reported, not promised). */
JPLG_API int32_t     jplg_panner_triangulation_is_tabulated(jplg_panner p);
/* number of speaker triangles / copies them (vertex numbering = non-LFE channel order, then dummies) */
JPLG_API uint32_t    jplg_panner_triangles(jplg_panner p, uint8_t* out_tris /* [n][3] or NULL */);

/* GetLUT (:320). 2D: `gains` receives 1024*L floats, `speakers` ignored (VBAPLUT2D.h:91-99).
 * 3D: `gains` receives 65536*3 floats and `speakers` 65536*3 bytes (VBAPLUT3D.h:163-171). Host pointers. */
JPLG_API jplg_status jplg_panner_get_lut(jplg_panner p, float* gains, uint8_t* speakers);
/* Replace the LUT contents (host pointers, same shapes as get_lut). Lets a caller run the kernels on a
 * LUT built elsewhere (used by the parity tests to decouple kernel parity from LUT-construction parity). */
JPLG_API jplg_status jplg_panner_set_lut(jplg_panner p, const float* gains, const uint8_t* speakers);

/* GetSpeakerGains(direction, outGains) batched (:329,441-445). dirs [n][3] normalised, non-zero;
 * out [n][L]. Unlike the reference's 3D query (which writes only 3 entries, VBAPLUT3D.h:1001-1009)
 * every row is fully written (zeros elsewhere). */
JPLG_API jplg_status jplg_panner_point_gains(jplg_panner p, const float* dirs, uint32_t n, float* out, int32_t mem);

/* InitializeSourceLayout(ChannelMap, SourceLayout&) (:324; VBAPanning2D.h:150-193; VBAPanning3D.h:242-310) */
JPLG_API jplg_status jplg_source_layout_create(jplg_panner p, uint32_t source_mask, jplg_source_layout* out);
JPLG_API void        jplg_source_layout_destroy(jplg_source_layout l);
JPLG_API uint32_t    jplg_source_layout_num_channels(jplg_source_layout l);        /* ChannelGroups.size() (LFE skipped) */
JPLG_API uint32_t    jplg_source_layout_num_virtual_sources(jplg_source_layout l); /* GetNumVirtualSources */
/* info[0]=channels, [1]=VS per channel, [2]=rings, [3]=samples per ring (3D; 0 for 2D) */
JPLG_API void        jplg_source_layout_info(jplg_source_layout l, uint32_t info[4], float* phi_term, float* offset_term, float* vs_weight);
/* channel-group quaternions [channels][4] = {x,y,z,w} (PannerBase.h:255-262) */
JPLG_API void        jplg_source_layout_groups(jplg_source_layout l, float* out_quats, uint32_t* out_channel_ids);

/* ProcessVBAPData(sourceLayout, PanUpdateData, outChannelMixMap) batched (:369,486-500).
 * params [n][5] = { dir.x, dir.y, dir.z, focus, spread } ; out [n][channels*L]. */
JPLG_API jplg_status jplg_panner_mdap(jplg_panner p, jplg_source_layout l, const float* params, uint32_t n, float* out, int32_t mem);
/* ProcessVBAPData(sourceLayout, PanUpdateDataWithOrientation, outChannelMixMap) batched (:375,501-515): the pan rotation is
 * QuatFromUpAndForward(Orientation.Up, Orientation.Forward) instead of QuatLookAt(SourceDirection).
 * params [n][11] = { dir.xyz, focus, spread, up.xyz, forward.xyz } ; out [n][channels*L]. */
JPLG_API jplg_status jplg_panner_mdap_oriented(jplg_panner p, jplg_source_layout l, const float* params, uint32_t n, float* out, int32_t mem);

/* ProcessVirtualSources(span<VirtualSource>, outGains) (:385,731-742): ONE gain vector from n weighted
 * directions. vs [n][4] = { dir.x, dir.y, dir.z, weight }; out [L] is ACCUMULATED INTO and then L2-normalised,
 * like the reference (caller zeroes it). `batches` independent sets of n (vs [batches][n][4], out [batches][L]). */
JPLG_API jplg_status jplg_panner_process_virtual_sources(jplg_panner p, const float* vs, uint32_t n, uint32_t batches, float* out, int32_t mem);
/* NormalizeWeights (:389,879-892), host helper, in place on [n][4] records */
JPLG_API void        jplg_normalize_weights(float* vs, uint32_t n);
/* GetSpreadFromSourceSize (:69-74) */
JPLG_API float       jplg_spread_from_source_size(float source_size, float distance);

/* ---------------------------------------------------------------------------------------------- */
/* Renderer: N x DirectSoundEffect  (replaces: inc/Auralization/DirectSound.h:41-125,               */
/*                                              src/Auralization/DirectSound.cpp:34-239)            */

typedef struct jplg_renderer_config {
    float    sample_rate;        /* DirectSoundEffect ctor `sampleRate` */
    uint32_t num_sources;        /* number of effects in the batch */
    uint32_t source_channels;    /* `numSourceChannels` (same for every source of one renderer) */
    uint32_t output_channels;    /* `numOutputChannels` = L */
    uint32_t ring_size;          /* 0 = reference value bit_ceil(2*fs) (DirectSound.cpp:79); else a power of two */
    uint32_t max_block_frames;   /* largest numFrames of process_block; 0 = 512 */
    int32_t  mode;               /* JPLG_MODE_PARITY / JPLG_MODE_THROUGHPUT */
    float    split_f1, split_f2, split_f3; /* 0 = reference defaults 176/775/3408 Hz (inc/FrequencyBands.h:46) */
} jplg_renderer_config;

/* Allocates rings/state for cfg.num_sources effects. Initial parameters like the ctor:
 * band_gains [N][4] (b1..b4), delay_seconds [N], mix_maps [N][C*L]; host pointers, any may be NULL
 * (gains 1, delay 0, mix 0). */
JPLG_API jplg_status jplg_renderer_create(jplg_context ctx, const jplg_renderer_config* cfg,
                                          const float* band_gains, const float* delay_seconds, const float* mix_maps,
                                          jplg_renderer* out);
JPLG_API void        jplg_renderer_destroy(jplg_renderer r);

/* UpdateParameters(filterGains, delayTimeSeconds, channelMixMap) for all N sources (DirectSound.cpp:231-239).
 * The new targets become visible at the next process_block (the reference picks them up at its next
 * 480-frame chunk, DirectSound.cpp:136-142). Any pointer may be NULL = leave unchanged. */
JPLG_API jplg_status jplg_renderer_update(jplg_renderer r, const float* band_gains, const float* delay_seconds,
                                          const float* mix_maps, int32_t mem);

/* ProcessInterleaved for all N sources (DirectSound.cpp:95-229).
 * in : [N][frames][C] (each source's block is interleaved exactly as the reference's `input` span)
 * bus: [frames][L] interleaved, ACCUMULATED INTO like the reference's `output` span; all N sources add
 *      into the same bus (the host loop "for each effect: ProcessInterleaved(in_i, bus)"). */
/* jplg_panner_mdap + jplg_renderer_update in one launch, for sources whose pan parameters change every block (replaces the pair
 * VBAPanner::ProcessVBAPData, inc/Panning/PannerBase.h:315 -> DirectSoundEffect::UpdateParameters, src/Auralization/DirectSound.cpp:231-239, per
 * source): pan_params [N][5], band_gains [N][4], delay_seconds [N], all JPLG_MEM_DEVICE. The mix maps never leave the device: the panning kernel
 * writes them, the band gains and the delays straight into the renderer's NEXT set of targets, which the next jplg_renderer_process_block adopts
 * (it waits for them on the device). The work is enqueued on the PANNER's context stream — with a JPLG_CONTEXT_BACKGROUND panner context it runs
 * while the renderer's context still renders the previous block, which keeps reading the current targets. Same values, bit for bit, as the two
 * separate calls. 3D panners; the panner's channel count must be the renderer's output_channels, and the renderer's source_channels either the
 * layout's panned channel count or the channel count of the layout's source map: in that case (a 5.1 source rendered with all six channels) every
 * panned channel group lands on the row of its channel in ascending EChannel order and the rows the panner skips (LFE, PannerBase.h:937) get zero
 * gains, which is the mix map an engine assembles from the reference's per-channel gains. */
JPLG_API jplg_status jplg_renderer_update_panned(jplg_renderer r, jplg_panner p, jplg_source_layout l, const float* pan_params,
                                                 const float* band_gains, const float* delay_seconds, int32_t mem);
JPLG_API jplg_status jplg_renderer_process_block(jplg_renderer r, const float* in, float* bus, uint32_t frames, int32_t mem);
/* Offline / throughput rendering: issues the ring pushes of the NEXT block (device pointer `in`, already resident) on a side stream, so that
 * they overlap the render of the block enqueued before this call: the copy is HBM-bound and runs on the SMs the render leaves free. The two
 * touch disjoint ring slots as long as the ring holds the longest delay in use plus two blocks of pushes plus one chunk; the call checks that against
 * the delay bound it knows (jplg_renderer_set_max_delay_seconds below) and fails otherwise (true for the reference ring size bit_ceil(2 fs) up to 1.7 s). The next jplg_renderer_process_block must get the same `in` / `frames`
 * (JPLG_MEM_DEVICE); it waits for these pushes instead of issuing its own (any other block simply overwrites them: the ring position only advances
 * in process_block). The side stream starts them once the render enqueued last is fully resident, and ranks above JPLG_CONTEXT_BACKGROUND
 * contexts, so they neither delay that render's placement nor trail behind background panning work. A real-time engine, whose next block does
 * not exist yet, does not need this call. */
JPLG_API jplg_status jplg_renderer_push_ahead(jplg_renderer r, const float* in, uint32_t frames);
/* Upper bound of the target delays (seconds) for renderers whose delays only exist on the device (jplg_renderer_update with JPLG_MEM_DEVICE,
 * jplg_renderer_update_panned): host-side updates are tracked by the library itself. jplg_renderer_push_ahead refuses to run without a known bound, and
 * when longest delay + two blocks of pushes + one 480-frame chunk does not fit the ring (the pushed-ahead block would overwrite samples the render of the
 * current block still reads). Declaring a bound smaller than the delays actually in use is a caller error. */
JPLG_API jplg_status jplg_renderer_set_max_delay_seconds(jplg_renderer r, float seconds);
/* Makes `waiting`'s stream (a JPLG_CONTEXT_BACKGROUND context on the same device) wait until every CTA of the k_render launch enqueued last by
 * jplg_renderer_process_block is resident on the device. Work enqueued on `waiting` afterwards (the next block's jplg_panner_mdap, the scene
 * front-end) runs beside that render on the SMs it leaves free instead of occupying SMs its CTAs need. No-op before the first block. */
JPLG_API jplg_status jplg_renderer_wait_render_started(jplg_renderer r, jplg_context waiting);
/* JPLG_MEM_HOST_ASYNC blocks complete in submission order; this waits (on the host) for the oldest block not yet waited for: its `bus`
 * is then valid. At most two blocks are in flight (a third submission first waits for the oldest). */
JPLG_API jplg_status jplg_renderer_wait_block(jplg_renderer r);
/* Same, but also writes every source's own contribution: per_source [N][frames][L] (overwritten). */
JPLG_API jplg_status jplg_renderer_process_block_debug(jplg_renderer r, const float* in, float* bus, float* per_source,
                                                       uint32_t frames, int32_t mem);
/* JPLG_MODE_THROUGHPUT renderers render small scenes TIME-PARALLEL: FourBandCrossover::ProcessBlock (inc/Auralization/CrossoverFilter.h:505-535)
 * is a linear recurrence with coefficients shared by every source, so a block is cut into
 * 16-frame tiles that are filtered from zero state in parallel, stitched by a scan over the tiles' 24-float filter states (s' = M s + e, M fixed
 * per sample rate) and re-run from their true start states (jplg_scan_kernels.cu); the delay smoother and the Thiran all-pass stay sequential
 * three-operation chains, bit for bit the reference's (their rounding is part of the result: integer-delay statics never forget a state error). It costs twice the filter arithmetic, so by default
 * (policy -1) it is used only while the scene leaves the machine under-filled (<= 3,072 source-channels, whole 16-frame tiles); 0 = never, 1 = whenever a block qualifies. Same state arrays as the sequential kernel: the policy may change between blocks.
 * Results agree with JPLG_MODE_PARITY to rounding (bar 1e-5 * max(|ref|, rms, 1e-3)); PARITY renderers ignore the policy. */
JPLG_API jplg_status jplg_renderer_set_time_parallel(jplg_renderer r, int32_t policy);
/* Blocks rendered through the time-parallel path so far (tests / benchmarks check which path ran). */
JPLG_API uint64_t    jplg_renderer_time_parallel_blocks(jplg_renderer r);
/* Device pointer of the renderer's partial bus for the LAST block ([frames][L], not accumulated into
 * anything): the buffer a multi-GPU caller reduces across ranks. */
JPLG_API float*      jplg_renderer_device_bus(jplg_renderer r);

/* Checkpoint / parity seeding: per-source-channel state, host pointers.
 * state [N*C][JPLG_STATE_FLOATS]: [0] smoothed delay (samples), [1] Thiran y[n-1], [2..25] crossover state
 * (X2.S1{s1,s2} X2.S2{s1,s2} X13.S1 lanes0,1{s1,s2} X13.S2 lanes0..3{s1,s2} allpass lanes0..3{s1,s2}),
 * [26..29] current band gains b1..b4, [30..31] reserved, [32..63] current mix gains (first L used).
 * rings [N*C][ring_size]: raw ring, the sample pushed at time t lives at index t & (ring_size-1);
 * `*ring_pos` = number of samples pushed so far. */
#define JPLG_STATE_FLOATS 64
JPLG_API uint32_t    jplg_renderer_ring_size(jplg_renderer r);
JPLG_API jplg_status jplg_renderer_state_export(jplg_renderer r, float* state, float* rings /* may be NULL */, uint64_t* ring_pos);
JPLG_API jplg_status jplg_renderer_state_import(jplg_renderer r, const float* state, const float* rings /* may be NULL */, uint64_t ring_pos);

/* ---------------------------------------------------------------------------------------------- */
/* Components (batched value types)                                                                 */

/* FourBandCrossover x n (replaces inc/Auralization/CrossoverFilter.h:416-535) */
JPLG_API jplg_status jplg_crossover_bank_create(jplg_context ctx, uint32_t n, float sample_rate,
                                                float f1, float f2, float f3, int32_t mode, jplg_crossover_bank* out); /* Prepare :425 */
JPLG_API void        jplg_crossover_bank_destroy(jplg_crossover_bank b);
JPLG_API jplg_status jplg_crossover_bank_reset(jplg_crossover_bank b);                                                /* Reset :433 */
/* coefficients as 27 floats: X2{g,d,-(R+g)}, X13{g[4],d[4],-(R+g)[4]}, allpass{g[4],d[4],-(R+g)[4]} */
JPLG_API void        jplg_crossover_bank_coeffs(jplg_crossover_bank b, float out[27]);
/* ProcessBlock(in, startGains, endGains, out) (:505-535): in/out [n][frames]; gains [n][4]. In place allowed. */
JPLG_API jplg_status jplg_crossover_bank_process_ramp(jplg_crossover_bank b, const float* in, const float* start_gains,
                                                      const float* end_gains, float* out, uint32_t frames, int32_t mem);
/* ProcessBlock(in, outBands) (:462-470): out_bands [n][frames][4] in band order b1..b4 */
JPLG_API jplg_status jplg_crossover_bank_split(jplg_crossover_bank b, const float* in, float* out_bands, uint32_t frames, int32_t mem);

/* DelayLine + DelayTap x n (replaces inc/Auralization/DelayLine.h:43-68,129-230; FractionalDelay.h:65-136) */
enum { JPLG_INTERP_NEAREST = 0, JPLG_INTERP_LINEAR = 1, JPLG_INTERP_THIRAN1 = 2 };
JPLG_API jplg_status jplg_delay_bank_create(jplg_context ctx, uint32_t n, uint32_t min_size, jplg_delay_bank* out);   /* DelayLine(minSize) :142 */
JPLG_API void        jplg_delay_bank_destroy(jplg_delay_bank b);
JPLG_API uint32_t    jplg_delay_bank_size(jplg_delay_bank b);                                                         /* GetSize :171 */
JPLG_API jplg_status jplg_delay_bank_clear(jplg_delay_bank b);                                                        /* Clear :175 */
/* Push (:183-197) `frames` samples per line: in [n][frames] */
JPLG_API jplg_status jplg_delay_bank_push(jplg_delay_bank b, const float* in, uint32_t frames, int32_t mem);
/* For each line: for k in [0,taps): tap.SetDelay(delays[line][k]); out[line][k] = tap.Process(line) (:53-67),
 * the tap's interpolator state carried across k (and across calls for Thiran). */
JPLG_API jplg_status jplg_delay_bank_read(jplg_delay_bank b, const float* delays, uint32_t taps, int32_t interp, float* out, int32_t mem);
/* DelayTapXFade x n (:73-122). For each line: for k in [0,taps): tap.SetCrossfade(alphas[line][k]); tap.SetDelay(delays[line][k]);
 * out[line][k] = tap.Process(line); the tap's previous / target delay pair is carried across k and across calls (both start at 0). */
JPLG_API jplg_status jplg_delay_bank_read_xfade(jplg_delay_bank b, const float* delays, const float* alphas, uint32_t taps, float* out, int32_t mem);

/* ChannelMixing free functions (replaces inc/Auralization/ChannelMixing.h:32-136), batched over `n`
 * independent (dest, source) pairs laid out back to back. */
JPLG_API jplg_status jplg_mix_add_and_apply_gain_ramp(jplg_context ctx, float* dest, const float* source, uint32_t n,
                                                      uint32_t dest_channel, uint32_t source_channel,
                                                      uint32_t dest_num_channels, uint32_t source_num_channels,
                                                      uint32_t frames, const float* gain_start, const float* gain_end, int32_t mem); /* :89-107 */
JPLG_API jplg_status jplg_mix_apply_gain(jplg_context ctx, float* dest, uint32_t n, uint32_t frames, const float* gains, int32_t mem); /* :50-63 */
/* Add (:32-48): dest[pair][f][dest_channel] += source[pair][f][source_channel]. The reference's mono-source overload (:41-48) is
 * source_channel = 0, source_num_channels = 1. */
JPLG_API jplg_status jplg_mix_add(jplg_context ctx, float* dest, const float* source, uint32_t n, uint32_t dest_channel, uint32_t source_channel,
                                  uint32_t dest_num_channels, uint32_t source_num_channels, uint32_t frames, int32_t mem);
/* DownmixToMono (:129-136): dest [n][frames] += source [n][frames][num_channels] summed in channel order; dest is not cleared and
 * nothing is normalised, like the reference. (The reference's numSamples argument is frames * num_channels.) */
JPLG_API jplg_status jplg_mix_downmix_to_mono(jplg_context ctx, float* dest, const float* source, uint32_t n, uint32_t num_channels,
                                              uint32_t frames, int32_t mem);

/* ---------------------------------------------------------------------------------------------- */
/* ReverbBus x n  (replaces: inc/Auralization/LateReverb.h:34-117, src/Auralization/LateReverb.cpp:31-134;           */
/*                 FDN inc/Auralization/ReverbDesign.h:311-349 Householder, :388-448 Hadamard, :451-492 mixers,        */
/*                 :573-704 MultiChannelMixedFeedback<16, AttenuationFilter>)                                          */

/* Prepare(sampleRate, splits) for n buses (LateReverb.cpp:42-48); f1..f3 = 0 -> reference defaults. output_channels =
 * channel count of the `output` span of ProcessInterleaved (1..16). Initial decay slope -60 dB/s (RT60 = 1 s). */
JPLG_API jplg_status jplg_reverb_bank_create(jplg_context ctx, uint32_t n, float sample_rate, float f1, float f2, float f3,
                                             uint32_t output_channels, uint32_t max_block_frames, int32_t mode, jplg_reverb_bank* out);
JPLG_API void        jplg_reverb_bank_destroy(jplg_reverb_bank b);
/* SetRT60(newRT60Seconds) (:31-35): rt60 [n][4] seconds per band, clamped to [0.03, 10]. Visible at the next process. */
JPLG_API jplg_status jplg_reverb_bank_set_rt60(jplg_reverb_bank b, const float* rt60_seconds, int32_t mem);
/* SetDecaySlope(newDecaySlope) (:37-40): [n][4] dB/s, clamped to [-2000, -6] */
JPLG_API jplg_status jplg_reverb_bank_set_decay_slope(jplg_reverb_bank b, const float* slope_db_per_second, int32_t mem);
/* AttenuationFilter::DecayGains of every FDN channel = dBToGain(slope * delaySeconds) (LateReverb.h:64-68): host out [n][16][4] */
JPLG_API jplg_status jplg_reverb_bank_decay_gains(jplg_reverb_bank b, float* out);
/* ProcessInterleaved(input, output, numFrames) for all n buses (:53-134). in [n][frames][in_channels] interleaved; multichannel input is down-mixed like
 * the reference (DownmixToMono + ApplyGain(1 / in_channels), :80-92) for blocks of at most 480 frames — beyond one chunk the reference over-runs its scratch
 * array (:83), so longer multichannel blocks are refused. out [n][frames][output_channels] interleaved, ACCUMULATED INTO like the reference's `output` span. */
JPLG_API jplg_status jplg_reverb_bank_process(jplg_reverb_bank b, const float* in, uint32_t in_channels, float* out, uint32_t frames, int32_t mem);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it). jplg_reverb_bank_process with DEVICE pointers only (`in`, `out` as above, JPLG_MEM_DEVICE implied) that
 * allocates nothing: the mono down-mix of a multichannel input goes to a file-scope device array of JPLG_REVERB_NO_ALLOC_MAX_FLOATS floats (960 KiB, held
 * by every GPU that loads the library until the process exits) instead of the context's growable scratch. instances * frames beyond that size is refused
 * with JPLG_ERR_INVALID_ARGUMENT (jplg_last_error says so); mono input does not use the scratch and has no such limit. Same kernels and results as
 * jplg_reverb_bank_process. The array is shared by every bank and context on a device and keeps the last call's contents (nothing in it needs clearing:
 * the down-mix overwrites every element the FDN kernel reads); calls that overlap on the GPU (different contexts' streams) race on it. The first call on
 * a device looks the array's address up (cudaGetSymbolAddress): make it once outside a graph capture. */
#define JPLG_REVERB_NO_ALLOC_MAX_FLOATS 245760u
JPLG_API jplg_status jplg_reverb_bank_process_no_alloc(jplg_reverb_bank b, const float* in, uint32_t in_channels, float* out, uint32_t frames);

/* ---------------------------------------------------------------------------------------------- */
/* ERBus x B  (replaces: inc/Auralization/EarlyReflectionsBus.h:45-249,                              */
/*             src/Auralization/EarlyReflectionsBus.cpp:60-225 Prepare / ProcessInterleaved, :239-417 SetTaps) */

/* Prepare(sampleRate, numChannels) for num_buses buses (:60-81). Every bus owns one mono ring (ring_size 0 = the reference's
 * bit_ceil(2 * fs)) and max_taps_per_bus tap slots (rounded up to a power of two in [8, 128]; taps that are still fading
 * out occupy a slot until the SetTaps call after their last chunk). */
JPLG_API jplg_status jplg_erbus_bank_create(jplg_context ctx, uint32_t num_buses, uint32_t max_taps_per_bus, float sample_rate,
                                            uint32_t output_channels, uint32_t ring_size, uint32_t max_block_frames, int32_t mode,
                                            jplg_erbus_bank* out);
JPLG_API void        jplg_erbus_bank_destroy(jplg_erbus_bank b);
/* SetTaps(span<const ERUpdateData>) of bus `bus` (:239-406), host pointers: filter_gains [n][4], delay_seconds [n] (ERUpdateData::Delay),
 * ids [n], gains [n][L]. Ids not in the list fade out over the next chunk and are then removed; new ids fade in from gain 0;
 * known ids get new targets. Takes effect at the next process call, like the reference's RealtimeObject hand-over. */
JPLG_API jplg_status jplg_erbus_bank_set_taps(jplg_erbus_bank b, uint32_t bus, uint32_t n, const float* filter_gains,
                                              const float* delay_seconds, const uint32_t* ids, const float* gains);
JPLG_API uint32_t    jplg_erbus_bank_num_taps(jplg_erbus_bank b, uint32_t bus);       /* GetNumTaps (incl. taps fading out) */
JPLG_API uint32_t    jplg_erbus_bank_slots_per_bus(jplg_erbus_bank b);
JPLG_API int32_t     jplg_erbus_bank_tap_slot(jplg_erbus_bank b, uint32_t bus, uint32_t id);  /* row of per_tap for this id, -1 if unknown */
/* ProcessInterleaved(input, output, numInputChannels, numFrames) for all buses (:83-225). in [B][frames][in_channels]
 * (down-mixed to mono per bus), out [B][frames][L] interleaved, ACCUMULATED INTO like the reference's `output` span. */
JPLG_API jplg_status jplg_erbus_bank_process(jplg_erbus_bank b, const float* in, uint32_t in_channels, float* out, uint32_t frames, int32_t mem);
/* Same, and every tap slot's own contribution: per_tap [B * slots_per_bus][frames][L] (overwritten; dead slots give zeros). */
JPLG_API jplg_status jplg_erbus_bank_process_debug(jplg_erbus_bank b, const float* in, uint32_t in_channels, float* out, float* per_tap,
                                                   uint32_t frames, int32_t mem);

/* ---------------------------------------------------------------------------------------------- */
/* Scene front-end: the per-source body of SpatialManager::AdvanceSimulation batched               */
/* (replaces: inc/SpatialManager.h:374-420 and the services it calls)                              */
/*   DirectPathService::ProcessDirectPath / EvaluateDistance / EvaluateDirection                   */
/*                                     inc/Services/DirectPathService.h:232-364                    */
/*   AttenuationBaseModel / AttenuationCurve ::Evaluate   inc/DistanceAttenuation.h:66-119,172-215 */
/*   GetSpreadFromSourceSize           inc/SpatialManager.h:617-623 (inc/Panning/PannerBase.h:69-74) */
/*   Impl::SanitizeSourceDir           inc/Services/PanningService.h:480-495                       */
/*   AirAbsorption::CacheParameters / ComputeForDistance    inc/AirAbsorption.h:179-234            */
/*   dBToGain(simd)                    inc/Math/DecibelsAndGain.h:40-50                            */
/*   SpeedOfSound                      inc/SpeedOfSound.h:24-47                                    */
/* Positions are 9 floats: Location, Orientation.Up, Orientation.Forward (inc/Math/Position.h:31-83). */

/* one AttenuationFunction */
typedef struct jplg_attenuation_function {
    int32_t  kind;          /* 0: AttenuationBaseModel, 1: AttenuationCurve */
    int32_t  model;         /* AttenuationModel: 0 None, 1 Inverse, 2 Linear, 3 Exponential (base model only) */
    float    min_distance, max_distance, rolloff;   /* DistanceAttenuationParameters */
    uint32_t first_point, num_points;               /* curve: its range in `points` */
} jplg_attenuation_function;

/* AirAbsorptionCache (inc/AirAbsorption.h:51-87) */
typedef struct jplg_air_absorption_cache {
    float    band_centers[4], lower_third_centers[4];
    float    center_db[4], lower_third_db[4], weighted_db[4], decay_slope[4];
    float    distance_to_knee[4], linearized_loss_factor[4];
    uint32_t has_knee[4];
} jplg_air_absorption_cache;

/* Host builders (no device needed; the reference also runs these once per environment on the CPU). */
/* AirAbsorption::CacheParameters(frequencyBands, params, nyquist); bands NULL = cBandCenters (inc/FrequencyBands.h:68). */
JPLG_API void  jplg_air_absorption_cache_build(const float* bands, float temperature_c, float humidity_percent, float pressure_kpa,
                                               float nyquist, jplg_air_absorption_cache* out);
JPLG_API void  jplg_default_band_centers(float out[4]);
JPLG_API float jplg_speed_of_sound_for_temperature(float temperature_c);   /* SpeedOfSound::ForTemperature */

typedef struct jplg_frontend_config {
    const jplg_attenuation_function* functions;  /* attenuation function table (may be NULL / 0) */
    uint32_t num_functions;
    const float* points;                         /* [num_points][3] = Distance, Value, Curve::EType (inc/Math/Curves.h:40-52) */
    uint32_t num_points;
    int32_t  sort_points;                        /* 1: sort every curve's points by distance first (AttenuationCurve::SortPoints) */
    const jplg_air_absorption_cache* air;        /* NULL = cDefaultAirAbsCache (20 C, 50 %, 101.325 kPa, default bands) */
    float    speed_of_sound;                     /* 0 = SpeedOfSound::cReference (343.2 m/s); delay = distance * (1 / c) */
    int32_t  spread_flavour;                     /* 0: SpatialManager::GetSpreadFromSourceSize, 1: the free function of PannerBase.h */
} jplg_frontend_config;

/* Arrays of one call; inputs and outputs may each be NULL (= absent / not wanted). All arrays live where `mem` says. */
typedef struct jplg_frontend_io {
    uint32_t n;                      /* sources */
    uint32_t num_listeners;
    const float*    sources;         /* [n][9] world-space Position of every source */
    const float*    listeners;       /* [num_listeners][9] */
    const uint32_t* listener_index;  /* [n], NULL = listener 0 */
    const float*    distance_in;     /* [n] distances to evaluate when sources == NULL (EvaluateDistance / ComputeForDistance on their own) */
    const float*    dot_in;          /* [n] direction cosines when sources == NULL (ProcessAngleAttenuation on its own) */
    const float*    cones;           /* [n][2] AttenuationCone: InnerAngle, OuterAngle (radians) */
    const float*    object_sizes;    /* [n] SourceData::ObjectSize; < 0 = none (spread stays as given) */
    const float*    focus_spread;    /* [n][2] PanEffectParameters; NULL = focus / spread below */
    float           focus, spread;
    const uint32_t* function_index;  /* [n] attenuation function per source, NULL = function 0 */
    uint32_t        pan_stride;      /* 5: PanUpdateData rows, 11: PanUpdateDataWithOrientation rows */
    float* direct;        /* (16-byte aligned) [n][4] Distance, DirectionDot, InvDirectionDot, cone factor (ProcessAngleAttenuationImpl(InvDirectionDot, cone)) */
    float* relative;      /* [n][9] DirectPathResult::Position */
    float* pan_params;    /* [n][pan_stride]: what jplg_panner_mdap / jplg_panner_mdap_oriented consume */
    float* distance_attenuation; /* [n] AttenuationFunction::Evaluate(Distance) */
    float* air_db;        /* [n][4] AirAbsorption::ComputeForDistance */
    float* band_gains;    /* [n][4] dBToGain(-air_db): the filterGains of jplg_renderer_update */
    float* delay_seconds; /* [n] Distance / speed of sound: the delaySeconds of jplg_renderer_update */
} jplg_frontend_io;

JPLG_API jplg_status jplg_frontend_create(jplg_context ctx, const jplg_frontend_config* cfg, jplg_frontend* out);
JPLG_API void        jplg_frontend_destroy(jplg_frontend f);
/* One kernel launch for all n sources. JPLG_MEM_DEVICE: asynchronous on the context stream (outputs can feed
 * jplg_panner_mdap / jplg_renderer_update directly); JPLG_MEM_HOST: staged and synchronous. */
JPLG_API jplg_status jplg_frontend_process(jplg_frontend f, const jplg_frontend_io* io, int32_t mem);
/* the fallback direction SanitizeSourceDir substitutes for straight-up / straight-down sources */
JPLG_API void        jplg_sanitize_fallback_dir(float out[3]);

/* ---------------------------------------------------------------------------------------------- */
/* Multi-GPU: peer-memory combine of the per-GPU partial buses (no counterpart in the reference,   */
/* which is a single process). One process per GPU; the ROOT rank (0) owns a gather  */
/* buffer, every other rank maps it through CUDA IPC and stores its bus into its slot over NVLink; */
/* the root sums the slots in rank order (deterministic). Nothing but the 64-byte handle has to    */
/* travel through the host application (e.g. torch.distributed broadcast).                          */

#define JPLG_IPC_HANDLE_BYTES 64
/* rank 0: allocates the gather buffer for `world` ranks x two blocks in flight x max_floats, returns the IPC handle to distribute */
JPLG_API jplg_status jplg_bus_exchange_create_root(jplg_context ctx, uint32_t world, uint32_t max_floats, jplg_bus_exchange* out,
                                                   uint8_t handle_out[JPLG_IPC_HANDLE_BYTES]);
/* rank != 0: maps the root's buffer (the root's GPU must be peer-accessible from this context's GPU) */
JPLG_API jplg_status jplg_bus_exchange_open_peer(jplg_context ctx, uint32_t world, uint32_t rank, uint32_t max_floats,
                                                 const uint8_t handle[JPLG_IPC_HANDLE_BYTES], jplg_bus_exchange* out);
JPLG_API void        jplg_bus_exchange_destroy(jplg_bus_exchange x);
/* Once per block and rank with the same `seq` (0, 1, 2, ...) on every rank. `device_bus`: this rank's partial
 * bus (n floats, 16-byte aligned, e.g. jplg_renderer_device_bus). Rank != 0: waits (on the device) until the root has consumed block seq-2,
 * stores the bus into the root's slot and publishes it. Rank 0: waits for every rank's block `seq`, then device_out[i] (+)= own[i] + sum over
 * ranks in rank order (`accumulate` != 0 adds into device_out like the renderer's bus), and acknowledges.
 * The call copies `device_bus` into a private staging slot on the context stream and runs the peer traffic on a side stream, so the context
 * stream continues with the next block while the root waits for the other ranks; device_out is complete after jplg_bus_exchange_wait. */
JPLG_API jplg_status jplg_bus_exchange_submit(jplg_bus_exchange x, const float* device_bus, uint32_t n, uint64_t seq, float* device_out,
                                              int32_t accumulate);
/* makes the context stream wait (on the device, not the host) for every block submitted so far */
JPLG_API jplg_status jplg_bus_exchange_wait(jplg_bus_exchange x);
/* 0 = healthy; != 0 after a device-side wait timed out (a rank died or stalled for longer than the time-out): nothing hangs and nothing stale is
 * summed — a sender that timed out stores / publishes nothing, a gather that timed out (or runs after an error) leaves device_out untouched and only
 * acknowledges. The state is sticky until jplg_bus_exchange_clear_error. This call may read the root's header (a blocking 4-byte copy);
 * jplg_bus_exchange_poll_error only reads this rank's own host-mapped error word (cheap enough to call after every block). */
JPLG_API int32_t     jplg_bus_exchange_error(jplg_bus_exchange x);
JPLG_API int32_t     jplg_bus_exchange_poll_error(jplg_bus_exchange x);
/* device-side wait time-out of this rank's exchange kernels, default 2,000 ms; set it before the first submit or between blocks */
JPLG_API jplg_status jplg_bus_exchange_set_timeout_ms(jplg_bus_exchange x, uint32_t milliseconds);
/* collective recovery: every rank drains its streams and clears its error word, the root also zeroes the header (flags, acknowledgement); call it on
 * every rank between two host-side barriers, then restart `seq` at 0 */
JPLG_API jplg_status jplg_bus_exchange_clear_error(jplg_bus_exchange x);

#ifdef __cplusplus
}
#endif
#endif /* JPLG_H */
