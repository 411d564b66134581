// C-ABI layer (include/jplg.h): handles, host<->device staging, launch sequencing. No compute happens on the host:
// every data-path call fails with JPLG_ERR_NO_DEVICE when the context has no GPU.
#include "../../include/jplg.h"
#include "jplg_host.hpp"
#include "jplg_kernels.hpp"

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace jplg;

static jplg_trace_fn gTrace = nullptr;

struct jplg_context_t
{
	int device = -1;              // -1: host-only context (builders work, data-path calls fail loudly)
	cudaStream_t stream = nullptr;
	int numSMs = 0;
	uint32_t* dRsq = nullptr;
	bool rsqVerified = false;
	bool sticky = false;
	uint64_t launches = 0;
	std::string lastError;
	// growable device scratch for JPLG_MEM_HOST calls
	void* scratch[4] = { nullptr, nullptr, nullptr, nullptr };
	size_t scratchBytes[4] = { 0, 0, 0, 0 };
	std::mutex mtx;
	// optional per-kernel timing (CUDA events on the launching stream)
	struct Timed { int which; cudaEvent_t a, b; };
	bool timing = false;
	bool timedThis = false;
	uint32_t timingMask = 0;   // bit (1 << JPLG_K_*) = that kernel is bracketed by events
	std::vector<Timed> pending;
	std::vector<cudaEvent_t> freeEvents;
	double totalMs[JPLG_K_COUNT] = {};
	uint64_t counts[JPLG_K_COUNT] = {};
	// host-path pipelining (JPLG_MEM_HOST process_block): slices of sources are copied / pushed / rendered on auxiliary streams
	// so the PCIe transfer of slice k+1 overlaps the render of slice k. Created on first use.
	static constexpr int kAux = 4, kMaxSlices = 16;
	cudaStream_t aux[kAux] = {};
	cudaEvent_t evStart = nullptr, evCopied[kMaxSlices] = {}, evDone[kAux] = {};
	bool auxReady = false;
	// 0 (default): single stream (copy, then compute); 1: sliced DMA + concurrent slices; 2: like 1, but pinned input is read in place by k_push (zero-copy).
	// Measured at 16,384 sources (r01b): process_block 797 / 818 / 1072 us for modes 0 / 1 / 2 against a 610 us PCIe floor (55 GB/s): the render of the
	// last slice is as long as the render of the whole block (latency-bound per warp), so slicing only hides ~100 us and its event traffic costs as much.
	int hostPipeline = 0;
};

static void TimeBegin(jplg_context ctx, int which, cudaStream_t onStream = nullptr)
{
	const cudaStream_t ts = onStream ? onStream : ctx->stream;
	if (!ctx->timing || !((ctx->timingMask >> which) & 1u)) { ctx->timedThis = false; return; }
	ctx->timedThis = true;
	cudaEvent_t ev[2];
	for (auto& e : ev)
	{
		if (!ctx->freeEvents.empty()) { e = ctx->freeEvents.back(); ctx->freeEvents.pop_back(); }
		else cudaEventCreate(&e);
	}
	cudaEventRecord(ev[0], ts);
	ctx->pending.push_back({ which, ev[0], ev[1] });
}
static void TimeEnd(jplg_context ctx, cudaStream_t onStream = nullptr)
{
	if (!ctx->timing || !ctx->timedThis || ctx->pending.empty()) return;
	cudaEventRecord(ctx->pending.back().b, onStream ? onStream : ctx->stream);
}
static void TimeCollect(jplg_context ctx)
{
	if (ctx->pending.empty()) return;
	cudaStreamSynchronize(ctx->stream);
	if (ctx->auxReady) for (auto& a : ctx->aux) cudaStreamSynchronize(a);
	for (auto& t : ctx->pending)
	{
		float ms = 0.0f;
		if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) { ctx->totalMs[t.which] += ms; ctx->counts[t.which]++; }
		else cudaGetLastError();
		ctx->freeEvents.push_back(t.a); ctx->freeEvents.push_back(t.b);
	}
	ctx->pending.clear();
}
// launch + count + (optionally) time one kernel
#define JPLG_LAUNCH(ctx, which, expr) JPLG_LAUNCH_ON(ctx, which, nullptr, expr)
#define JPLG_LAUNCH_ON(ctx, which, onStream, expr) do { TimeBegin((ctx), (which), (onStream)); cudaError_t _le = (expr); TimeEnd((ctx), (onStream)); (ctx)->launches++; \
	if ((ctx)->timing && (ctx)->pending.size() > 4096) TimeCollect((ctx)); \
	if (_le != cudaSuccess) { (ctx)->sticky = true; return Fail((ctx), JPLG_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_le)); } } while (0)

static jplg_status Fail(jplg_context ctx, jplg_status st, const std::string& msg)
{
	if (ctx) ctx->lastError = msg;
	if (gTrace) gTrace(msg.c_str());
	return st;
}

#define JPLG_CUDA_TRY(ctx, expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { (ctx)->sticky = true; \
	return Fail((ctx), JPLG_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); } } while (0)
#define JPLG_NEED_DEVICE(ctx) do { if (!(ctx) || (ctx)->device < 0) return Fail((ctx), JPLG_ERR_NO_DEVICE, \
	"no CUDA device in this context: the library has no CPU fallback"); if ((ctx)->sticky) return Fail((ctx), JPLG_ERR_CUDA, "context is in a sticky CUDA error state: " + (ctx)->lastError); } while (0)

static jplg_status Scratch(jplg_context ctx, int slot, size_t bytes, void** out)
{
	if (bytes > ctx->scratchBytes[slot])
	{
		if (ctx->scratch[slot]) { JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(ctx->scratch[slot]); ctx->scratch[slot] = nullptr; ctx->scratchBytes[slot] = 0; }
		const size_t want = bytes + bytes / 4 + 256;
		if (cudaMalloc(&ctx->scratch[slot], want) != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "device scratch allocation failed"); }
		ctx->scratchBytes[slot] = want;
	}
	*out = ctx->scratch[slot];
	return JPLG_OK;
}

extern "C" {

void jplg_set_trace_callback(jplg_trace_fn fn) { gTrace = fn; }
const char* jplg_version(void) { return "jplspatial_b200 0.1 (sm_100a)"; }
int32_t jplg_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

//==================================================================================================
jplg_status jplg_context_create(int32_t device, jplg_context* out) { return jplg_context_create_ex(device, 0u, out); }

jplg_status jplg_context_create_ex(int32_t device, uint32_t flags, jplg_context* out)
{
	if (!out || (flags & ~uint32_t(JPLG_CONTEXT_BACKGROUND))) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	auto* ctx = new jplg_context_t();
	static uint32_t sTable[2048];
	static bool sVerified = false, sBuilt = false;
	static std::mutex sMtx;
	{
		std::lock_guard<std::mutex> lk(sMtx);
		if (!sBuilt) { sVerified = BuildRsqrtTable(sTable); sBuilt = true; }
	}
	ctx->rsqVerified = sVerified;
	if (const char* hp = std::getenv("JPLG_HOST_PIPELINE")) ctx->hostPipeline = std::atoi(hp); // measurement knob: 0 / 1 / 2 (see jplg_context_t)
	if (device < 0) { ctx->device = -1; *out = ctx; return JPLG_OK; }
	if (device >= jplg_device_count()) { delete ctx; if (gTrace) gTrace("jplg_context_create: no such CUDA device"); return JPLG_ERR_NO_DEVICE; }
	ctx->device = device;
	cudaError_t e = cudaSetDevice(device);
	int prioLeast = 0, prioGreatest = 0;
	if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prioLeast, &prioGreatest);
	// audio-rate contexts dispatch ahead of background ones (and of the library's own copy / push-ahead side streams, which stay at the default)
	if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, (flags & JPLG_CONTEXT_BACKGROUND) ? prioLeast : prioGreatest);
	if (e == cudaSuccess) e = cudaDeviceGetAttribute(&ctx->numSMs, cudaDevAttrMultiProcessorCount, device);
	if (e == cudaSuccess) e = cudaMalloc(&ctx->dRsq, sizeof(sTable));
	if (e == cudaSuccess) e = cudaMemcpy(ctx->dRsq, sTable, sizeof(sTable), cudaMemcpyHostToDevice);
	if (e != cudaSuccess) { if (gTrace) gTrace(cudaGetErrorString(e)); delete ctx; return JPLG_ERR_CUDA; }
	*out = ctx;
	return JPLG_OK;
}

void jplg_context_destroy(jplg_context ctx)
{
	if (!ctx) return;
	if (ctx->device >= 0)
	{
		cudaSetDevice(ctx->device);
		if (ctx->stream) cudaStreamSynchronize(ctx->stream);
		TimeCollect(ctx);
		for (auto& e : ctx->freeEvents) cudaEventDestroy(e);
		if (ctx->auxReady)
		{
			for (auto& a : ctx->aux) { cudaStreamSynchronize(a); cudaStreamDestroy(a); }
			cudaEventDestroy(ctx->evStart);
			for (auto& e : ctx->evCopied) cudaEventDestroy(e);
			for (auto& e : ctx->evDone) cudaEventDestroy(e);
		}
		for (auto& s : ctx->scratch) if (s) cudaFree(s);
		if (ctx->dRsq) cudaFree(ctx->dRsq);
		if (ctx->stream) cudaStreamDestroy(ctx->stream);
	}
	delete ctx;
}
const char* jplg_last_error(jplg_context ctx) { return ctx ? ctx->lastError.c_str() : "null context"; }
void* jplg_context_stream(jplg_context ctx) { return ctx ? static_cast<void*>(ctx->stream) : nullptr; }
jplg_status jplg_context_synchronize(jplg_context ctx)
{
	JPLG_NEED_DEVICE(ctx);
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	return JPLG_OK;
}
int32_t jplg_context_rsqrt_model_verified(jplg_context ctx) { return ctx && ctx->rsqVerified ? 1 : 0; }
uint64_t jplg_context_launch_count(jplg_context ctx) { return ctx ? ctx->launches : 0; }
void jplg_context_set_kernel_timing(jplg_context ctx, int32_t enabled)
{
	if (!ctx || ctx->device < 0) return;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	if (!enabled) TimeCollect(ctx);
	ctx->timing = enabled != 0;
	ctx->timingMask = enabled == 1 ? 0xffffffffu : static_cast<uint32_t>(enabled) >> 1;   // 1 = all kernels, else (mask << 1)
}
void jplg_context_reset_kernel_timing(jplg_context ctx)
{
	if (!ctx || ctx->device < 0) return;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	TimeCollect(ctx);
	for (int i = 0; i < JPLG_K_COUNT; ++i) { ctx->totalMs[i] = 0.0; ctx->counts[i] = 0; }
}
jplg_status jplg_context_kernel_time(jplg_context ctx, int32_t which, double* total_ms, uint64_t* launches)
{
	if (!ctx || which < 0 || which >= JPLG_K_COUNT) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	TimeCollect(ctx);
	if (total_ms) *total_ms = ctx->totalMs[which];
	if (launches) *launches = ctx->counts[which];
	return JPLG_OK;
}

jplg_status jplg_context_probe_fp32(jplg_context ctx, double* lane_ops_per_second, double* kernel_ms)
{
	if (!ctx || !lane_ops_per_second) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	void* sink = nullptr;
	jplg_status st = Scratch(ctx, 0, 256, &sink); if (st) return st;
	cudaEvent_t a, b;
	JPLG_CUDA_TRY(ctx, cudaEventCreate(&a)); JPLG_CUDA_TRY(ctx, cudaEventCreate(&b));
	const uint32_t grid = static_cast<uint32_t>(ctx->numSMs), iters = 20000;
	double best = 1e30;
	for (int rep = 0; rep < 4; ++rep)     // first launch warms up; best of the following three
	{
		JPLG_CUDA_TRY(ctx, cudaEventRecord(a, ctx->stream));
		JPLG_CUDA_TRY(ctx, LaunchFp32Probe(static_cast<float*>(sink), grid, iters, ctx->stream)); ctx->launches++;
		JPLG_CUDA_TRY(ctx, cudaEventRecord(b, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaEventSynchronize(b));
		float ms = 0.0f; JPLG_CUDA_TRY(ctx, cudaEventElapsedTime(&ms, a, b));
		if (rep > 0 && ms < best) best = ms;
	}
	cudaEventDestroy(a); cudaEventDestroy(b);
	const double laneOps = double(grid) * 512.0 * double(iters) * 16.0 * 2.0;
	*lane_ops_per_second = laneOps / (best * 1e-3);
	if (kernel_ms) *kernel_ms = best;
	return JPLG_OK;
}

//==================================================================================================
uint32_t jplg_channel_map_num_channels(uint32_t mask) { return NumChannels(mask); }
int32_t jplg_channel_map_has_lfe(uint32_t mask) { return HasLFE(mask) ? 1 : 0; }
int32_t jplg_channel_map_has_top_channels(uint32_t mask) { return HasTopChannels(mask) ? 1 : 0; }
uint32_t jplg_channel_map_channel_index(uint32_t mask, uint32_t bit) { return ChannelIndex(mask, bit); }
uint32_t jplg_channel_map_from_num_channels(uint32_t n) { return MaskFromNumChannels(n); }

} // extern "C"

//==================================================================================================
struct jplg_panner_t
{
	jplg_context ctx = nullptr;
	PannerData host;
	uint4* dLut3d = nullptr;
	float* dLut2d = nullptr;
	PannerDev dev;
};

struct jplg_source_layout_t
{
	jplg_panner panner = nullptr;
	LayoutData host;
	MdapLayoutParams lp;
};

static jplg_status UploadLUT(jplg_panner p)
{
	jplg_context ctx = p->ctx;
	if (ctx->device < 0) return JPLG_OK;
	cudaSetDevice(ctx->device);
	if (p->host.is3D)
	{
		std::vector<uint4> cells(kLUT3DCells);
		for (uint32_t i = 0; i < kLUT3DCells; ++i)
		{
			uint32_t g[3]; std::memcpy(g, &p->host.lut3dGains[size_t(i) * 3], 12);
			const uint8_t* s = &p->host.lut3dSpeakers[size_t(i) * 3];
			cells[i] = make_uint4(g[0], g[1], g[2], uint32_t(s[0]) | (uint32_t(s[1]) << 8) | (uint32_t(s[2]) << 16));
		}
		if (!p->dLut3d) JPLG_CUDA_TRY(ctx, cudaMalloc(&p->dLut3d, sizeof(uint4) * kLUT3DCells));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(p->dLut3d, cells.data(), sizeof(uint4) * kLUT3DCells, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	else
	{
		const size_t bytes = p->host.lut2d.size() * sizeof(float);
		if (!p->dLut2d) JPLG_CUDA_TRY(ctx, cudaMalloc(&p->dLut2d, bytes));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(p->dLut2d, p->host.lut2d.data(), bytes, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	p->dev.is3D = p->host.is3D;
	p->dev.numChannels = p->host.numChannels;
	p->dev.lut3d = p->dLut3d;
	p->dev.lut2d = p->dLut2d;
	p->dev.rsqTable = ctx->dRsq;
	return JPLG_OK;
}

extern "C" {

jplg_status jplg_panner_create_ex(jplg_context ctx, uint32_t target_mask, int32_t lut_variant, jplg_panner* out)
{
	if (!ctx || !out) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	if (lut_variant == JPLG_LUT_KB_655 || lut_variant == JPLG_LUT_KB_65)
		return Fail(ctx, JPLG_ERR_UNSUPPORTED, lut_variant == JPLG_LUT_KB_655
			? "LUT variant KB_655 combines KB_851's triplet indirection (undefined with an LFE, VBAPLUT3D.h:618-620) with KB_786's gain packing: use one of those"
			: "LUT variant KB_65 returns NaN rows in the reference (gains recomputed per query from a triangle table indexed like KB_851's): not built");
	if (lut_variant < JPLG_LUT_KB_983 || lut_variant > JPLG_LUT_KB_65) return JPLG_ERR_INVALID_ARGUMENT;
	auto* p = new jplg_panner_t();
	p->ctx = ctx;
	std::string err;
	if (!BuildPanner(target_mask, p->host, err, lut_variant))
	{
		delete p;
		if (err.find("LUT variant") != std::string::npos || err.find("LUT storage") != std::string::npos) return Fail(ctx, JPLG_ERR_UNSUPPORTED, "VBAPannerBase: " + err);
		return Fail(ctx, err.find("triangulate") != std::string::npos ? JPLG_ERR_TRIANGULATION : JPLG_ERR_INVALID_CHANNEL_MAP, "VBAPannerBase: " + err);
	}
	const jplg_status st = UploadLUT(p);
	if (st != JPLG_OK) { delete p; return st; }
	*out = p;
	return JPLG_OK;
}
jplg_status jplg_panner_create(jplg_context ctx, uint32_t target_mask, jplg_panner* out) { return jplg_panner_create_ex(ctx, target_mask, JPLG_LUT_KB_983, out); }
void jplg_panner_destroy(jplg_panner p)
{
	if (!p) return;
	if (p->ctx->device >= 0) { cudaSetDevice(p->ctx->device); cudaStreamSynchronize(p->ctx->stream); if (p->dLut3d) cudaFree(p->dLut3d); if (p->dLut2d) cudaFree(p->dLut2d); }
	delete p;
}
int32_t jplg_panner_is_3d(jplg_panner p) { return p && p->host.is3D ? 1 : 0; }
uint32_t jplg_panner_num_channels(jplg_panner p) { return p ? p->host.numChannels : 0; }
uint32_t jplg_panner_channel_mask(jplg_panner p) { return p ? p->host.mask : 0; }
float jplg_panner_shortest_aperture(jplg_panner p) { return p ? p->host.shortestAperture : 0.0f; }
int32_t jplg_panner_triangulation_is_tabulated(jplg_panner p) { return p && p->host.trisTabulated ? 1 : 0; }
uint32_t jplg_panner_triangles(jplg_panner p, uint8_t* out_tris)
{
	if (!p) return 0;
	if (out_tris && !p->host.tris.empty()) std::memcpy(out_tris, p->host.tris.data(), p->host.tris.size());
	return static_cast<uint32_t>(p->host.tris.size() / 3);
}

jplg_status jplg_panner_get_lut(jplg_panner p, float* gains, uint8_t* speakers)
{
	if (!p || !gains) return JPLG_ERR_INVALID_ARGUMENT;
	if (p->host.is3D)
	{
		std::memcpy(gains, p->host.lut3dGains.data(), p->host.lut3dGains.size() * sizeof(float));
		if (speakers) std::memcpy(speakers, p->host.lut3dSpeakers.data(), p->host.lut3dSpeakers.size());
	}
	else std::memcpy(gains, p->host.lut2d.data(), p->host.lut2d.size() * sizeof(float));
	return JPLG_OK;
}
jplg_status jplg_panner_set_lut(jplg_panner p, const float* gains, const uint8_t* speakers)
{
	if (!p || !gains) return JPLG_ERR_INVALID_ARGUMENT;
	if (p->host.is3D)
	{
		if (!speakers) return JPLG_ERR_INVALID_ARGUMENT;
		std::memcpy(p->host.lut3dGains.data(), gains, p->host.lut3dGains.size() * sizeof(float));
		std::memcpy(p->host.lut3dSpeakers.data(), speakers, p->host.lut3dSpeakers.size());
	}
	else std::memcpy(p->host.lut2d.data(), gains, p->host.lut2d.size() * sizeof(float));
	return UploadLUT(p);
}

jplg_status jplg_panner_point_gains(jplg_panner p, const float* dirs, uint32_t n, float* out, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!p || (n && (!dirs || !out))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = p->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (n == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const uint32_t L = p->host.numChannels;
	const float* dDirs = dirs; float* dOut = out;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *b;
		jplg_status st = Scratch(ctx, 0, size_t(n) * 3 * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, size_t(n) * L * 4, &b); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, dirs, size_t(n) * 3 * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDirs = static_cast<float*>(a); dOut = static_cast<float*>(b);
	}
	JPLG_LAUNCH(ctx, JPLG_K_POINT_GAINS, LaunchPointGains(p->dev, dDirs, n, dOut, ctx->stream));
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, size_t(n) * L * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

jplg_status jplg_source_layout_create(jplg_panner p, uint32_t source_mask, jplg_source_layout* out)
{
	if (!p || !out) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	auto* l = new jplg_source_layout_t();
	l->panner = p;
	std::string err;
	if (!BuildSourceLayout(p->host, source_mask, l->host, err)) { delete l; return Fail(p->ctx, JPLG_ERR_INVALID_CHANNEL_MAP, "SourceLayout: " + err); }
	MdapLayoutParams& lp = l->lp;
	std::memset(&lp, 0, sizeof(lp));
	lp.numChannels = l->host.numChannels; lp.numVS = l->host.numVS; lp.numRings = l->host.numRings; lp.samplesPerRing = l->host.samplesPerRing;
	const uint32_t numSIMDOps = l->host.samplesPerRing >> 2;
	lp.halfSamples = (numSIMDOps & ~1u) >> 1; lp.tail = numSIMDOps & 1u;
	lp.phiTerm = l->host.phiTerm; lp.offsetTerm = l->host.offsetTerm; lp.vsWeight = l->host.vsWeight;
	std::memcpy(lp.quats, l->host.quats, sizeof(lp.quats));
	if (l->host.is3D)
	{
		if (l->host.azSin.size() > kMaxAzimuthFloats || l->host.numRings > 24)
		{
			delete l;
			return Fail(p->ctx, JPLG_ERR_UNSUPPORTED, "SourceLayout: spread-cap dimensions exceed the kernel's tables (rings > 24 or samples per ring > 128)");
		}
		for (size_t i = 0; i < l->host.azSin.size(); ++i) { lp.azSin[i] = l->host.azSin[i]; lp.azCos[i] = l->host.azCos[i]; }
	}
	*out = l;
	return JPLG_OK;
}
void jplg_source_layout_destroy(jplg_source_layout l) { delete l; }
uint32_t jplg_source_layout_num_channels(jplg_source_layout l) { return l ? l->host.numChannels : 0; }
uint32_t jplg_source_layout_num_virtual_sources(jplg_source_layout l) { return l ? l->host.numVS : 0; }
void jplg_source_layout_info(jplg_source_layout l, uint32_t info[4], float* phi_term, float* offset_term, float* vs_weight)
{
	if (!l) return;
	info[0] = l->host.numChannels; info[1] = l->host.numVS; info[2] = l->host.numRings; info[3] = l->host.samplesPerRing;
	if (phi_term) *phi_term = l->host.phiTerm;
	if (offset_term) *offset_term = l->host.offsetTerm;
	if (vs_weight) *vs_weight = l->host.vsWeight;
}
void jplg_source_layout_groups(jplg_source_layout l, float* out_quats, uint32_t* out_channel_ids)
{
	if (!l) return;
	for (uint32_t i = 0; i < l->host.numChannels; ++i)
	{
		if (out_quats) std::memcpy(out_quats + 4 * i, l->host.quats[i], 16);
		if (out_channel_ids) out_channel_ids[i] = l->host.channelIds[i];
	}
}

// JPLG_MEM_HOST_ASYNC with page-locked caller buffers (cudaHostAlloc / cudaHostRegister): under unified addressing such a buffer has a device alias,
// so the small per-block arrays (pan parameters, band gains, delays, mix maps, the bus) are read / written by the kernels themselves over PCIe
// instead of queueing behind the 33.5 MB audio transfer on the one host-to-device copy engine (measured: 0.15 ms of a 0.86 ms step).
// Returns nullptr for pageable memory (the staged copy path is used then).
static float* MappedHostAliasRaw(const void* hostPtr)
{
	if (!hostPtr) return nullptr;
	cudaPointerAttributes at{};
	if (cudaPointerGetAttributes(&at, hostPtr) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	if (at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
	return static_cast<float*>(at.devicePointer);
}
static const float* MappedHostAlias(const float* p) { return MappedHostAliasRaw(p); }

static jplg_status MdapImpl(jplg_panner p, jplg_source_layout l, const float* params, uint32_t stride, uint32_t n, float* out, int32_t mem)
{
	if (!p || !l || l->panner != p || (n && (!params || !out))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = p->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (n == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t outFloats = size_t(n) * l->host.numChannels * p->host.numChannels;
	const float* dParams = params; float* dOut = out;
	if (mem == JPLG_MEM_HOST || mem == JPLG_MEM_HOST_ASYNC)
	{
		void *a, *b;
		jplg_status st = Scratch(ctx, 1, outFloats * 4, &b); if (st) return st;
		dOut = static_cast<float*>(b);      // the kernel rescales its own output in place: staged on the device, one D2H copy
		const float* alias = mem == JPLG_MEM_HOST_ASYNC ? MappedHostAlias(params) : nullptr;
		if (alias) dParams = alias;
		else
		{
			st = Scratch(ctx, 0, size_t(n) * stride * 4, &a); if (st) return st;
			JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, params, size_t(n) * stride * 4, cudaMemcpyHostToDevice, ctx->stream));
			dParams = static_cast<float*>(a);
		}
	}
	JPLG_LAUNCH(ctx, JPLG_K_MDAP, LaunchMdap(p->dev, l->lp, dParams, stride, n, dOut, ctx->numSMs, ctx->stream));
	if (mem == JPLG_MEM_HOST || mem == JPLG_MEM_HOST_ASYNC)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, outFloats * 4, cudaMemcpyDeviceToHost, ctx->stream));
		if (mem == JPLG_MEM_HOST) JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}
jplg_status jplg_panner_mdap(jplg_panner p, jplg_source_layout l, const float* params, uint32_t n, float* out, int32_t mem)
{
	return MdapImpl(p, l, params, 5, n, out, mem);
}
jplg_status jplg_panner_mdap_oriented(jplg_panner p, jplg_source_layout l, const float* params, uint32_t n, float* out, int32_t mem)
{
	return MdapImpl(p, l, params, 11, n, out, mem);
}

jplg_status jplg_panner_process_virtual_sources(jplg_panner p, const float* vs, uint32_t n, uint32_t batches, float* out, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!p || (batches && (!out || (n && !vs)))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = p->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (batches == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const uint32_t L = p->host.numChannels;
	const float* dVs = vs; float* dOut = out;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *b;
		jplg_status st = Scratch(ctx, 0, size_t(batches) * n * 16 + 16, &a); if (st) return st;
		st = Scratch(ctx, 1, size_t(batches) * L * 4, &b); if (st) return st;
		if (n) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, vs, size_t(batches) * n * 16, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b, out, size_t(batches) * L * 4, cudaMemcpyHostToDevice, ctx->stream)); // accumulated into
		dVs = static_cast<float*>(a); dOut = static_cast<float*>(b);
	}
	JPLG_LAUNCH(ctx, JPLG_K_OTHER, LaunchProcessVS(p->dev, dVs, n, batches, dOut, ctx->stream));
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, size_t(batches) * L * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

void jplg_normalize_weights(float* vs, uint32_t n)
{
	// VBAPannerBase::NormalizeWeights (inc/Panning/PannerBase.h:879-892)
	float sum = 0.0f;
	for (uint32_t i = 0; i < n; ++i) { const float w = vs[4 * i + 3]; sum += (w < 0 ? -w : w); }
	const float invSum = sum == 0.0f ? 0.0f : 1.0f / sum;
	for (uint32_t i = 0; i < n; ++i) vs[4 * i + 3] *= invSum;
}
float jplg_spread_from_source_size(float source_size, float distance)
{
	// GetSpreadFromSourceSize (inc/Panning/PannerBase.h:69-74)
	return distance <= 0.0f ? 1.0f : atanf((source_size * 0.5f) / distance) * (0.318309886183790671538f * 2.0f);
}

} // extern "C"

//==================================================================================================
// Renderer
struct jplg_renderer_t
{
	jplg_context ctx = nullptr;
	jplg_renderer_config cfg{};
	uint32_t numSC = 0, ringSize = 0, maxFrames = 0, numCTAs = 0;
	uint64_t pos = 0;
	float alpha = 0.0f, coef[27] = {};
	float *state = nullptr, *curMix = nullptr, *tgtDelay = nullptr, *tgtBand = nullptr, *tgtMix = nullptr, *rings = nullptr;
	float *partial = nullptr, *deviceBus = nullptr, *stageIn = nullptr, *stageBus = nullptr, *stagePerSource = nullptr;
	float *stageParams = nullptr; // update staging: band [N][4] | delay [N] | mix [N][C*L]
	size_t stagePerSourceFloats = 0;
	// JPLG_MEM_HOST_ASYNC: two input staging buffers (slot 0 = stageIn) fed by the copy stream, so the PCIe transfer of block k+1 overlaps the
	// render of block k; evPushed[s] = k_push has consumed staging slot s, evDone[s] = the block's bus is back in host memory
	float* stageIn2 = nullptr;
	cudaEvent_t evCopiedA[2] = {}, evPushedA[2] = {}, evDoneA[2] = {};
	bool asyncReady = false;
	uint64_t asyncSubmitted = 0, asyncWaited = 0;
	// jplg_renderer_push_ahead: the next block's ring pushes, issued early on a side stream
	const float* aheadIn = nullptr;
	uint32_t aheadFrames = 0;
	cudaEvent_t evAhead = nullptr;
	cudaEvent_t evStarted = nullptr;   // launch-completion event of the last whole-grid k_render: fires when all its CTAs are resident
	bool startedRecorded = false;
	// jplg_renderer_update_panned: a second set of target arrays, written (possibly by another context's stream) while the render of the
	// current block still reads the active set; the next process_block swaps the sets. evSetFree[id] = the last render that read set id is done.
	float *tgt2Delay = nullptr, *tgt2Band = nullptr, *tgt2Mix = nullptr;
	int activeSet = 0;
	bool pendingPanned = false, setFreeRecorded[2] = { false, false };
	cudaEvent_t evUpdated = nullptr, evSetFree[2] = {};
	// upper bound of the target delays in samples, as far as the HOST knows it: the maximum of host-side updates, or what the caller declared with
	// jplg_renderer_set_max_delay_seconds for delays that only exist on the device (jplg_renderer_update(.., JPLG_MEM_DEVICE), update_panned).
	// jplg_renderer_push_ahead needs it: the next block's pushes and the current render touch disjoint ring slots only while
	// maxDelay + 2 * TotalPushes + 480 < ringSize.
	float maxDelaySamples = 0.0f, declaredMaxDelaySamples = -1.0f;
	bool delayBoundKnown = true;
	// time-parallel throughput path (jplg_scan_kernels.cu): -1 = chosen by size, 0 = never, 1 = whenever the block qualifies
	int32_t scanPolicy = -1;
	float *scanM = nullptr, *scanScratch = nullptr;
	uint32_t* scanCounters = nullptr;
	uint64_t scanBlocks = 0;
};

// targets in the reference's layouts -> SoA per source-channel (UpdateParameters, src/Auralization/DirectSound.cpp:231-239)
__global__ void k_update_targets(const float* __restrict__ band, const float* __restrict__ delaySec, const float* __restrict__ mix,
								 uint32_t numSC, uint32_t C, uint32_t L, float sampleRate,
								 float* __restrict__ tgtBand, float* __restrict__ tgtDelay, float* __restrict__ tgtMix)
{
	const uint32_t sc = blockIdx.x * blockDim.x + threadIdx.x;
	if (sc >= numSC) return;
	const uint32_t src = sc / C, k = sc - src * C;
	if (band) for (uint32_t b = 0; b < 4; ++b) tgtBand[size_t(b) * numSC + sc] = band[size_t(src) * 4 + b];
	if (delaySec) tgtDelay[sc] = delaySec[src] * sampleRate;
	if (mix) for (uint32_t c = 0; c < L; ++c) tgtMix[size_t(c) * numSC + sc] = mix[(size_t(src) * C + k) * L + c];
}
// constructor semantics: current = target (DirectSound.cpp:48-63)
__global__ void k_init_current(uint32_t numSC, uint32_t L, const float* __restrict__ tgtBand, const float* __restrict__ tgtDelay, const float* __restrict__ tgtMix,
							   float* __restrict__ state, float* __restrict__ curMix)
{
	const uint32_t sc = blockIdx.x * blockDim.x + threadIdx.x;
	if (sc >= numSC) return;
	state[size_t(ST_DELAY) * numSC + sc] = tgtDelay[sc];
	for (uint32_t b = 0; b < 4; ++b) state[size_t(ST_BAND + b) * numSC + sc] = tgtBand[size_t(b) * numSC + sc];
	for (uint32_t c = 0; c < L; ++c) curMix[size_t(c) * numSC + sc] = tgtMix[size_t(c) * numSC + sc];
}
__global__ void k_fill(float* p, size_t n, float v)
{
	const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
	if (i < n) p[i] = v;
}
// state blob [numSC][64] (AoS) <-> device SoA state[32][numSC] + curMix[L][numSC]
__global__ void k_state_export(const float* __restrict__ state, const float* __restrict__ curMix, uint32_t numSC, uint32_t L, float* __restrict__ blob)
{
	const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
	if (idx >= size_t(numSC) * 64) return;
	const uint32_t sc = static_cast<uint32_t>(idx / 64), k = static_cast<uint32_t>(idx % 64);
	float v = 0.0f;
	if (k < 32) v = state[size_t(k) * numSC + sc]; else if (k - 32 < L) v = curMix[size_t(k - 32) * numSC + sc];
	blob[idx] = v;
}
__global__ void k_state_import(float* __restrict__ state, float* __restrict__ curMix, uint32_t numSC, uint32_t L, const float* __restrict__ blob)
{
	const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
	if (idx >= size_t(numSC) * 64) return;
	const uint32_t sc = static_cast<uint32_t>(idx / 64), k = static_cast<uint32_t>(idx % 64);
	if (k < 32) state[size_t(k) * numSC + sc] = blob[idx]; else if (k - 32 < L) curMix[size_t(k - 32) * numSC + sc] = blob[idx];
}

static void FreeRenderer(jplg_renderer r)
{
	float** ptrs[] = { &r->state, &r->curMix, &r->tgtDelay, &r->tgtBand, &r->tgtMix, &r->rings, &r->partial, &r->deviceBus, &r->stageIn, &r->stageBus, &r->stagePerSource, &r->stageParams };
	for (float** p : ptrs) if (*p) { cudaFree(*p); *p = nullptr; }
	if (r->asyncReady)
	{
		if (r->ctx->auxReady) cudaStreamSynchronize(r->ctx->aux[0]);
		if (r->stageIn2) { cudaFree(r->stageIn2); r->stageIn2 = nullptr; }
		for (int i = 0; i < 2; ++i) { cudaEventDestroy(r->evCopiedA[i]); cudaEventDestroy(r->evPushedA[i]); cudaEventDestroy(r->evDoneA[i]); }
		r->asyncReady = false;
	}
	if (r->evStarted) { cudaEventDestroy(r->evStarted); r->evStarted = nullptr; }
	if (r->evUpdated) { cudaEventDestroy(r->evUpdated); r->evUpdated = nullptr; }
	for (auto& e : r->evSetFree) if (e) { cudaEventDestroy(e); e = nullptr; }
	for (float** q : { &r->tgt2Delay, &r->tgt2Band, &r->tgt2Mix, &r->scanM, &r->scanScratch }) if (*q) { cudaFree(*q); *q = nullptr; }
	if (r->scanCounters) { cudaFree(r->scanCounters); r->scanCounters = nullptr; }
	if (r->evAhead) { if (r->ctx->auxReady) cudaStreamSynchronize(r->ctx->aux[1]); cudaEventDestroy(r->evAhead); r->evAhead = nullptr; }
}

// delays the host never sees (device-resident updates): known only through the caller's declared bound
static void NoteDeviceDelays(jplg_renderer r)
{
	if (r->declaredMaxDelaySamples >= 0.0f) { r->maxDelaySamples = r->declaredMaxDelaySamples; r->delayBoundKnown = true; }
	else r->delayBoundKnown = false;
}

static jplg_status RendererUpdateLocked(jplg_renderer r, const float* band, const float* delay, const float* mix, int32_t mem)
{
	jplg_context ctx = r->ctx;
	const uint32_t N = r->cfg.num_sources, C = r->cfg.source_channels, L = r->cfg.output_channels;
	const float *dBand = band, *dDelay = delay, *dMix = mix;
	if (mem == JPLG_MEM_HOST || mem == JPLG_MEM_HOST_ASYNC)
	{
		float* base = r->stageParams;
		const bool tryAlias = mem == JPLG_MEM_HOST_ASYNC;   // page-locked buffers are read by k_update_targets directly (MappedHostAlias)
		const float* al;
		if (band) { if (tryAlias && (al = MappedHostAlias(band))) dBand = al; else { JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(base, band, size_t(N) * 4 * 4, cudaMemcpyHostToDevice, ctx->stream)); dBand = base; } }
		if (delay) { if (tryAlias && (al = MappedHostAlias(delay))) dDelay = al; else { JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(base + size_t(N) * 4, delay, size_t(N) * 4, cudaMemcpyHostToDevice, ctx->stream)); dDelay = base + size_t(N) * 4; } }
		if (mix) { if (tryAlias && (al = MappedHostAlias(mix))) dMix = al; else { JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(base + size_t(N) * 5, mix, size_t(N) * C * L * 4, cudaMemcpyHostToDevice, ctx->stream)); dMix = base + size_t(N) * 5; } }
	}
	if (delay)
	{
		if (mem == JPLG_MEM_DEVICE) NoteDeviceDelays(r);
		else
		{
			float mx = 0.0f;
			for (uint32_t i = 0; i < N; ++i) mx = std::max(mx, delay[i] * r->cfg.sample_rate);
			r->maxDelaySamples = mx; r->delayBoundKnown = true;      // a complete set of host-side delays replaces whatever was there
		}
	}
	float *wBand = r->tgtBand, *wDelay = r->tgtDelay, *wMix = r->tgtMix;
	if (r->pendingPanned)
	{
		// the next block will use the set jplg_renderer_update_panned has written: this update refines that one, after it
		JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, r->evUpdated, 0));
		wBand = r->tgt2Band; wDelay = r->tgt2Delay; wMix = r->tgt2Mix;
	}
	TimeBegin(ctx, JPLG_K_UPDATE);
	k_update_targets<<<(r->numSC + 127) / 128, 128, 0, ctx->stream>>>(dBand, dDelay, dMix, r->numSC, C, L, r->cfg.sample_rate, wBand, wDelay, wMix);
	TimeEnd(ctx);
	JPLG_CUDA_TRY(ctx, cudaGetLastError()); ctx->launches++;
	if (mem == JPLG_MEM_HOST) JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream)); // host buffers are caller-owned: safe to reuse on return
	return JPLG_OK;
}

extern "C" {

jplg_status jplg_renderer_create(jplg_context ctx, const jplg_renderer_config* cfgIn, const float* band_gains, const float* delay_seconds,
								 const float* mix_maps, jplg_renderer* out)
{
	if (!ctx || !cfgIn || !out) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	JPLG_NEED_DEVICE(ctx);
	jplg_renderer_config cfg = *cfgIn;
	if (cfg.num_sources == 0 || cfg.source_channels == 0 || cfg.source_channels > 128 || cfg.output_channels == 0 || cfg.output_channels > kMaxChannelsDev || !(cfg.sample_rate > 0.0f))
		return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_renderer_create: bad configuration");
	if (cfg.max_block_frames == 0) cfg.max_block_frames = 512;
	uint32_t ring = cfg.ring_size;
	if (ring == 0) { const uint32_t minSize = static_cast<uint32_t>(cfg.sample_rate * 2.0f); ring = 1; while (ring < minSize) ring <<= 1; } // std::bit_ceil (DirectSound.cpp:79)
	if (ring & (ring - 1)) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_renderer_create: ring_size must be a power of two");
	if (ring < TotalPushes(cfg.max_block_frames) + kChunkFrames + 4) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_renderer_create: ring_size too small for max_block_frames");
	if (cfg.split_f1 == 0.0f) cfg.split_f1 = 176.0f;
	if (cfg.split_f2 == 0.0f) cfg.split_f2 = 775.0f;
	if (cfg.split_f3 == 0.0f) cfg.split_f3 = 3408.0f;
	cfg.ring_size = ring;

	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	auto* r = new jplg_renderer_t();
	r->ctx = ctx; r->cfg = cfg; r->ringSize = ring; r->maxFrames = cfg.max_block_frames;
	r->numSC = cfg.num_sources * cfg.source_channels;
	r->numCTAs = RenderNumCTAs(r->numSC, cfg.source_channels);
	r->alpha = ExpAlpha(kChunkFrames);                                    // CreateExpSmoothing(.., cScratchSize) (DirectSound.cpp:53)
	CrossoverCoefficients(cfg.sample_rate, cfg.split_f1, cfg.split_f2, cfg.split_f3, r->coef); // Filter.Prepare(sampleRate, {}) (:83)
	const uint32_t N = cfg.num_sources, C = cfg.source_channels, L = cfg.output_channels;
	const size_t nsc = r->numSC;
	auto alloc = [&](float** p, size_t floats) { return cudaMalloc(reinterpret_cast<void**>(p), floats * sizeof(float)) == cudaSuccess; };
	bool ok = alloc(&r->state, nsc * kStateFloats) && alloc(&r->curMix, nsc * L) && alloc(&r->tgtDelay, nsc) && alloc(&r->tgtBand, nsc * 4) && alloc(&r->tgtMix, nsc * L)
		&& alloc(&r->rings, nsc * ring) && alloc(&r->partial, size_t(r->numCTAs) * r->maxFrames * L) && alloc(&r->deviceBus, size_t(r->maxFrames) * L)
		&& alloc(&r->stageIn, nsc * r->maxFrames) && alloc(&r->stageBus, size_t(r->maxFrames) * L) && alloc(&r->stageParams, size_t(N) * (5 + size_t(C) * L));
	if (!ok) { cudaGetLastError(); FreeRenderer(r); delete r; return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "jplg_renderer_create: device allocation failed (rings need num_sources*channels*ring_size*4 bytes)"); }
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(r->state, 0, nsc * kStateFloats * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(r->rings, 0, nsc * ring * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(r->tgtDelay, 0, nsc * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(r->tgtMix, 0, nsc * L * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(r->deviceBus, 0, size_t(r->maxFrames) * L * 4, ctx->stream));
	k_fill<<<static_cast<uint32_t>((nsc * 4 + 255) / 256), 256, 0, ctx->stream>>>(r->tgtBand, nsc * 4, 1.0f); ctx->launches++;
	jplg_status st = RendererUpdateLocked(r, band_gains, delay_seconds, mix_maps, JPLG_MEM_HOST);
	if (st != JPLG_OK) { FreeRenderer(r); delete r; return st; }
	k_init_current<<<(r->numSC + 127) / 128, 128, 0, ctx->stream>>>(r->numSC, L, r->tgtBand, r->tgtDelay, r->tgtMix, r->state, r->curMix); ctx->launches++;
	JPLG_CUDA_TRY(ctx, cudaGetLastError());
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	*out = r;
	return JPLG_OK;
}

void jplg_renderer_destroy(jplg_renderer r)
{
	if (!r) return;
	cudaSetDevice(r->ctx->device);
	cudaStreamSynchronize(r->ctx->stream);
	FreeRenderer(r);
	delete r;
}

jplg_status jplg_renderer_update(jplg_renderer r, const float* band_gains, const float* delay_seconds, const float* mix_maps, int32_t mem)
{
	if (!r) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(r->ctx);
	std::lock_guard<std::mutex> lk(r->ctx->mtx);
	cudaSetDevice(r->ctx->device);
	return RendererUpdateLocked(r, band_gains, delay_seconds, mix_maps, mem);
}

jplg_status jplg_renderer_update_panned(jplg_renderer r, jplg_panner p, jplg_source_layout l, const float* pan_params, const float* band_gains,
										const float* delay_seconds, int32_t mem)
{
	if (!r || !p || !l || l->panner != p || !pan_params || !band_gains || !delay_seconds) return JPLG_ERR_INVALID_ARGUMENT;
	if (mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;       // a device-side fusion: host callers use jplg_panner_mdap + jplg_renderer_update
	jplg_context rc = r->ctx, ctx = p->ctx;                             // the work runs on the PANNER's context stream
	JPLG_NEED_DEVICE(rc);
	JPLG_NEED_DEVICE(ctx);
	if (rc->device != ctx->device) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "update_panned: panner and renderer on different devices");
	if (!p->host.is3D) return Fail(ctx, JPLG_ERR_UNSUPPORTED, "update_panned: 3D panners only (use jplg_panner_mdap + jplg_renderer_update)");
	// the renderer's channels per source: the layout's panned channels, or every channel of the layout's source map (a 5.1 source rendered with its
	// LFE: the panner skips it, PannerBase.h:937, and its mix row stays zero)
	uint8_t rows[kMaxSourceChannelsDev] = {};
	{
		const uint32_t C = r->cfg.source_channels;
		uint32_t full = 0; for (uint32_t v = l->host.sourceMask; v; v &= v - 1) ++full;
		bool okRows = p->host.numChannels == r->cfg.output_channels;
		if (l->host.numChannels == C) for (uint32_t k = 0; k < C; ++k) rows[k] = static_cast<uint8_t>(k);
		else if (full == C && C <= kMaxSourceChannelsDev)
		{
			uint32_t idx = 0, k = 0;
			for (uint32_t v = l->host.sourceMask; v; v &= v - 1, ++idx)
			{
				const uint32_t bit = v & (0u - v);
				if (bit == 0x8u /* LFE */ || bit >= (1u << 13) /* top channels are not panned as source channels */) continue;
				if (k < kMaxSourceChannelsDev) rows[k] = static_cast<uint8_t>(idx);
				++k;
			}
			okRows = okRows && k == l->host.numChannels;
		}
		else okRows = false;
		if (!okRows) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "update_panned: source layout / panner channel counts do not match the renderer");
	}
	std::unique_lock<std::mutex> lkA(rc->mtx, std::defer_lock), lkB(ctx->mtx, std::defer_lock);
	if (rc == ctx) lkA.lock(); else std::lock(lkA, lkB);
	cudaSetDevice(ctx->device);
	const uint32_t N = r->cfg.num_sources, L = r->cfg.output_channels;
	const size_t nsc = r->numSC;
	if (!r->tgt2Mix)
	{
		if (cudaMalloc(reinterpret_cast<void**>(&r->tgt2Delay), nsc * 4) != cudaSuccess || cudaMalloc(reinterpret_cast<void**>(&r->tgt2Band), nsc * 16) != cudaSuccess
			|| cudaMalloc(reinterpret_cast<void**>(&r->tgt2Mix), nsc * L * 4) != cudaSuccess)
		{ cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "update_panned: target set allocation failed"); }
		JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evUpdated, cudaEventDisableTiming));
		for (auto& e : r->evSetFree) JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	}
	const int w = r->activeSet ^ 1;
	// the render that last read the set about to be overwritten (two blocks back) has to be done; the one reading the active set may still run
	if (r->setFreeRecorded[w]) JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, r->evSetFree[w], 0));
	void* scratch = nullptr;
	jplg_status st = Scratch(ctx, 1, size_t(N) * l->host.numChannels * L * 4, &scratch); if (st) return st;
	MdapFuse fz;
	fz.tgtMix = r->tgt2Mix; fz.tgtBand = r->tgt2Band; fz.tgtDelay = r->tgt2Delay;
	fz.band = band_gains; fz.delaySec = delay_seconds; fz.sampleRate = r->cfg.sample_rate; fz.numSC = r->numSC;
	fz.rendC = r->cfg.source_channels; std::memcpy(fz.rows, rows, sizeof(rows));
	fz.unpannedRows = (1u << fz.rendC) - 1u;
	for (uint32_t k = 0; k < l->host.numChannels; ++k) fz.unpannedRows &= ~(1u << rows[k]);
	JPLG_LAUNCH(ctx, JPLG_K_MDAP, LaunchMdap(p->dev, l->lp, pan_params, 5, N, static_cast<float*>(scratch), ctx->numSMs, ctx->stream, &fz));
	JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evUpdated, ctx->stream));
	r->pendingPanned = true;
	NoteDeviceDelays(r);
	return JPLG_OK;
}

static jplg_status EnsureAux(jplg_context ctx)
{
	if (ctx->auxReady) return JPLG_OK;
	// The side streams rank between a rendering context's stream (greatest priority) and a background context's (least): the next block's ring pushes
	// (aux[1]) then take the SMs a render leaves free BEFORE the background panning kernels do and are over in a third of the render's duration;
	// spread over all of it, sharing those SMs, their ring writes slow the render's ring reads down by 9 % (measured).
	int prioLeast = 0, prioGreatest = 0;
	JPLG_CUDA_TRY(ctx, cudaDeviceGetStreamPriorityRange(&prioLeast, &prioGreatest));
	const int prioSide = prioLeast - 1 >= prioGreatest ? prioLeast - 1 : prioLeast;
	for (auto& a : ctx->aux) JPLG_CUDA_TRY(ctx, cudaStreamCreateWithPriority(&a, cudaStreamNonBlocking, prioSide));
	JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->evStart, cudaEventDisableTiming));
	for (auto& e : ctx->evCopied) JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	for (auto& e : ctx->evDone) JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	ctx->auxReady = true;
	return JPLG_OK;
}

// JPLG_MODE_THROUGHPUT renders small scenes time-parallel (jplg_scan_kernels.cu): the scan does the crossover arithmetic twice, so it pays only
// while (source-channels x tiles) threads still find idle schedulers. Cross-over measured on a B200 (profiles/r02_scan_ab.txt); the policy can be
// pinned with jplg_renderer_set_time_parallel.
constexpr size_t kScanAutoMaxSC = 3072;
static bool UseScanPath(jplg_renderer r, const RenderParams& rp)
{
	if (r->cfg.mode != JPLG_MODE_THROUGHPUT || r->scanPolicy == 0 || !RenderCanScan(rp)) return false;
	if (r->scanPolicy > 0) return true;
	return size_t(rp.numSC) <= kScanAutoMaxSC;
}

static jplg_status LaunchScan(jplg_renderer r, const RenderParams& rp, float* busAcc)
{
	jplg_context ctx = r->ctx;
	if (!r->scanM)
	{
		float M[24 * 24];
		CrossoverTransition(r->coef, kScanTile, M);
		if (cudaMalloc(reinterpret_cast<void**>(&r->scanM), sizeof(M)) != cudaSuccess
			|| cudaMalloc(reinterpret_cast<void**>(&r->scanScratch), ScanScratchFloats(r->numSC, r->maxFrames) * sizeof(float)) != cudaSuccess
			|| cudaMalloc(reinterpret_cast<void**>(&r->scanCounters), (size_t(r->numCTAs) + r->maxFrames / kScanTile + 1) * sizeof(uint32_t)) != cudaSuccess)
		{ cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "process_block: time-parallel scratch allocation failed"); }
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(r->scanM, M, sizeof(M), cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemsetAsync(r->scanCounters, 0, (size_t(r->numCTAs) + r->maxFrames / kScanTile + 1) * sizeof(uint32_t), ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));   // M is a stack array
	}
	ScanParams sp{};
	sp.rp = rp; sp.M = r->scanM; sp.scratch = r->scanScratch; sp.counters = r->scanCounters; sp.push = 1u;
	sp.deviceBus = r->deviceBus; sp.busAcc = busAcc;
	JPLG_LAUNCH(ctx, JPLG_K_RENDER, LaunchRenderScan(sp, ctx->stream));
	ctx->launches += 1;   // k_scan_front, k_scan_mix
	r->scanBlocks++;
	return JPLG_OK;
}

static jplg_status ProcessBlock(jplg_renderer r, const float* in, float* bus, float* perSource, uint32_t frames, int32_t mem)
{
	if (!r || (frames && (!in || !bus))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = r->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (frames == 0) return JPLG_OK;
	if (frames > r->maxFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "process_block: frames exceeds max_block_frames");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const uint32_t N = r->cfg.num_sources, C = r->cfg.source_channels, L = r->cfg.output_channels;
	const float* dIn = in; float* dBus = bus; float* dPer = perSource;
	const size_t perFloats = size_t(N) * frames * L;
	if (r->aheadIn && !(mem == JPLG_MEM_DEVICE && r->aheadIn == in && r->aheadFrames == frames && !perSource))
	{
		// a different block than the one pushed ahead: its pushes went to the slots this block is about to write (the ring position only
		// advances in process_block), so let them land first and push this block over them
		JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, r->evAhead, 0));
		r->aheadIn = nullptr;
	}
	const bool pipelinedHost = mem == JPLG_MEM_HOST && ctx->hostPipeline > 0 && r->numCTAs >= 32 && !perSource;
	int asyncSlot = -1;
	float* busAlias = nullptr;
	if (mem == JPLG_MEM_HOST_ASYNC)
	{
		if (perSource) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "process_block_debug: JPLG_MEM_HOST_ASYNC is not supported");
		jplg_status ast = EnsureAux(ctx); if (ast != JPLG_OK) return ast;
		if (!r->asyncReady)
		{
			if (cudaMalloc(reinterpret_cast<void**>(&r->stageIn2), size_t(r->numSC) * r->maxFrames * 4) != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "async staging allocation failed"); }
			for (int i = 0; i < 2; ++i)
			{
				JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evCopiedA[i], cudaEventDisableTiming));
				JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evPushedA[i], cudaEventDisableTiming));
				JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evDoneA[i], cudaEventDisableTiming));
			}
			r->asyncReady = true;
		}
		if (r->asyncSubmitted - r->asyncWaited >= 2)   // both slots in flight: wait for the older block (its host bus is then valid too)
		{
			JPLG_CUDA_TRY(ctx, cudaEventSynchronize(r->evDoneA[r->asyncWaited & 1]));
			r->asyncWaited++;
		}
		asyncSlot = static_cast<int>(r->asyncSubmitted & 1);
		float* stage = asyncSlot ? r->stageIn2 : r->stageIn;
		cudaStream_t copy = ctx->aux[0];
		if (r->asyncSubmitted >= 2) JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(copy, r->evPushedA[asyncSlot], 0));   // the push of block k-2 has read this slot
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(stage, in, size_t(N) * frames * C * 4, cudaMemcpyHostToDevice, copy));
		JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evCopiedA[asyncSlot], copy));
		JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, r->evCopiedA[asyncSlot], 0));
		busAlias = MappedHostAliasRaw(bus);    // page-locked bus: k_bus_reduce accumulates into it over PCIe, no staged copies either way
		if (!busAlias) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(r->stageBus, bus, size_t(frames) * L * 4, cudaMemcpyHostToDevice, ctx->stream));
		dIn = stage; dBus = busAlias ? busAlias : r->stageBus;
	}
	if (mem == JPLG_MEM_HOST)
	{
		if (!pipelinedHost) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(r->stageIn, in, size_t(N) * frames * C * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(r->stageBus, bus, size_t(frames) * L * 4, cudaMemcpyHostToDevice, ctx->stream));
		dIn = r->stageIn; dBus = r->stageBus;
		if (perSource)
		{
			if (perFloats > r->stagePerSourceFloats)
			{
				JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
				if (r->stagePerSource) cudaFree(r->stagePerSource);
				r->stagePerSource = nullptr; r->stagePerSourceFloats = 0;
				if (cudaMalloc(reinterpret_cast<void**>(&r->stagePerSource), perFloats * 4) != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "per-source debug buffer allocation failed"); }
				r->stagePerSourceFloats = perFloats;
			}
			dPer = r->stagePerSource;
		}
	}
	if (r->pendingPanned)
	{
		// targets written by jplg_renderer_update_panned (maybe on another context's stream): they become the active set once they are complete
		JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, r->evUpdated, 0));
		std::swap(r->tgtDelay, r->tgt2Delay); std::swap(r->tgtBand, r->tgt2Band); std::swap(r->tgtMix, r->tgt2Mix);
		r->activeSet ^= 1;
		r->pendingPanned = false;
	}
	bool scanReduced = false;
	RenderParams rp{};
	rp.numSC = r->numSC; rp.srcChannels = C; rp.L = L; rp.ringSize = r->ringSize; rp.ringMask = r->ringSize - 1; rp.pos = r->pos; rp.frames = frames;
	rp.alpha = r->alpha; std::memcpy(rp.coef, r->coef, sizeof(rp.coef));
	rp.state = r->state; rp.curMix = r->curMix; rp.tgtDelay = r->tgtDelay; rp.tgtBand = r->tgtBand; rp.tgtMix = r->tgtMix; rp.rings = r->rings;
	rp.in = dIn; rp.partial = r->partial; rp.perSource = dPer; rp.numCTAs = r->numCTAs;
	const uint32_t srcPerCTA = RenderSourcesPerCTA(C);
	// ---- host path, pipelined: slices of CTAs; the transfer of slice k+1 overlaps push + render of slice k on other streams
	uint32_t numSlices = 1;
	if (pipelinedHost) numSlices = 8;
	if (numSlices > 1)
	{
		jplg_status ast = EnsureAux(ctx); if (ast != JPLG_OK) return ast;
		bool zeroCopy = false;
		if (ctx->hostPipeline == 2)
		{
			cudaPointerAttributes pa{};
			if (cudaPointerGetAttributes(&pa, in) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer != nullptr) { zeroCopy = true; rp.in = static_cast<const float*>(pa.devicePointer); }
			else cudaGetLastError();
		}
		JPLG_CUDA_TRY(ctx, cudaEventRecord(ctx->evStart, ctx->stream));
		for (auto& a : ctx->aux) JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(a, ctx->evStart, 0));
		const uint32_t ctasPerSlice = (r->numCTAs + numSlices - 1) / numSlices;
		for (uint32_t k = 0; k < numSlices; ++k)
		{
			const uint32_t c0 = k * ctasPerSlice, c1 = std::min(r->numCTAs, c0 + ctasPerSlice);
			if (c0 >= c1) break;
			const uint32_t s0 = c0 * srcPerCTA, s1 = std::min(N, c1 * srcPerCTA);
			cudaStream_t cs = ctx->aux[1 + k % (jplg_context_t::kAux - 1)];
			if (!zeroCopy)
			{
				// the (re-issued) staged copy of this slice runs on the copy stream aux[0]; the first, whole-block copy above is skipped
				JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(r->stageIn + size_t(s0) * frames * C, in + size_t(s0) * frames * C, size_t(s1 - s0) * frames * C * 4, cudaMemcpyHostToDevice, ctx->aux[0]));
				JPLG_CUDA_TRY(ctx, cudaEventRecord(ctx->evCopied[k], ctx->aux[0]));
				JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(cs, ctx->evCopied[k], 0));
			}
			JPLG_LAUNCH_ON(ctx, JPLG_K_PUSH, cs, LaunchPush(rp, s0, s1 - s0, cs));
			JPLG_LAUNCH_ON(ctx, JPLG_K_RENDER, cs, LaunchRender(rp, r->cfg.mode, c0, c1 - c0, cs));
		}
		for (int a = 1; a < jplg_context_t::kAux; ++a)
		{
			JPLG_CUDA_TRY(ctx, cudaEventRecord(ctx->evDone[a], ctx->aux[a]));
			JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->evDone[a], 0));
		}
	}
	else if (mem == JPLG_MEM_DEVICE && r->aheadIn == in && r->aheadFrames == frames && !perSource)
	{
		// this block's pushes were issued ahead of time (jplg_renderer_push_ahead): only wait for them
		JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, r->evAhead, 0));
		r->aheadIn = nullptr;
		if (!r->evStarted) JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evStarted, cudaEventDisableTiming));
		JPLG_LAUNCH(ctx, JPLG_K_RENDER, LaunchRender(rp, r->cfg.mode, 0, r->numCTAs, ctx->stream, r->evStarted));
		r->startedRecorded = true;
	}
	else
	{
		// real-time order: this block's input enters the rings now. Mono renderers let k_render's own push role do it (no separate kernel, the copy
		// overlaps the start of the filters); everything else launches the push kernel first.
		const bool scan = UseScanPath(r, rp);
		rp.fusedPush = (scan || RenderCanFusePush(rp)) ? 1u : 0u;    // the time-parallel path pushes inside its own first kernel
		if (!rp.fusedPush)
		{
			JPLG_LAUNCH(ctx, JPLG_K_PUSH, LaunchPush(rp, 0, N, ctx->stream));
			if (asyncSlot >= 0) JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evPushedA[asyncSlot], ctx->stream));
		}
		if (!r->evStarted) JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evStarted, cudaEventDisableTiming));
		if (scan)
		{
			jplg_status sst = LaunchScan(r, rp, dBus); if (sst != JPLG_OK) return sst;
			scanReduced = true;
			JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evStarted, ctx->stream));   // small grids: "resident" and "done" are the same thing here
		}
		else JPLG_LAUNCH(ctx, JPLG_K_RENDER, LaunchRender(rp, r->cfg.mode, 0, r->numCTAs, ctx->stream, r->evStarted));
		r->startedRecorded = true;
		if (rp.fusedPush && asyncSlot >= 0) JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evPushedA[asyncSlot], ctx->stream));   // the staging slot was read by the render
	}
	if (r->evSetFree[r->activeSet]) { JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evSetFree[r->activeSet], ctx->stream)); r->setFreeRecorded[r->activeSet] = true; }
	if (!scanReduced) JPLG_LAUNCH(ctx, JPLG_K_BUS_REDUCE, LaunchBusReduce(rp, r->deviceBus, dBus, ctx->stream));   // the time-parallel path reduces inside k_scan_mix
	r->pos += TotalPushes(frames);
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(bus, dBus, size_t(frames) * L * 4, cudaMemcpyDeviceToHost, ctx->stream));
		if (perSource) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(perSource, dPer, perFloats * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	if (asyncSlot >= 0)
	{
		if (!busAlias) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(bus, dBus, size_t(frames) * L * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evDoneA[asyncSlot], ctx->stream));
		r->asyncSubmitted++;
	}
	return JPLG_OK;
}

jplg_status jplg_renderer_push_ahead(jplg_renderer r, const float* in, uint32_t frames)
{
	if (!r || !in || frames == 0) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = r->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (frames > r->maxFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "push_ahead: frames exceeds max_block_frames");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	if (r->aheadIn) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "push_ahead: a block is already pushed ahead");
	if (r->ringSize < 4u * TotalPushes(frames)) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "push_ahead: ring too small to hold a block ahead");
	// the render of the current block reads samples up to (longest delay + this block's pushes + one chunk) old; the pushes issued here land
	// TotalPushes further on: both fit the ring without meeting only below this bound
	if (!r->delayBoundKnown)
		return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "push_ahead: the target delays were set on the device (jplg_renderer_update with JPLG_MEM_DEVICE / jplg_renderer_update_panned); "
			"declare their upper bound with jplg_renderer_set_max_delay_seconds first");
	if (static_cast<double>(r->maxDelaySamples) + 2.0 * TotalPushes(frames) + kChunkFrames >= static_cast<double>(r->ringSize))
		return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "push_ahead: longest delay + two blocks of pushes + one chunk does not fit the ring: the pushed-ahead block would overwrite samples the current render reads");
	jplg_status ast = EnsureAux(ctx); if (ast != JPLG_OK) return ast;
	if (!r->evAhead) JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&r->evAhead, cudaEventDisableTiming));
	RenderParams rp{};
	rp.numSC = r->numSC; rp.srcChannels = r->cfg.source_channels; rp.L = r->cfg.output_channels; rp.ringSize = r->ringSize; rp.ringMask = r->ringSize - 1;
	rp.pos = r->pos; rp.frames = frames; rp.rings = r->rings; rp.in = in;
	cudaStream_t side = ctx->aux[1];
	// not before the render enqueued last has all its CTAs resident: the pushes then only take SMs that grid leaves free
	if (r->startedRecorded) JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(side, r->evStarted, 0));
	JPLG_LAUNCH_ON(ctx, JPLG_K_PUSH, side, LaunchPush(rp, 0, r->cfg.num_sources, side));
	JPLG_CUDA_TRY(ctx, cudaEventRecord(r->evAhead, side));
	r->aheadIn = in; r->aheadFrames = frames;
	return JPLG_OK;
}

jplg_status jplg_renderer_wait_render_started(jplg_renderer r, jplg_context waiting)
{
	if (!r || !waiting) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = r->ctx;
	JPLG_NEED_DEVICE(ctx);
	JPLG_NEED_DEVICE(waiting);
	if (waiting->device != ctx->device) return Fail(waiting, JPLG_ERR_INVALID_ARGUMENT, "wait_render_started: contexts on different devices");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	if (r->startedRecorded) JPLG_CUDA_TRY(waiting, cudaStreamWaitEvent(waiting->stream, r->evStarted, 0));
	return JPLG_OK;
}

jplg_status jplg_renderer_wait_block(jplg_renderer r)
{
	if (!r) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = r->ctx;
	JPLG_NEED_DEVICE(ctx);
	cudaEvent_t ev;
	{
		std::lock_guard<std::mutex> lk(ctx->mtx);
		if (r->asyncWaited >= r->asyncSubmitted) return JPLG_OK;   // nothing in flight
		ev = r->evDoneA[r->asyncWaited & 1];
	}
	cudaSetDevice(ctx->device);
	JPLG_CUDA_TRY(ctx, cudaEventSynchronize(ev));      // not under the lock: another thread may enqueue the next block meanwhile
	{
		std::lock_guard<std::mutex> lk(ctx->mtx);
		r->asyncWaited++;
	}
	return JPLG_OK;
}

jplg_status jplg_renderer_process_block(jplg_renderer r, const float* in, float* bus, uint32_t frames, int32_t mem)
{
	return ProcessBlock(r, in, bus, nullptr, frames, mem);
}
jplg_status jplg_renderer_process_block_debug(jplg_renderer r, const float* in, float* bus, float* per_source, uint32_t frames, int32_t mem)
{
	if (!per_source) return JPLG_ERR_INVALID_ARGUMENT;
	return ProcessBlock(r, in, bus, per_source, frames, mem);
}
jplg_status jplg_renderer_set_max_delay_seconds(jplg_renderer r, float seconds)
{
	if (!r || !(seconds >= 0.0f)) return JPLG_ERR_INVALID_ARGUMENT;
	std::lock_guard<std::mutex> lk(r->ctx->mtx);
	r->declaredMaxDelaySamples = seconds * r->cfg.sample_rate;
	if (!r->delayBoundKnown || r->declaredMaxDelaySamples > r->maxDelaySamples) { r->maxDelaySamples = std::max(r->maxDelaySamples, r->declaredMaxDelaySamples); r->delayBoundKnown = true; }
	return JPLG_OK;
}
jplg_status jplg_renderer_set_time_parallel(jplg_renderer r, int32_t policy)
{
	if (!r || policy < -1 || policy > 1) return JPLG_ERR_INVALID_ARGUMENT;
	std::lock_guard<std::mutex> lk(r->ctx->mtx);
	r->scanPolicy = policy;
	return JPLG_OK;
}
uint64_t jplg_renderer_time_parallel_blocks(jplg_renderer r) { return r ? r->scanBlocks : 0; }
float* jplg_renderer_device_bus(jplg_renderer r) { return r ? r->deviceBus : nullptr; }
uint32_t jplg_renderer_ring_size(jplg_renderer r) { return r ? r->ringSize : 0; }

jplg_status jplg_renderer_state_export(jplg_renderer r, float* state, float* rings, uint64_t* ring_pos)
{
	if (!r || !state) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = r->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	void* blob; jplg_status st = Scratch(ctx, 2, size_t(r->numSC) * 64 * 4, &blob); if (st) return st;
	k_state_export<<<static_cast<uint32_t>((size_t(r->numSC) * 64 + 255) / 256), 256, 0, ctx->stream>>>(r->state, r->curMix, r->numSC, r->cfg.output_channels, static_cast<float*>(blob)); ctx->launches++;
	JPLG_CUDA_TRY(ctx, cudaGetLastError());
	JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(state, blob, size_t(r->numSC) * 64 * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (rings) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(rings, r->rings, size_t(r->numSC) * r->ringSize * 4, cudaMemcpyDeviceToHost, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	if (ring_pos) *ring_pos = r->pos;
	return JPLG_OK;
}
jplg_status jplg_renderer_state_import(jplg_renderer r, const float* state, const float* rings, uint64_t ring_pos)
{
	if (!r || !state) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = r->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	void* blob; jplg_status st = Scratch(ctx, 2, size_t(r->numSC) * 64 * 4, &blob); if (st) return st;
	JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(blob, state, size_t(r->numSC) * 64 * 4, cudaMemcpyHostToDevice, ctx->stream));
	k_state_import<<<static_cast<uint32_t>((size_t(r->numSC) * 64 + 255) / 256), 256, 0, ctx->stream>>>(r->state, r->curMix, r->numSC, r->cfg.output_channels, static_cast<const float*>(blob)); ctx->launches++;
	JPLG_CUDA_TRY(ctx, cudaGetLastError());
	if (rings) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(r->rings, rings, size_t(r->numSC) * r->ringSize * 4, cudaMemcpyHostToDevice, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	r->pos = ring_pos;
	return JPLG_OK;
}

} // extern "C"

//==================================================================================================
// Components
struct jplg_crossover_bank_t { jplg_context ctx; uint32_t n; int mode; float coef[27]; float* state; };
struct jplg_delay_bank_t { jplg_context ctx; uint32_t n, ringSize; uint64_t pos; float* rings; float* y1; };

extern "C" {

jplg_status jplg_crossover_bank_create(jplg_context ctx, uint32_t n, float sample_rate, float f1, float f2, float f3, int32_t mode, jplg_crossover_bank* out)
{
	if (!ctx || !out || n == 0 || !(sample_rate > 0.0f)) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	auto* b = new jplg_crossover_bank_t{ ctx, n, mode, {}, nullptr };
	if (f1 == 0.0f) f1 = 176.0f; if (f2 == 0.0f) f2 = 775.0f; if (f3 == 0.0f) f3 = 3408.0f;
	CrossoverCoefficients(sample_rate, f1, f2, f3, b->coef);
	if (ctx->device >= 0)
	{
		cudaSetDevice(ctx->device);
		if (cudaMalloc(reinterpret_cast<void**>(&b->state), size_t(n) * 24 * 4) != cudaSuccess) { cudaGetLastError(); delete b; return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "crossover bank allocation failed"); }
		cudaMemsetAsync(b->state, 0, size_t(n) * 24 * 4, ctx->stream);
	}
	*out = b;
	return JPLG_OK;
}
void jplg_crossover_bank_destroy(jplg_crossover_bank b)
{
	if (!b) return;
	if (b->state) { cudaSetDevice(b->ctx->device); cudaStreamSynchronize(b->ctx->stream); cudaFree(b->state); }
	delete b;
}
jplg_status jplg_crossover_bank_reset(jplg_crossover_bank b)
{
	if (!b) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(b->ctx);
	cudaSetDevice(b->ctx->device);
	JPLG_CUDA_TRY(b->ctx, cudaMemsetAsync(b->state, 0, size_t(b->n) * 24 * 4, b->ctx->stream));
	return JPLG_OK;
}
void jplg_crossover_bank_coeffs(jplg_crossover_bank b, float out[27]) { if (b && out) std::memcpy(out, b->coef, sizeof(b->coef)); }

jplg_status jplg_crossover_bank_process_ramp(jplg_crossover_bank b, const float* in, const float* start_gains, const float* end_gains, float* out, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || !in || !start_gains || !end_gains || !out) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nf = size_t(b->n) * frames;
	const float *dIn = in, *dG0 = start_gains, *dG1 = end_gains; float* dOut = out;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *g, *o;
		jplg_status st = Scratch(ctx, 0, nf * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, size_t(b->n) * 8 * 4, &g); if (st) return st;
		st = Scratch(ctx, 2, nf * 4, &o); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, in, nf * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(g, start_gains, size_t(b->n) * 16, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(static_cast<float*>(g) + size_t(b->n) * 4, end_gains, size_t(b->n) * 16, cudaMemcpyHostToDevice, ctx->stream));
		dIn = static_cast<float*>(a); dG0 = static_cast<float*>(g); dG1 = dG0 + size_t(b->n) * 4; dOut = static_cast<float*>(o);
	}
	XoverParams xp{}; xp.n = b->n; xp.frames = frames; std::memcpy(xp.coef, b->coef, sizeof(xp.coef)); xp.state = b->state;
	JPLG_CUDA_TRY(ctx, LaunchXoverRamp(xp, dIn, dG0, dG1, dOut, b->mode, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, nf * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

jplg_status jplg_crossover_bank_split(jplg_crossover_bank b, const float* in, float* out_bands, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || !in || !out_bands) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nf = size_t(b->n) * frames;
	const float* dIn = in; float* dOut = out_bands;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *o;
		jplg_status st = Scratch(ctx, 0, nf * 4, &a); if (st) return st;
		st = Scratch(ctx, 2, nf * 16, &o); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, in, nf * 4, cudaMemcpyHostToDevice, ctx->stream));
		dIn = static_cast<float*>(a); dOut = static_cast<float*>(o);
	}
	XoverParams xp{}; xp.n = b->n; xp.frames = frames; std::memcpy(xp.coef, b->coef, sizeof(xp.coef)); xp.state = b->state;
	JPLG_CUDA_TRY(ctx, LaunchXoverSplit(xp, dIn, dOut, b->mode, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out_bands, dOut, nf * 16, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

jplg_status jplg_delay_bank_create(jplg_context ctx, uint32_t n, uint32_t min_size, jplg_delay_bank* out)
{
	if (!ctx || !out || n == 0) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	JPLG_NEED_DEVICE(ctx);
	uint32_t ring = 1; while (ring < min_size) ring <<= 1; // std::bit_ceil (inc/Utilities/AbstractIndex.h:235)
	cudaSetDevice(ctx->device);
	auto* b = new jplg_delay_bank_t{ ctx, n, ring, 0, nullptr, nullptr };
	if (cudaMalloc(reinterpret_cast<void**>(&b->rings), size_t(n) * ring * 4) != cudaSuccess || cudaMalloc(reinterpret_cast<void**>(&b->y1), size_t(n) * 12) != cudaSuccess)   // y1 [n] | DelayTapXFade {last, new} delays [2][n]
	{
		cudaGetLastError(); if (b->rings) cudaFree(b->rings); delete b;
		return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "delay bank allocation failed");
	}
	cudaMemsetAsync(b->rings, 0, size_t(n) * ring * 4, ctx->stream);
	cudaMemsetAsync(b->y1, 0, size_t(n) * 12, ctx->stream);
	*out = b;
	return JPLG_OK;
}
void jplg_delay_bank_destroy(jplg_delay_bank b)
{
	if (!b) return;
	cudaSetDevice(b->ctx->device); cudaStreamSynchronize(b->ctx->stream);
	cudaFree(b->rings); cudaFree(b->y1);
	delete b;
}
uint32_t jplg_delay_bank_size(jplg_delay_bank b) { return b ? b->ringSize : 0; }
jplg_status jplg_delay_bank_clear(jplg_delay_bank b)
{
	if (!b) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(b->ctx);
	cudaSetDevice(b->ctx->device);
	JPLG_CUDA_TRY(b->ctx, cudaMemsetAsync(b->rings, 0, size_t(b->n) * b->ringSize * 4, b->ctx->stream));
	b->pos = 0; // mIndex.Reset()
	return JPLG_OK;
}
jplg_status jplg_delay_bank_push(jplg_delay_bank b, const float* in, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || (frames && !in)) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (frames == 0) return JPLG_OK;
	if (frames > b->ringSize) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "delay_bank_push: frames exceeds ring size");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const float* dIn = in;
	if (mem == JPLG_MEM_HOST)
	{
		void* a; jplg_status st = Scratch(ctx, 0, size_t(b->n) * frames * 4, &a); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, in, size_t(b->n) * frames * 4, cudaMemcpyHostToDevice, ctx->stream));
		dIn = static_cast<float*>(a);
	}
	JPLG_CUDA_TRY(ctx, LaunchDelayPush(b->rings, b->n, b->ringSize, b->pos, dIn, frames, ctx->stream)); ctx->launches++;
	b->pos += frames;
	if (mem == JPLG_MEM_HOST) JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	return JPLG_OK;
}
jplg_status jplg_delay_bank_read(jplg_delay_bank b, const float* delays, uint32_t taps, int32_t interp, float* out, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || (taps && (!delays || !out)) || interp < 0 || interp > 2) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (taps == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nt = size_t(b->n) * taps;
	const float* dDelays = delays; float* dOut = out;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *o;
		jplg_status st = Scratch(ctx, 0, nt * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, nt * 4, &o); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, delays, nt * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDelays = static_cast<float*>(a); dOut = static_cast<float*>(o);
	}
	JPLG_CUDA_TRY(ctx, LaunchDelayRead(b->rings, b->n, b->ringSize, b->pos, dDelays, taps, interp, b->y1, dOut, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, nt * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

jplg_status jplg_delay_bank_read_xfade(jplg_delay_bank b, const float* delays, const float* alphas, uint32_t taps, float* out, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || (taps && (!delays || !alphas || !out))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (taps == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nt = size_t(b->n) * taps;
	const float* dDelays = delays; const float* dAlphas = alphas; float* dOut = out;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *c, *o;
		jplg_status st = Scratch(ctx, 0, nt * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, nt * 4, &o); if (st) return st;
		st = Scratch(ctx, 2, nt * 4, &c); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, delays, nt * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(c, alphas, nt * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDelays = static_cast<float*>(a); dAlphas = static_cast<float*>(c); dOut = static_cast<float*>(o);
	}
	JPLG_CUDA_TRY(ctx, LaunchDelayReadXFade(b->rings, b->n, b->ringSize, b->pos, dDelays, dAlphas, taps, reinterpret_cast<int32_t*>(b->y1 + b->n), dOut, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, nt * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

jplg_status jplg_mix_add_and_apply_gain_ramp(jplg_context ctx, float* dest, const float* source, uint32_t n, uint32_t dest_channel, uint32_t source_channel,
											 uint32_t dest_num_channels, uint32_t source_num_channels, uint32_t frames, const float* gain_start, const float* gain_end, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!ctx || !dest || !source || !gain_start || !gain_end || dest_channel >= dest_num_channels || source_channel >= source_num_channels) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(ctx);
	if (n == 0 || frames == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nd = size_t(n) * frames * dest_num_channels, ns = size_t(n) * frames * source_num_channels;
	float* dDest = dest; const float* dSrc = source; const float *dG0 = gain_start, *dG1 = gain_end;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *s, *g;
		jplg_status st = Scratch(ctx, 0, nd * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, ns * 4, &s); if (st) return st;
		st = Scratch(ctx, 2, size_t(n) * 8, &g); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, dest, nd * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(s, source, ns * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(g, gain_start, size_t(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(static_cast<float*>(g) + n, gain_end, size_t(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDest = static_cast<float*>(a); dSrc = static_cast<float*>(s); dG0 = static_cast<float*>(g); dG1 = dG0 + n;
	}
	JPLG_CUDA_TRY(ctx, LaunchMixRamp(dDest, dSrc, n, dest_channel, source_channel, dest_num_channels, source_num_channels, frames, dG0, dG1, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(dest, dDest, nd * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

// Add (ChannelMixing.h:32-48). source_num_channels == 1 and source_channel == 0 is the reference's second overload (mono source).
jplg_status jplg_mix_add(jplg_context ctx, float* dest, const float* source, uint32_t n, uint32_t dest_channel, uint32_t source_channel,
						 uint32_t dest_num_channels, uint32_t source_num_channels, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;
	if (!ctx || !dest || !source || dest_channel >= dest_num_channels || source_channel >= source_num_channels) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(ctx);
	if (n == 0 || frames == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nd = size_t(n) * frames * dest_num_channels, ns = size_t(n) * frames * source_num_channels;
	float* dDest = dest; const float* dSrc = source;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *s;
		jplg_status st = Scratch(ctx, 0, nd * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, ns * 4, &s); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, dest, nd * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(s, source, ns * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDest = static_cast<float*>(a); dSrc = static_cast<float*>(s);
	}
	JPLG_CUDA_TRY(ctx, LaunchMixAdd(dDest, dSrc, n, dest_channel, source_channel, dest_num_channels, source_num_channels, frames, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(dest, dDest, nd * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

// DownmixToMono (ChannelMixing.h:129-136): dest [n][frames] += sum over channels of source [n][frames][num_channels]
jplg_status jplg_mix_downmix_to_mono(jplg_context ctx, float* dest, const float* source, uint32_t n, uint32_t num_channels, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;
	if (!ctx || !dest || !source) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(ctx);
	if (n == 0 || frames == 0 || num_channels == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nd = size_t(n) * frames, ns = nd * num_channels;
	float* dDest = dest; const float* dSrc = source;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *s;
		jplg_status st = Scratch(ctx, 0, nd * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, ns * 4, &s); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, dest, nd * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(s, source, ns * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDest = static_cast<float*>(a); dSrc = static_cast<float*>(s);
	}
	JPLG_CUDA_TRY(ctx, LaunchDownmixToMono(dDest, dSrc, n, num_channels, frames, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(dest, dDest, nd * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

jplg_status jplg_mix_apply_gain(jplg_context ctx, float* dest, uint32_t n, uint32_t frames, const float* gains, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!ctx || !dest || !gains) return JPLG_ERR_INVALID_ARGUMENT;
	JPLG_NEED_DEVICE(ctx);
	if (n == 0 || frames == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nd = size_t(n) * frames;
	float* dDest = dest; const float* dG = gains;
	if (mem == JPLG_MEM_HOST)
	{
		void *a, *g;
		jplg_status st = Scratch(ctx, 0, nd * 4, &a); if (st) return st;
		st = Scratch(ctx, 1, size_t(n) * 4, &g); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, dest, nd * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(g, gains, size_t(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
		dDest = static_cast<float*>(a); dG = static_cast<float*>(g);
	}
	JPLG_CUDA_TRY(ctx, LaunchApplyGain(dDest, n, frames, dG, ctx->stream)); ctx->launches++;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(dest, dDest, nd * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

//==================================================================================================
// ReverbBus bank (LateReverb.h:34-117, LateReverb.cpp:31-134)
struct jplg_reverb_bank_t
{
	jplg_context ctx; uint32_t n, L, maxFrames; int mode; float sampleRate, invSampleRate; float coef[27];
	float *state = nullptr, *decay = nullptr, *slope = nullptr, *rings = nullptr, *stageIn = nullptr, *stageOut = nullptr;
	uint64_t pos = 0; bool slopeDirty = true;
};
static void FreeReverb(jplg_reverb_bank b)
{
	for (float* p : { b->state, b->decay, b->slope, b->rings, b->stageIn, b->stageOut }) if (p) cudaFree(p);
}
jplg_status jplg_reverb_bank_create(jplg_context ctx, uint32_t n, float sample_rate, float f1, float f2, float f3,
									uint32_t output_channels, uint32_t max_block_frames, int32_t mode, jplg_reverb_bank* out)
{
	if (!ctx || !out || n == 0 || !(sample_rate > 0.0f)) return JPLG_ERR_INVALID_ARGUMENT;
	// numOutChannels <= cFDNChannels is asserted by the reference (LateReverb.cpp:66)
	if (output_channels == 0 || output_channels > 16) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_create: output_channels must be 1..16 (FDN order)");
	if (max_block_frames == 0) max_block_frames = 512;
	if (max_block_frames > kFdnMaxFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_create: max_block_frames must not exceed 1024 (shortest FDN delay is 1609 samples)");
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	auto* b = new jplg_reverb_bank_t{};
	b->ctx = ctx; b->n = n; b->L = output_channels; b->maxFrames = max_block_frames; b->mode = mode; b->sampleRate = sample_rate;
	b->invSampleRate = 1.0f / sample_rate;                                 // MultiChannelFeedbackBase::Prepare (ReverbDesign.h:598)
	if (f1 == 0.0f) f1 = 176.0f; if (f2 == 0.0f) f2 = 775.0f; if (f3 == 0.0f) f3 = 3408.0f;
	CrossoverCoefficients(sample_rate, f1, f2, f3, b->coef);
	const size_t nCh = size_t(n) * 16;
	auto alloc = [&](float** p, size_t floats) { return cudaMalloc(reinterpret_cast<void**>(p), floats * 4) == cudaSuccess; };
	const bool ok = alloc(&b->state, nCh * 24) && alloc(&b->decay, nCh * 4) && alloc(&b->slope, size_t(n) * 4) && alloc(&b->rings, nCh * kFdnRingSize)
		&& alloc(&b->stageIn, size_t(n) * max_block_frames) && alloc(&b->stageOut, size_t(n) * max_block_frames * output_channels);
	if (!ok) { cudaGetLastError(); FreeReverb(b); delete b; return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "jplg_reverb_bank_create: device allocation failed (rings need n*16*8192*4 bytes)"); }
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(b->state, 0, nCh * 24 * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(b->rings, 0, nCh * kFdnRingSize * 4, ctx->stream));
	k_fill<<<static_cast<uint32_t>((size_t(n) * 4 + 255) / 256), 256, 0, ctx->stream>>>(b->slope, size_t(n) * 4, -60.0f); ctx->launches++;   // mDecaySlopeDbPerSecond{ -60 / 1 } (LateReverb.h:116)
	JPLG_CUDA_TRY(ctx, cudaGetLastError());
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	*out = b;
	return JPLG_OK;
}
void jplg_reverb_bank_destroy(jplg_reverb_bank b)
{
	if (!b) return;
	cudaSetDevice(b->ctx->device);
	cudaStreamSynchronize(b->ctx->stream);
	FreeReverb(b);
	delete b;
}
// slope = clamp(...) per the two setters; `rt60` selects SetRT60 (LateReverb.cpp:31-35) or SetDecaySlope (:37-40)
__global__ void k_reverb_set_slope(const float* __restrict__ v, size_t count, int rt60, float* __restrict__ slope)
{
	const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const float cMinRT60 = 0.03f, cMaxRT60 = 10.0f;
	if (rt60) slope[i] = -60.0f / fminf(fmaxf(v[i], cMinRT60), cMaxRT60);
	else slope[i] = fminf(fmaxf(v[i], -60.0f / cMinRT60), -60.0f / cMaxRT60);
}
static jplg_status ReverbSet(jplg_reverb_bank b, const float* v, int rt60, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || !v) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t count = size_t(b->n) * 4;
	const float* dV = v;
	if (mem == JPLG_MEM_HOST)
	{
		void* a; jplg_status st = Scratch(ctx, 0, count * 4, &a); if (st) return st;
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(a, v, count * 4, cudaMemcpyHostToDevice, ctx->stream));
		dV = static_cast<float*>(a);
	}
	k_reverb_set_slope<<<static_cast<uint32_t>((count + 255) / 256), 256, 0, ctx->stream>>>(dV, count, rt60, b->slope); ctx->launches++;
	JPLG_CUDA_TRY(ctx, cudaGetLastError());
	if (mem == JPLG_MEM_HOST) JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	b->slopeDirty = true;
	return JPLG_OK;
}
jplg_status jplg_reverb_bank_set_rt60(jplg_reverb_bank b, const float* rt60_seconds, int32_t mem) { return ReverbSet(b, rt60_seconds, 1, mem); }
jplg_status jplg_reverb_bank_set_decay_slope(jplg_reverb_bank b, const float* slope_db_per_second, int32_t mem) { return ReverbSet(b, slope_db_per_second, 0, mem); }

static jplg_status ReverbRefreshDecay(jplg_reverb_bank b)
{
	if (!b->slopeDirty) return JPLG_OK;
	jplg_context ctx = b->ctx;
	// mFDN.UpdateFilterParameters(slope) at the top of every ProcessInterleaved (LateReverb.cpp:55); recomputed only when the slope changed
	JPLG_LAUNCH(ctx, JPLG_K_OTHER, LaunchFdnDecay(b->slope, b->n, b->invSampleRate, b->decay, b->mode, ctx->stream));
	b->slopeDirty = false;
	return JPLG_OK;
}
jplg_status jplg_reverb_bank_decay_gains(jplg_reverb_bank b, float* out /* host [n][16][4] */)
{
	if (!b || !out) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	jplg_status st = ReverbRefreshDecay(b); if (st) return st;
	const size_t nCh = size_t(b->n) * 16;
	std::vector<float> soa(nCh * 4);
	JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(soa.data(), b->decay, nCh * 16, cudaMemcpyDeviceToHost, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	for (size_t i = 0; i < nCh; ++i) for (int k = 0; k < 4; ++k) out[i * 4 + k] = soa[size_t(k) * nCh + i];
	return JPLG_OK;
}
jplg_status jplg_reverb_bank_process(jplg_reverb_bank b, const float* in, uint32_t in_channels, float* out, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || (frames && (!in || !out))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (frames == 0) return JPLG_OK;
	if (in_channels == 0) return JPLG_ERR_INVALID_ARGUMENT;
	// Multichannel input: DownmixToMono + ApplyGain(1 / numIn) like the reference (LateReverb.cpp:80-92), for blocks of at most one 480-frame chunk. From
	// its second chunk on the reference assigns `framesProcessed` (not `framesToProcess`) elements to a 480-element scratch array (:83) and over-runs
	// it: there is no defined behaviour to match for longer multichannel blocks, so those are refused.
	if (in_channels != 1 && frames > kChunkFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_process: multichannel input is supported for blocks of at most "
		"480 frames (beyond that the reference's down-mix over-runs its scratch array, LateReverb.cpp:83)");
	if (frames > b->maxFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_process: frames exceeds max_block_frames");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const size_t nIn = size_t(b->n) * frames, nOut = nIn * b->L;
	const float* dIn = in; float* dOut = out;
	if (mem == JPLG_MEM_HOST)
	{
		if (in_channels == 1) { JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->stageIn, in, nIn * 4, cudaMemcpyHostToDevice, ctx->stream)); dIn = b->stageIn; }
		else
		{
			void* raw = nullptr;
			jplg_status sst = Scratch(ctx, 0, nIn * in_channels * 4, &raw); if (sst) return sst;
			JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(raw, in, nIn * in_channels * 4, cudaMemcpyHostToDevice, ctx->stream));
			dIn = static_cast<const float*>(raw);
		}
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->stageOut, out, nOut * 4, cudaMemcpyHostToDevice, ctx->stream));
		dOut = b->stageOut;
	}
	if (in_channels != 1)
	{
		void* mono = nullptr;
		jplg_status sst = Scratch(ctx, 1, nIn * 4, &mono); if (sst) return sst;
		JPLG_LAUNCH(ctx, JPLG_K_OTHER, LaunchReverbDownmix(dIn, b->n, frames, in_channels, static_cast<float*>(mono), ctx->stream));
		dIn = static_cast<const float*>(mono);
	}
	jplg_status st = ReverbRefreshDecay(b); if (st) return st;
	FdnParams fp{}; fp.n = b->n; fp.L = b->L; fp.frames = frames; fp.pos = b->pos; std::memcpy(fp.coef, b->coef, sizeof(fp.coef));
	fp.hadamardScale = FdnHadamardScale();
	fp.state = b->state; fp.decay = b->decay; fp.rings = b->rings; fp.in = dIn; fp.out = dOut;
	JPLG_LAUNCH(ctx, JPLG_K_FDN, LaunchFdn(fp, b->mode, ctx->stream));
	b->pos += frames;
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, nOut * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}

// jplg_reverb_bank_process for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): DEVICE pointers
// only, and the mono down-mix of a multichannel input goes to a file-scope device array (jplg_reverb_kernels.cu) instead of the context's growable
// scratch. No cudaMalloc / cudaFree / allocating routine on this path; n * frames beyond the array's size is refused. Same kernels, same order, same results.
jplg_status jplg_reverb_bank_process_no_alloc(jplg_reverb_bank b, const float* in, uint32_t in_channels, float* out, uint32_t frames)
{
	if (!b || (frames && (!in || !out))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (frames == 0) return JPLG_OK;
	if (in_channels == 0) return JPLG_ERR_INVALID_ARGUMENT;
	if (in_channels != 1 && frames > kChunkFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_process_no_alloc: multichannel input is supported for blocks of at most "
		"480 frames (beyond that the reference's down-mix over-runs its scratch array, LateReverb.cpp:83)");
	if (frames > b->maxFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_process_no_alloc: frames exceeds max_block_frames");
	const size_t nIn = size_t(b->n) * frames;
	if (in_channels != 1 && nIn > kReverbNoAllocMaxFloats)
		return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_reverb_bank_process_no_alloc: instances * frames = " + std::to_string(nIn) + " exceeds the static scratch of "
			+ std::to_string(kReverbNoAllocMaxFloats) + " floats (JPLG_REVERB_NO_ALLOC_MAX_FLOATS); use jplg_reverb_bank_process or a smaller bank / block");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const float* dIn = in;
	if (in_channels != 1)
	{
		float* mono = nullptr;
		JPLG_CUDA_TRY(ctx, ReverbNoAllocScratch(ctx->device, &mono));
		JPLG_LAUNCH(ctx, JPLG_K_OTHER, LaunchReverbDownmix(in, b->n, frames, in_channels, mono, ctx->stream));
		dIn = mono;
	}
	jplg_status st = ReverbRefreshDecay(b); if (st) return st;
	FdnParams fp{}; fp.n = b->n; fp.L = b->L; fp.frames = frames; fp.pos = b->pos; std::memcpy(fp.coef, b->coef, sizeof(fp.coef));
	fp.hadamardScale = FdnHadamardScale();
	fp.state = b->state; fp.decay = b->decay; fp.rings = b->rings; fp.in = dIn; fp.out = out;
	JPLG_LAUNCH(ctx, JPLG_K_FDN, LaunchFdn(fp, b->mode, ctx->stream));
	b->pos += frames;
	return JPLG_OK;
}

//==================================================================================================
// ERBus bank (EarlyReflectionsBus.h:45-249, EarlyReflectionsBus.cpp:83-417): B buses x up to T taps, every tap = smoothed fractional
// delay + four-band filter + gain-ramped pan reading the bus' ONE mono ring. The taps run through k_render (scPerRing = T,
// plain pushes, one mix group per bus); this layer restates SetTaps' add / update / fade-out bookkeeping on slots.
struct ErTap { uint32_t id, slot; int state; };          // state: 1 Rendering, 2 FadingOut, 0 Stopped (realtime side finished the fade)
struct jplg_erbus_bank_t
{
	jplg_context ctx; uint32_t B, T, L, ringSize, maxFrames; int mode; float sampleRate, alpha, coef[27];
	float *state = nullptr, *curMix = nullptr, *tgtMix = nullptr, *tgtDelay = nullptr, *tgtBand = nullptr, *rings = nullptr, *busOut = nullptr,
		  *stageIn = nullptr, *stageOut = nullptr, *stagePer = nullptr;
	uint8_t* flags = nullptr; uint32_t* initList = nullptr;
	size_t stagePerFloats = 0, stageInFloats = 0;
	uint64_t pos = 0;
	// host mirrors (SoA like the device arrays) + per-bus tap tables
	std::vector<float> hMix, hDelay, hBand; std::vector<uint8_t> hFlags; std::vector<uint32_t> hInit;
	std::vector<std::vector<ErTap>> taps; std::vector<std::vector<uint32_t>> freeSlots;
	bool dirty = true;
};
static void FreeErBus(jplg_erbus_bank b)
{
	for (float* p : { b->state, b->curMix, b->tgtMix, b->tgtDelay, b->tgtBand, b->rings, b->busOut, b->stageIn, b->stageOut, b->stagePer }) if (p) cudaFree(p);
	if (b->flags) cudaFree(b->flags);
	if (b->initList) cudaFree(b->initList);
}
static uint32_t BitCeil(uint32_t v) { uint32_t p = 1; while (p < v) p <<= 1; return p; }

jplg_status jplg_erbus_bank_create(jplg_context ctx, uint32_t num_buses, uint32_t max_taps_per_bus, float sample_rate, uint32_t output_channels,
								   uint32_t ring_size, uint32_t max_block_frames, int32_t mode, jplg_erbus_bank* out)
{
	if (!ctx || !out || num_buses == 0 || max_taps_per_bus == 0 || !(sample_rate > 0.0f)) return JPLG_ERR_INVALID_ARGUMENT;
	if (output_channels == 0 || output_channels > kMaxChannelsDev) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_erbus_bank_create: output_channels must be 1..32");
	if (max_taps_per_bus > 128) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_erbus_bank_create: at most 128 taps per bus (one CTA renders a whole bus)");
	if (ring_size && (ring_size & (ring_size - 1))) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_erbus_bank_create: ring_size must be a power of two");
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	auto* b = new jplg_erbus_bank_t{};
	b->ctx = ctx; b->B = num_buses; b->L = output_channels; b->mode = mode; b->sampleRate = sample_rate;
	// slots per bus: a power of two so that 128 / T buses share a CTA; at least 8 so every bus gets its own mix subset
	b->T = std::max(8u, BitCeil(max_taps_per_bus));
	b->ringSize = ring_size ? ring_size : BitCeil(static_cast<uint32_t>(sample_rate * 2.0f));   // DelayLineType(sampleRate * 2) (EarlyReflectionsBus.cpp:74)
	b->maxFrames = max_block_frames ? max_block_frames : 512;
	b->alpha = ExpAlpha(kChunkFrames);                                   // CreateExpSmoothing(d, d, cScratchSize) (:452)
	CrossoverCoefficients(sample_rate, 176.0f, 775.0f, 3408.0f, b->coef); // Filter.Prepare(mSampleRate) (:450)
	const size_t nsc = size_t(b->B) * b->T, L = b->L;
	auto alloc = [&](float** p, size_t floats) { return cudaMalloc(reinterpret_cast<void**>(p), floats * 4) == cudaSuccess; };
	const bool ok = alloc(&b->state, nsc * kStateFloats) && alloc(&b->curMix, nsc * L) && alloc(&b->tgtMix, nsc * L) && alloc(&b->tgtDelay, nsc) && alloc(&b->tgtBand, nsc * 4)
		&& alloc(&b->rings, size_t(b->B) * b->ringSize) && alloc(&b->busOut, size_t(b->B) * b->maxFrames * L) && alloc(&b->stageOut, size_t(b->B) * b->maxFrames * L)
		&& cudaMalloc(reinterpret_cast<void**>(&b->flags), nsc) == cudaSuccess && cudaMalloc(reinterpret_cast<void**>(&b->initList), nsc * 4) == cudaSuccess;
	if (!ok) { cudaGetLastError(); FreeErBus(b); delete b; return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "jplg_erbus_bank_create: device allocation failed"); }
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(b->state, 0, nsc * kStateFloats * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(b->curMix, 0, nsc * L * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaMemsetAsync(b->rings, 0, size_t(b->B) * b->ringSize * 4, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	b->hMix.assign(nsc * L, 0.0f); b->hDelay.assign(nsc, 0.0f); b->hBand.assign(nsc * 4, 1.0f); b->hFlags.assign(nsc, 0);
	b->taps.resize(b->B); b->freeSlots.resize(b->B);
	for (uint32_t i = 0; i < b->B; ++i) for (uint32_t t = b->T; t-- > 0;) b->freeSlots[i].push_back(t);
	*out = b;
	return JPLG_OK;
}
void jplg_erbus_bank_destroy(jplg_erbus_bank b)
{
	if (!b) return;
	cudaSetDevice(b->ctx->device);
	cudaStreamSynchronize(b->ctx->stream);
	FreeErBus(b);
	delete b;
}
uint32_t jplg_erbus_bank_num_taps(jplg_erbus_bank b, uint32_t bus) { return (b && bus < b->B) ? static_cast<uint32_t>(b->taps[bus].size()) : 0; }
uint32_t jplg_erbus_bank_slots_per_bus(jplg_erbus_bank b) { return b ? b->T : 0; }

// ERBus::SetTaps (EarlyReflectionsBus.cpp:239-406) for one bus; host pointers.
jplg_status jplg_erbus_bank_set_taps(jplg_erbus_bank b, uint32_t bus, uint32_t n, const float* filter_gains, const float* delay_seconds,
									 const uint32_t* ids, const float* gains)
{
	if (!b || bus >= b->B || (n && (!filter_gains || !delay_seconds || !ids || !gains))) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	const size_t nsc = size_t(b->B) * b->T; const uint32_t L = b->L;
	auto& taps = b->taps[bus];
	auto slotIndex = [&](uint32_t slot) { return size_t(bus) * b->T + slot; };
	// 1.-2. taps whose fade-out has been rendered are erased and their slot is released (:246-322)
	for (size_t i = 0; i < taps.size();)
	{
		if (taps[i].state == 0) { b->hFlags[slotIndex(taps[i].slot)] = 0; b->freeSlots[bus].push_back(taps[i].slot); taps.erase(taps.begin() + i); }
		else ++i;
	}
	auto inNew = [&](uint32_t id) { for (uint32_t k = 0; k < n; ++k) if (ids[k] == id) return true; return false; };
	// 3. taps that are no longer requested fade out: target channel gains 0, one more chunk is rendered (:329-343)
	for (auto& t : taps)
	{
		if (inNew(t.id) || t.state == 2) continue;
		const size_t sc = slotIndex(t.slot);
		for (uint32_t c = 0; c < L; ++c) b->hMix[size_t(c) * nsc + sc] = 0.0f;
		t.state = 2; b->hFlags[sc] = 1 | 2;
	}
	// 4.-6. update existing taps, create new ones (:352-400)
	uint32_t needed = 0;
	for (uint32_t k = 0; k < n; ++k) { bool found = false; for (auto& t : taps) if (t.id == ids[k]) found = true; if (!found) ++needed; }
	if (needed > b->freeSlots[bus].size())
		return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_erbus_bank_set_taps: more simultaneous taps (incl. taps still fading out) than slots per bus");
	for (uint32_t k = 0; k < n; ++k)
	{
		ErTap* hit = nullptr;
		for (auto& t : taps) if (t.id == ids[k]) { hit = &t; break; }
		if (hit && hit->state == 2) continue;   // a fading tap cannot be restarted (the reference asserts this path dead, :366-376); it is re-created once erased
		if (!hit)
		{
			const uint32_t slot = b->freeSlots[bus].back(); b->freeSlots[bus].pop_back();
			taps.push_back({ ids[k], slot, 1 });
			hit = &taps.back();
			b->hInit.push_back(static_cast<uint32_t>(slotIndex(slot)));   // AllocateRTData: current = target delay / filter gains, channel gains 0 (:421-455)
		}
		const size_t sc = slotIndex(hit->slot);
		for (uint32_t q = 0; q < 4; ++q) b->hBand[size_t(q) * nsc + sc] = filter_gains[4 * size_t(k) + q];
		for (uint32_t c = 0; c < L; ++c) b->hMix[size_t(c) * nsc + sc] = gains[size_t(k) * L + c];
		b->hDelay[sc] = delay_seconds[k] * b->sampleRate;                 // newER.Delay * mSampleRate (:383,393)
		b->hFlags[sc] = 1;
	}
	b->dirty = true;
	return JPLG_OK;
}

// new slots: the realtime state AllocateRTData builds (EarlyReflectionsBus.cpp:421-455)
__global__ void k_er_init_slots(const uint32_t* __restrict__ slots, uint32_t n, uint32_t numSC, uint32_t L, const float* __restrict__ tgtDelay,
								const float* __restrict__ tgtBand, float* __restrict__ state, float* __restrict__ curMix)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t sc = slots[i];
	for (uint32_t k = 0; k < kStateFloats; ++k) state[size_t(k) * numSC + sc] = 0.0f;             // Filter.Prepare, CreateTap: zero state
	state[size_t(ST_DELAY) * numSC + sc] = tgtDelay[sc];                                           // CreateExpSmoothing(delay, delay, ..)
	for (uint32_t q = 0; q < 4; ++q) state[size_t(ST_BAND + q) * numSC + sc] = tgtBand[size_t(q) * numSC + sc];   // FilterGains = filterGains
	for (uint32_t c = 0; c < L; ++c) curMix[size_t(c) * numSC + sc] = 0.0f;                        // ChannelGains filled with 0: the tap fades in
}
// ERBus::ProcessInterleaved head (EarlyReflectionsBus.cpp:100-122): down-mix to mono, ApplyGain(1 / numInputChannels), push
__global__ void k_er_push(float* __restrict__ rings, const float* __restrict__ in, uint32_t numBuses, uint32_t inCh, uint32_t frames,
						  uint32_t ringSize, uint64_t pos)
{
	const uint64_t idx = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
	if (idx >= uint64_t(numBuses) * frames) return;
	const uint32_t bus = static_cast<uint32_t>(idx / frames), f = static_cast<uint32_t>(idx - uint64_t(bus) * frames);
	const float* x = in + (size_t(bus) * frames + f) * inCh;
	float sample = 0.0f;
	for (uint32_t ch = 0; ch < inCh; ++ch) sample += x[ch];
	sample *= 1.0f / static_cast<float>(inCh);
	rings[size_t(bus) * ringSize + (static_cast<uint32_t>(pos + f) & (ringSize - 1))] = sample;
}
__global__ void k_accumulate(float* __restrict__ dst, const float* __restrict__ src, size_t n)
{
	const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
	if (i < n) dst[i] += src[i];
}

static jplg_status ErProcess(jplg_erbus_bank b, const float* in, uint32_t in_channels, float* out, float* perTap, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!b || (frames && (!in || !out)) || in_channels == 0) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = b->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (frames == 0) return JPLG_OK;
	if (frames > b->maxFrames) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_erbus_bank_process: frames exceeds max_block_frames");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	const uint32_t numSC = b->B * b->T, L = b->L;
	const size_t nIn = size_t(b->B) * frames * in_channels, nOut = size_t(b->B) * frames * L, nPer = size_t(numSC) * frames * L;
	// ---- SafeERsRead: the audio thread picks up the latest tap data (:141-143)
	if (b->dirty)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->tgtMix, b->hMix.data(), size_t(numSC) * L * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->tgtDelay, b->hDelay.data(), size_t(numSC) * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->tgtBand, b->hBand.data(), size_t(numSC) * 16, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->flags, b->hFlags.data(), numSC, cudaMemcpyHostToDevice, ctx->stream));
		if (!b->hInit.empty())
		{
			const uint32_t n = static_cast<uint32_t>(b->hInit.size());
			JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->initList, b->hInit.data(), size_t(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
			k_er_init_slots<<<(n + 127) / 128, 128, 0, ctx->stream>>>(b->initList, n, numSC, L, b->tgtDelay, b->tgtBand, b->state, b->curMix); ctx->launches++;
			JPLG_CUDA_TRY(ctx, cudaGetLastError());
		}
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));   // the host mirrors may be rewritten by the next set_taps
		b->hInit.clear();
		b->dirty = false;
	}
	const float* dIn = in; float* dOut = out; float* dPer = perTap;
	if (mem == JPLG_MEM_HOST)
	{
		if (nIn > b->stageInFloats)
		{
			if (b->stageIn) cudaFree(b->stageIn);
			b->stageIn = nullptr; b->stageInFloats = 0;
			if (cudaMalloc(reinterpret_cast<void**>(&b->stageIn), nIn * 4) != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "erbus input staging allocation failed"); }
			b->stageInFloats = nIn;
		}
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->stageIn, in, nIn * 4, cudaMemcpyHostToDevice, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(b->stageOut, out, nOut * 4, cudaMemcpyHostToDevice, ctx->stream));
		dIn = b->stageIn; dOut = b->stageOut;
		if (perTap)
		{
			if (nPer > b->stagePerFloats)
			{
				if (b->stagePer) cudaFree(b->stagePer);
				b->stagePer = nullptr; b->stagePerFloats = 0;
				if (cudaMalloc(reinterpret_cast<void**>(&b->stagePer), nPer * 4) != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "per-tap debug buffer allocation failed"); }
				b->stagePerFloats = nPer;
			}
			dPer = b->stagePer;
		}
	}
	{
		const uint64_t total = uint64_t(b->B) * frames;
		TimeBegin(ctx, JPLG_K_ER_PUSH);
		k_er_push<<<static_cast<uint32_t>((total + 255) / 256), 256, 0, ctx->stream>>>(b->rings, dIn, b->B, in_channels, frames, b->ringSize, b->pos);
		TimeEnd(ctx); ctx->launches++;
		JPLG_CUDA_TRY(ctx, cudaGetLastError());
	}
	RenderParams rp{};
	rp.numSC = numSC; rp.srcChannels = 1; rp.L = L; rp.ringSize = b->ringSize; rp.ringMask = b->ringSize - 1; rp.pos = b->pos; rp.frames = frames;
	rp.alpha = b->alpha; std::memcpy(rp.coef, b->coef, sizeof(rp.coef));
	rp.state = b->state; rp.curMix = b->curMix; rp.tgtDelay = b->tgtDelay; rp.tgtBand = b->tgtBand; rp.tgtMix = b->tgtMix; rp.rings = b->rings;
	rp.in = nullptr; rp.partial = b->busOut; rp.perSource = dPer;
	rp.numCTAs = RenderNumCTAs(numSC, 1);
	rp.scPerRing = b->T; rp.plainPush = 1; rp.groupsPerCTA = 128 / b->T; rp.numGroups = b->B; rp.flags = b->flags;
	JPLG_LAUNCH(ctx, JPLG_K_RENDER, LaunchRender(rp, b->mode, 0, rp.numCTAs, ctx->stream));
	k_accumulate<<<static_cast<uint32_t>((nOut + 255) / 256), 256, 0, ctx->stream>>>(dOut, b->busOut, nOut); ctx->launches++;   // the bus adds into `output`
	JPLG_CUDA_TRY(ctx, cudaGetLastError());
	b->pos += frames;
	// a tap that was fading out has rendered its last chunk: realtime state Stopped (:218-220); SetTaps erases it next time
	for (uint32_t bus = 0; bus < b->B; ++bus)
		for (auto& t : b->taps[bus])
			if (t.state == 2) { t.state = 0; b->hFlags[size_t(bus) * b->T + t.slot] = 0; b->dirty = true; }
	if (mem == JPLG_MEM_HOST)
	{
		JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, nOut * 4, cudaMemcpyDeviceToHost, ctx->stream));
		if (perTap) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(perTap, dPer, nPer * 4, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}
jplg_status jplg_erbus_bank_process(jplg_erbus_bank b, const float* in, uint32_t in_channels, float* out, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	return ErProcess(b, in, in_channels, out, nullptr, frames, mem);
}
jplg_status jplg_erbus_bank_process_debug(jplg_erbus_bank b, const float* in, uint32_t in_channels, float* out, float* per_tap, uint32_t frames, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!per_tap) return JPLG_ERR_INVALID_ARGUMENT;
	return ErProcess(b, in, in_channels, out, per_tap, frames, mem);
}
// slot (row of per_tap) of the tap with this id, or -1
int32_t jplg_erbus_bank_tap_slot(jplg_erbus_bank b, uint32_t bus, uint32_t id)
{
	if (!b || bus >= b->B) return -1;
	for (auto& t : b->taps[bus]) if (t.id == id) return static_cast<int32_t>(bus * b->T + t.slot);
	return -1;
}

#include "jplg_api_frontend.inl"
#include "jplg_api_exchange.inl"

} // extern "C"