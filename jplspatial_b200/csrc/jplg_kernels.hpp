// Internal interface between the C-ABI layer (jplg_api.cu) and the kernels.
#pragma once

#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>

#define JPLG_MDAP_WARPS 4

namespace jplg
{
	constexpr uint32_t kMaxChannelsDev = 32;       // VBAPBaseTraits::MAX_CHANNELS
	constexpr uint32_t kMaxSourceChannelsDev = 8;  // MAX_SOURCE_CHANNELS
	constexpr uint32_t kMaxAzimuthFloats = 64;

	struct PannerDev
	{
		bool is3D = false;
		uint32_t numChannels = 0;
		const uint4* lut3d = nullptr;    // 65536 cells {g0,g1,g2, spk0|spk1<<8|spk2<<16} (repacked KB_983 LUT, 1 MiB, L2 resident)
		const float* lut2d = nullptr;    // [1024][L]
		const uint32_t* rsqTable = nullptr; // 2048-entry RSQRTPS model
	};

	struct MdapLayoutParams
	{
		uint32_t numChannels, numVS, numRings, samplesPerRing, halfSamples, tail;
		float phiTerm, offsetTerm, vsWeight;
		uint32_t paramStride;   // 5 or 11 floats per source (set by LaunchMdap)
		float quats[kMaxSourceChannelsDev][4];
		float azSin[kMaxAzimuthFloats], azCos[kMaxAzimuthFloats];
	};

	cudaError_t LaunchPointGains(const PannerDev& p, const float* dirs, uint32_t n, float* out, cudaStream_t st);
	// Optional fused epilogue of k_mdap_3d (jplg_renderer_update_panned): the normalised mix maps go straight into a renderer's target arrays
	// (SoA [L][numSC]; the layout's channel count is the renderer's source channel count) together with the band gains and the delay in samples,
	// i.e. what k_update_targets would make of them. tgtMix == nullptr: plain MDAP.
	struct MdapFuse
	{
		float* tgtMix = nullptr; float* tgtBand = nullptr; float* tgtDelay = nullptr;
		const float* band = nullptr; const float* delaySec = nullptr;
		float sampleRate = 0.0f; uint32_t numSC = 0;
		// renderer channels per source (>= the layout's panned channels) and the renderer row of every panned channel group: a 5.1 source is
		// rendered with 6 channels, its LFE (skipped by the panner, PannerBase.h:937) keeps the all-zero mix row it was created with
		uint32_t rendC = 0; uint8_t rows[kMaxSourceChannelsDev] = {};
		uint32_t unpannedRows = 0;   // bit k: renderer row k has no panned channel group; its targets are written as zeros
	};
	cudaError_t LaunchMdap(const PannerDev& p, const MdapLayoutParams& lp, const float* params, uint32_t paramStride, uint32_t n, float* out, int numSMs, cudaStream_t st,
						   const MdapFuse* fuse = nullptr);
	cudaError_t LaunchProcessVS(const PannerDev& p, const float* vs, uint32_t n, uint32_t batches, float* out, cudaStream_t st);

	//======================================================================
	// Renderer
	constexpr uint32_t kStateFloats = 32;   // device-side DSP state floats per source-channel (JPLG_STATE_FLOATS = 64 adds the 32 current mix gains)
	constexpr uint32_t kChunkFrames = 480;  // DirectSoundEffect::cScratchSize (inc/Auralization/DirectSound.h:90)

	// state float indices (SoA: state[k * numSC + sc])
	enum : uint32_t
	{
		ST_DELAY = 0,      // SmoothedValue::Current (samples)
		ST_THIRAN_Y1 = 1,  // Thiran1stInterpolator::mY1
		ST_FILTER = 2,     // 24 floats: X2.S1{s1,s2} X2.S2{s1,s2} X13.S1 lane0{s1,s2} lane1{s1,s2} X13.S2 lanes0..3{s1,s2} AP lanes0..3{s1,s2}
		ST_BAND = 26,      // current band gains b1..b4 (mFilterGains)
		ST_SPARE = 30
	};

	struct RenderParams
	{
		uint32_t numSC;        // source-channels = sources * channels per source
		uint32_t srcChannels;  // channels per source (C)
		uint32_t L;            // output channels
		uint32_t ringSize, ringMask;
		uint64_t pos;          // ring pushes before this block
		uint32_t frames;
		float alpha;           // exp smoothing coefficient for cScratchSize ticks
		float coef[27];        // crossover coefficients (see jplg_host.hpp)
		float* state;          // [kStateFloats][numSC]
		float* curMix;         // [L][numSC] current channel gains (mCurrentGains)
		const float* tgtDelay; // [numSC] target delay in samples
		const float* tgtBand;  // [4][numSC]
		const float* tgtMix;   // [L][numSC]
		float* rings;          // [numSC][ringSize], sample pushed at time t lives at t & ringMask
		const float* in;       // [numSources][frames][C]
		float* partial;        // [numCTAs][frames][L] per-CTA partial buses
		float* perSource;      // debug: [numSources][frames][L] or nullptr
		uint32_t numCTAs;
		uint32_t ctaBase;      // first CTA of this launch (set by LaunchRender)
		// ---- generalisations used by the ERBus bank (all zero / null for N x DirectSoundEffect)
		uint32_t scPerRing;    // consecutive source-channels that read the SAME ring (ER taps of one bus); 0 or 1 = one ring each
		uint32_t plainPush;    // 1: a chunk pushes only its own frames (ERBus, EarlyReflectionsBus.cpp:119-122); 0: DirectSoundEffect's quirk
		uint32_t groupsPerCTA; // mix groups per CTA (ER: buses per CTA = 128 / taps per bus), each with its own partial bus; 0 or 1 = one
		uint32_t numGroups;    // total number of partial buses (groups) when groupsPerCTA > 1
		uint32_t fusedPush;    // 1: k_render's push role writes this block's input into the rings itself (mono sources; set by LaunchRender's caller via RenderCanFusePush)
		const uint8_t* flags;  // per source-channel: bit 0 = live, bit 1 = fading out (rendered for the first chunk of the block only); null = all live
	};

	uint32_t RenderNumCTAs(uint32_t numSC, uint32_t srcChannels);
	uint32_t TotalPushes(uint32_t frames); // ring pushes one block of `frames` causes (reference quirk, see k_push)
	uint32_t RenderSourcesPerCTA(uint32_t srcChannels);
	bool RenderCanFusePush(const RenderParams& rp);   // conditions of k_render's push role (mono, 4-aligned block / ring position, input 16-byte aligned)
	cudaError_t LaunchPush(const RenderParams& rp, uint32_t srcBegin, uint32_t srcCount, cudaStream_t st);
	cudaError_t LaunchRender(const RenderParams& rp, int mode, uint32_t ctaBegin, uint32_t ctaCount, cudaStream_t st, cudaEvent_t started = nullptr);
	// bus[frames][L] (+)= sum over CTAs of partial; also writes the plain sum to `deviceBus`
	cudaError_t LaunchBusReduce(const RenderParams& rp, float* deviceBus, float* busAccumulate, cudaStream_t st);
	cudaError_t LaunchFp32Probe(float* sink, uint32_t grid, uint32_t iters, cudaStream_t st);
	cudaError_t LaunchStateTranspose(const float* in, float* out, uint32_t rows, uint32_t cols, cudaStream_t st);

	// ---- time-parallel throughput path (jplg_scan_kernels.cu): the block as a scan over kScanTile-frame tiles
	constexpr uint32_t kScanTile = 16;      // 480 = 30 tiles: a tile never straddles one of DirectSoundEffect's chunks
	struct ScanParams
	{
		RenderParams rp;
		const float* M;         // [24][24] zero-input state transition of the crossover over kScanTile frames (host: CrossoverTransition)
		float* scratch;         // ScanScratchFloats(numSC, maxFrames) floats
		uint32_t* counters;     // [numCTAs + maxFrames / kScanTile], zero between launches: tiles done per source group, source groups done per tile
		float* deviceBus;       // [frames][L] the block's bus (k_scan_mix reduces the partial buses itself: no LaunchBusReduce on this path)
		float* busAcc;          // caller's bus, accumulated into (may be null)
		uint32_t push;          // 1: k_scan_front also pushes the block's input into the rings (0: already pushed)
		// carved out of `scratch` by LaunchRenderScan
		float* xth;             // [frames][numSC] filtered, band-gained source signals
		float* tileGain;        // [nTiles][numSC][L + 4] gain ramps at the tile starts
		uint32_t nTiles;
	};
	size_t ScanScratchFloats(uint32_t numSC, uint32_t maxFrames);
	bool RenderCanScan(const RenderParams& rp);   // plain N x DirectSoundEffect, whole tiles
	cudaError_t LaunchRenderScan(const ScanParams& sp, cudaStream_t st);

	// components
	struct XoverParams { uint32_t n, frames; float coef[27]; float* state; /* [24][n] */ };
	cudaError_t LaunchXoverRamp(const XoverParams& xp, const float* in, const float* g0, const float* g1, float* out, int mode, cudaStream_t st);
	cudaError_t LaunchXoverSplit(const XoverParams& xp, const float* in, float* outBands, int mode, cudaStream_t st);
	cudaError_t LaunchDelayPush(float* rings, uint32_t n, uint32_t ringSize, uint64_t pos, const float* in, uint32_t frames, cudaStream_t st);
	cudaError_t LaunchDelayRead(const float* rings, uint32_t n, uint32_t ringSize, uint64_t pos, const float* delays, uint32_t taps,
								int interp, float* thiranY1, float* out, cudaStream_t st);
	cudaError_t LaunchDelayReadXFade(const float* rings, uint32_t n, uint32_t ringSize, uint64_t pos, const float* delays, const float* alphas,
									 uint32_t taps, int32_t* lastNew, float* out, cudaStream_t st);
	cudaError_t LaunchMixRamp(float* dest, const float* src, uint32_t n, uint32_t destCh, uint32_t srcCh, uint32_t nDest, uint32_t nSrc,
							  uint32_t frames, const float* g0, const float* g1, cudaStream_t st);
	cudaError_t LaunchApplyGain(float* dest, uint32_t n, uint32_t frames, const float* gains, cudaStream_t st);
	cudaError_t LaunchMixAdd(float* dest, const float* src, uint32_t n, uint32_t destCh, uint32_t srcCh, uint32_t nDest, uint32_t nSrc, uint32_t frames, cudaStream_t st);
	cudaError_t LaunchDownmixToMono(float* dest, const float* src, uint32_t n, uint32_t numChannels, uint32_t frames, cudaStream_t st);
	//======================================================================
	// ReverbBus bank (jplg_reverb_kernels.cu)
	constexpr uint32_t kFdnRingSize = 8192;   // power of two above the longest FDN delay + 1 (4259); the reference sizes each line bit_ceil(delay + 1)
	constexpr uint32_t kFdnMaxFrames = 1024;  // a call must not be longer than the shortest FDN delay (1609): no feedback dependency inside a call
	struct FdnParams
	{
		uint32_t n, L, frames;
		uint64_t pos;           // samples pushed into every ring so far
		float coef[27];         // crossover coefficients (jplg_host.hpp)
		float hadamardScale;    // Hadamard::MixInPlace scalingFactor (see FdnHadamardScale)
		float* state;           // [24][n*16] crossover state, SoA
		const float* decay;     // [4][n*16] per-channel decay gains b1..b4
		float* rings;           // [n*16][kFdnRingSize]
		const float* in;        // [n][frames] mono
		float* out;             // [n][frames][L], accumulated into
	};
	// Hadamard::MixInPlace's `static constexpr Sample scalingFactor = Math::Sqrt(1.0 / Size)` (ReverbDesign.h:443-446) is evaluated at compile time
	// by the reference's constexpr fallback on a DOUBLE (Quake-style seed + 2 Newton steps on 1/sqrt, no polish: inc/Math/Math.h:213-262), then
	// narrowed to float: 0x3e7fffb8 = 0.249998927 for Size = 16, not 0.25. Restated on the host so the value is derived, not pasted.
	float FdnHadamardScale();
	cudaError_t LaunchFdnDecay(const float* slope, uint32_t n, float invSampleRate, float* decay, int mode, cudaStream_t st);
	cudaError_t LaunchFdn(const FdnParams& fp, int mode, cudaStream_t st);
	cudaError_t LaunchReverbDownmix(const float* in, uint32_t n, uint32_t frames, uint32_t numIn, float* mono, cudaStream_t st);   // [n][frames][numIn] -> [n][frames]
	// jplg_reverb_bank_process_no_alloc: device address of the file-scope mono down-mix scratch (jplg_reverb_kernels.cu), resolved once per device
	// with cudaGetSymbolAddress; `device` must be current. kReverbNoAllocMaxFloats = JPLG_REVERB_NO_ALLOC_MAX_FLOATS (include/jplg.h).
	constexpr size_t kReverbNoAllocMaxFloats = 245760;
	cudaError_t ReverbNoAllocScratch(int device, float** out);
	//======================================================================
	// Scene front-end (jplg_frontend_kernels.cu): the per-source body of SpatialManager::AdvanceSimulation
	// (inc/SpatialManager.h:374-420) plus the parameter producers that sit between it and stages 1-3.
	struct AttFunctionDev     // one AttenuationFunction (inc/DistanceAttenuation.h:41-162)
	{
		int32_t kind;          // 0: AttenuationBaseModel, 1: AttenuationCurve
		int32_t model;         // AttenuationModel
		float minDistance, maxDistance, rolloff;
		uint32_t firstPoint, numPoints; // curve: range in the point table (sorted by distance)
	};
	struct AirCacheDev        // the fields of AirAbsorptionCache that ComputeForDistance reads (inc/AirAbsorption.h:212-234)
	{
		float centerDb[4], lowerThirdDb[4], distanceToKnee[4], linFactor[4];
		uint32_t hasKnee[4];
	};
	struct FrontendArgs
	{
		uint32_t n;
		// ---- inputs (null = absent)
		const float* sources;          // [n][9] Position: location, up, forward
		const float* listeners;        // [m][9]
		const uint32_t* listenerIndex; // [n] (null: listener 0)
		const float* distanceIn;       // [n] used instead of the direct path when sources == null
		const float* dotIn;            // [n] cone input (InvDirectionDot) when sources == null
		const float* cones;            // [n][2] inner, outer angle (radians)
		const float* sizes;            // [n] object size; < 0 = the source has none (spread unchanged)
		const float* focusSpread;      // [n][2] PanEffectParameters (null: focusConst / spreadConst)
		float focusConst, spreadConst;
		const uint32_t* attIndex;      // [n] index into att (null: 0)
		const AttFunctionDev* att;     // attenuation function table
		uint32_t numAtt;
		const float* attPoints;        // [p][3] distance, value, Curve::EType
		AirCacheDev air;
		float invSpeedOfSound;         // SpeedOfSound::cInvReference or 1 / ForTemperature(T)
		float fallbackDir[3];          // SanitizeSourceDir's fallback (PanningService.h:487)
		int32_t spreadFlavour;         // 0: SpatialManager::GetSpreadFromSourceSize (SpatialManager.h:617-623); 1: the free function (PannerBase.h:69-74)
		// ---- outputs (null = not wanted)
		float* direct;     // [n][4] distance, directionDot, invDirectionDot, cone factor
		float* rel;        // [n][9] DirectPathResult::Position (direction, up, forward in the listener's frame)
		float* pan;        // [n][panStride] {sanitised dir, focus, spread} (+ {up, forward} when panStride == 11): the input of LaunchMdap
		uint32_t panStride;
		float* distAtt;    // [n]
		float* airDb;      // [n][4] ComputeForDistance
		float* bandGains;  // [n][4] dBToGain(-airDb) (simd flavour): DirectSoundEffect's filterGains
		float* delay;      // [n] seconds
	};
	cudaError_t LaunchFrontend(const FrontendArgs& fa, int numSMs, cudaStream_t st);
	//======================================================================
	// Peer-memory bus exchange (jplg_exchange_kernels.cu)
	constexpr uint32_t kMaxWorld = 16;
	struct ExchangeHeader
	{
		unsigned long long ack;                 // blocks consumed by the root
		uint32_t done, error;
		unsigned long long flags[2][kMaxWorld]; // (seq + 1) last published by a rank into that slot
	};
	constexpr size_t kExchangeDataOffset = 512;   // bytes: header padded so the slot data is 256-byte aligned
	static_assert(sizeof(ExchangeHeader) <= kExchangeDataOffset, "exchange header too large");
	// spinLimitCycles: device-side wait time-out (SM clock cycles); localError: this rank's host-mapped error word
	cudaError_t LaunchBusSend(ExchangeHeader* hdr, float* slotData, const float* bus, uint32_t n, uint32_t rank, uint64_t seq, long long spinLimitCycles,
							  uint32_t* localError, cudaStream_t st);
	cudaError_t LaunchBusGather(ExchangeHeader* hdr, const float* data, uint32_t world, uint32_t maxFloats, const float* own, uint32_t n, uint64_t seq,
								float* out, int accumulate, long long spinLimitCycles, uint32_t* localError, cudaStream_t st);
}
