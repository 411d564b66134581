// Time-parallel render path of JPLG_MODE_THROUGHPUT: stages K2 + K3 + K4 of N x DirectSoundEffect as a chunked linear-recurrence scan over
// kScanTile-frame tiles (src/Auralization/DirectSound.cpp:95-229; BASELINE.json north_star stage 3).
//
// k_render (jplg_render_kernels.cu) walks a block sample by sample inside one thread per source-channel: 512 dependent steps of ~200 operations,
// which leaves a small scene (a few thousand source-channels: fewer source-warps than the machine has schedulers) latency-bound on a handful of
// SMs. Here only the two recurrences whose ROUNDING is part of the result stay sequential, as three-operation chains; everything else is spread
// over (source-channel, tile) threads. k_scan_front: one CTA = 4 source-channels x 32 tile slots, phases separated by __syncthreads:
//
//   P0  all threads: DelayLine::Push of the block (with the reference's chunk quirk, DirectSound.cpp:107-128; see k_push): loads before P1, stores after it
//   P1  warp 0, lanes 0-3 = the four source-channels in lock step, sequential: SmoothedValue::Tick for every frame (inc/Utilities/SmoothedValue.h:73-77).
//       The smoothed delay decides the integer / fraction split of every tap (at 40,000 samples of delay one float ulp is 0.004 samples), so it is the
//       reference's recurrence bit for bit, not a closed form. Warps 1-3 meanwhile run the reference's gain-ramp recurrences (`gain += delta` per frame,
//       ChannelMixing.h:99-107, CrossoverFilter.h:518-531 — a constant increment rounds the same way step after step, the accumulated ramp drifts up to
//       480 x 0.5 ulp from start + n * delta, so the closed form is outside the bar) and leave the ramp values at the tile starts.
//   P2  per (sc, tile), parallel: DelayTap::SetDelay -> Thiran coefficient (1 - f) / (1 + f) -> GetReadWindow<2> (DelayLine.h:53-59,199-213;
//       FractionalDelay.h:105-110): { a[n], x[n], x[n-1] } as one float4 per sample
//   P3  lanes 0-3 of warp 0, sequential: y[n] = a[n] (x[n] - y[n-1]) + x[n-1] (FractionalDelay.h:117-124), bit for bit: for a static source on an
//       exactly-integer delay a = 1, the all-pass has its pole ON the unit circle and a differently-rounded state never decays —
//       an affine scan of this stage (y_end = P y_start + Q per tile) drifted past the bar within 20 blocks (r02 measurement).
//   P4  per (sc, tile), parallel: the FourBandCrossover (CrossoverFilter.h:505-535) on the tile's delayed samples from ZERO state: the zero-state
//       end state e (24 floats)
//   P5  one warp per sc: the 24-state scan over the tiles, s[t+1] = M s[t] + e[t], with M = the crossover's zero-input transition over
//       kScanTile frames (fixed per sample rate, built on the host in double: CrossoverTransition); leaves every tile's true START state
//   P6  per (sc, tile), parallel: re-runs the tile's crossover from its start state with the ramped band gains -> Y [frame][sc]
//
// k_scan_mix: per (128 sc, tile) CTA, the gain-ramped mix of the CTA's source-channels into a partial bus (thread mapping of k_render's mix role); the
// last CTA of a tile to finish reduces the tile's partial buses over the source groups in a fixed order (no k_bus_reduce launch on this path), the last
// tile-CTA of a source group commits mCurrentGains = targets (DirectSound.cpp:221). Launched as a programmatic dependent of k_scan_front.
//
// With one or two warps per scheduler every instruction costs its full latency, so these kernels are as long as their longest per-thread instruction
// sequence: the design rule behind the float4 taps, the lock-step recurrence lanes and the ramps in warps of their own (profiles/r02v_scan_pipe_timeline.txt).
// The crossover is stable (its state errors decay), so the scan's re-association there stays at rounding level. Twice the filter arithmetic of
// the sequential kernel (zero-state pass + final pass) plus the 24 x 24 scan: it only pays while the machine is under-filled; the launcher's
// caller (jplg_api.cu) selects it by size. Bar: 1e-5 * max(|ref|, rms, 1e-3) per sample against oracle/_ref, tests/test_gpu_scan.py.
#include "jplg_kernels.hpp"
#include "jplg_render_common.cuh"
#include <cstdlib>
#include <algorithm>

namespace jplg
{
	namespace
	{
		constexpr uint32_t kT = kScanTile;
		constexpr uint32_t kScanThreads = 128;
		constexpr uint32_t kFrontSC = 4;                               // source-channels per k_scan_front CTA (one per warp in the sequential phases)
		constexpr uint32_t kFrontSlots = kScanThreads / kFrontSC;      // tile slots per source-channel

		struct ScanTile { uint32_t n0, chunkStart, chunkLen, t0; };
		__device__ __forceinline__ ScanTile scan_tile(uint32_t t, uint32_t F)
		{
			ScanTile ti;
			ti.n0 = t * kT;
			const uint32_t k = ti.n0 / kChunkFrames;            // 480 = 30 * 16: a tile never straddles a chunk
			ti.chunkStart = k * kChunkFrames;
			ti.chunkLen = min(F - ti.chunkStart, kChunkFrames);
			ti.t0 = ti.n0 - ti.chunkStart;
			return ti;
		}
	}

	// dynamic shared memory: float4 taps [4][nTiles * 17 + 2] | e [4][nTiles][25] + 32 floats | a 32-float staging row per scan warp
#ifdef JPLG_PROF
	__device__ unsigned long long g_frontProf[16];
	#define FRONT_STAMP(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_frontProf[i] = clock64(); } while (0)
	extern "C" __attribute__((visibility("default"))) void jplg_debug_front_prof(unsigned long long* out) { cudaMemcpyFromSymbol(out, g_frontProf, sizeof(g_frontProf)); }
#else
	#define FRONT_STAMP(i)
#endif
	__global__ void __launch_bounds__(kScanThreads)
	k_scan_front(const ScanParams sp)
	{
		extern __shared__ __align__(16) float sF[];
		const RenderParams& rp = sp.rp;
		const uint32_t tid = threadIdx.x, F = rp.frames, nT = sp.nTiles, C = rp.srcChannels, L = rp.L, nG = L + 4;
		const size_t nSC = rp.numSC;
		const uint32_t scBase = blockIdx.x * kFrontSC;
		// One float4 per (source-channel q, frame n) at q * tapRow + (n / 16) * 17 + n % 16 (tiles 17 apart, rows + 2: a quarter-warp's STS.128 in P2 is
		// conflict-free). Its life: P1 leaves the smoothed delay in .x; P2 replaces it with the taps { a, x[n], x[n-1], . }, so that P3 fetches a sample's
		// operands with ONE LDS.128 (every memory instruction in a single-warp recurrence costs ~2 cycles of its chain); P3 leaves the delayed sample in .w.
		const uint32_t tapRow = nT * (kT + 1) + 2u;
		float4* sTap = reinterpret_cast<float4*>(sF);
		float* sE = reinterpret_cast<float*>(sTap + kFrontSC * tapRow);      // state k of (q, tile t) at (q * nT + t) * 25 + k + 8 q
		const uint32_t pos32 = static_cast<uint32_t>(rp.pos);
		FRONT_STAMP(0);
		asm volatile("griddepcontrol.launch_dependents;" ::: "memory");     // k_scan_mix may be scheduled now: its CTAs wait (griddepcontrol.wait) for this grid to finish

		// ---- P0 (loads): ring pushes of the CTA's source-channels (chunk k pushes in[480k .. F) at push offset PushesThroughChunk(F, k)). A thread keeps
		// frames f = tid + 128 j of all four source-channels in registers; the loads are in flight while P1 runs, the stores follow it.
		const uint32_t numChunks = (F + kChunkFrames - 1) / kChunkFrames;
		const bool pushInRegs = sp.push && F <= 4 * kScanThreads;
		float pv[kFrontSC][4];
		if (pushInRegs)
		{
			#pragma unroll
			for (uint32_t qq = 0; qq < kFrontSC; ++qq)
			{
				const uint32_t s0 = min(scBase + qq, rp.numSC - 1u), src = s0 / C, ch = s0 - src * C;
				const float* in = rp.in + size_t(src) * F * C + ch;
				#pragma unroll
				for (uint32_t j = 0; j < 4; ++j) { const uint32_t f = tid + kScanThreads * j; pv[qq][j] = f < F ? __ldcs(in + size_t(f) * C) : 0.0f; }
			}
		}
		else if (sp.push)
		{
			for (uint32_t qq = 0; qq < kFrontSC && scBase + qq < rp.numSC; ++qq)      // long blocks: plain element-wise push
			{
				const uint32_t s0 = scBase + qq, src = s0 / C, ch = s0 - src * C;
				const float* in = rp.in + size_t(src) * F * C + ch;
				float* ring = rp.rings + size_t(s0) * rp.ringSize;
				for (uint32_t f = tid; f < F; f += kScanThreads)
				{
					const float v = in[size_t(f) * C];
					for (uint32_t k = 0; k < numChunks && k * kChunkFrames <= f; ++k) ring[(pos32 + PushesThroughChunk(F, k) + (f - k * kChunkFrames)) & rp.ringMask] = v;
				}
			}
		}
		// ---- P1: warp 0, lanes 0-3 = the four source-channels in lock step: the smoothed delay of every frame (three dependent operations per frame).
		// Warps 1-3 meanwhile run the reference's gain-ramp recurrences (`gain += delta` per frame, ChannelMixing.h:99-107, CrossoverFilter.h:518-531) for the
		// L mix gains + 4 band gains of the four source-channels and leave them at every tile start of the first chunk — in warps of their own, because
		// independent work interleaved INTO a recurrence's instruction stream slows it (measured: + 5 k cycles with the ramps in the delay warp).
		if (tid < 32)
		{
			const uint32_t s1 = scBase + tid;
			if (tid < kFrontSC && s1 < rp.numSC)
			{
				float cur = rp.state[ST_DELAY * nSC + s1];
				const float tgt = rp.tgtDelay[s1], alpha = rp.alpha;
				float4* o = sTap + tid * tapRow;
				for (uint32_t t = 0; t < nT; ++t)
				{
					#pragma unroll
					for (uint32_t i = 0; i < kT; ++i)
					{
						cur = alpha * (tgt - cur) + cur;     // same expression as role D of k_render (the smoother does not restart at chunk boundaries)
						o[t * (kT + 1) + i].x = cur;
					}
				}
				rp.state[ST_DELAY * nSC + s1] = cur;
			}
		}
		else
		{
			const uint32_t chunk0 = min(F, kChunkFrames), tilesRamped = (chunk0 + kT - 1) / kT;
			float g[2], dg[2]; float* gp[2];
			#pragma unroll
			for (uint32_t h = 0; h < 2; ++h)
			{
				const uint32_t v = (tid - 32u) + 96u * h, qq = v / nG, j = v - qq * nG, s1 = scBase + qq;
				float c0 = 0.0f, tg = 0.0f;
				gp[h] = nullptr;
				if (qq < kFrontSC && s1 < rp.numSC)
				{
					if (j < L) { c0 = rp.curMix[j * nSC + s1]; tg = rp.tgtMix[j * nSC + s1]; }
					else { c0 = rp.state[(ST_BAND + j - L) * nSC + s1]; tg = rp.tgtBand[(j - L) * nSC + s1]; }
					gp[h] = sp.tileGain + size_t(s1) * nG + j;        // + t * nSC * nG: [tile][sc][L + 4]
				}
				g[h] = c0; dg[h] = (tg - c0) / static_cast<float>(chunk0);
			}
			for (uint32_t t = 0; t < tilesRamped; ++t)
			{
				if (gp[0]) gp[0][size_t(t) * nSC * nG] = g[0];
				if (gp[1]) gp[1][size_t(t) * nSC * nG] = g[1];
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i) { g[0] += dg[0]; g[1] += dg[1]; }
			}
		}
		// ---- P0 (stores)
		if (pushInRegs)
		{
			#pragma unroll
			for (uint32_t qq = 0; qq < kFrontSC; ++qq)
			{
				if (scBase + qq >= rp.numSC) break;
				float* ring = rp.rings + size_t(scBase + qq) * rp.ringSize;
				#pragma unroll
				for (uint32_t j = 0; j < 4; ++j)
				{
					const uint32_t f = tid + kScanThreads * j;
					if (f >= F) continue;
					for (uint32_t k = 0; k < numChunks && k * kChunkFrames <= f; ++k)
						ring[(pos32 + PushesThroughChunk(F, k) + (f - k * kChunkFrames)) & rp.ringMask] = pv[qq][j];
				}
			}
		}
		__syncthreads();
		FRONT_STAMP(1);
		// ---- P2: taps. thread = (tile slot, source-channel q = tid & 3): the four source-channels' rows interleave across lanes
		const uint32_t q = tid & (kFrontSC - 1u), slot = tid / kFrontSC, sc = scBase + q;
		const bool scValid = sc < rp.numSC;
		if (scValid)
		{
			const float* ring = rp.rings + size_t(sc) * rp.ringSize;
			for (uint32_t t = slot; t < nT; t += kFrontSlots)
			{
				const ScanTile ti = scan_tile(t, F);
				const uint32_t pe = pos32 + PushesThroughChunk(F, ti.chunkStart / kChunkFrames + 1);   // ring push count after this chunk's pushes
				float4 tap[kT];
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i)
				{
					const float d = static_cast<float>(ti.chunkLen - (ti.t0 + i)) + sTap[q * tapRow + t * (kT + 1) + i].x;   // SetDelay(float(framesToProcess - s) + delay)
					const uint32_t D = static_cast<uint32_t>(d);
					const float fr = d - static_cast<float>(static_cast<int32_t>(D));
					const uint32_t r = pe - 1u - D;
					tap[i] = make_float4(thiran_coef(fr), ring[r & rp.ringMask], ring[(r - 1u) & rp.ringMask], 0.0f);
				}
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i) sTap[q * tapRow + t * (kT + 1) + i] = tap[i];
			}
		}
		__syncthreads();
		FRONT_STAMP(2);
		// ---- P3: the Thiran recurrence, lanes 0-3 of warp 0 = the four source-channels in lock step
		if (tid < kFrontSC && scBase + tid < rp.numSC)
		{
			const uint32_t s2 = scBase + tid;
			float y1 = rp.state[ST_THIRAN_Y1 * nSC + s2];
			float4* tp = sTap + tid * tapRow;
			for (uint32_t t = 0; t < nT; ++t)
			{
				const uint32_t n0 = t * (kT + 1);
				float4 v[kT];
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i) v[i] = tp[n0 + i];
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i)
				{
					y1 = v[i].x * (v[i].y - y1) + v[i].z;     // Thiran1stInterpolator::Process: never fused in the reference
					tp[n0 + i].w = y1;
				}
			}
			rp.state[ST_THIRAN_Y1 * nSC + s2] = y1;
		}
		__syncthreads();
		FRONT_STAMP(3);
		// ---- P4: zero-state response of every tile
		if (scValid)
		{
			for (uint32_t t = slot; t < nT; t += kFrontSlots)
			{
				float st[24];
				#pragma unroll
				for (int k = 0; k < 24; ++k) st[k] = 0.0f;
				const float4* xi = sTap + q * tapRow + t * (kT + 1);
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i) { float bands[4]; xover_sample<1>(rp.coef, st, xi[i].w, bands); }
				float* e = sE + (q * nT + t) * 25 + 8 * q;
				#pragma unroll
				for (int k = 0; k < 24; ++k) e[k] = st[k];
			}
		}
		__syncthreads();
		FRONT_STAMP(4);
		// ---- P5: warp w = source-channel w: the scan over the tiles, s[t+1] = M s[t] + e[t]; lane r < 24 owns row r of M and state r (lanes 24-31 shadow
		// row 23: same addresses, same values, no divergence). The state vector goes through a 16-byte aligned staging row: 6 broadcast LDS.128 per step
		// instead of 24 shuffles. (Tried: the scan blocked in groups of four tiles with M^4 — 3 + nT/4 + 3 dependent steps instead of nT, but three times
		// the products in the same single warp: 1,024 sources 0.0421 -> 0.0468 ms per block, profiles/r02zz_scan_width_ab.txt.)
		{
			const uint32_t w = tid >> 5, lane = tid & 31u, s5 = scBase + w;
			if (s5 < rp.numSC)
			{
				const uint32_t r = min(lane, 23u);
				float m[24];
				#pragma unroll
				for (int k = 0; k < 24; ++k) m[k] = sp.M[r * 24 + k];
				float s = rp.state[(ST_FILTER + r) * nSC + s5];
				float* e = sE + w * nT * 25 + 8 * w + r;
				float* stage = sE + kFrontSC * nT * 25 + 32 + 32 * w;
				auto matvec = [](const float (&mm)[24], const float* row, float add)
				{
					const float4* v4 = reinterpret_cast<const float4*>(row);
					float a0 = add, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
					#pragma unroll
					for (int k = 0; k < 6; ++k)
					{
						const float4 v = v4[k];
						a0 = __fmaf_rn(mm[4 * k], v.x, a0);
						a1 = __fmaf_rn(mm[4 * k + 1], v.y, a1);
						a2 = __fmaf_rn(mm[4 * k + 2], v.z, a2);
						a3 = __fmaf_rn(mm[4 * k + 3], v.w, a3);
					}
					return (a0 + a1) + (a2 + a3);
				};
				{
					for (uint32_t t = 0; t < nT; ++t)
					{
						const float et = e[t * 25];
						e[t * 25] = s; stage[lane] = s;           // the tile's start state replaces its zero-state response in place
						__syncwarp();
						s = matvec(m, stage, et);
						__syncwarp();      // every lane has read the staging row before the next step overwrites it
					}
				}
				if (lane < 24) rp.state[(ST_FILTER + lane) * nSC + s5] = s;
			}
		}
		__syncthreads();
		FRONT_STAMP(5);
		// ---- P6: every tile again from its true start state, with the ramped band gains (CrossoverFilter.h:505-535) -> Y [frame][sc]
		float tgB[4] = { 0.0f, 0.0f, 0.0f, 0.0f }, curB[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
		if (scValid)
		{
			#pragma unroll
			for (int b = 0; b < 4; ++b) { tgB[b] = rp.tgtBand[b * nSC + sc]; curB[b] = rp.state[(ST_BAND + b) * nSC + sc]; }
		}
		__syncthreads();      // every thread has read the block-start band gains: slot 0 commits mFilterGains = target now (DirectSound.cpp:225), under the pass
		if (scValid)
		{
			if (slot == 0)
			{
				#pragma unroll
				for (int b = 0; b < 4; ++b) rp.state[(ST_BAND + b) * nSC + sc] = tgB[b];
			}
			for (uint32_t t = slot; t < nT; t += kFrontSlots)
			{
				const ScanTile ti = scan_tile(t, F);
				const bool ramping = ti.chunkStart == 0;      // later chunks sit at the target (mFilterGains = target after a chunk, DirectSound.cpp:225)
				const float chunkF = static_cast<float>(ti.chunkLen);
				float st[24], band[4], inc[4];
				const float* e = sE + (q * nT + t) * 25 + 8 * q;
				#pragma unroll
				for (int k = 0; k < 24; ++k) st[k] = e[k];
				const float* gt = sp.tileGain + (size_t(t) * nSC + sc) * nG;     // written by this CTA's P5 (same CTA: ordered by the barrier above)
				#pragma unroll
				for (int b = 0; b < 4; ++b)
				{
					const float b0 = ramping ? curB[b] : tgB[b];
					inc[b] = (tgB[b] - b0) / chunkF;
					band[b] = ramping ? gt[L + b] : tgB[b];
				}
				const float4* xi = sTap + q * tapRow + t * (kT + 1);
				float* yo = sp.xth + size_t(ti.n0) * nSC + sc;
				#pragma unroll
				for (uint32_t i = 0; i < kT; ++i)
				{
					float b[4];
					xover_sample<1>(rp.coef, st, xi[i].w, b);
					// (bands * gains).reduce() with the bands interleaved { b1, b3, b2, b4 } (CrossoverFilter.h:516-531, SIMD.h:938-945)
					yo[size_t(i) * nSC] = (b[0] * band[0] + b[1] * band[2]) + (b[2] * band[1] + b[3] * band[3]);
					#pragma unroll
					for (int k = 0; k < 4; ++k) band[k] += inc[k];
				}
			}
		}
		FRONT_STAMP(6);
			FRONT_STAMP(15);
	}

	//======================================================================
	// k_scan_front_pipe: the same six phases as k_scan_front, but as STAGES of a pipeline over the tiles instead of barrier-separated phases. The
	// two sequential recurrences (3 of the 32 us of k_scan_front each leave 124 of its 128 threads idle) get a warp of their own and trail each other by a
	// quarter of the block; the parallel stages run behind them quarter by quarter:
	//   warps 0-3  "tile" warps: thread = (tile slot, source-channel): P0 push, then for its tile P2 taps -> P4 zero-state pass -> P6 final pass
	//   warp 4     P1 delay smoother, lanes 0-3 = the CTA's four source-channels in lock step
	//   warp 5     P3 Thiran recurrence, lanes 0-3
	//   warps 6-9  P5 state scan + gain ramps, one warp per source-channel
	// Hand-offs are progress counters in shared memory (st.release / ld.acquire at CTA scope): tiles done by P1, P3, P5 (per source-channel), quarters
	// (8 tiles) done by P2 and P4 per tile warp.
	namespace
	{
		constexpr uint32_t kPipeThreads = 320;
#ifdef JPLG_PROF
		__device__ unsigned long long g_scanProf[16];
		__device__ unsigned long long g_scanCta[3 * 1024];      // per CTA: globaltimer at start / end, SM id
		__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
		__device__ __forceinline__ uint32_t smid() { uint32_t v; asm volatile("mov.u32 %0, %%smid;" : "=r"(v)); return v; }
		#define SCAN_STAMP(i) do { if (blockIdx.x == 0 && lane == 0) g_scanProf[i] = clock64(); } while (0)
#else
		#define SCAN_STAMP(i)
#endif
		__device__ __forceinline__ uint32_t ld_acq(const uint32_t* p)
		{
			uint32_t v; asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(p))) : "memory"); return v;
		}
		__device__ __forceinline__ void st_rel(uint32_t* p, uint32_t v)
		{
			asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(p))), "r"(v) : "memory");
		}
		__device__ __forceinline__ void wait_ge(const uint32_t* p, uint32_t want) { while (ld_acq(p) < want) __nanosleep(100); }   // backing off: spinning tile warps otherwise take issue slots from the two recurrence warps they wait for
	}

	// dynamic shared memory: as k_scan_front, + flags: dlyDone, thDone, scanDone[4], tapDone[16], zsrDone[16] (uint32)
	__global__ void __launch_bounds__(kPipeThreads)
	k_scan_front_pipe(const ScanParams sp)
	{
		extern __shared__ __align__(16) float sF[];
		const RenderParams& rp = sp.rp;
		const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
		const uint32_t F = rp.frames, nT = sp.nTiles, C = rp.srcChannels, L = rp.L, nG = L + 4;
		const size_t nSC = rp.numSC;
		const uint32_t scBase = blockIdx.x * kFrontSC;
		const uint32_t rowS = nT * (kT + 1) + 8u, arrS = kFrontSC * rowS;
		float* sDly = sF;
		float* sA = sDly + arrS;
		float* sX = sA + arrS;
		float* sX1 = sX + arrS;
		float* sE = sX1 + arrS;
		uint32_t* sFlag = reinterpret_cast<uint32_t*>(sE + kFrontSC * nT * 25 + 32);
		uint32_t* dlyDone = sFlag;          // tiles whose smoothed delays are in sDly
		uint32_t* thDone = sFlag + 1;       // tiles whose delayed samples are in sA
		uint32_t* scanDone = sFlag + 2;     // [4] per source-channel: tiles whose start state (and gain ramps) are published
		uint32_t* tapDone = sFlag + 8;      // [16] per (round, tile warp): the quarter's taps are in sA / sX / sX1
		uint32_t* zsrDone = sFlag + 24;     // [16] per (round, tile warp): the quarter's zero-state end states are in sE
#ifdef JPLG_PROF
		if (tid == 0 && blockIdx.x < 1024) { g_scanCta[3 * blockIdx.x] = gtime(); g_scanCta[3 * blockIdx.x + 2] = smid(); if (blockIdx.x == 0) g_scanProf[12] = clock64(); }
#endif
		if (tid < 40) sFlag[tid] = 0u;
		__syncthreads();
		if (warp == 0) SCAN_STAMP(0);
		const uint32_t pos32 = static_cast<uint32_t>(rp.pos);
		const uint32_t rounds = (nT + kFrontSlots - 1) / kFrontSlots;

		if (warp < 4)
		{
			//============================================================== tile warps
			if (sp.push)
			{
				const uint32_t numChunks = (F + kChunkFrames - 1) / kChunkFrames;
				for (uint32_t f0 = 0; f0 < F; f0 += 4 * kScanThreads)
				{
					float v[kFrontSC][4];
					#pragma unroll
					for (uint32_t qq = 0; qq < kFrontSC; ++qq)
					{
						const uint32_t s0 = min(scBase + qq, rp.numSC - 1u), src = s0 / C, ch = s0 - src * C;
						const float* in = rp.in + size_t(src) * F * C + ch;
						#pragma unroll
						for (uint32_t j = 0; j < 4; ++j) { const uint32_t f = f0 + tid + kScanThreads * j; v[qq][j] = f < F ? __ldcs(in + size_t(f) * C) : 0.0f; }
					}
					#pragma unroll
					for (uint32_t qq = 0; qq < kFrontSC; ++qq)
					{
						if (scBase + qq >= rp.numSC) break;
						float* ring = rp.rings + size_t(scBase + qq) * rp.ringSize;
						#pragma unroll
						for (uint32_t j = 0; j < 4; ++j)
						{
							const uint32_t f = f0 + tid + kScanThreads * j;
							if (f >= F) continue;
							for (uint32_t k = 0; k < numChunks && k * kChunkFrames <= f; ++k)
								ring[(pos32 + PushesThroughChunk(F, k) + (f - k * kChunkFrames)) & rp.ringMask] = v[qq][j];
						}
					}
				}
				__threadfence_block();
				asm volatile("bar.sync 1, 128;" ::: "memory");      // the four tile warps: every push of the CTA's rings is visible before any tap reads them
			}
			if (warp == 0) SCAN_STAMP(1);
			const uint32_t q = tid & (kFrontSC - 1u), slot = tid / kFrontSC, sc = scBase + q;
			const bool scValid = sc < rp.numSC;
			const float* ring = rp.rings + size_t(scValid ? sc : 0) * rp.ringSize;
			float tgB[4] = { 0, 0, 0, 0 }, curB[4] = { 0, 0, 0, 0 };
			if (scValid)
			{
				#pragma unroll
				for (int b = 0; b < 4; ++b) { tgB[b] = rp.tgtBand[b * nSC + sc]; curB[b] = rp.state[(ST_BAND + b) * nSC + sc]; }
			}
			for (uint32_t r = 0; r < rounds; ++r)
			{
				const uint32_t t = slot + r * kFrontSlots;
				const bool has = t < nT && scValid;
				const uint32_t tLast = min(8u * warp + 7u + r * kFrontSlots, nT - 1u);      // the last tile of this warp's quarter
				const ScanTile ti = scan_tile(min(t, nT - 1u), F);
				// ---- P2: taps of the tile
				wait_ge(dlyDone, tLast + 1u);
				if (has)
				{
					const uint32_t pe = pos32 + PushesThroughChunk(F, ti.chunkStart / kChunkFrames + 1);
					// (loops of this kernel are unrolled by 4, not 16: every tile thread runs its code once per tile, so fully unrolled bodies made the
					// kernel wait on instruction fetch — 15 k cycles between CTA entry and the first barrier, `no_instruction` the top stall of P4 / P6)
					#pragma unroll 4
					for (uint32_t i = 0; i < kT; ++i)
					{
						const float d = static_cast<float>(ti.chunkLen - (ti.t0 + i)) + sDly[q * rowS + t * (kT + 1) + i];
						const uint32_t D = static_cast<uint32_t>(d);
						const float fr = d - static_cast<float>(static_cast<int32_t>(D));
						const uint32_t rd = pe - 1u - D;
						const uint32_t o = q * rowS + t * (kT + 1) + i;
						sA[o] = thiran_coef(fr);
						sX[o] = ring[rd & rp.ringMask];
						sX1[o] = ring[(rd - 1u) & rp.ringMask];
					}
				}
				__syncwarp();
				if (lane == 0) st_rel(tapDone + 4 * r + warp, 1u);
				if (warp == 0) SCAN_STAMP(2); if (warp == 3) SCAN_STAMP(3);
				// ---- P4: zero-state pass of the tile
				wait_ge(thDone, tLast + 1u);
				if (has)
				{
					float st[24];
					#pragma unroll
					for (int k = 0; k < 24; ++k) st[k] = 0.0f;
					const float* xi = sA + q * rowS + t * (kT + 1);
					#pragma unroll 4
					for (uint32_t i = 0; i < kT; ++i) { float bands[4]; xover_sample<1>(rp.coef, st, xi[i], bands); }
					float* e = sE + (q * nT + t) * 25 + 8 * q;
					#pragma unroll
					for (int k = 0; k < 24; ++k) e[k] = st[k];
				}
				__syncwarp();
				if (lane == 0) st_rel(zsrDone + 4 * r + warp, 1u);
				if (warp == 0) SCAN_STAMP(4); if (warp == 3) SCAN_STAMP(5);
				// ---- P6: final pass of the tile from its true start state
				wait_ge(scanDone + q, min(t, nT - 1u) + 1u);
				if (has)
				{
					const bool ramping = ti.chunkStart == 0;
					const float chunkF = static_cast<float>(ti.chunkLen);
					float st[24], band[4], inc[4];
					const float* e = sE + (q * nT + t) * 25 + 8 * q;
					#pragma unroll
					for (int k = 0; k < 24; ++k) st[k] = e[k];
					const float* gt = sp.tileGain + (size_t(t) * nSC + sc) * nG;
					#pragma unroll
					for (int b = 0; b < 4; ++b)
					{
						const float b0 = ramping ? curB[b] : tgB[b];
						inc[b] = (tgB[b] - b0) / chunkF;
						band[b] = ramping ? gt[L + b] : tgB[b];
					}
					const float* xi = sA + q * rowS + t * (kT + 1);
					float* yo = sp.xth + size_t(ti.n0) * nSC + sc;
					#pragma unroll 4
					for (uint32_t i = 0; i < kT; ++i)
					{
						float b[4];
						xover_sample<1>(rp.coef, st, xi[i], b);
						yo[size_t(i) * nSC] = (b[0] * band[0] + b[1] * band[2]) + (b[2] * band[1] + b[3] * band[3]);
						#pragma unroll
						for (int k = 0; k < 4; ++k) band[k] += inc[k];
					}
				}
			}
		}
		else if (warp == 4)
		{
			//============================================================== P1: delay smoother
			const uint32_t q = lane, sc = scBase + q;
			const bool on = q < kFrontSC && sc < rp.numSC;
			float cur = on ? rp.state[ST_DELAY * nSC + sc] : 0.0f;
			const float tgt = on ? rp.tgtDelay[sc] : 0.0f, alpha = rp.alpha;
			float* o = sDly + min(q, kFrontSC - 1u) * rowS;
			for (uint32_t t = 0; t < nT; ++t)
			{
				if (on)
				{
					#pragma unroll
					for (uint32_t i = 0; i < kT; ++i)
					{
						cur = alpha * (tgt - cur) + cur;
						o[t * (kT + 1) + i] = cur;
					}
				}
				__syncwarp();
				if (lane == 0) st_rel(dlyDone, t + 1u);
			}
			if (on) rp.state[ST_DELAY * nSC + sc] = cur;
			SCAN_STAMP(6);
		}
		else if (warp == 5)
		{
			//============================================================== P3: Thiran recurrence
			const uint32_t q = lane, sc = scBase + q;
			const bool on = q < kFrontSC && sc < rp.numSC;
			float y1 = on ? rp.state[ST_THIRAN_Y1 * nSC + sc] : 0.0f;
			const uint32_t qq = min(q, kFrontSC - 1u);
			float* a = sA + qq * rowS; const float* x = sX + qq * rowS; const float* x1 = sX1 + qq * rowS;
			for (uint32_t t = 0; t < nT; ++t)
			{
				if ((t & 7u) == 0) wait_ge(tapDone + 4 * (t / kFrontSlots) + ((t % kFrontSlots) >> 3), 1u);
				if (on)
				{
					const uint32_t n0 = t * (kT + 1);
					float va[kT], vx[kT], v1[kT];
					#pragma unroll
					for (uint32_t i = 0; i < kT; ++i) { va[i] = a[n0 + i]; vx[i] = x[n0 + i]; v1[i] = x1[n0 + i]; }
					#pragma unroll
					for (uint32_t i = 0; i < kT; ++i)
					{
						y1 = va[i] * (vx[i] - y1) + v1[i];
						a[n0 + i] = y1;
					}
				}
				__syncwarp();
				if (lane == 0) st_rel(thDone, t + 1u);
			}
			if (on) rp.state[ST_THIRAN_Y1 * nSC + sc] = y1;
			SCAN_STAMP(7);
		}
		else
		{
			//============================================================== P5: state scan + gain ramps, warp per source-channel
			const uint32_t w = warp - 6u, s5 = scBase + w;
			if (s5 < rp.numSC)
			{
				const uint32_t r = min(lane, 23u);
				float m[24];
				#pragma unroll
				for (int k = 0; k < 24; ++k) m[k] = sp.M[r * 24 + k];
				float s = rp.state[(ST_FILTER + r) * nSC + s5];
				const uint32_t chunk0 = min(F, kChunkFrames);
				float g[2], dg[2];
				#pragma unroll
				for (uint32_t h = 0; h < 2; ++h)
				{
					const uint32_t j = lane + 32u * h;
					float cur = 0.0f, tg = 0.0f;
					if (j < L) { cur = rp.curMix[j * nSC + s5]; tg = rp.tgtMix[j * nSC + s5]; }
					else if (j < nG) { cur = rp.state[(ST_BAND + j - L) * nSC + s5]; tg = rp.tgtBand[(j - L) * nSC + s5]; }
					g[h] = cur; dg[h] = (tg - cur) / static_cast<float>(chunk0);
				}
				float* e = sE + w * nT * 25 + 8 * w + r;
				for (uint32_t t = 0; t < nT; ++t)
				{
					if ((t & 7u) == 0) wait_ge(zsrDone + 4 * (t / kFrontSlots) + ((t % kFrontSlots) >> 3), 1u);
					const float et = e[t * 25];
					if (lane < 24) e[t * 25] = s;
					if (t * kT < chunk0)
					{
						float* gt = sp.tileGain + (size_t(t) * nSC + s5) * nG;
						if (lane < nG) gt[lane] = g[0];
						if (lane + 32u < nG) gt[lane + 32u] = g[1];
						#pragma unroll
						for (uint32_t i = 0; i < kT; ++i) { g[0] += dg[0]; g[1] += dg[1]; }
					}
					__threadfence_block();
					__syncwarp();
					if (lane == 0) st_rel(scanDone + w, t + 1u);      // tile t's start state and ramp values are published
					float a0 = et, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
					#pragma unroll
					for (int k = 0; k < 24; k += 4)
					{
						a0 = __fmaf_rn(m[k], __shfl_sync(0xffffffffu, s, k), a0);
						a1 = __fmaf_rn(m[k + 1], __shfl_sync(0xffffffffu, s, k + 1), a1);
						a2 = __fmaf_rn(m[k + 2], __shfl_sync(0xffffffffu, s, k + 2), a2);
						a3 = __fmaf_rn(m[k + 3], __shfl_sync(0xffffffffu, s, k + 3), a3);
					}
					s = (a0 + a1) + (a2 + a3);
				}
				if (lane < 24) rp.state[(ST_FILTER + lane) * nSC + s5] = s;
				if (w == 0) SCAN_STAMP(8);
			}
			else if (lane == 0) st_rel(scanDone + w, nT);     // no source-channel here: nobody must wait for it
		}
		if (warp == 0) SCAN_STAMP(9); if (warp == 3) SCAN_STAMP(10);
		__syncthreads();       // every tile has read the block-start band gains
		if (warp == 0) SCAN_STAMP(11);
#ifdef JPLG_PROF
		if (tid == 0 && blockIdx.x < 1024) { g_scanCta[3 * blockIdx.x + 1] = gtime(); if (blockIdx.x == 0) g_scanProf[13] = clock64(); }
#endif
		if (tid < kFrontSC && scBase + tid < rp.numSC)
		{
			#pragma unroll
			for (int b = 0; b < 4; ++b) rp.state[(ST_BAND + b) * nSC + scBase + tid] = rp.tgtBand[b * nSC + scBase + tid];
		}
	}
#ifdef JPLG_PROF
	extern "C" __attribute__((visibility("default"))) void jplg_debug_scan_prof(unsigned long long* out) { cudaMemcpyFromSymbol(out, g_scanProf, sizeof(g_scanProf)); }
	extern "C" __attribute__((visibility("default"))) void jplg_debug_scan_cta(unsigned long long* out) { cudaMemcpyFromSymbol(out, g_scanCta, sizeof(g_scanCta)); }
#endif

	__device__ __forceinline__ void cp_async16(float* dstShared, const float* src)
	{
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(dstShared))), "l"(src) : "memory");
	}
	__device__ __forceinline__ void cp_async4(float* dstShared, const float* src)
	{
		asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(dstShared))), "l"(src) : "memory");
	}

	// k_scan_mix: CTA = (128 source-channels, one tile), kMixThreads threads. With one or two warps per scheduler every instruction costs its full latency
	// (r02y: 1,934 instructions per warp in 14.9 k cycles with 128 threads), so the work is spread thin: staging by (source-channel, quarter), the mix by
	// (output channel pair, subset of TWO sources), the cross-subset reduce by output element. The last CTA of a tile to finish also sums the tile's
	// partial buses over the source groups in a fixed order into the device bus and the caller's bus (no k_bus_reduce launch on this path); the last tile-CTA
	// of a source group commits mCurrentGains = targets (DirectSound.cpp:221).
	// dynamic shared memory (floats): Y [kT][128] | P [Jused][nPairs][kT + 1] f2 | tg, cur, gt [128][Lp] (target gains, block-start gains, ramp values at
	// the tile start: staged with 4-byte cp.async, so the scattered loads of a thread are all in flight at once).
	// DEBUG: also writes every source's own contribution (jplg_renderer_process_block_debug), channel after channel like the reference's
	// `dest += source * gain`.
	namespace { constexpr uint32_t kMixThreads = 256, kMixSC = 128; }
#ifdef JPLG_PROF
	__device__ unsigned long long g_mixProf[16];
	#define MIX_STAMP(i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_mixProf[i] = clock64(); } while (0)
	extern "C" __attribute__((visibility("default"))) void jplg_debug_mix_prof(unsigned long long* out) { cudaMemcpyFromSymbol(out, g_mixProf, sizeof(g_mixProf)); }
#else
	#define MIX_STAMP(i)
#endif
	template<bool DEBUG>
	__global__ void __launch_bounds__(kMixThreads)
	k_scan_mix(const ScanParams sp)
	{
		extern __shared__ __align__(16) float sScan[];
		__shared__ uint32_t sLast, sLastOfTile;
		const RenderParams& rp = sp.rp;
		MIX_STAMP(0);
		const uint32_t tid = threadIdx.x, cta = blockIdx.x, t = blockIdx.y;
		const uint32_t C = rp.srcChannels, L = rp.L, nG = L + 4;
		const uint32_t srcPerCTA = kMixSC / C, scPerCTA = srcPerCTA * C, scBase = cta * scPerCTA;
		const uint32_t numSCInCTA = min(scPerCTA, rp.numSC - scBase);
		const size_t nSC = rp.numSC;
		const uint32_t F = rp.frames;
		const ScanTile ti = scan_tile(t, F);
		const float chunkF = static_cast<float>(ti.chunkLen);
		const bool ramping = ti.chunkStart == 0;      // later chunks sit at the target (mCurrentGains = target after a chunk, DirectSound.cpp:221)
		const uint32_t nPairs = (L + 1) / 2, Lp = 2 * nPairs;
		const uint32_t J = kMixThreads / nPairs, pStride = (kT + 1) * nPairs;
		const uint32_t srcPerSubset = (srcPerCTA + J - 1) / J, Jused = (srcPerCTA + srcPerSubset - 1) / srcPerSubset;
		float* sY = sScan;
		f2* sP = reinterpret_cast<f2*>(sScan + kT * kMixSC);
		float* sTg = reinterpret_cast<float*>(sP + Jused * pStride);
		float* sCur = sTg + kMixSC * Lp;
		float* sGt = sCur + kMixSC * Lp;

		// Launched with programmatic stream serialisation: the CTAs are resident while k_scan_front still runs. What that kernel does not write (the
		// target and block-start gains) is staged before griddepcontrol.wait, under its run time; its outputs (Y, the ramp values at the tile start) after.
		const uint32_t col = tid & (kMixSC - 1u), part = tid / kMixSC;     // staging: thread = (source-channel, half): every second output channel
		if (col < numSCInCTA)
		{
			const uint32_t sc = scBase + col;
			for (uint32_t c = part; c < L; c += kMixThreads / kMixSC)
			{
				cp_async4(sTg + col * Lp + c, rp.tgtMix + c * nSC + sc);
				if (ramping) cp_async4(sCur + col * Lp + c, rp.curMix + c * nSC + sc);
			}
			if (part == 0 && L < Lp) { sTg[col * Lp + L] = 0.0f; sCur[col * Lp + L] = 0.0f; sGt[col * Lp + L] = 0.0f; }
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
		asm volatile("griddepcontrol.wait;" ::: "memory");
		if (((nSC | scBase) & 3u) == 0)
		{
			// Y tile: rows of up to 128 consecutive source-channels, 16 bytes per copy (a last quad may reach past a ragged group: still inside Y, columns unused)
			for (uint32_t e = tid; e < kT * (kMixSC / 4); e += kMixThreads)
			{
				const uint32_t i = e / (kMixSC / 4), c4 = (e - i * (kMixSC / 4)) * 4;
				if (c4 < numSCInCTA) cp_async16(sY + i * kMixSC + c4, sp.xth + size_t(ti.n0 + i) * nSC + scBase + c4);
			}
		}
		else if (col < numSCInCTA)
		{
			const float* yi = sp.xth + size_t(ti.n0) * nSC + scBase + col;
			#pragma unroll
			for (uint32_t i = part; i < kT; i += kMixThreads / kMixSC) cp_async4(sY + i * kMixSC + col, yi + size_t(i) * nSC);
		}
		if (ramping && col < numSCInCTA)
		{
			const float* gt = sp.tileGain + (size_t(t) * nSC + scBase + col) * nG;
			for (uint32_t c = part; c < L; c += kMixThreads / kMixSC) cp_async4(sGt + col * Lp + c, gt + c);
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
		asm volatile("cp.async.wait_group 0;" ::: "memory");
		__syncthreads();
		MIX_STAMP(1);

		// ---- K4: AddAndApplyGainRamp (ChannelMixing.h:89-107) for every (source-channel, output channel) of the tile, re-associated across sources:
		// thread = (output channel pair, source subset), kT packed frame accumulators, fixed-order cross-subset reduce (the mix role of k_render)
		const uint32_t mixP = tid % nPairs, mixJ = tid / nPairs;
		const uint32_t c0 = 2 * mixP, c1 = c0 + 1;
		const bool has1 = c1 < L, mixActive = mixJ < Jused;
		const uint32_t subBegin = min(mixJ * srcPerSubset * C, numSCInCTA), subEnd = min((mixJ + 1) * srcPerSubset * C, numSCInCTA);
		if (mixActive)
		{
			f2 acc[kT];
			#pragma unroll
			for (uint32_t i = 0; i < kT; ++i) acc[i] = 0ull;
			const f2 one = pk(1.0f, 1.0f);
			for (uint32_t s = subBegin; s < subEnd; ++s)
			{
				const float2 tg = *reinterpret_cast<const float2*>(sTg + s * Lp + c0);
				f2 g = pk(tg.x, tg.y), dg = 0ull;
				if (ramping)
				{
					const float2 a = *reinterpret_cast<const float2*>(sCur + s * Lp + c0), gs = *reinterpret_cast<const float2*>(sGt + s * Lp + c0);
					dg = pk((tg.x - a.x) / chunkF, (tg.y - a.y) / chunkF);       // delta = (gainEnd - gainStart) / numFrames (ChannelMixing.h:99)
					g = pk(gs.x, gs.y);                                            // the reference's accumulated ramp at the tile start (k_scan_front P1)
				}
				if (DEBUG)
				{
					const uint32_t gsc = scBase + s, src = gsc / C, k = gsc - src * C;
					float* ps = rp.perSource + (size_t(src) * F + ti.n0) * L;
					for (uint32_t i = 0; i < kT; ++i)
					{
						const float y = sY[i * kMixSC + s];
						const f2 v = mul2(pk(y, y), g);
						float* o = ps + size_t(i) * L;
						o[c0] = (k == 0 ? 0.0f : o[c0]) + lo(v);
						if (has1) o[c1] = (k == 0 ? 0.0f : o[c1]) + hi(v);
						acc[i] = fma2(v, one, acc[i]);
						g = fma2(g, one, dg);
					}
				}
				else
				{
					#pragma unroll
					for (uint32_t i = 0; i < kT; ++i)
					{
						const float y = sY[i * kMixSC + s];
						acc[i] = fma2(mul2(pk(y, y), g), one, acc[i]);     // product rounded first, as in the per-source outputs (same bus either way)
						g = fma2(g, one, dg);
					}
				}
			}
			f2* row = sP + mixJ * pStride + mixP * (kT + 1);
			#pragma unroll
			for (uint32_t i = 0; i < kT; ++i) row[i] = acc[i];
		}
		__syncthreads();
		MIX_STAMP(2);
		float* pb = rp.partial + (size_t(cta) * F + ti.n0) * L;
		if (tid < kT * Lp)
		{
			const float* sPf = reinterpret_cast<const float*>(sP);
			const uint32_t i = tid / Lp, c = tid - i * Lp;
			const float* src = sPf + ((c >> 1) * (kT + 1) + i) * 2 + (c & 1u);
			float s0 = 0.0f, s1 = 0.0f;      // two interleaved chains, combined at the end: fixed order
			uint32_t j = 0;
			for (; j + 1 < Jused; j += 2, src += 4 * pStride) { s0 += src[0]; s1 += src[2 * pStride]; }
			if (j < Jused) s0 += src[0];
			if (c < L) pb[i * L + c] = s0 + s1;
		}
		// ---- last arrivers: of this tile (over the source groups) -> bus reduction; of this source group (over the tiles) -> current = target
		__threadfence();
		__syncthreads();
		MIX_STAMP(3);
		if (tid == 0)
		{
			sLast = (atomicAdd(&sp.counters[cta], 1u) == sp.nTiles - 1u) ? 1u : 0u;
			sLastOfTile = (atomicAdd(&sp.counters[rp.numCTAs + t], 1u) == rp.numCTAs - 1u) ? 1u : 0u;
		}
		__syncthreads();
		if (sLastOfTile)
		{
			__threadfence();
			if (tid < kT * L)
			{
				// out[o] = sum over the source groups in ascending order (deterministic); also accumulated into the caller's bus like k_bus_reduce does
				const size_t o = size_t(ti.n0) * L + tid;
				float sum = 0.0f;
				for (uint32_t x0 = 0; x0 < rp.numCTAs; x0 += 8)      // eight loads in flight, added in ascending order
				{
					float v[8];
					#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) v[k] = x0 + k < rp.numCTAs ? __ldcg(rp.partial + size_t(x0 + k) * F * L + o) : 0.0f;
					#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) if (x0 + k < rp.numCTAs) sum += v[k];
				}
				sp.deviceBus[o] = sum;
				if (sp.busAcc) sp.busAcc[o] += sum;
			}
			if (tid == 0) sp.counters[rp.numCTAs + t] = 0u;
		}
		if (sLast)
		{
			__threadfence();
			if (tid < numSCInCTA)
			{
				const uint32_t sc = scBase + tid;
				for (uint32_t cb = 0; cb < L; cb += 8)          // eight loads in flight before the first store (the two arrays may alias as far as the compiler knows)
				{
					float v[8];
					#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) v[k] = cb + k < L ? __ldcg(rp.tgtMix + (cb + k) * nSC + sc) : 0.0f;
					#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) if (cb + k < L) rp.curMix[(cb + k) * nSC + sc] = v[k];
				}
			}
			if (tid == 0) sp.counters[cta] = 0u;
		}
		MIX_STAMP(15);
	}

	size_t ScanScratchFloats(uint32_t numSC, uint32_t maxFrames)
	{
		const size_t nT = (maxFrames + kScanTile - 1) / kScanTile;
		// Y [maxFrames][numSC]; tileGain [nT][numSC][L + 4] (L <= 32)
		return size_t(numSC) * (size_t(maxFrames) + nT * (kMaxChannelsDev + 4));
	}

	static size_t ScanFrontSmem(uint32_t frames)
	{
		const size_t nT = frames / kScanTile;
		// k_scan_front: float4 tap rows, zero-state / start states, a 24-float staging row per scan warp; k_scan_front_pipe: four float rows (+ 8) instead of the
		// float4 rows (its flags follow, + 256 bytes at the launch): the larger of the two
		const size_t phases = size_t(4) * kFrontSC * (nT * (kScanTile + 1) + 2), pipe = size_t(4) * kFrontSC * (nT * (kScanTile + 1) + 8);
		return (std::max(phases, pipe) + size_t(kFrontSC) * nT * 25 + 32 + 32 * kFrontSC) * sizeof(float);
	}

	bool RenderCanScan(const RenderParams& rp)   // per-source debug output: a multichannel source's channels must sit in one subset (they do: subsets hold whole sources)
	{
		const bool er = rp.scPerRing > 1 || rp.plainPush || rp.groupsPerCTA > 1 || rp.flags;
		return !er && rp.frames >= kScanTile && rp.frames % kScanTile == 0 && rp.numSC > 0 && ScanFrontSmem(rp.frames) <= 200 * 1024
			   && TotalPushes(rp.frames) <= rp.ringSize;
	}

	cudaError_t LaunchRenderScan(const ScanParams& spIn, cudaStream_t st)
	{
		ScanParams sp = spIn;
		const RenderParams& rp = sp.rp;
		sp.nTiles = rp.frames / kScanTile;
		const size_t nSC = rp.numSC;
		sp.xth = sp.scratch;
		sp.tileGain = sp.scratch + size_t(rp.frames) * nSC;
		const uint32_t nPairs = (rp.L + 1) / 2, J = kMixThreads / nPairs, srcPerCTA = kMixSC / rp.srcChannels;
		const uint32_t srcPerSubset = (srcPerCTA + J - 1) / J, Jused = (srcPerCTA + srcPerSubset - 1) / srcPerSubset;
		const size_t dynFront = ScanFrontSmem(rp.frames);
		const size_t dynMix = size_t(kT) * kMixSC * sizeof(float) + size_t(Jused) * (kT + 1) * nPairs * sizeof(f2) + size_t(3) * kMixSC * 2 * nPairs * sizeof(float);
		static bool attr = false;
		if (!attr)
		{
			cudaFuncSetAttribute(k_scan_front, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
			cudaFuncSetAttribute(k_scan_front_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024);
			cudaFuncSetAttribute(k_scan_mix<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
			cudaFuncSetAttribute(k_scan_mix<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
			attr = true;
		}
		// JPLG_SCAN_PIPE=1 (experiment switch): the warp-pipelined k_scan_front_pipe instead of the barrier-separated phases. Same results; measured
		// the same speed (r02v, 1,024 sources: 0.0567 vs 0.0561 ms per block): with ten warps per CTA the two recurrence warps advance 2.3 - 3.4 x
		// slower than their 4.15-cycle dependent-issue latency allows (profiles/r02v_scan_pipe_timeline.txt), which eats what the overlap buys.
		static const bool pipe = [] { const char* e = std::getenv("JPLG_SCAN_PIPE"); return e && e[0] == '1'; }();
		if (!pipe) k_scan_front<<<(rp.numSC + kFrontSC - 1) / kFrontSC, kScanThreads, dynFront, st>>>(sp);
		else k_scan_front_pipe<<<(rp.numSC + kFrontSC - 1) / kFrontSC, kPipeThreads, dynFront + 256, st>>>(sp);
		{
			// programmatic dependent launch: the mix grid's launch latency overlaps the front kernel's run (both are a few microseconds on this path)
			cudaLaunchConfig_t cfg = {};
			cfg.gridDim = dim3(rp.numCTAs, sp.nTiles); cfg.blockDim = dim3(kMixThreads); cfg.dynamicSmemBytes = dynMix; cfg.stream = st;
			cudaLaunchAttribute at[1];
			at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
			at[0].val.programmaticStreamSerializationAllowed = pipe ? 0 : 1;     // (k_scan_front_pipe does not signal)
			cfg.attrs = at; cfg.numAttrs = 1;
			if (rp.perSource) cudaLaunchKernelEx(&cfg, k_scan_mix<true>, sp);
			else cudaLaunchKernelEx(&cfg, k_scan_mix<false>, sp);
		}
		return cudaGetLastError();
	}
}
