// Batched ReverbBus for sm_100a: n x 16th-order feedback delay network with four-band decay filters
// (inc/Auralization/LateReverb.h:34-117, src/Auralization/LateReverb.cpp:53-134, inc/Auralization/ReverbDesign.h:311-349,388-492,573-704).
//
//   k_fdn_decay   AttenuationFilter::UpdateParameters for every (instance, FDN channel): DecayGains = dBToGain(slope * delaySeconds)
//                 (LateReverb.h:64-68; dBToGain = SIMD exp, inc/Math/DecibelsAndGain.h:40-56, inc/Math/SIMDMath.h:270-316)
//   k_fdn         one block of ReverbBus::ProcessInterleaved for n instances. Thread = (instance, FDN channel); the 16 channels of an
//                 instance sit in 16 consecutive lanes, so the Householder feedback mix and the Hadamard output mix are warp shuffles.
//                 Every FDN delay (1609..4259 samples) is longer than a block, so inside a block the only time recurrences are the 16
//                 per-channel crossover filters; delayed samples are staged from the rings, and the new ring samples / output frames are
//                 staged back, through shared memory with coalesced global accesses.
//                 The stage-in is double-buffered (cp.async: tile k+1 arrives while tile k runs); the crossover runs its ten SVFs after X2 as five
//                 packed FP32x2 filters (xover_process_sample2).
//
// Arithmetic modes as in jplg_render_kernels.cu: PARITY = every reference FMA as mul + add (the SSE2 baseline semantic), THROUGHPUT = fused.
// Compile with -fmad=false.
#include "jplg_kernels.hpp"
#include "jplg_render_common.cuh"   // mad, svf, packed FP32x2 helpers
#include <cstring>
#include <mutex>

namespace jplg
{
	namespace
	{
		// FourBandCrossover::ProcessSample(sample, gains) (CrossoverFilter.h:440-459): SplitSample -> { b1, b2, b3, b4 } * gains, reduce()
		// = (b1*g1 + b2*g2) + (b3*g3 + b4*g4) (simd::reduce, inc/Math/SIMD.h:938-945). st: 24 state floats in ST_FILTER order.
		template<int MODE> __device__ __forceinline__ float xover_process_sample(const float* __restrict__ cf, float* st, const float x, const float g[4])
		{
			const float minusR = -1.41421356237309504880f;
			const SvfP p2{ cf[0], cf[1], cf[2] };
			float yL, yB, yH, yL2, yB2, yH2;
			svf<MODE>(p2, x, st[0], st[1], yL, yB, yH);
			svf<MODE>(p2, yL, st[2], st[3], yL2, yB2, yH2);
			const float low = yL2;
			const float high = mad<MODE>(minusR, yB, yL + yH - yL2);
			float aL, aB, aH, bL, bB, bH;
			svf<MODE>(SvfP{ cf[3], cf[7], cf[11] }, low, st[4], st[5], aL, aB, aH);
			svf<MODE>(SvfP{ cf[4], cf[8], cf[12] }, high, st[6], st[7], bL, bB, bH);
			float o[4], t0, t1;
			svf<MODE>(SvfP{ cf[3], cf[7], cf[11] }, aL, st[8], st[9], o[0], t0, t1);
			svf<MODE>(SvfP{ cf[4], cf[8], cf[12] }, bL, st[10], st[11], o[1], t0, t1);
			svf<MODE>(SvfP{ cf[5], cf[9], cf[13] }, aH, st[12], st[13], t0, t1, o[2]);
			svf<MODE>(SvfP{ cf[6], cf[10], cf[14] }, bH, st[14], st[15], t0, t1, o[3]);
			float b[4];   // interleaved { b1, b3, b2, b4 }
			#pragma unroll
			for (int k = 0; k < 4; ++k)
			{
				float pL, pB, pH;
				svf<MODE>(SvfP{ cf[15 + k], cf[19 + k], cf[23 + k] }, o[k], st[16 + 2 * k], st[17 + 2 * k], pL, pB, pH);
				b[k] = mad<MODE>(minusR, pB, pH + pL);
			}
			return (b[0] * g[0] + b[2] * g[1]) + (b[1] * g[2] + b[3] * g[3]);
		}

		// The same sample on FP32x2 lanes (FMUL2 / FFMA2): X2's two SVFs stay scalar (they are a dependent chain of one signal), the ten SVFs
		// after it run as five packed ones, lane pairs as in k_render's filter roles: X13 stage 1 { low, high }, stage 2 { aL, bL } and
		// { aH, bH }, the all-pass pairs { o0, o1 } and { o2, o3 }. Every lane performs the scalar sequence above operation for operation
		// (P2<MODE>::svf), so the result is bit-identical; the instruction count per sample drops from ~155 to ~95.
		struct XoverState2 { float x2[4]; f2 s3a, s3b, s4a1, s4a2, s4b1, s4b2, s5a1, s5a2, s5b1, s5b2; };
		__device__ __forceinline__ void xover_pack_state(const float* st, XoverState2& p)
		{
			#pragma unroll
			for (int k = 0; k < 4; ++k) p.x2[k] = st[k];
			p.s3a = pk(st[4], st[6]); p.s3b = pk(st[5], st[7]);
			p.s4a1 = pk(st[8], st[10]); p.s4a2 = pk(st[9], st[11]);
			p.s4b1 = pk(st[12], st[14]); p.s4b2 = pk(st[13], st[15]);
			p.s5a1 = pk(st[16], st[18]); p.s5a2 = pk(st[17], st[19]);
			p.s5b1 = pk(st[20], st[22]); p.s5b2 = pk(st[21], st[23]);
		}
		__device__ __forceinline__ void xover_unpack_state(const XoverState2& p, float* st)
		{
			#pragma unroll
			for (int k = 0; k < 4; ++k) st[k] = p.x2[k];
			st[4] = lo(p.s3a); st[6] = hi(p.s3a); st[5] = lo(p.s3b); st[7] = hi(p.s3b);
			st[8] = lo(p.s4a1); st[10] = hi(p.s4a1); st[9] = lo(p.s4a2); st[11] = hi(p.s4a2);
			st[12] = lo(p.s4b1); st[14] = hi(p.s4b1); st[13] = lo(p.s4b2); st[15] = hi(p.s4b2);
			st[16] = lo(p.s5a1); st[18] = hi(p.s5a1); st[17] = lo(p.s5a2); st[19] = hi(p.s5a2);
			st[20] = lo(p.s5b1); st[22] = hi(p.s5b1); st[21] = lo(p.s5b2); st[23] = hi(p.s5b2);
		}
		template<int MODE> __device__ __forceinline__ float xover_process_sample2(const float* __restrict__ cf, const PackedConsts& K, XoverState2& p, const float x, const f2 gA /*{g1,g3}*/, const f2 gB /*{g2,g4}*/)
		{
			const float minusR = -1.41421356237309504880f;
			const P2<MODE> M{ K };
			const SvfP p2{ cf[0], cf[1], cf[2] };
			float yL, yB, yH, yL2, yB2, yH2;
			svf<MODE>(p2, x, p.x2[0], p.x2[1], yL, yB, yH);
			svf<MODE>(p2, yL, p.x2[2], p.x2[3], yL2, yB2, yH2);
			const float high = mad<MODE>(minusR, yB, yL + yH - yL2);
			f2 qL, qB, qH, tL, tH, u0, u1;
			M.svf(PC_X13A, pk(yL2, high), p.s3a, p.s3b, qL, qB, qH);
			M.svf(PC_X13A, qL, p.s4a1, p.s4a2, tL, u0, u1);
			M.svf(PC_X13B, qH, p.s4b1, p.s4b2, u0, u1, tH);
			f2 pL, pB, pH;
			M.svf(PC_APA, tL, p.s5a1, p.s5a2, pL, pB, pH);
			const f2 bA = M.mad(K.v[PC_MINUS_R], pB, M.add(pH, pL));        // { b1, b3 }
			M.svf(PC_APB, tH, p.s5b1, p.s5b2, pL, pB, pH);
			const f2 bB = M.mad(K.v[PC_MINUS_R], pB, M.add(pH, pL));        // { b2, b4 }
			const f2 sum = M.add(mul2(bA, gA), mul2(bB, gB));               // { b1*g1 + b2*g2, b3*g3 + b4*g4 }
			return lo(sum) + hi(sum);
		}

		// JPL::exp(simd) per lane (inc/Math/SIMDMath.h:296-316) with the SSE2 floor (inc/Math/SIMD.h:1861-1879)
		template<int MODE> __device__ __forceinline__ float exp_ref(float x)
		{
			x = fminf(fmaxf(x, -88.3762626647949f), 88.3762626647949f);
			const float v = mad<MODE>(x, 1.44269504088896341f, 0.5f);
			const float tr = static_cast<float>(static_cast<int32_t>(v));     // cvttps2dq -> cvtdq2ps
			const float n = (v == 0.0f) ? v : tr - (tr > v ? 1.0f : 0.0f);
			const float nC2 = n * -2.12194440e-4f;
			x = x - mad<MODE>(n, 0.693359375f, nC2);
			const float x2 = x * x;
			float y = 1.9875691500e-4f;
			y = mad<MODE>(y, x, 1.3981999507e-3f);
			y = mad<MODE>(y, x, 8.3334519073e-3f);
			y = mad<MODE>(y, x, 4.1665795894e-2f);
			y = mad<MODE>(y, x, 1.6666665459e-1f);
			y = mad<MODE>(y, x, 5.0000001201e-1f);
			y = mad<MODE>(y, x2, x + 1.0f);
			const float pow2n = __int_as_float((static_cast<int32_t>(n) + 0x7f) << 23);
			return y * pow2n;
		}
	}

	// cPrimeDelayLengths[0..15] (ReverbDesign.h:578-581)
	__constant__ uint32_t cFdnDelays[16] = { 2927, 2593, 2273, 3697, 1877, 3877, 2477, 3461, 1609, 3779, 3541, 4259, 1669, 3539, 3637, 4013 };

	template<int MODE>
	__global__ void k_fdn_decay(const float* __restrict__ slope /*[n][4]*/, uint32_t n, float invSampleRate, float* __restrict__ decay /*[4][n*16]*/)
	{
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n * 16u) return;
		const uint32_t inst = i >> 4, c = i & 15u;
		const float delaySeconds = static_cast<float>(cFdnDelays[c]) * invSampleRate;      // ReverbDesign.h:625
		const float ln10div20 = 2.302585092994046f * 0.05f;                                  // DecibelsAndGain.h:47 (float product, folded like the reference's static)
		#pragma unroll
		for (int k = 0; k < 4; ++k)
		{
			const float dbLossPerLoop = slope[4 * size_t(inst) + k] * delaySeconds;           // LateReverb.h:66
			decay[size_t(k) * n * 16u + i] = exp_ref<MODE>(ln10div20 * dbLossPerLoop);
		}
	}

	__device__ __forceinline__ void cp_async4_fdn(float* dstShared, const float* src)
	{
		asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(dstShared))), "l"(src) : "memory");
	}
	constexpr uint32_t kFdnTile = 32;        // frames per shared-memory tile
	constexpr uint32_t kFdnInstPerCTA = 8;   // 8 instances x 16 channels = 128 threads
	constexpr uint32_t kFdnRow = kFdnTile + 1;

	// ReverbBus::ProcessInterleaved, mono input (LateReverb.cpp:53-134), for `n` instances. The 480-frame chunking of the reference
	// has no arithmetic effect here (the filter parameters are updated once per call, the FDN runs sample by sample).
	template<int MODE>
	__global__ void __launch_bounds__(kFdnInstPerCTA * 16)
	k_fdn(const FdnParams fp, const PackedConsts K)
	{
#ifdef JPLG_FDN_NO_PREFETCH
		__shared__ float sDelA[1][kFdnInstPerCTA * 16][kFdnRow];
		__shared__ float sInA[1][kFdnInstPerCTA][kFdnTile];
		__shared__ float sOut[kFdnInstPerCTA][kFdnTile][16];   // FDN output mix per frame (first L used)
#else
		// two stage-in buffers: tile k+1's delayed samples and inputs are fetched (cp.async) while tile k is computed and staged out. Every FDN delay (>= 1,609
		// samples) exceeds the longest block (1,024 frames), so a block never reads a sample it pushes itself: the prefetch cannot overtake a push.
		extern __shared__ __align__(16) float sFdn[];
		float (*sDelA)[kFdnInstPerCTA * 16][kFdnRow] = reinterpret_cast<float (*)[kFdnInstPerCTA * 16][kFdnRow]>(sFdn);
		float (*sInA)[kFdnInstPerCTA][kFdnTile] = reinterpret_cast<float (*)[kFdnInstPerCTA][kFdnTile]>(sFdn + 2 * kFdnInstPerCTA * 16 * kFdnRow);
		float (*sOut)[kFdnTile][16] = reinterpret_cast<float (*)[kFdnTile][16]>(sFdn + 2 * kFdnInstPerCTA * 16 * kFdnRow + 2 * kFdnInstPerCTA * kFdnTile);
#endif

		const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
		const uint32_t c = tid & 15u;                          // FDN channel
		const uint32_t instBase = blockIdx.x * kFdnInstPerCTA;
		const uint32_t inst = instBase + (tid >> 4);
		const bool active = inst < fp.n;
		const uint32_t L = fp.L;
		const size_t nCh = size_t(fp.n) * 16u, ch = size_t(inst) * 16u + c;
		const uint32_t pos32 = static_cast<uint32_t>(fp.pos);
		constexpr uint32_t ringMask = kFdnRingSize - 1;

		float st[24], g[4];
		#pragma unroll
		for (int k = 0; k < 24; ++k) st[k] = active ? fp.state[size_t(k) * nCh + ch] : 0.0f;
		#pragma unroll
		for (int k = 0; k < 4; ++k) g[k] = active ? fp.decay[size_t(k) * nCh + ch] : 0.0f;

#ifndef JPLG_FDN_SCALAR
		XoverState2 ps; xover_pack_state(st, ps);
		const f2 gA = pk(g[0], g[2]), gB = pk(g[1], g[3]);
#endif
		const uint32_t oFrame0 = tid / L, oChan0 = tid - oFrame0 * L, oStepFrame = (kFdnInstPerCTA * 16) / L, oStepChan = (kFdnInstPerCTA * 16) - oStepFrame * L;
#ifndef JPLG_FDN_NO_PREFETCH
		// ---- stage in (asynchronous): warp w fetches rows [32w, 32w+32): delayed[c][n] = pushed[n - 1 - D_c] (GetReadWindow<1>(D) before this frame's push),
		// 32 + 2 copies of 4 bytes per thread in flight; rows / frames outside the bank or the block are zero-filled with plain stores
		auto prefetch = [&](uint32_t t0n, uint32_t buf)
		{
			const uint32_t lenN = min(kFdnTile, fp.frames - t0n);
			#pragma unroll 8
			for (uint32_t r = 0; r < 32; ++r)
			{
				const uint32_t row = warp * 32 + r, rInst = instBase + (row >> 4), rc = row & 15u;
				float* d = &sDelA[buf][row][lane];
				if (rInst < fp.n && lane < lenN)
					cp_async4_fdn(d, fp.rings + (size_t(rInst) * 16u + rc) * kFdnRingSize + ((pos32 + t0n + lane - 1u - cFdnDelays[rc]) & ringMask));
				else *d = 0.0f;
			}
			for (uint32_t e = tid; e < kFdnInstPerCTA * kFdnTile; e += blockDim.x)
			{
				const uint32_t i = e / kFdnTile, f = e - i * kFdnTile;
				if (instBase + i < fp.n && f < lenN) cp_async4_fdn(&sInA[buf][i][f], fp.in + size_t(instBase + i) * fp.frames + t0n + f);
			}
			asm volatile("cp.async.commit_group;" ::: "memory");
		};
		if (fp.frames) prefetch(0, 0);
#endif
		uint32_t tileIdx = 0;
		for (uint32_t t0 = 0; t0 < fp.frames; t0 += kFdnTile, ++tileIdx)
		{
			const uint32_t len = min(kFdnTile, fp.frames - t0);
#ifdef JPLG_FDN_NO_PREFETCH
			float (*sDel)[kFdnRow] = sDelA[0];
			float (*sIn)[kFdnTile] = sInA[0];
			// ---- stage in: warp w loads rows [32w, 32w+32): delayed[c][n] = pushed[n - 1 - D_c] (GetReadWindow<1>(D) before this frame's push).
			// Eight independent loads are issued before the first dependent shared-memory store (DRAM latency is paid once per batch).
			for (uint32_t r0 = 0; r0 < 32; r0 += 8)
			{
				float v[8];
				#pragma unroll
				for (uint32_t r = 0; r < 8; ++r)
				{
					const uint32_t row = warp * 32 + r0 + r, rInst = instBase + (row >> 4), rc = row & 15u;
					v[r] = 0.0f;
					if (rInst < fp.n && lane < len)
						v[r] = __ldcs(fp.rings + (size_t(rInst) * 16u + rc) * kFdnRingSize + ((pos32 + t0 + lane - 1u - cFdnDelays[rc]) & ringMask));
				}
				#pragma unroll
				for (uint32_t r = 0; r < 8; ++r) sDel[warp * 32 + r0 + r][lane] = v[r];
			}
			for (uint32_t e = tid; e < kFdnInstPerCTA * kFdnTile; e += blockDim.x)
			{
				const uint32_t i = e / kFdnTile, f = e - i * kFdnTile;
				if (instBase + i < fp.n && f < len) sIn[i][f] = fp.in[size_t(instBase + i) * fp.frames + t0 + f];
			}
			__syncthreads();
#else
			float (*sDel)[kFdnRow] = sDelA[tileIdx & 1u];
			float (*sIn)[kFdnTile] = sInA[tileIdx & 1u];
			asm volatile("cp.async.wait_group 0;" ::: "memory");
			__syncthreads();            // this tile has landed; the other buffer and sOut are free (the stage-out of the tile before is done)
			if (t0 + kFdnTile < fp.frames) prefetch(t0 + kFdnTile, (tileIdx + 1u) & 1u);
#endif

			// ---- compute, pass 1 (no time recurrence): the feedback mix and the output mix only need the delayed samples, so every
			//      frame of the tile is independent here. x[f] overwrites the delayed sample in place.
			for (uint32_t f = 0; f < len; ++f)
			{
				const float d = sDel[tid][f];
				// Householder::MixInPlace, 16-channel special case (ReverbDesign.h:311-349): mix4 inside each group of 4, then across the groups
				float s = d + __shfl_xor_sync(0xffffffffu, d, 1);
				s = s + __shfl_xor_sync(0xffffffffu, s, 2);
				float m = 2.0f * d - s;                                 // FMA(2, v, -sum): 2v is exact, one rounding in either build
				float s2 = m + __shfl_xor_sync(0xffffffffu, m, 4);
				s2 = s2 + __shfl_xor_sync(0xffffffffu, s2, 8);
				m = (2.0f * m - s2) * 0.25f;
				// FDNInput::InjectNormalized: every channel gets monoInput * sqrt(1/16) (ReverbDesign.h:455-459)
				sDel[tid][f] = m + sIn[tid >> 4][f] * 0.25f;
				// FDNOutputMixer::Mix (ReverbDesign.h:480-492): Hadamard of the delayed samples (RecursiveUnscaled :391-437), * scalingFactor
				float h = d;
				#pragma unroll
				for (uint32_t bit = 1; bit < 16; bit <<= 1)
				{
					const float o = __shfl_xor_sync(0xffffffffu, h, bit);
					h = (c & bit) ? o - h : h + o;
				}
				sOut[tid >> 4][f][c] = h * fp.hadamardScale;
			}
			// ---- pass 2: AttenuationFilter::Process per channel, frames in order (the only time recurrence); unrolled so the filter
			//      stages of neighbouring frames overlap. The result is the sample Push()ed into the ring (ReverbDesign.h:696-699).
#ifdef JPLG_FDN_SCALAR
			if (len == kFdnTile)
			{
				#pragma unroll 8
				for (uint32_t f = 0; f < kFdnTile; ++f) sDel[tid][f] = xover_process_sample<MODE>(fp.coef, st, sDel[tid][f], g);
			}
			else for (uint32_t f = 0; f < len; ++f) sDel[tid][f] = xover_process_sample<MODE>(fp.coef, st, sDel[tid][f], g);
#else
			if (len == kFdnTile)
			{
				#pragma unroll 8
				for (uint32_t f = 0; f < kFdnTile; ++f) sDel[tid][f] = xover_process_sample2<MODE>(fp.coef, K, ps, sDel[tid][f], gA, gB);
			}
			else for (uint32_t f = 0; f < len; ++f) sDel[tid][f] = xover_process_sample2<MODE>(fp.coef, K, ps, sDel[tid][f], gA, gB);
#endif
			__syncthreads();

			// ---- stage out: push the new samples (ring index = push count), fold the mixed FDN channels into the output frames
			for (uint32_t r = 0; r < 32; ++r)
			{
				const uint32_t row = warp * 32 + r, rInst = instBase + (row >> 4), rc = row & 15u;
				if (rInst < fp.n && lane < len)
					fp.rings[(size_t(rInst) * 16u + rc) * kFdnRingSize + ((pos32 + t0 + lane) & ringMask)] = sDel[row][lane];
			}
			// an instance's tile of output frames is one contiguous run of len * L floats: thread -> elements tid, tid + 128, ... of every
			// instance ((frame, channel) advance incrementally, no divisions); the 8 instances' read-modify-writes are independent
			const uint32_t perInst = len * L;
			uint32_t ff = oFrame0, oc = oChan0;
			for (uint32_t idx = tid; idx < perInst; idx += kFdnInstPerCTA * 16)
			{
				float prev[kFdnInstPerCTA];
				#pragma unroll
				for (uint32_t i = 0; i < kFdnInstPerCTA; ++i)
					if (instBase + i < fp.n) prev[i] = fp.out[(size_t(instBase + i) * fp.frames + t0) * L + idx];
				#pragma unroll
				for (uint32_t i = 0; i < kFdnInstPerCTA; ++i)
				{
					if (instBase + i >= fp.n) continue;
					float acc = prev[i];                                 // out[i % numOut] += fdn[i], i ascending
					for (uint32_t k = oc; k < 16; k += L) acc += sOut[i][ff][k];
					fp.out[(size_t(instBase + i) * fp.frames + t0) * L + idx] = acc;
				}
				ff += oStepFrame; oc += oStepChan;
				if (oc >= L) { oc -= L; ++ff; }
			}
#ifdef JPLG_FDN_NO_PREFETCH
			__syncthreads();
#endif
		}

#ifndef JPLG_FDN_SCALAR
		xover_unpack_state(ps, st);
#endif
		if (active)
		{
			#pragma unroll
			for (int k = 0; k < 24; ++k) fp.state[size_t(k) * nCh + ch] = st[k];
		}
	}

	float FdnHadamardScale()
	{
		const double x = 1.0 / 16.0;
		uint64_t i; std::memcpy(&i, &x, 8);
		i = 0x5fe6eb50c7b537a9ull - (i >> 1);              // InvSqrtInitialGuess<double> (Math.h:221-226)
		double y; std::memcpy(&y, &i, 8);
		const double halfx = 0.5 * x;
		for (int k = 0; k < 2; ++k) y = y * (1.5 - halfx * y * y);   // InvSqrtNewtonRefine<double, 2> (:236-242)
		return static_cast<float>(x * y);                  // SqrtConstevalFallback (:255), narrowed by `constexpr Sample`
	}

	// ReverbBus::ProcessInterleaved's multichannel input stage (src/Auralization/LateReverb.cpp:80-92): scratch = 0, DownmixToMono adds the channels in
	// order (inc/Auralization/ChannelMixing.h:129-136), ApplyGain scales by 1 / numInChannels (:50-63). One thread per frame of one instance.
	__global__ void k_reverb_downmix(const float* __restrict__ in, uint64_t total, uint32_t numIn, float norm, float* __restrict__ mono)
	{
		const uint64_t idx = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
		if (idx >= total) return;
		const float* s = in + idx * numIn;
		float d = 0.0f;
		for (uint32_t ch = 0; ch < numIn; ++ch) d += s[ch];
		mono[idx] = d * norm;
	}
	cudaError_t LaunchReverbDownmix(const float* in, uint32_t n, uint32_t frames, uint32_t numIn, float* mono, cudaStream_t st)
	{
		const uint64_t total = uint64_t(n) * frames;
		if (total == 0) return cudaSuccess;
		k_reverb_downmix<<<static_cast<uint32_t>((total + 255) / 256), 256, 0, st>>>(in, total, numIn, 1.0f / static_cast<float>(numIn), mono);
		return cudaGetLastError();
	}

	//======================================================================
	// Scratch of jplg_reverb_bank_process_no_alloc (include/jplg.h): the mono down-mix of a multichannel reverb input, [n][frames], in a file-scope device
	// array instead of the context's cudaMalloc'ed scratch. 245,760 floats = 960 KiB, held by every GPU that loads the library until the process exits.
	// k_reverb_downmix overwrites every element k_fdn later reads, so nothing in it has to start at zero; all calls share it (calls that overlap on
	// the GPU race). In HOST code the array's name is a host-side stand-in, not a device address: the device address comes from cudaGetSymbolAddress,
	// looked up once per device with that device current.
	static __device__ __align__(256) float g_reverbNoAllocMono[kReverbNoAllocMaxFloats];

	cudaError_t ReverbNoAllocScratch(int device, float** out)
	{
		constexpr int kMaxDevices = 64;
		static std::mutex mtx;
		static float* addr[kMaxDevices] = {};
		*out = nullptr;
		if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
		std::lock_guard<std::mutex> lk(mtx);
		if (!addr[device])
		{
			int current = -1;
			cudaError_t e = cudaGetDevice(&current); if (e != cudaSuccess) return e;
			if (current != device) return cudaErrorInvalidDevice;      // the caller makes the context's device current first
			void* p = nullptr;
			e = cudaGetSymbolAddress(&p, g_reverbNoAllocMono); if (e != cudaSuccess) return e;
			addr[device] = static_cast<float*>(p);
		}
		*out = addr[device];
		return cudaSuccess;
	}

	cudaError_t LaunchFdnDecay(const float* slope, uint32_t n, float invSampleRate, float* decay, int mode, cudaStream_t st)
	{
		if (n == 0) return cudaSuccess;
		const uint32_t total = n * 16u, grid = (total + 255) / 256;
		if (mode == 0) k_fdn_decay<0><<<grid, 256, 0, st>>>(slope, n, invSampleRate, decay); else k_fdn_decay<1><<<grid, 256, 0, st>>>(slope, n, invSampleRate, decay);
		return cudaGetLastError();
	}

	cudaError_t LaunchFdn(const FdnParams& fp, int mode, cudaStream_t st)
	{
		if (fp.n == 0 || fp.frames == 0) return cudaSuccess;
		const uint32_t grid = (fp.n + kFdnInstPerCTA - 1) / kFdnInstPerCTA;
		const PackedConsts K = MakePackedConsts(fp.coef);
#ifdef JPLG_FDN_NO_PREFETCH
		const size_t dyn = 0;
#else
		const size_t dyn = (size_t(2) * kFdnInstPerCTA * 16 * kFdnRow + size_t(2) * kFdnInstPerCTA * kFdnTile + size_t(kFdnInstPerCTA) * kFdnTile * 16) * sizeof(float);   // 52 KB
		static bool attr[64] = {};                        // function attributes are per device
		int dev = 0; cudaGetDevice(&dev);
		if (dev >= 0 && dev < 64 && !attr[dev])
		{
			cudaFuncSetAttribute(k_fdn<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(dyn));
			cudaFuncSetAttribute(k_fdn<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(dyn));
			attr[dev] = true;
		}
#endif
		if (mode == 0) k_fdn<0><<<grid, kFdnInstPerCTA * 16, dyn, st>>>(fp, K); else k_fdn<1><<<grid, kFdnInstPerCTA * 16, dyn, st>>>(fp, K);
		return cudaGetLastError();
	}
}
