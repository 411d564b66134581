// K1: batched VBAP / MDAP gain evaluation for sm_100a.
//   k_point_gains_{2d,3d}  one thread per query         (GetSpeakerGains, inc/Panning/PannerBase.h:441-445)
//   k_mdap_3d<G>           G = 8 or 32 lanes per source (ProcessVBAPData, PannerBase.h:486-728, 3D panner)
//   k_mdap_2d              one warp per source           (same, 2D panner)
//   k_process_vs           one thread per virtual-source set (ProcessVirtualSources, PannerBase.h:731-804)
// The direction -> LUT-index arithmetic is bit-exact with the reference's SSE2 build (jplg_device_math.cuh);
// the per-speaker accumulation is done in the reference's own order, so gains are bit-exact too.
// Compile with -fmad=false.
#include "jplg_kernels.hpp"
#include "jplg_device_math.cuh"
#include <cstdlib>

namespace jplg
{
	using namespace dev;

	//======================================================================
	__global__ void k_point_gains_3d(const uint4* __restrict__ lut, const float* __restrict__ dirs, uint32_t n, uint32_t L, float* __restrict__ out)
	{
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n) return;
		const float x = dirs[3 * i], y = dirs[3 * i + 1], z = dirs[3 * i + 2];
		const uint32_t code = oct_encode_scalar(x, y, z);
		const uint4 cell = __ldg(&lut[code]);
		float* row = out + size_t(i) * L;
		for (uint32_t c = 0; c < L; ++c) row[c] = 0.0f;
		// outGains[speaker[k]] = gain[k] in slot order (inc/Panning/VBAPLUT3D.h:1001-1009)
		const uint32_t s0 = cell.w & 0xffu, s1 = (cell.w >> 8) & 0xffu, s2 = (cell.w >> 16) & 0xffu;
		if (s0 < L) row[s0] = u2f(cell.x);
		if (s1 < L) row[s1] = u2f(cell.y);
		if (s2 < L) row[s2] = u2f(cell.z);
	}

	__global__ void k_point_gains_2d(const float* __restrict__ lut, const float* __restrict__ dirs, uint32_t n, uint32_t L, float* __restrict__ out)
	{
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n) return;
		// LUTQuery2D::GainsFor (inc/Panning/VBAPLUT2D.h:193-205): zero Y, normalise (exact 1/sqrt), diamond index, copy a row
		float x = dirs[3 * i], z = dirs[3 * i + 2];
		const float y = 0.0f;
		const float il = invsqrt_exact(x * x + y * y + z * z);
		x *= il; z *= il;
		const uint32_t pos = lut2d_pos_scalar(x, z);
		const float* src = lut + size_t(pos) * L;
		float* row = out + size_t(i) * L;
		for (uint32_t c = 0; c < L; ++c) row[c] = __ldg(&src[c]);
	}

	//======================================================================
	struct MdapCommon
	{
		Q4 qPan;
		float focus, tSlerp;
		F3 target;
	};

	// stride 5: PanUpdateData {dir, focus, spread}; stride 11: PanUpdateDataWithOrientation {dir, focus, spread, up, forward}
	// (PannerBase.h:486-500 vs :501-515: only the pan quaternion differs)
	__device__ __forceinline__ MdapCommon mdap_prologue(const float* __restrict__ params, uint32_t src, uint32_t stride)
	{
		const float* p = params + size_t(src) * stride;
		MdapCommon m;
		m.target = { p[0], p[1], p[2] };
		const float focusIn = p[3], spreadIn = p[4];
		m.focus = (focusIn == 1.0f) ? 0.9999f : focusIn;          // PannerBase.h:589
		const float spread = (spreadIn == 0.0f) ? 0.0001f : spreadIn; // :646
		m.tSlerp = 1.0f - spread;                                  // :648
		m.qPan = stride == 11 ? quat_from_up_and_forward(F3{ p[5], p[6], p[7] }, F3{ p[8], p[9], p[10] })  // :509 QuatFromUpAndForward(Orientation.Up, Orientation.Forward)
		                      : quat_from_up_and_forward(F3{ 0.0f, 1.0f, 0.0f }, m.target);                // :495 QuatLookAt(SourceDirection, world up)
		return m;
	}

	// G lanes per source. Targets with <= 16 channels take G = 8 (r02): four sources share a warp, so the per-source work every lane repeats (pan quaternion,
	// per-channel basis, ring sines) is amortised over twice as many sources as with 16 lanes, the 72 = 9 x 8 virtual sources of mono -> 7.1.4 fill every lane slot
	// (16 lanes: 4.5 -> 5 trips, 10 % idle), and each lane adds two columns of the gain matrix. Wider targets take one warp per source (G = 32).
	// ROWS = virtual sources per accumulation batch (rows of the per-group gain matrix), STRIDE = floats between matrix rows: odd, so that the scatter of a
	// batch (row = lane, column = speaker) and the column sums (row after row, column = lane) both spread over the banks (r01e: 1.5 conflicts per shared-memory
	// instruction with the [32][16] matrix).
	// Shared memory per group: ring sin/cos (2*24 floats) + gain matrix + mixBuffer.
	// ROWS for G = 8 (same-box A/B, profiles/r02q_mdap_rows_ab.txt): 8-row batches need 22 KB of shared memory per CTA instead of 40 KB (9 instead of 5 resident
	// CTAs per SM: the kernel waits on its LUT gathers, more warps hide them) and run a whole-machine launch 17 % faster (config 2: 91.4 -> 76.2 us); the fused
	// update that runs BESIDE a render on the SMs it leaves free is 6 % slower that way (config 3 step 0.1275 -> 0.1344 ms), so it keeps 24-row batches.
	template<int G, int ROWS8 = 24> struct MdapShape { static constexpr uint32_t kRows = G == 8 ? uint32_t(ROWS8) : 32u, kCols = G == 32 ? 32u : 16u, kStride = kCols + 1u, kGroupFloats = 48u + kRows * kStride + kMaxChannelsDev; };
	template<int G, int ROWS8>
	__global__ void __launch_bounds__(JPLG_MDAP_WARPS * 32)
	k_mdap_3d(const uint4* __restrict__ lut, const uint32_t* __restrict__ rsqTable, MdapLayoutParams lp,
			  const float* __restrict__ params, uint32_t n, uint32_t L, float* __restrict__ out, const MdapFuse fz)
	{
		using Shape = MdapShape<G, ROWS8>;
		constexpr uint32_t kRows = Shape::kRows, kStride = Shape::kStride, kGroupFloats = Shape::kGroupFloats;
		constexpr uint32_t kColsPerLane = Shape::kCols / G;      // columns a lane sums: 2 for G = 8, else 1
		extern __shared__ __align__(16) unsigned char smemRaw[];
		uint32_t* sRsq = reinterpret_cast<uint32_t*>(smemRaw);                       // 2048
		float* sAzSin = reinterpret_cast<float*>(sRsq + 2048);                        // 32 (max 24 used: halfAndTail*4 <= 16)
		float* sAzCos = sAzSin + 32;
		float* sGroupBase = sAzCos + 32;
		constexpr uint32_t kGroups = JPLG_MDAP_WARPS * 32 / G;
		const uint32_t grp = threadIdx.x / G, gl = threadIdx.x % G;
		// per group: ring sin/cos [24+24] | gain matrix [kRows][kStride] (row = virtual source of the current batch, column = speaker; all zero
		// between batches) | mixBuffer [32]
		float* sRingS = sGroupBase + grp * kGroupFloats;
		float* sRingC = sRingS + 24;
		float* sMat = sRingC + 24;
		float* sMix = sMat + kRows * kStride;                                         // [32] mixBuffer

		for (uint32_t i = threadIdx.x; i < kGroups * kRows * kStride; i += blockDim.x) sGroupBase[(i / (kRows * kStride)) * kGroupFloats + 48 + i % (kRows * kStride)] = 0.0f;
		for (uint32_t i = threadIdx.x; i < 2048; i += blockDim.x) sRsq[i] = rsqTable[i];
		if (threadIdx.x < 32) { sAzSin[threadIdx.x] = lp.azSin[threadIdx.x]; sAzCos[threadIdx.x] = lp.azCos[threadIdx.x]; }
		const uint32_t R = lp.numRings, H = lp.halfSamples;
		const uint32_t packsHalf = R * H;
		// virtual source -> (ring, azimuth slot, mirrored) of GenerateSpreadCap's pack order (VBAPanning3D.h:449-518): the same for every source,
		// so the two integer divisions per virtual source are done once per CTA. Entry = ring | azimuth float index << 8 | mirror << 16.
		uint32_t* sVsIdx = reinterpret_cast<uint32_t*>(sGroupBase + kGroups * kGroupFloats);
		for (uint32_t v = threadIdx.x; v < lp.numVS; v += blockDim.x)
		{
			const uint32_t pack = v >> 2, l4 = v & 3u;
			uint32_t sci, azi, mirror = 0;
			if (pack < packsHalf) { sci = pack / H; azi = pack - sci * H; }
			else if (pack < 2 * packsHalf) { const uint32_t p2 = pack - packsHalf; sci = p2 / H; azi = p2 - sci * H; mirror = 1; }
			else { sci = pack - 2 * packsHalf; azi = H; }
			sVsIdx[v] = sci | ((azi * 4 + l4) << 8) | (mirror << 16);
		}
		__syncthreads();

		const float nChF = static_cast<float>(lp.numChannels);

		for (uint32_t srcBase = blockIdx.x * kGroups; srcBase < n; srcBase += gridDim.x * kGroups)
		{
			const bool valid = srcBase + grp < n;                       // a warp's groups run the same trip counts; the idle one only skips its stores
			const uint32_t src = valid ? srcBase + grp : n - 1;
			const MdapCommon m = mdap_prologue(params, src, lp.paramStride);

			// GenerateSpreadCap ring centres (inc/Panning/VBAPanning3D.h:414-448): rings in groups of 4 through the
			// SIMD SinCos, the remaining (R & 3) rings through libm
			for (uint32_t r = gl; r < R; r += G)
			{
				const float capSpread = 1.0f - m.focus;
				const float phiTermS = capSpread * lp.phiTerm;
				const float rampv = static_cast<float>((r & 3u) + 1u);
				float phi = phiTermS * rampv;
				const float delta = 4.0f * phiTermS;
				for (uint32_t g = 0; g < (r >> 2); ++g) phi += delta;
				float s, c;
				if (r < (R & ~3u)) sincos_simd(phi, s, c); else sincos_libm(phi, s, c);
				sRingS[r] = s; sRingC[r] = c;
			}
			#pragma unroll
			for (uint32_t k = 0; k < kColsPerLane; ++k) if (gl + k * G < L) sMix[gl + k * G] = 0.0f;
			__syncwarp();

			float* outSrc = out + size_t(src) * lp.numChannels * L;

			for (uint32_t ch = 0; ch < lp.numChannels; ++ch)
			{
				const Q4 qc{ lp.quats[ch][0], lp.quats[ch][1], lp.quats[ch][2], lp.quats[ch][3] };
				const B3 basis = quat_to_basis(quat_mul(m.qPan, qc)); // PannerBase.h:593-601

				// outGains[speakers[s]] += gains[s] in entry order (PannerBase.h:846-849), per speaker == column. A batch of kRows virtual sources
				// scatters its (vsWeight * g) * g values into the zeroed gain matrix (the three speakers of a cell are distinct: a triangle, the dummy
				// replaced by an index not in it, VBAPLUT3D.h:803-815), then every column is added row after row by its lane: the entries of other
				// speakers contribute +0.0f, which leaves a non-negative running sum unchanged bit for bit, so the additions that matter happen in the
				// reference's order. 2 instructions per entry instead of a 14-instruction compare / select scan (ncu r01d: that scan was 33 % of
				// the kernel's issued instructions).
				float acc[kColsPerLane];
				#pragma unroll
				for (uint32_t k = 0; k < kColsPerLane; ++k) acc[k] = 0.0f;
				for (uint32_t vb = 0; vb < lp.numVS; vb += kRows)
				{
					uint32_t mine[kRows / G];      // speakers this lane wrote (to clear them afterwards), 3 x 8 bits each
					#pragma unroll
					for (uint32_t it = 0; it < kRows / G; ++it)
					{
						const uint32_t row = it * G + gl, v = vb + row;
						mine[it] = 0xffffffffu;
						if (v < lp.numVS)
						{
							const uint32_t vi = sVsIdx[v], sci = vi & 0xffu, az = (vi >> 8) & 0xffu;
							const float sinPhi = sRingS[sci], cosPhi = sRingC[sci];
							F3 d;
							d.x = sinPhi * sAzCos[az];
							d.y = sinPhi * sAzSin[az];
							if (vi >> 16) d.y = 0.0f - d.y;
							d.z = cosPhi;
							d = basis_transform_simd(basis, d);                 // RotateChannelCap (:895-901)
							d = slerp_simd(sRsq, d, m.target, m.tSlerp);        // SlerpChannelCap (:904-919)
							const uint32_t code = oct_encode_simd(d.x, d.y, d.z);
							const uint4 cell = __ldg(&lut[code]);
							// (vsWeight * g) * g (PannerBase.h:836-839)
							const float g0 = u2f(cell.x), g1 = u2f(cell.y), g2 = u2f(cell.z);
							const uint32_t s0 = cell.w & 0xffu, s1 = (cell.w >> 8) & 0xffu, s2 = (cell.w >> 16) & 0xffu;
							float* mrow = sMat + row * kStride;
							if (s0 < L) mrow[s0] = (lp.vsWeight * g0) * g0;
							if (s1 < L) mrow[s1] = (lp.vsWeight * g1) * g1;
							if (s2 < L) mrow[s2] = (lp.vsWeight * g2) * g2;
							mine[it] = cell.w & 0xffffffu;
						}
					}
					__syncwarp();
					{
						const uint32_t rows = min(kRows, lp.numVS - vb);
						#pragma unroll
						for (uint32_t k = 0; k < kColsPerLane; ++k)
						{
							if (gl + k * G >= L) continue;
							const float* col = sMat + gl + k * G;
							if (rows == kRows)
							{
								#pragma unroll
								for (uint32_t e = 0; e < kRows; ++e) acc[k] += col[e * kStride];
							}
							else
								for (uint32_t e = 0; e < rows; ++e) acc[k] += col[e * kStride];
						}
					}
					__syncwarp();
					#pragma unroll
					for (uint32_t it = 0; it < kRows / G; ++it)
					{
						if (mine[it] != 0xffffffffu)
						{
							float* mrow = sMat + (it * G + gl) * kStride;
							const uint32_t s0 = mine[it] & 0xffu, s1 = (mine[it] >> 8) & 0xffu, s2 = (mine[it] >> 16) & 0xffu;
							if (s0 < L) mrow[s0] = 0.0f;
							if (s1 < L) mrow[s1] = 0.0f;
							if (s2 < L) mrow[s2] = 0.0f;
						}
					}
					// the next batch's writes of a lane only touch rows it owns, which it has just cleared; readers of this batch are behind the barrier above
				}
				#pragma unroll
				for (uint32_t k = 0; k < kColsPerLane; ++k)
				{
					const uint32_t c = gl + k * G;
					if (c < L)
					{
						if (valid) outSrc[ch * L + c] = acc[k]; // pre-normalisation value; rescaled below
						sMix[c] += acc[k];                       // mixBuffer[i] += gains[i] (:683-686)
					}
				}
				__syncwarp();
			}

			// Normalise for the number of input channels (:694-713)
			#pragma unroll
			for (uint32_t k = 0; k < kColsPerLane; ++k)
			{
				const uint32_t c = gl + k * G;
				if (c < L && valid)
				{
					const float g = sMix[c];
					const float norm = g > 0.0f ? invsqrt_exact(g * nChF) : 0.0f;
					for (uint32_t ch = 0; ch < lp.numChannels; ++ch)
					{
						const float v = outSrc[ch * L + c] * norm;
						outSrc[ch * L + c] = v;
						if (fz.tgtMix) fz.tgtMix[size_t(c) * fz.numSC + size_t(src) * fz.rendC + fz.rows[ch]] = v;   // k_update_targets' transpose
					}
				}
			}
			if (fz.tgtMix && valid)
			{
				if (fz.unpannedRows)
				{
					#pragma unroll
					for (uint32_t k = 0; k < kColsPerLane; ++k)
					{
						const uint32_t c = gl + k * G;
						if (c < L)
							for (uint32_t ch = 0; ch < fz.rendC; ++ch)
								if ((fz.unpannedRows >> ch) & 1u) fz.tgtMix[size_t(c) * fz.numSC + size_t(src) * fz.rendC + ch] = 0.0f;
					}
				}
				// UpdateParameters (src/Auralization/DirectSound.cpp:231-239) for this source's channels: band gains, delay in samples
				if (gl < 4)
				{
					const float b = fz.band[size_t(src) * 4 + gl];
					for (uint32_t ch = 0; ch < fz.rendC; ++ch) fz.tgtBand[size_t(gl) * fz.numSC + size_t(src) * fz.rendC + ch] = b;
				}
				if (gl == 4)
				{
					const float d = fz.delaySec[src] * fz.sampleRate;
					for (uint32_t ch = 0; ch < fz.rendC; ++ch) fz.tgtDelay[size_t(src) * fz.rendC + ch] = d;
				}
			}
			__syncwarp();
		}
	}

	// 2D panner MDAP: one warp per source; lanes cover (channel, virtual source) pairs (<= 32 in total)
	__global__ void __launch_bounds__(JPLG_MDAP_WARPS * 32)
	k_mdap_2d(const float* __restrict__ lut, const uint32_t* __restrict__ rsqTable, MdapLayoutParams lp,
			  const float* __restrict__ params, uint32_t n, uint32_t L, float* __restrict__ out)
	{
		extern __shared__ __align__(16) unsigned char smemRaw[];
		uint32_t* sRsq = reinterpret_cast<uint32_t*>(smemRaw);
		uint32_t* sPosBase = sRsq + 2048;
		const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
		uint32_t* sPos = sPosBase + warp * 32;
		for (uint32_t i = threadIdx.x; i < 2048; i += blockDim.x) sRsq[i] = rsqTable[i];
		__syncthreads();

		const uint32_t numVS = lp.numVS, packs = numVS >> 2;
		const uint32_t toMirror = packs & ~1u, tail = packs & 1u, H = toMirror >> 1, halfAndTail = H + tail;
		const uint32_t totalVS = numVS * lp.numChannels; // <= 32 (reference buffer capacity)
		const float nChF = static_cast<float>(lp.numChannels);

		for (uint32_t src = blockIdx.x * JPLG_MDAP_WARPS + warp; src < n; src += gridDim.x * JPLG_MDAP_WARPS)
		{
			const MdapCommon m = mdap_prologue(params, src, lp.paramStride);
			if (lane < totalVS)
			{
				const uint32_t ch = lane / numVS, v = lane - ch * numVS;
				const uint32_t pack = v >> 2, l4 = v & 3u;
				// GenerateSpreadCap (inc/Panning/VBAPanning2D.h:196-246)
				const uint32_t gpack = pack < halfAndTail ? pack : pack - halfAndTail;
				const bool mirror = pack >= halfAndTail;
				const float spread = 1.0f - m.focus;
				const float phiTerm = spread * lp.phiTerm;
				const float offset = (-spread) * lp.offsetTerm;
				const float rampv = 0.5f + static_cast<float>(l4);
				float phi = phiTerm * rampv + offset;
				const float delta = 4.0f * phiTerm;
				for (uint32_t g = 0; g < gpack; ++g) phi += delta;
				F3 d; float s, c;
				sincos_simd(phi, s, c);
				d.x = mirror ? (0.0f - s) : s; d.y = 0.0f; d.z = c;
				const Q4 qc{ lp.quats[ch][0], lp.quats[ch][1], lp.quats[ch][2], lp.quats[ch][3] };
				const B3 basis = quat_to_basis(quat_mul(m.qPan, qc));
				d = basis_transform_simd(basis, d);
				d = slerp_simd(sRsq, d, m.target, m.tSlerp);
				sPos[lane] = lut2d_pos_simd(sRsq, d.x, d.z);           // LUT2D::CartesianToLUTPosition(simd) (VBAPLUT2D.h:376-385)
			}
			__syncwarp();
			float* outSrc = out + size_t(src) * lp.numChannels * L;
			if (lane < L)
			{
				// ProcessVirtualSourcesImplSIMD 2D branch (PannerBase.h:852-875): per pack, per speaker: (w*g*g).reduce()
				float mix = 0.0f;
				float gains[kMaxSourceChannelsDev];
				for (uint32_t ch = 0; ch < lp.numChannels; ++ch)
				{
					float acc = 0.0f;
					for (uint32_t pk = 0; pk < packs; ++pk)
					{
						float t[4];
						#pragma unroll
						for (uint32_t l4 = 0; l4 < 4; ++l4)
						{
							const float g = __ldg(&lut[size_t(sPos[ch * numVS + pk * 4 + l4]) * L + lane]);
							t[l4] = (lp.vsWeight * g) * g;
						}
						acc += (t[0] + t[1]) + (t[2] + t[3]);
					}
					gains[ch] = acc;
					mix += acc;
				}
				const float norm = mix > 0.0f ? invsqrt_exact(mix * nChF) : 0.0f;
				for (uint32_t ch = 0; ch < lp.numChannels; ++ch) outSrc[ch * L + lane] = gains[ch] * norm;
			}
			__syncwarp();
		}
	}

	//======================================================================
	// ProcessVirtualSources: bAccumulatePow=false path + Algo::NormalizeL2 (PannerBase.h:731-804). One thread per set.
	__global__ void k_process_vs(const uint4* __restrict__ lut3d, const float* __restrict__ lut2d, int is3D,
								 const float* __restrict__ vs, uint32_t n, uint32_t batches, uint32_t L, float* __restrict__ out)
	{
		const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
		if (b >= batches) return;
		float* o = out + size_t(b) * L;
		const float* v = vs + size_t(b) * n * 4;
		for (uint32_t i = 0; i < n; ++i)
		{
			const float x = v[4 * i], y = v[4 * i + 1], z = v[4 * i + 2], w = v[4 * i + 3];
			if (is3D)
			{
				const uint4 cell = __ldg(&lut3d[oct_encode_scalar(x, y, z)]);
				const uint32_t s0 = cell.w & 0xffu, s1 = (cell.w >> 8) & 0xffu, s2 = (cell.w >> 16) & 0xffu;
				if (s0 < L) o[s0] += w * u2f(cell.x);
				if (s1 < L) o[s1] += w * u2f(cell.y);
				if (s2 < L) o[s2] += w * u2f(cell.z);
			}
			else
			{
				float xx = x, zz = z;
				const float il = invsqrt_exact(xx * xx + 0.0f * 0.0f + zz * zz);
				xx *= il; zz *= il;
				const float* row = lut2d + size_t(lut2d_pos_scalar(xx, zz)) * L;
				for (uint32_t s = 0; s < L; ++s) o[s] += w * __ldg(&row[s]);
			}
		}
		float sum2 = 0.0f;
		for (uint32_t s = 0; s < L; ++s) sum2 = sum2 + o[s] * o[s];
		const float inv = invsqrt_exact(sum2);
		for (uint32_t s = 0; s < L; ++s) o[s] *= inv;
	}

	//======================================================================
	// launchers
	cudaError_t LaunchPointGains(const PannerDev& p, const float* dirs, uint32_t n, float* out, cudaStream_t st)
	{
		if (n == 0) return cudaSuccess;
		const uint32_t block = 128, grid = (n + block - 1) / block;
		if (p.is3D) k_point_gains_3d<<<grid, block, 0, st>>>(p.lut3d, dirs, n, p.numChannels, out);
		else k_point_gains_2d<<<grid, block, 0, st>>>(p.lut2d, dirs, n, p.numChannels, out);
		return cudaGetLastError();
	}

	// JPLG_MDAP_LANES=16 (experiment switch): the r01 mapping, 16 lanes per source for targets with <= 16 channels
	static uint32_t MdapLanesPerSource(const PannerDev& p)
	{
		static const bool lanes16 = [] { const char* e = std::getenv("JPLG_MDAP_LANES"); return e && e[0] == '1' && e[1] == '6'; }();
		return (p.is3D && p.numChannels <= 16) ? (lanes16 ? 16u : 8u) : 32u;
	}
	static uint32_t MdapGroupFloats(uint32_t G, bool rows8) { return G == 8 ? (rows8 ? MdapShape<8, 8>::kGroupFloats : MdapShape<8, 24>::kGroupFloats) : G == 16 ? MdapShape<16>::kGroupFloats : MdapShape<32>::kGroupFloats; }

	size_t MdapSmemBytes(const PannerDev& p, const MdapLayoutParams& lp, bool rows8)
	{
		const uint32_t groups = JPLG_MDAP_WARPS * 32 / MdapLanesPerSource(p);
		if (p.is3D) return (2048 + 64) * 4 + size_t(groups) * MdapGroupFloats(MdapLanesPerSource(p), rows8) * 4 + size_t(lp.numVS) * 4;
		return (2048 + JPLG_MDAP_WARPS * 32) * 4;
	}

	cudaError_t LaunchMdap(const PannerDev& p, const MdapLayoutParams& lpIn, const float* params, uint32_t paramStride, uint32_t n, float* out, int numSMs, cudaStream_t st,
						   const MdapFuse* fuse)
	{
		const MdapFuse fz = fuse ? *fuse : MdapFuse{};
		if (fz.tgtMix && !p.is3D) return cudaErrorNotSupported;
		MdapLayoutParams lp = lpIn;
		lp.paramStride = paramStride;
		if (n == 0) return cudaSuccess;
		static const int rowsEnv = [] { const char* e = std::getenv("JPLG_MDAP_ROWS8"); return e ? std::atoi(e) : -1; }();
		bool rows8 = fz.tgtMix == nullptr;     // plain (foreground, whole machine) launches: 8-row batches; the fused background update: 24
		if (rowsEnv >= 0) rows8 = rowsEnv != 0;
		const size_t smem = MdapSmemBytes(p, lp, rows8);
		const uint32_t groups = JPLG_MDAP_WARPS * 32 / MdapLanesPerSource(p);
		uint32_t grid = (n + groups - 1) / groups;
		// JPLG_MDAP_CTAS_PER_SM=n, JPLG_MDAP_ROWS8=0/1 (experiment switches): grid cap, batch rows of the 8-lane kernel whatever the launch kind
		static const uint32_t capPerSM = [] { const char* e = std::getenv("JPLG_MDAP_CTAS_PER_SM"); const int v = e ? std::atoi(e) : 0; return v > 0 ? uint32_t(v) : 8u; }();
		const uint32_t maxGrid = static_cast<uint32_t>(numSMs) * capPerSM; // persistent-ish: amortise the 8 KB rsqrt table load
		if (grid > maxGrid) grid = maxGrid;
		if (p.is3D)
		{
			auto kernel = groups == JPLG_MDAP_WARPS ? k_mdap_3d<32, 24> : (groups == JPLG_MDAP_WARPS * 2 ? k_mdap_3d<16, 24> : (rows8 ? k_mdap_3d<8, 8> : k_mdap_3d<8, 24>));
			cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
			kernel<<<grid, JPLG_MDAP_WARPS * 32, smem, st>>>(p.lut3d, p.rsqTable, lp, params, n, p.numChannels, out, fz);
		}
		else
		{
			k_mdap_2d<<<grid, JPLG_MDAP_WARPS * 32, smem, st>>>(p.lut2d, p.rsqTable, lp, params, n, p.numChannels, out);
		}
		return cudaGetLastError();
	}

	cudaError_t LaunchProcessVS(const PannerDev& p, const float* vs, uint32_t n, uint32_t batches, float* out, cudaStream_t st)
	{
		if (batches == 0) return cudaSuccess;
		const uint32_t block = 64, grid = (batches + block - 1) / block;
		k_process_vs<<<grid, block, 0, st>>>(p.lut3d, p.lut2d, p.is3D ? 1 : 0, vs, n, batches, p.numChannels, out);
		return cudaGetLastError();
	}
}
