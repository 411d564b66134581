// Scene front-end (SURVEY.md 8f.2 + 8f.3): the per-source body of SpatialManager::AdvanceSimulation batched on the device, so the
// parameters of stages 1-3 (pan parameters, band gains, delay) are produced next to the kernels that consume them, without a host
// round trip. One thread per source; 60-170 B of traffic per source, so the kernel is launch-latency / HBM-latency bound and tiny.
//
// Reference, per source (inc/SpatialManager.h:374-420):
//   DirectPathService::ProcessDirectPath            inc/Services/DirectPathService.h:280-338
//   DirectPathService::EvaluateDistance  -> AttenuationFunction::Evaluate          inc/DistanceAttenuation.h:66-119,172-215, inc/Math/Curves.h:58-85
//   DirectPathService::EvaluateDirection -> ProcessAngleAttenuationImpl            DirectPathService.h:340-364
//   GetSpreadFromSourceSize                                                         SpatialManager.h:617-623 (PannerBase.h:69-74)
//   PanningService::EvaluateDirection -> Impl::SanitizeSourceDir                    inc/Services/PanningService.h:480-495,527-549
// and the producers an engine calls with the cached distance:
//   AirAbsorption::ComputeForDistance + dBToGain(simd)                              inc/AirAbsorption.h:212-234, inc/Math/DecibelsAndGain.h:40-50
//   distance * SpeedOfSound::cInvReference                                          inc/SpeedOfSound.h:32-33
//
// Arithmetic: every scalar float operation of the reference is executed as the same IEEE binary32 operation in the same order
// (-fmad=false). Where the reference calls libm (cosf of the cone angles, atanf, powf / logf / sinf / cosf of a curve segment) the
// device evaluates the function in double and rounds once: equal to glibc's result except where glibc itself is not correctly
// rounded (tests allow 2 ulp there and say so).
#include "jplg_kernels.hpp"
#include "jplg_device_math.cuh"

namespace jplg
{
	using namespace dev;

	namespace fe
	{
		// JPL::exp(simd), one lane, SSE2 semantic (inc/Math/SIMDMath.h:296-316; floor inc/Math/SIMD.h:1861-1879)
		__device__ __forceinline__ float exp_simd_lane(float x)
		{
			x = fminf(fmaxf(x, -88.3762626647949f), 88.3762626647949f);
			const float v = x * 1.44269504088896341f + 0.5f;
			const float tr = static_cast<float>(static_cast<int32_t>(v));
			const float n = (v == 0.0f) ? v : tr - (tr > v ? 1.0f : 0.0f);
			const float nC2 = n * -2.12194440e-4f;
			x = x - (n * 0.693359375f + nC2);
			const float x2 = x * x;
			float y = 1.9875691500e-4f;
			y = y * x + 1.3981999507e-3f;
			y = y * x + 8.3334519073e-3f;
			y = y * x + 4.1665795894e-2f;
			y = y * x + 1.6666665459e-1f;
			y = y * x + 5.0000001201e-1f;
			y = y * x2 + (x + 1.0f);
			return y * __int_as_float((static_cast<int32_t>(n) + 0x7f) << 23);
		}

		__device__ __forceinline__ float libm1(double v) { return static_cast<float>(v); }
		__device__ __forceinline__ float clampf(float v, float lo, float hi) { return v < lo ? lo : (hi < v ? hi : v); } // std::clamp

		// Curve::EvaluateCurve (inc/Math/Curves.h:58-85)
		__device__ float evaluate_curve(int type, float t)
		{
			const float halfPi = 3.14159274f * 0.5f;
			switch (type)
			{
			case 0: return 1.0f;
			case 1: return t;
			case 2: return libm1(sin(static_cast<double>(t * halfPi)));
			case 3: return 1.0f - libm1(cos(static_cast<double>(t * halfPi)));
			case 4: return t * t * (3.0f - 2.0f * t);
			case 5: { const float inv = 1.0f / libm1(log(static_cast<double>(1.41f))); return libm1(log(static_cast<double>(1.0f + 0.41f * t))) * inv; }
			case 6: { const float inv = 1.0f / libm1(log(static_cast<double>(3.0f))); return libm1(log(static_cast<double>(1.0f + 2.0f * t))) * inv; }
			case 7: return (libm1(pow(static_cast<double>(1.41f), static_cast<double>(t))) - 1.0f) * (1.0f / 0.41f);
			case 8: return (libm1(pow(3.0, static_cast<double>(t))) - 1.0f) * 0.5f;
			default: return t;
			}
		}

		// AttenuationBaseModel::Evaluate (inc/DistanceAttenuation.h:66-119)
		__device__ float evaluate_model(const AttFunctionDev& f, float distance)
		{
			const bool valid = f.minDistance < f.maxDistance;
			const float dc = clampf(distance, f.minDistance, f.maxDistance);
			switch (f.model)
			{
			case 1: return valid ? f.minDistance / (f.minDistance + f.rolloff * (dc - f.minDistance)) : 1.0f;
			case 2: return valid ? 1.0f - f.rolloff * (dc - f.minDistance) / (f.maxDistance - f.minDistance) : 1.0f;
			case 3: return valid ? libm1(pow(static_cast<double>(dc / f.minDistance), static_cast<double>(-f.rolloff))) : 1.0f;
			default: return 1.0f;
			}
		}

		// AttenuationCurve::Evaluate / EvaluateFunction / ComputeAttenuation (inc/DistanceAttenuation.h:172-215)
		__device__ float evaluate_points(const float* __restrict__ pts, uint32_t count, float distance)
		{
			if (count == 0) return 1.0f;
			if (distance <= pts[0]) return pts[1];                                   // before the first point: EvaluateFunction(front, d, nullptr) = Value
			if (distance >= pts[3 * (count - 1)]) return pts[3 * (count - 1) + 1];   // beyond the last point
			uint32_t lo = 0, len = count;                                            // std::upper_bound(distance, value < point.Distance)
			while (len > 0)
			{
				const uint32_t half = len >> 1, mid = lo + half;
				if (distance < pts[3 * mid]) len = half; else { lo = mid + 1; len -= half + 1; }
			}
			const float* prev = pts + 3 * (lo - 1);
			if (lo == count) return prev[1];
			const float* next = pts + 3 * lo;
			const int type = static_cast<int>(prev[2]);
			if (type == 0) return prev[1];                                           // Constant
			const float amplitude = next[1] - prev[1];
			float t = (distance - prev[0]) / (next[0] - prev[0]);
			t = clampf(t, 0.0f, 1.0f);
			return prev[1] + amplitude * evaluate_curve(type, t);
		}

		// DirectPathService::ProcessAngleAttenuationImpl (DirectPathService.h:340-364)
		__device__ float angle_attenuation(float azimuthCos, float innerAngle, float outerAngle)
		{
			const float cutoffInner = libm1(cos(static_cast<double>(innerAngle * 0.5f)));
			const float cutoffOuter = libm1(cos(static_cast<double>(outerAngle * 0.5f)));
			if (azimuthCos > cutoffInner) return 0.0f;
			if (azimuthCos > cutoffOuter) return (cutoffInner - azimuthCos) / (cutoffInner - cutoffOuter);
			return 1.0f;
		}
	}

	__global__ void __launch_bounds__(128) k_scene_frontend(const FrontendArgs fa)
	{
		using namespace fe;
		for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < fa.n; i += gridDim.x * blockDim.x)
		{
			float distance, invDirectionDot, directionDot = 0.0f;
			F3 dir{ 0.0f, 0.0f, -1.0f };
			if (fa.sources)
			{
				// ---- ProcessDirectPath (DirectPathService.h:284-337)
				const float* s = fa.sources + size_t(i) * 9;
				const float* l = fa.listeners + size_t(fa.listenerIndex ? fa.listenerIndex[i] : 0u) * 9;
				const F3 sLoc{ s[0], s[1], s[2] }, sUp{ s[3], s[4], s[5] }, sFwd{ s[6], s[7], s[8] };
				const F3 lLoc{ l[0], l[1], l[2] }, lUp{ l[3], l[4], l[5] };
				const F3 lZ{ -l[6], -l[7], -l[8] };                       // ToBasisUnsafe: forward flipped (inc/Math/Position.h:58-68)
				const B3 lb{ cross3(lUp, lZ), lUp, lZ };
				const F3 relPos{ sLoc.x - lLoc.x, sLoc.y - lLoc.y, sLoc.z - lLoc.z };
				F3 p{ dot3(relPos, lb.X), dot3(relPos, lb.Y), dot3(relPos, lb.Z) };   // Basis::InverseTransform
				if (absbits(p.x) <= 1e-6f && absbits(p.z) <= 1e-6f)     // straight above / below / on top: nudge forward (:291-294)
				{
					p.x += 0.0f * 1e-5f; p.y += 0.0f * 1e-5f; p.z += -1.0f * 1e-5f;
				}
				distance = sqrtf(p.x * p.x + p.y * p.y + p.z * p.z);
				dir = { p.x / distance, p.y / distance, p.z / distance };
				directionDot = dir.x * 0.0f + dir.y * 0.0f + dir.z * -1.0f;
				const F3 nd{ -dir.x, -dir.y, -dir.z };
				const F3 l2s{ dot3(nd, lb.X), dot3(nd, lb.Y), dot3(nd, lb.Z) };
				invDirectionDot = dot3(sFwd, l2s);
				if (fa.rel || fa.panStride == 11)
				{
					// source orientation in the listener's frame: listenerBasis.InverseTransform(source basis) (:307)
					const F3 sZ{ -sFwd.x, -sFwd.y, -sFwd.z };
					const F3 relUp{ dot3(sUp, lb.X), dot3(sUp, lb.Y), dot3(sUp, lb.Z) };
					const F3 relFwd{ dot3(sZ, lb.X), dot3(sZ, lb.Y), dot3(sZ, lb.Z) };   // .Forward = basis.Z (:334)
					if (fa.rel)
					{
						float* o = fa.rel + size_t(i) * 9;
						o[0] = dir.x; o[1] = dir.y; o[2] = dir.z; o[3] = relUp.x; o[4] = relUp.y; o[5] = relUp.z; o[6] = relFwd.x; o[7] = relFwd.y; o[8] = relFwd.z;
					}
					if (fa.pan && fa.panStride == 11)
					{
						float* o = fa.pan + size_t(i) * 11;
						o[5] = relUp.x; o[6] = relUp.y; o[7] = relUp.z; o[8] = relFwd.x; o[9] = relFwd.y; o[10] = relFwd.z;
					}
				}
			}
			else
			{
				distance = fa.distanceIn ? fa.distanceIn[i] : 0.0f;
				invDirectionDot = fa.dotIn ? fa.dotIn[i] : 0.0f;
			}

			// ---- EvaluateDirection: cone factor of the source's cone towards the listener (SpatialManager.h:404)
			if (fa.direct)
			{
				// one 16-byte store per source (the rows are 16-byte aligned: cudaMalloc / pinned staging); without a cone the last column keeps 0
				float cone = 0.0f;
				if (fa.cones) { const float2 c = reinterpret_cast<const float2*>(fa.cones)[i]; cone = angle_attenuation(invDirectionDot, c.x, c.y); }
				reinterpret_cast<float4*>(fa.direct)[i] = make_float4(fa.sources ? distance : 0.0f, directionDot, fa.sources ? invDirectionDot : 0.0f, cone);
			}

			// ---- EvaluateDistance (SpatialManager.h:401)
			if (fa.distAtt)
			{
				float a = 1.0f;
				const uint32_t k = fa.attIndex ? fa.attIndex[i] : 0u;
				if (k < fa.numAtt)
				{
					const AttFunctionDev f = fa.att[k];
					a = f.kind == 0 ? evaluate_model(f, distance) : evaluate_points(fa.attPoints + size_t(f.firstPoint) * 3, f.numPoints, distance);
				}
				fa.distAtt[i] = a;
			}

			// ---- pan parameters: spread from the object's size (SpatialManager.h:407-412), sanitised direction (PanningService.h:480-495)
			if (fa.pan)
			{
				float focus = fa.focusConst, spread = fa.spreadConst;
				if (fa.focusSpread) { focus = fa.focusSpread[2 * size_t(i)]; spread = fa.focusSpread[2 * size_t(i) + 1]; }
				if (fa.sizes)
				{
					const float size = fa.sizes[i];
					if (size >= 0.0f)
					{
						if (distance <= 0.0f) spread = 1.0f;
						else if (fa.spreadFlavour == 0) spread = libm1(atan(static_cast<double>((0.5f * size) / distance))) * 0.318309886f;
						else spread = libm1(atan(static_cast<double>((size * 0.5f) / distance))) * (0.318309886f * 2.0f);
					}
				}
				F3 d = dir;
				const float dy = absbits(dir.y) - 1.0f;
				if (absbits(dy) <= 1e-6f) d = { fa.fallbackDir[0], fa.fallbackDir[1], fa.fallbackDir[2] };
				d.y = copysignf(d.y, dir.y);
				float* o = fa.pan + size_t(i) * fa.panStride;
				o[0] = d.x; o[1] = d.y; o[2] = d.z; o[3] = focus; o[4] = spread;
			}

			// ---- air absorption -> band gains (AirAbsorption.h:212-234); SSE2 build: Math::FMA = multiply, then add
			if (fa.airDb || fa.bandGains)
			{
				float loss[4], gain[4];
				#pragma unroll
				for (int b = 0; b < 4; ++b)
				{
					const float lin = distance * fa.air.lowerThirdDb[b] + fa.air.linFactor[b];
					const float centre = distance * fa.air.centerDb[b];
					loss[b] = (fa.air.hasKnee[b] && distance >= fa.air.distanceToKnee[b]) ? lin : centre;
					gain[b] = exp_simd_lane((2.302585092994046f * 0.05f) * (0.0f - loss[b]));
				}
				if (fa.airDb) reinterpret_cast<float4*>(fa.airDb)[i] = make_float4(loss[0], loss[1], loss[2], loss[3]);
				if (fa.bandGains) reinterpret_cast<float4*>(fa.bandGains)[i] = make_float4(gain[0], gain[1], gain[2], gain[3]);
			}
			if (fa.delay) fa.delay[i] = distance * fa.invSpeedOfSound;
		}
	}

	cudaError_t LaunchFrontend(const FrontendArgs& fa, int numSMs, cudaStream_t st)
	{
		if (fa.n == 0) return cudaSuccess;
		const uint32_t want = (fa.n + 127u) / 128u;
		const uint32_t cap = static_cast<uint32_t>(numSMs > 0 ? numSMs : 148) * 16u;   // grid-stride beyond 16 CTAs per SM
		k_scene_frontend<<<want < cap ? want : cap, 128, 0, st>>>(fa);
		return cudaGetLastError();
	}
}
