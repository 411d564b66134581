// Device-side restatement of the reference's per-query math, lane by lane.
// The reference evaluates these with 4-wide SSE2 vectors; every lane is independent, so one CUDA thread
// evaluating one lane with the same sequence of individually rounded IEEE-754 operations produces the
// same bits. This translation unit MUST be compiled with -fmad=false (no implicit contraction) and without
// --use_fast_math; fused operations appear only through explicit __fmaf_rn in throughput mode.
// Reference paths: inc/ = Spatialization/include/JPLSpatial/.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace jplg { namespace dev {

struct F3 { float x, y, z; };
struct Q4 { float x, y, z, w; };
struct B3 { F3 X, Y, Z; };

__device__ __forceinline__ uint32_t f2u(float f) { return __float_as_uint(f); }
__device__ __forceinline__ float u2f(uint32_t u) { return __uint_as_float(u); }
__device__ __forceinline__ float absbits(float v) { return u2f(f2u(v) & 0x7fffffffu); } // abs() (inc/Math/SIMD.h:1830-1839)
__device__ __forceinline__ float minps(float a, float b) { return a < b ? a : b; }      // _mm_min_ps lane
__device__ __forceinline__ float maxps(float a, float b) { return a > b ? a : b; }      // _mm_max_ps lane
__device__ __forceinline__ float clampps(float v, float lo, float hi) { return maxps(minps(v, hi), lo); } // :1841-1844
__device__ __forceinline__ float sign2simd(float v) { return v < 0.0f ? -1.0f : 1.0f; } // Math::Sign2(simd) (:1788-1791)
__device__ __forceinline__ float sign2scalar(float v) { return copysignf(1.0f, v); }    // Math::Sign2 scalar (inc/Math/Math.h:131-137)
__device__ __forceinline__ int32_t cvtt(float v) { return __float2int_rz(v); }           // _mm_cvttps_epi32 lane (valid range only)

// _mm_rsqrt_ps model: 2048-entry table extracted from the host CPU + exact exponent scaling (see jplg_host.cpp)
__device__ __forceinline__ float rsqrt_est(const uint32_t* __restrict__ table, float x)
{
	const uint32_t u = f2u(x);
	const uint32_t E = (u >> 23) & 0xffu;
	const uint32_t par = E & 1u;
	const uint32_t idx = (par << 10) | ((u >> 13) & 0x3ffu);
	const int32_t half = (static_cast<int32_t>(E) - (par ? 127 : 128)) / 2;
	return u2f(table[idx] - (static_cast<uint32_t>(half) << 23));
}
// Math::InvSqrtFast = estimate + 1 Newton step, SSE2 flavour (inc/Math/SIMD.h:1709-1735)
__device__ __forceinline__ float invsqrt_fast(const uint32_t* __restrict__ table, float x)
{
	const float y = rsqrt_est(table, x);
	const float y2 = y * y;
	const float term = 1.5f - ((0.5f * x) * y2);
	return y * term;
}
__device__ __forceinline__ float invsqrt_exact(float x) { return 1.0f / sqrtf(x); } // Math::InvSqrt scalar runtime path

// Math::SinCos(simd) (inc/Math/SIMDMath.h:488-534)
__device__ __forceinline__ void sincos_simd(float xin, float& outS, float& outC)
{
	uint32_t sinSign = f2u(xin) & 0x80000000u;
	float x = absbits(xin);
	const uint32_t q = static_cast<uint32_t>(cvtt(1.27323954473516f * x + 1.0f)) & ~1u;
	const float fq = static_cast<float>(static_cast<int32_t>(q));
	x = fq * -0.78515625f + x;
	x = fq * -2.4187564849853515625e-4f + x;
	x = fq * -3.77489497744594108e-8f + x;
	const float x2 = x * x;
	float tc = 2.443315711809948e-5f * x2 + -1.388731625493765e-3f;
	tc = tc * x2 + 4.166664568298827e-2f;
	tc = tc * (x2 * x2) + ((0.0f - (0.5f * x2)) + 1.0f);
	float ts = -1.9515295891e-4f * x2 + 8.3321608736e-3f;
	ts = ts * x2 + -1.6666654611e-1f;
	ts = ts * (x2 * x) + x;
	const bool m = (q & 2u) == 0u;
	const float s = m ? ts : tc;
	const float c = m ? tc : ts;
	const uint32_t bit1 = q << 30;
	const uint32_t bit2 = (q << 29) & 0x80000000u;
	sinSign ^= bit2;
	outS = u2f(f2u(s) ^ sinSign);
	outC = u2f(f2u(c) ^ (bit1 ^ bit2));
}

// asin / acos (inc/Math/SIMDMath.h:667-714)
__device__ __forceinline__ float asin_simd(float in)
{
	const uint32_t sign = f2u(in) & 0x80000000u;
	float a = absbits(in);
	a = minps(a, 1.0f);
	const bool o = a < 0.5f;
	const float x2 = o ? (a * a) : ((1.0f - a) * 0.5f);
	const float x = o ? a : sqrtf(x2);
	float y = 0.4197454825e-1f;
	y = y * x2 + 0.2424046025e-1f;
	y = y * x2 + 0.4547423869e-1f;
	y = y * x2 + 0.7495029271e-1f;
	y = y * x2 + 0.1666677296e+0f;
	y = y * (x * x2) + x;
	const float halfPi = 3.14159265358979323846f * 0.5f;
	const float r = o ? y : (halfPi - (y + y));
	return u2f(f2u(r) ^ sign);
}
__device__ __forceinline__ float acos_simd(float in) { return (3.14159265358979323846f * 0.5f) - asin_simd(in); }

__device__ __forceinline__ float dot3(const F3& a, const F3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ F3 cross3(const F3& a, const F3& b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }

// Math::Slerp(Vec3Pack&, const Vec3Pack&, simd t), one lane (inc/Math/Vec3Math.h:206-307)
__device__ __forceinline__ F3 slerp_simd(const uint32_t* __restrict__ rsq, F3 v0, const F3& v1, float t)
{
	const float EPS = 1.1920928955078125e-7f * 100.0f;
	const float oneMinusEps = 1.0f - EPS;
	const float minusOnePlusEps = -1.0f + EPS;
	const float dot = clampps(dot3(v0, v1), -1.0f, 1.0f);
	const bool isParallel = dot > oneMinusEps;
	const bool isOpposite = dot < minusOnePlusEps;

	// The reference evaluates all three cases in every SIMD lane and selects; a lane's result only depends on the case it selects, so the
	// unselected ones are skipped here (bit-identical, and the parallel / opposite cases are rare)
	F3 mulA = v0, addC;
	float mulB;
	if (isParallel)
	{
		mulA = F3{ v1.x - v0.x, v1.y - v0.y, v1.z - v0.z };
		mulB = t;
		addC = v0;
	}
	else if (isOpposite)
	{
		const float dAxisX = v0.x * 1.0f + v0.y * 0.0f + v0.z * 0.0f;
		const bool isClose = absbits(dAxisX) > oneMinusEps;
		const F3 axis{ isClose ? 0.0f : 1.0f, isClose ? 1.0f : 0.0f, 0.0f };
		const float dAxis = dot3(v0, axis);
		F3 proj{ axis.x - v0.x * dAxis, axis.y - v0.y * dAxis, axis.z - v0.z * dAxis };
		const float ilp = invsqrt_fast(rsq, dot3(proj, proj));
		proj.x *= ilp; proj.y *= ilp; proj.z *= ilp;
		float sinA, cosA;
		sincos_simd(t * 3.14159265358979323846f, sinA, cosA);
		mulB = cosA;
		addC = F3{ proj.x * sinA, proj.y * sinA, proj.z * sinA };
	}
	else
	{
		const float invSinTheta = invsqrt_fast(rsq, 1.0f - dot * dot);
		const float theta = acos_simd(dot);
		float sinT, cosT;
		sincos_simd(t * theta, sinT, cosT);
		const float w1 = sinT * invSinTheta;
		mulB = cosT - dot * w1;
		addC = F3{ v1.x * w1, v1.y * w1, v1.z * w1 };
	}
	F3 r{ mulA.x * mulB + addC.x, mulA.y * mulB + addC.y, mulA.z * mulB + addC.z };
	const float il = invsqrt_fast(rsq, dot3(r, r));
	r.x *= il; r.y *= il; r.z *= il;
	return r;
}

// Octahedron16Bit::Encode(simd,simd,simd), one lane (inc/Math/DirectionEncoding.h:346-386)
__device__ __forceinline__ uint32_t oct_encode_simd(float x, float y, float z)
{
	const float L1 = absbits(x) + absbits(y) + absbits(z);
	const float inv = 1.0f / L1;
	float px = x * inv;
	float py = y * inv;
	if (z < 0.0f)
	{
		const float t = px;
		px = (1.0f - absbits(py)) * sign2simd(t);
		py = (1.0f - absbits(t)) * sign2simd(py);
	}
	const float h = 0.5f / 255.0f;
	const float lo = (0.0f - 1.0f) + h, hi = 1.0f - h;
	px = clampps(px, lo, hi);
	py = clampps(py, lo, hi);
	const float v0 = px * 0.5f + 0.5f;
	const float v1 = py * 0.5f + 0.5f;
	const uint32_t dx = static_cast<uint32_t>(cvtt(v0 * 255.0f)); // floor(SSE2) == trunc for the positive range reached here
	const uint32_t dy = static_cast<uint32_t>(cvtt(v1 * 255.0f));
	return ((dy << 8) | dx) & 0xffffu;
}
// Octahedron16Bit::Encode(Vec3) scalar (inc/Math/DirectionEncoding.h:292-343)
__device__ __forceinline__ uint32_t oct_encode_scalar(float x, float y, float z)
{
	const float L1 = fabsf(x) + fabsf(y) + fabsf(z);
	const float inv = 1.0f / L1;
	float px = x * inv;
	float py = y * inv;
	if (z < 0.0f)
	{
		const float t = px;
		px = (1.0f - fabsf(py)) * sign2scalar(t);
		py = (1.0f - fabsf(t)) * sign2scalar(py);
	}
	const float h = 0.5f / 255.0f;
	const float lo = -1.0f + h, hi = 1.0f - h;
	px = (px < lo) ? lo : ((hi < px) ? hi : px); // std::clamp
	py = (py < lo) ? lo : ((hi < py) ? hi : py);
	const float v0 = px * 0.5f + 0.5f;
	const float v1 = py * 0.5f + 0.5f;
	const uint32_t dx = static_cast<uint32_t>(static_cast<int32_t>(floorf(v0 * 255.0f)));
	const uint32_t dy = static_cast<uint32_t>(static_cast<int32_t>(floorf(v1 * 255.0f)));
	return ((dy << 8) | dx) & 0xffffu;
}

// ToDiamond scalar (inc/Math/DirectionEncoding.h:197-209) and LUT2D::CartesianToLUTPosition(float,float) (inc/Panning/VBAPLUT2D.h:346-351)
__device__ __forceinline__ uint32_t lut2d_pos_scalar(float x, float y)
{
	x = x / (fabsf(x) + fabsf(y));
	const float q = sign2scalar(y) * 0.25f;
	const float p = (-q) * x + (0.5f + q);
	return static_cast<uint32_t>(__float2int_rz(1024.0f * p + 0.5f)) & 1023u;
}
// SIMD flavour (VBAPLUT2D.h:376-385, DirectionEncoding.h:231-243), one lane
__device__ __forceinline__ uint32_t lut2d_pos_simd(const uint32_t* __restrict__ rsq, float x, float y)
{
	const float il = invsqrt_fast(rsq, x * x + y * y);
	float xn = x * il;
	const float yn = y * il;
	xn = xn / (absbits(xn) + absbits(yn));
	const float q = sign2simd(yn) * 0.25f;
	const float p = (0.0f - q) * xn + (0.5f + q);
	return static_cast<uint32_t>(cvtt(p * 1024.0f + 0.5f)) & 1023u;
}

// Math::QuatFromUpAndForward(up, forward) = Quat::FromBasis(Basis::FromUpAndForward(up, forward)); QuatLookAt(direction, up) is the
// same call with forward = direction (inc/Math/MinimalQuat.h:84-144, inc/Math/MinimalBasis.h:60-71); scalar IEEE
__device__ __forceinline__ Q4 quat_from_up_and_forward(const F3& up, const F3& forward)
{
	F3 Y = up;
	const F3 Z = forward;
	const float dYZ = dot3(Y, Z);
	const float ad = dYZ < 0 ? -dYZ : dYZ;
	const float diff = ad - 1.0f;
	if ((diff < 0 ? -diff : diff) <= 1e-6f)
	{
		const float az = Z.z < 0 ? -Z.z : Z.z;
		Y = (az < 0.999f) ? F3{ 0, 0, 1 } : F3{ 1, 0, 0 };
	}
	F3 X = cross3(Y, Z);
	{
		const float il = invsqrt_exact(dot3(X, X));
		X.x *= il; X.y *= il; X.z *= il;
	}
	Y = cross3(Z, X);
	const float r00 = X.x, r01 = Y.x, r02 = Z.x;
	const float r10 = X.y, r11 = Y.y, r12 = Z.y;
	const float r20 = X.z, r21 = Y.z, r22 = Z.z;
	const float tr = r00 + r11 + r22;
	float qw, qx, qy, qz;
	if (tr > 0.0f)
	{
		const float s = sqrtf(tr + 1.0f) * 2.0f;
		const float invS = 1.0f / s;
		qw = 0.25f * s; qx = (r21 - r12) * invS; qy = (r02 - r20) * invS; qz = (r10 - r01) * invS;
	}
	else if (r00 > r11 && r00 > r22)
	{
		const float s = sqrtf(1.0f + r00 - r11 - r22) * 2.0f;
		const float invS = 1.0f / s;
		qw = (r21 - r12) * invS; qx = 0.25f * s; qy = (r01 + r10) * invS; qz = (r02 + r20) * invS;
	}
	else if (r11 > r22)
	{
		const float s = sqrtf(1.0f + r11 - r00 - r22) * 2.0f;
		const float invS = 1.0f / s;
		qw = (r02 - r20) * invS; qx = (r01 + r10) * invS; qy = 0.25f * s; qz = (r12 + r21) * invS;
	}
	else
	{
		const float s = sqrtf(1.0f + r22 - r00 - r11) * 2.0f;
		const float invS = 1.0f / s;
		qw = (r10 - r01) * invS; qx = (r02 + r20) * invS; qy = (r12 + r21) * invS; qz = 0.25f * s;
	}
	return { qx, qy, qz, qw };
}

// Quat operator* (inc/Math/MinimalQuat.h:147-153)
__device__ __forceinline__ Q4 quat_mul(const Q4& a, const Q4& b)
{
	const F3 av{ a.x, a.y, a.z }, bv{ b.x, b.y, b.z };
	const F3 c = cross3(av, bv);
	Q4 r;
	r.x = (bv.x * a.w + av.x * b.w) + c.x;
	r.y = (bv.y * a.w + av.y * b.w) + c.y;
	r.z = (bv.z * a.w + av.z * b.w) + c.z;
	r.w = a.w * b.w - dot3(av, bv);
	return r;
}
// Quat::ToBasis (inc/Math/MinimalQuat.h:212-222)
__device__ __forceinline__ B3 quat_to_basis(const Q4& q)
{
	const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
	const float xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
	const float wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
	B3 b;
	b.X = { 1.0f - 2.0f * (yy + zz), 2.0f * (xy + wz), 2.0f * (xz - wy) };
	b.Y = { 2.0f * (xy - wz), 1.0f - 2.0f * (xx + zz), 2.0f * (yz + wx) };
	b.Z = { 2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (xx + yy) };
	return b;
}
// Basis::Transform(simd&,simd&,simd&), one lane (inc/Math/MinimalBasis.h:117-128)
__device__ __forceinline__ F3 basis_transform_simd(const B3& b, const F3& v)
{
	F3 r;
	r.x = b.X.x * v.x + (b.Y.x * v.y + (b.Z.x * v.z));
	r.y = b.X.y * v.x + (b.Y.y * v.y + (b.Z.y * v.z));
	r.z = b.X.z * v.x + (b.Y.z * v.y + (b.Z.z * v.z));
	return r;
}

// libm sinf/cosf as the reference's scalar Math::SinCos calls them (inc/Math/Math.h:164-168). glibc's sinf/cosf
// evaluate in double precision with < 1 ulp(float) error; rounding an (almost) exact double result reproduces them
// except in the measure-zero band where the true value sits within ~1e-8 ulp of a float rounding boundary.
__device__ __forceinline__ void sincos_libm(float x, float& s, float& c)
{
	double ds, dc;
	sincos(static_cast<double>(x), &ds, &dc);
	s = static_cast<float>(ds);
	c = static_cast<float>(dc);
}

}} // namespace jplg::dev
