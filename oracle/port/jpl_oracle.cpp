/* TEST INFRASTRUCTURE — NOT PRODUCT CODE. See jpl_oracle.h for the status header.
 *
 * Plain CPU restatement of JPL Spatial's hot path, written from the reference's algorithm. It includes no reference
 * header and no product code. The reference evaluates most of this with 4-wide SSE2 vectors ("packs"); every pack lane
 * is independent, so the restatement runs the same sequence of individually rounded IEEE-754 operations lane by lane.
 * Compile with -ffp-contract=off (no FMA contraction: the parity target is the reference's SSE2 baseline build, where
 * Math::FMA(a,b,c) is a multiply followed by an add).
 *
 * Citations: "inc/" = /root/reference/Spatialization/include/JPLSpatial/, "src/" = .../src/Spatialization/.
 */
#include "jpl_oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>
#include <xmmintrin.h>

namespace
{
	constexpr float PI = 3.14159265358979323846f;
	constexpr float TWO_PI = 2.0f * PI;
	constexpr float SQRT2 = 1.41421356237309504880f;

	inline uint32_t bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
	inline float fromBits(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
	inline float fabsBits(float v) { return fromBits(bits(v) & 0x7fffffffu); }                 // simd abs: and-not of the sign (inc/Math/SIMD.h:1830-1839)
	inline float laneMin(float a, float b) { return a < b ? a : b; }                            // MINPS lane
	inline float laneMax(float a, float b) { return a > b ? a : b; }                            // MAXPS lane
	inline float laneClamp(float v, float lo, float hi) { return laneMax(laneMin(v, hi), lo); } // inc/Math/SIMD.h:1838-1841
	inline float signSimd(float v) { return v < 0.0f ? -1.0f : 1.0f; }                          // Math::Sign2(simd) (SIMD.h:1788-1791): -0 -> +1
	inline float signScalar(float v) { return std::copysign(1.0f, v); }                         // Math::Sign2 scalar (inc/Math/Math.h:131-137)
	inline int32_t truncInt(float v) { return _mm_cvttss_si32(_mm_set_ss(v)); }                 // CVTTPS2DQ lane (to_mask, SIMD.h:990)
	inline float rsqrtps(float x) { return _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(x))); }        // RSQRTPS lane (SIMD.h:1709)
	inline float invSqrtExact(float x) { return 1.0f / std::sqrt(x); }                          // Math::InvSqrt scalar (Math.h:283-289)

	/* Math::InvSqrtFast: estimate + one Newton step, SSE2 flavour (inc/Math/SIMD.h:1709-1735) */
	inline float invSqrtNR1(float x)
	{
		const float y = rsqrtps(x);
		const float yy = y * y;
		return y * (1.5f - ((0.5f * x) * yy));
	}
	inline float invSqrtNR2(float x) /* Math::InvSqrt(simd): two steps */
	{
		float y = rsqrtps(x);
		for (int k = 0; k < 2; ++k) { const float yy = y * y; y = y * (1.5f - ((0.5f * x) * yy)); }
		return y;
	}

	/* Math::SinCos(simd) (inc/Math/SIMDMath.h:449-535): Cephes-style, 3-term Cody-Waite reduction by pi/4 octants */
	void sinCosLane(float xin, float& s, float& c)
	{
		uint32_t signS = bits(xin) & 0x80000000u;
		float x = fabsBits(xin);
		const uint32_t q = uint32_t(truncInt(1.27323954473516f * x + 1.0f)) & ~1u;
		const float qf = float(int32_t(q));
		x = qf * -0.78515625f + x;
		x = qf * -2.4187564849853515625e-4f + x;
		x = qf * -3.77489497744594108e-8f + x;
		const float z = x * x;
		float pc = 2.443315711809948e-5f * z + -1.388731625493765e-3f;
		pc = pc * z + 4.166664568298827e-2f;
		pc = pc * (z * z) + ((0.0f - (0.5f * z)) + 1.0f);
		float ps = -1.9515295891e-4f * z + 8.3321608736e-3f;
		ps = ps * z + -1.6666654611e-1f;
		ps = ps * (z * x) + x;
		const bool keep = (q & 2u) == 0u;
		const float sv = keep ? ps : pc, cv = keep ? pc : ps;
		const uint32_t b1 = q << 30, b2 = (q << 29) & 0x80000000u;
		signS ^= b2;
		s = fromBits(bits(sv) ^ signS);
		c = fromBits(bits(cv) ^ (b1 ^ b2));
	}

	/* Math::asin / acos (simd) (inc/Math/SIMDMath.h:667-714) */
	float asinLane(float in)
	{
		const uint32_t sign = bits(in) & 0x80000000u;
		float a = laneMin(fabsBits(in), 1.0f);
		const bool small = a < 0.5f;
		const float z = small ? a * a : (1.0f - a) * 0.5f;
		const float x = small ? a : std::sqrt(z);
		float y = 0.4197454825e-1f;
		y = y * z + 0.2424046025e-1f;
		y = y * z + 0.4547423869e-1f;
		y = y * z + 0.7495029271e-1f;
		y = y * z + 0.1666677296e+0f;
		y = y * (x * z) + x;
		const float r = small ? y : (PI * 0.5f) - (y + y);
		return fromBits(bits(r) ^ sign);
	}
	inline float acosLane(float in) { return (PI * 0.5f) - asinLane(in); }

	/* round() SSE2 path (inc/Math/SIMD.h:1930-1937) and Math::tan (simd) (inc/Math/SIMDMath.h:617-665) */
	float tanLane(float x)
	{
		const float half = fromBits(bits(0.5f) | (bits(x * 0.63661977236758134308f) & 0x80000000u));
		const float n = float(truncInt(x * 0.63661977236758134308f + half));
		float r = n * -1.5703125f + x;
		r = n * -0.0004837512969970703125f + r;
		r = n * -7.549789948768648e-8f + r;
		const float z = r * r;
		float y = 9.38540185543e-3f * z + 3.11992232697e-3f;
		y = y * z + 2.44301354525e-2f;
		y = y * z + 5.34112807005e-2f;
		y = y * z + 1.33387994085e-1f;
		y = y * z + 3.33331568548e-1f;
		y = y * (z * r) + r;
		return (truncInt(n) & 1) ? 0.0f - (1.0f / y) : y;
	}

	struct Vec3 { float x, y, z; };
	struct Quat { float x, y, z, w; };
	struct Basis { Vec3 X, Y, Z; };
	inline float dot(const Vec3& a, const Vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
	inline Vec3 cross(const Vec3& a, const Vec3& b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
	inline Vec3 normalizedExact(Vec3 v) { const float i = invSqrtExact(dot(v, v)); return { v.x * i, v.y * i, v.z * i }; }

	/* Octahedron16Bit::Encode, simd flavour (inc/Math/DirectionEncoding.h:346-386) */
	uint32_t octEncodeSimd(float x, float y, float z)
	{
		const float n = (fabsBits(x) + fabsBits(y)) + fabsBits(z);
		const float inv = 1.0f / n;
		float px = x * inv, py = y * inv;
		if (z < 0.0f)
		{
			const float ox = px;
			px = (1.0f - fabsBits(py)) * signSimd(ox);
			py = (1.0f - fabsBits(ox)) * signSimd(py);
		}
		const float h = 0.5f / 255.0f;
		px = laneClamp(px, (0.0f - 1.0f) + h, 1.0f - h);
		py = laneClamp(py, (0.0f - 1.0f) + h, 1.0f - h);
		const uint32_t dx = uint32_t(truncInt((px * 0.5f + 0.5f) * 255.0f)); /* SSE2 floor == trunc on this positive range (SIMD.h:1873-1877) */
		const uint32_t dy = uint32_t(truncInt((py * 0.5f + 0.5f) * 255.0f));
		return ((dy << 8) | dx) & 0xffffu;
	}
	/* Octahedron16Bit::Encode, scalar flavour (:292-343) */
	uint32_t octEncodeScalar(float x, float y, float z)
	{
		const float n = std::fabs(x) + std::fabs(y) + std::fabs(z);
		const float inv = 1.0f / n;
		float px = x * inv, py = y * inv;
		if (z < 0.0f)
		{
			const float ox = px;
			px = (1.0f - std::fabs(py)) * signScalar(ox);
			py = (1.0f - std::fabs(ox)) * signScalar(py);
		}
		const float h = 0.5f / 255.0f;
		px = std::clamp(px, -1.0f + h, 1.0f - h);
		py = std::clamp(py, -1.0f + h, 1.0f - h);
		const uint32_t dx = uint32_t(int32_t(std::floor((px * 0.5f + 0.5f) * 255.0f)));
		const uint32_t dy = uint32_t(int32_t(std::floor((py * 0.5f + 0.5f) * 255.0f)));
		return ((dy << 8) | dx) & 0xffffu;
	}
	/* Octahedron16Bit::Decode, scalar (:389-436): cell centre, fold the lower hemisphere, exact normalise */
	Vec3 octDecodeScalar(uint32_t code)
	{
		const uint32_t dx = std::min(code & 255u, 254u), dy = std::min((code >> 8) & 255u, 254u);
		const float k = 1.0f / 255.0f;
		const float px = -1.0f + 2.0f * (float(dx) + 0.5f) * k;
		const float py = -1.0f + 2.0f * (float(dy) + 0.5f) * k;
		Vec3 d{ px, py, 1.0f - std::fabs(px) - std::fabs(py) };
		const float t = std::max(-d.z, 0.0f);
		d.x += d.x > 0.0f ? -t : t;
		d.y += d.y > 0.0f ? -t : t;
		return normalizedExact(d);
	}

	/* FromDiamond (:212-227) / ToDiamond scalar (:197-209) / simd (:231-243) */
	void fromDiamond(float p, float& ox, float& oy)
	{
		const float sg = signScalar(p - 0.5f);
		ox = (-sg) * (4.0f * p) + (sg * 2.0f + 1.0f);
		oy = (-sg) * std::fabs(ox) + sg;
		const float i = invSqrtExact(ox * ox + oy * oy);
		ox *= i; oy *= i;
	}
	float toDiamondScalar(float x, float y)
	{
		x = x / (std::fabs(x) + std::fabs(y));
		const float q = signScalar(y) * 0.25f;
		return (-q) * x + (0.5f + q);
	}
	/* LUT2D::CartesianToLUTPosition (inc/Panning/VBAPLUT2D.h:346-351 scalar, :376-385 simd) */
	uint32_t lut2dIndexScalar(float x, float y) { return uint32_t(int(1024.0f * toDiamondScalar(x, y) + 0.5f)) & 1023u; }
	uint32_t lut2dIndexSimd(float x, float y)
	{
		const float i = invSqrtNR1(x * x + y * y);
		float xn = x * i; const float yn = y * i;
		xn = xn / (fabsBits(xn) + fabsBits(yn));
		const float q = signSimd(yn) * 0.25f;
		const float p = (0.0f - q) * xn + (0.5f + q);
		return uint32_t(truncInt(p * 1024.0f + 0.5f)) & 1023u;
	}

	/* Math::Slerp(Vec3Pack&, const Vec3Pack&, simd) one lane (inc/Math/Vec3Math.h:206-307): all three branches evaluated, then selected */
	Vec3 slerpLane(Vec3 a, const Vec3& b, float t)
	{
		const float eps = 100.0f * std::numeric_limits<float>::epsilon();
		const float d = laneClamp(dot(a, b), -1.0f, 1.0f);
		const bool parallel = d > 1.0f - eps, opposite = d < -1.0f + eps;
		/* opposite: rotate about an axis orthogonal to a */
		const bool nearX = fabsBits(a.x * 1.0f + a.y * 0.0f + a.z * 0.0f) > 1.0f - eps;
		const Vec3 ax{ nearX ? 0.0f : 1.0f, nearX ? 1.0f : 0.0f, 0.0f };
		const float da = dot(a, ax);
		Vec3 pr{ ax.x - a.x * da, ax.y - a.y * da, ax.z - a.z * da };
		const float ip = invSqrtNR1(dot(pr, pr));
		pr = { pr.x * ip, pr.y * ip, pr.z * ip };
		float sA, cA; sinCosLane(t * PI, sA, cA);
		/* general */
		const float invSin = invSqrtNR1(1.0f - d * d);
		const float theta = acosLane(d);
		float sT, cT; sinCosLane(t * theta, sT, cT);
		const float w1 = sT * invSin, w0 = cT - d * w1;
		const Vec3 m = parallel ? Vec3{ b.x - a.x, b.y - a.y, b.z - a.z } : a;
		const float k = parallel ? t : (opposite ? cA : w0);
		const Vec3 add = parallel ? a : (opposite ? Vec3{ pr.x * sA, pr.y * sA, pr.z * sA } : Vec3{ b.x * w1, b.y * w1, b.z * w1 });
		Vec3 r{ m.x * k + add.x, m.y * k + add.y, m.z * k + add.z };
		const float ir = invSqrtNR1(dot(r, r));
		return { r.x * ir, r.y * ir, r.z * ir };
	}

	/* Math::QuatLookAt(dir, worldUp) = Quat::FromBasis(Basis::FromUpAndForward) (inc/Math/MinimalQuat.h:84-144, MinimalBasis.h:60-71) */
	Quat quatFromUpAndForward(Vec3 up, const Vec3& fwd);
	Quat quatLookAt(const Vec3& fwd) { return quatFromUpAndForward(Vec3{ 0, 1, 0 }, fwd); }
	/* Math::QuatFromUpAndForward(up, forward) (inc/Math/MinimalQuat.h:136-139) */
	Quat quatFromUpAndForward(Vec3 up, const Vec3& fwd)
	{
		if (std::fabs(std::fabs(dot(up, fwd)) - 1.0f) <= 1e-6f) up = std::fabs(fwd.z) < 0.999f ? Vec3{ 0, 0, 1 } : Vec3{ 1, 0, 0 };
		const Vec3 X = normalizedExact(cross(up, fwd));
		const Vec3 Y = cross(fwd, X);
		const Vec3 Z = fwd;
		const float m00 = X.x, m01 = Y.x, m02 = Z.x, m10 = X.y, m11 = Y.y, m12 = Z.y, m20 = X.z, m21 = Y.z, m22 = Z.z;
		const float tr = m00 + m11 + m22;
		Quat q;
		if (tr > 0.0f)
		{
			const float s = std::sqrt(tr + 1.0f) * 2.0f, is = 1.0f / s;
			q = { (m21 - m12) * is, (m02 - m20) * is, (m10 - m01) * is, 0.25f * s };
		}
		else if (m00 > m11 && m00 > m22)
		{
			const float s = std::sqrt(1.0f + m00 - m11 - m22) * 2.0f, is = 1.0f / s;
			q = { 0.25f * s, (m01 + m10) * is, (m02 + m20) * is, (m21 - m12) * is };
		}
		else if (m11 > m22)
		{
			const float s = std::sqrt(1.0f + m11 - m00 - m22) * 2.0f, is = 1.0f / s;
			q = { (m01 + m10) * is, 0.25f * s, (m12 + m21) * is, (m02 - m20) * is };
		}
		else
		{
			const float s = std::sqrt(1.0f + m22 - m00 - m11) * 2.0f, is = 1.0f / s;
			q = { (m02 + m20) * is, (m12 + m21) * is, 0.25f * s, (m10 - m01) * is };
		}
		return q;
	}
	Quat quatMul(const Quat& a, const Quat& b) /* inc/Math/MinimalQuat.h:147-153 */
	{
		const Vec3 av{ a.x, a.y, a.z }, bv{ b.x, b.y, b.z }, c = cross(av, bv);
		return { (bv.x * a.w + av.x * b.w) + c.x, (bv.y * a.w + av.y * b.w) + c.y, (bv.z * a.w + av.z * b.w) + c.z, a.w * b.w - dot(av, bv) };
	}
	Basis quatToBasis(const Quat& q) /* :212-222 */
	{
		const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z, xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z, wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
		return { { 1.0f - 2.0f * (yy + zz), 2.0f * (xy + wz), 2.0f * (xz - wy) },
				 { 2.0f * (xy - wz), 1.0f - 2.0f * (xx + zz), 2.0f * (yz + wx) },
				 { 2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (xx + yy) } };
	}
	Vec3 basisTransformSimd(const Basis& b, const Vec3& v) /* inc/Math/MinimalBasis.h:117-128 */
	{
		return { b.X.x * v.x + (b.Y.x * v.y + (b.Z.x * v.z)), b.X.y * v.x + (b.Y.y * v.y + (b.Z.y * v.z)), b.X.z * v.x + (b.Y.z * v.y + (b.Z.z * v.z)) };
	}

	/* ---- channel tables (inc/ChannelMap.h:38-79; inc/Panning/PannerBase.h:108-187) */
	enum : uint32_t { LFE = 1u << 3, FRONT_CENTER = 1u << 2, TOP_CENTER = 1u << 13, FIRST_TOP = 1u << 13, ALL = (1u << 22) - 1 };
	/* indexed by EChannel bit position: FL FR FC LFE SL SR FLC FRC BL BR BC WL WR | TC TFL TFC TFR TSL TSR TBL TBC TBR */
	const Vec3 kSpeakerDir[22] = {
		{ -0.7071f, 0, -0.7071f }, { 0.7071f, 0, -0.7071f }, { 0, 0, -1 }, { 0, 0, -1 }, { -1, 0, 0 }, { 1, 0, 0 },
		{ -0.3162f, 0, -0.9487f }, { 0.3162f, 0, -0.9487f }, { -0.7071f, 0, 0.7071f }, { 0.7071f, 0, 0.7071f }, { 0, 0, 1 },
		{ -0.866f, 0, -0.5f }, { 0.866f, 0, -0.5f }, { 0, 1, 0 },
		{ -0.5774f, 0.5774f, -0.5774f }, { 0, 0.7071f, -0.7071f }, { 0.5774f, 0.5774f, -0.5774f }, { -0.7071f, 0.7071f, 0 }, { 0.7071f, 0.7071f, 0 },
		{ -0.5774f, 0.5774f, 0.5774f }, { 0, 0.7071f, 0.7071f }, { 0.5774f, 0.5774f, 0.5774f } };
	const float kSpeakerAngle[22] = { -0.785398f, 0.785398f, 0, 0, -1.5708f, 1.5708f, -0.321719f, 0.321719f, -2.35619f, 2.35619f, 3.14159f,
		-1.0472f, 1.0472f, 3.14159f, -0.785398f, 0, 0.785398f, -1.5708f, 1.5708f, -2.35619f, 3.14159f, 2.35619f };
	/* cChannelPlaneRatios (inc/ChannelConversion.h:43-78) */
	const float kPlane[22][6] = {
		{ .5f, 0, .5f, 0, 0, 0 }, { 0, .5f, .5f, 0, 0, 0 }, { 0, 0, 1, 0, 0, 0 }, { 0, 0, 0, 0, 0, 0 }, { 1, 0, 0, 0, 0, 0 }, { 0, 1, 0, 0, 0, 0 },
		{ .25f, 0, .75f, 0, 0, 0 }, { 0, .25f, .75f, 0, 0, 0 }, { .5f, 0, 0, .5f, 0, 0 }, { 0, .5f, 0, .5f, 0, 0 }, { 0, 0, 0, 1, 0, 0 },
		{ 1, 0, 0, 0, 0, 0 }, { 0, 1, 0, 0, 0, 0 }, { 0, 0, 0, 0, 0, 1 }, { .33f, 0, .33f, 0, 0, .34f }, { 0, 0, .5f, 0, 0, .5f }, { 0, .33f, .33f, 0, 0, .34f },
		{ .5f, 0, 0, 0, 0, .5f }, { 0, .5f, 0, 0, 0, .5f }, { .33f, 0, 0, .33f, 0, .34f }, { 0, 0, 0, .5f, 0, .5f }, { 0, .33f, 0, .33f, 0, .34f } };

	inline uint32_t popcount(uint32_t v) { return uint32_t(__builtin_popcount(v)); }
	inline uint32_t bitPos(uint32_t bit) { return uint32_t(__builtin_ctz(bit)); }
	inline uint32_t roundUp4(uint32_t n) { return (n + 3u) & ~3u; } /* inc/Memory/Bits.h:52 */

	struct AngleId { float angle; uint32_t id; };
	/* VBAP::ChannelAngle list, sorted (inc/Panning/VBAPEx.h:239-295) */
	std::vector<AngleId> sortedAngles(uint32_t mask, bool skipLFE)
	{
		std::vector<AngleId> v;
		uint32_t rank = 0;
		for (uint32_t m = mask; m; m &= m - 1, ++rank)
		{
			const uint32_t bit = m & (0u - m);
			if ((skipLFE && bit == LFE) || bit >= FIRST_TOP) continue;
			uint32_t id = rank;
			if (skipLFE && (mask & LFE) && bit > LFE) --id;
			const float a = kSpeakerAngle[bitPos(bit)];
			v.push_back({ a < 0.0f ? a + TWO_PI : a, id });
		}
		std::stable_sort(v.begin(), v.end(), [](const AngleId& l, const AngleId& r) { return l.angle < r.angle; });
		return v;
	}
}

/* ================================================================================================================== */
struct orc_panner
{
	uint32_t mask = 0, L = 0;
	bool is3d = false;
	float aperture = 0.0f;
	std::vector<float> lut2d;            /* [1024][L] (inc/Panning/VBAPLUT2D.h:91-99) */
	std::vector<float> g3;               /* [65536][3] (inc/Panning/VBAPLUT3D.h:163-171) */
	std::vector<uint8_t> s3;             /* [65536][3] */
};

namespace
{
	/* LUTBuilder2D (inc/Panning/VBAPLUT2D.h:536-733) */
	bool build2d(orc_panner& p)
	{
		const uint32_t quad = 0x303u; /* FL | FR | BL | BR */
		const uint32_t work = p.L < 4 ? quad : p.mask; /* <4 target channels: build on quad, convert (:543,556-560) */
		const uint32_t lfeId = (work & LFE) ? popcount(work & (LFE - 1)) : 0xffffffffu;
		const std::vector<AngleId> A = sortedAngles(work, false);
		if (A.size() < 2) return false;
		/* ForEachChannelAnglePair (:674-701): neighbours in angle order, skipping the LFE entry, wrapping at the end */
		auto skip = [&](uint32_t i) { return i + (A[i].id == lfeId ? 1u : 0u); };
		std::vector<std::pair<uint32_t, uint32_t>> pr;
		for (uint32_t i = 0; i + 1 < A.size(); ++i)
		{
			const uint32_t a = skip(i);
			if (a + 1 >= A.size()) break;
			const uint32_t b = skip(a + 1);
			if (b >= A.size()) break;
			pr.emplace_back(a, b);
		}
		pr.emplace_back(uint32_t(A.size()) - 1, skip(0));
		/* ComputePairMatrices (:645-671): inverse of [[sin a1, -cos a1], [sin a2, -cos a2]] (column vectors), Mat2::TryInverse (inc/Math/MinimalMat.h:49-68) */
		struct Pair { float i00, i01, i10, i11; uint32_t id1, id2; };
		std::vector<Pair> P;
		for (auto [i1, i2] : pr)
		{
			const float s1 = std::sin(A[i1].angle), c1 = std::cos(A[i1].angle), s2 = std::sin(A[i2].angle), c2 = std::cos(A[i2].angle);
			const float c0x = s1, c0y = -c1, c1x = s2, c1y = -c2;
			const float det = c0x * c1y - c0y * c1x;
			Pair q{ 1, 0, 0, 1, A[i1].id, A[i2].id };
			if (std::fabs(det) >= 1e-6f)
			{
				const float inv = 1.0f / det;
				/* rows of the adjugate scaled by 1/det, stored back as columns */
				const float r0x = c1y * inv, r0y = (-c1x) * inv, r1x = (-c0y) * inv, r1y = c0x * inv;
				q.i00 = r0x; q.i10 = r1x; /* column 0 = { r0.x, r1.x } */
				q.i01 = r0y; q.i11 = r1y; /* column 1 = { r0.y, r1.y } */
			}
			P.push_back(q);
		}
		/* FindShortestAperture (:566-583) */
		float shortest = std::numeric_limits<float>::max();
		for (auto [i1, i2] : pr)
			shortest = std::min(shortest, A[i1].id == A.back().id ? TWO_PI - A[i1].angle + A[i2].angle : A[i2].angle - A[i1].angle);
		p.aperture = shortest;
		/* rectangular conversion weights quad -> target (inc/ChannelConversion.h:163-236) */
		std::vector<float> W; /* [out][in] */
		const uint32_t nIn = popcount(work);
		if (work != p.mask)
		{
			W.assign(size_t(p.L) * nIn, 0.0f);
			auto spatial = [](uint32_t bit) { return bit != LFE && bit != FRONT_CENTER; };
			auto weight = [](uint32_t a, uint32_t b) { float s = 0; for (int k = 0; k < 6; ++k) s = k == 0 ? kPlane[a][0] * kPlane[b][0] : s + kPlane[a][k] * kPlane[b][k]; return s; };
			uint32_t ii = 0;
			for (uint32_t mi = work; mi; mi &= mi - 1, ++ii)
			{
				uint32_t oi = 0;
				for (uint32_t mo = p.mask; mo; mo &= mo - 1, ++oi)
					if ((mi & (0u - mi)) == (mo & (0u - mo))) W[oi * nIn + ii] = 1.0f;
			}
			ii = 0;
			for (uint32_t mi = work; mi; mi &= mi - 1, ++ii)
			{
				const uint32_t cin = mi & (0u - mi);
				if (!spatial(cin) || (p.mask & cin)) continue;
				uint32_t oi = 0;
				for (uint32_t mo = p.mask; mo; mo &= mo - 1, ++oi)
				{
					const uint32_t cout = mo & (0u - mo);
					if (spatial(cout) && W[oi * nIn + ii] == 0.0f) W[oi * nIn + ii] = weight(bitPos(cin), bitPos(cout));
				}
			}
			uint32_t oi = 0;
			for (uint32_t mo = p.mask; mo; mo &= mo - 1, ++oi)
			{
				const uint32_t cout = mo & (0u - mo);
				if (!spatial(cout) || (work & cout)) continue;
				ii = 0;
				for (uint32_t mi = work; mi; mi &= mi - 1, ++ii)
				{
					const uint32_t cin = mi & (0u - mi);
					if (spatial(cin) && W[oi * nIn + ii] == 0.0f) W[oi * nIn + ii] = weight(bitPos(cout), bitPos(cin));
				}
			}
		}
		/* BuildForAllDirections (:619-642): the diamond parameter advances by repeated addition of 1/1024 */
		p.lut2d.assign(size_t(1024) * p.L, 0.0f);
		float dia = 0.0f;
		for (uint32_t pos = 0; pos < 1024; ++pos, dia += 1.0f / 1024.0f)
		{
			float dx, dy; fromDiamond(dia, dx, dy);
			float* row = &p.lut2d[size_t(pos) * p.L];
			for (const Pair& q : P) /* ComputeCellFor (:586-616): first pair with both gains >= 0 */
			{
				float g1 = q.i00 * dx + q.i01 * dy, g2 = q.i10 * dx + q.i11 * dy;
				if (g1 < 0.0f || g2 < 0.0f) continue;
				const float inv = invSqrtExact(g1 * g1 + g2 * g2);
				g1 *= inv; g2 *= inv;
				if (W.empty()) { row[q.id1] = g1; row[q.id2] = g2; }
				else
				{
					float e = 0.0f;
					for (uint32_t o = 0; o < p.L; ++o) row[o] = g1 * W[o * nIn + q.id1] + g2 * W[o * nIn + q.id2];
					for (uint32_t o = 0; o < p.L; ++o) e += row[o] * row[o];
					const float n = invSqrtExact(e);
					for (uint32_t o = 0; o < p.L; ++o) row[o] *= n;
				}
				break;
			}
		}
		return true;
	}

	/* LUTBuilder 3D (inc/Panning/VBAPLUT3D.h:354-424,466-487,522-716) on a given triangle list */
	bool build3d(orc_panner& p, const uint8_t* tris, uint32_t ntris)
	{
		std::vector<Vec3> V;
		uint32_t lfeRank = 0xffffffffu, rank = 0, tops = 0;
		for (uint32_t m = p.mask; m; m &= m - 1, ++rank)
		{
			const uint32_t bit = m & (0u - m);
			if (bit == LFE) { lfeRank = rank; continue; }
			V.push_back(kSpeakerDir[bitPos(bit)]);
			tops += bit >= FIRST_TOP;
		}
		const uint32_t real = uint32_t(V.size());
		V.push_back({ 0, -1, 0 });                                               /* dummy below (:389-399) */
		if ((tops == 4 || tops == 6) && !(p.mask & TOP_CENTER)) V.push_back({ 0, 1, 0 });
		if (!tris || ntris == 0) return false;
		struct M3 { Vec3 c0, c1, c2; };
		std::vector<M3> inv(ntris);
		float best = -std::numeric_limits<float>::max();
		for (uint32_t t = 0; t < ntris; ++t)
		{
			const Vec3& a = V[tris[3 * t]]; const Vec3& b = V[tris[3 * t + 1]]; const Vec3& c = V[tris[3 * t + 2]];
			/* Mat3::TryInverse (inc/Math/MinimalMat.h:92-117) */
			const Vec3 r0 = cross(b, c), r1 = cross(c, a), r2 = cross(a, b);
			const float det = dot(a, r0);
			if (std::fabs(det) >= 1e-6f)
			{
				const float k = 1.0f / det;
				const Vec3 R0{ r0.x * k, r0.y * k, r0.z * k }, R1{ r1.x * k, r1.y * k, r1.z * k }, R2{ r2.x * k, r2.y * k, r2.z * k };
				inv[t] = { { R0.x, R1.x, R2.x }, { R0.y, R1.y, R2.y }, { R0.z, R1.z, R2.z } };
			}
			else inv[t] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
			best = std::max({ best, dot(a, b), dot(b, c), dot(c, a) });          /* FindShortestAperture (:466-480): largest cosine */
		}
		p.aperture = best;
		p.g3.assign(size_t(65536) * 3, 0.0f);
		p.s3.assign(size_t(65536) * 3, 0);
		for (uint32_t dy = 0; dy < 255; ++dy)
			for (uint32_t dx = 0; dx < 255; ++dx)
			{
				const uint32_t code = (dy << 8) | dx;
				const Vec3 d = octDecodeScalar(code);
				bool hit = false;
				for (uint32_t t = 0; t < ntris && !hit; ++t) /* ComputeCellFor (:522-632): FIRST triangle with all gains >= -1e-6 */
				{
					const M3& m = inv[t];
					Vec3 g{ (m.c0.x * d.x + m.c1.x * d.y) + m.c2.x * d.z, (m.c0.y * d.x + m.c1.y * d.y) + m.c2.y * d.z, (m.c0.z * d.x + m.c1.z * d.y) + m.c2.z * d.z };
					if (g.x < -1e-6f || g.y < -1e-6f || g.z < -1e-6f) continue;
					if (std::fabs(dot(g, g)) <= 1e-6f) continue;
					uint32_t id[3] = { tris[3 * t], tris[3 * t + 1], tris[3 * t + 2] };
					int dummy = id[0] >= real ? 0 : id[1] >= real ? 1 : id[2] >= real ? 2 : -1;
					if (dummy == 0) g.x = 0; else if (dummy == 1) g.y = 0; else if (dummy == 2) g.z = 0;
					if (dummy >= 0) { uint32_t r = 0; while (id[0] == r || id[1] == r || id[2] == r) ++r; id[dummy] = r; } /* :803-815 */
					g = normalizedExact(g);
					float* gp = &p.g3[size_t(code) * 3]; gp[0] = g.x; gp[1] = g.y; gp[2] = g.z;
					for (int k = 0; k < 3; ++k) p.s3[size_t(code) * 3 + k] = uint8_t(id[k] + (id[k] >= lfeRank ? 1u : 0u)); /* SanitizeSpeakerIndex (:516-519) */
					hit = true;
				}
				if (!hit) for (int k = 0; k < 3; ++k) p.s3[size_t(code) * 3 + k] = 255;
			}
		return true;
	}

	struct Layout
	{
		uint32_t channels = 0, vs = 0, rings = 0, perRing = 0;
		float phi = 0, offset = 0, weight = 0;
		float quat[8][4]; uint32_t ids[8];
		std::vector<float> azSin, azCos;
	};

	/* SourceLayout::Initialize 2D (inc/Panning/VBAPanning2D.h:150-193), 3D (VBAPanning3D.h:242-346,534-612), InitializeBase (PannerBase.h:922-985) */
	bool makeLayout(const orc_panner& p, uint32_t src, Layout& o)
	{
		if (src == 0 || (src & ~uint32_t(ALL)) || src >= FIRST_TOP) return false;
		const uint32_t n = popcount(src) - ((src & LFE) ? 1u : 0u);
		if (n == 0 || n > 8) return false;
		if (!p.is3d)
		{
			uint32_t total = 4;
			if (p.aperture < std::numeric_limits<float>::max()) total = uint32_t(std::ceil(TWO_PI / std::max(p.aperture, 1e-6f)));
			o.vs = roundUp4(std::max(total / n, 1u));
			if (o.vs * n > 32) return false;
			o.phi = TWO_PI / (o.vs * n);
			o.offset = n > 1 ? o.phi * o.vs * 0.5f : 0.0f;
		}
		else
		{
			uint32_t total = 12, per = 24;
			if (p.aperture < std::numeric_limits<float>::max())
			{
				total = uint32_t(std::ceil(PI / std::max(std::acos(p.aperture), 1e-6f)));
				if (n <= 2) per = total << 1;
				else
				{
					/* MinPointsOnCapRing (VBAPanning3D.h:213-235) */
					const float al = 0.5f * (TWO_PI / n), s = std::sin(al), c = std::cos(al), s2 = s * s, c2 = c * c;
					if (!(s2 != 0.0)) per = 1;
					else
					{
						const float lim = std::acos(std::clamp((p.aperture - c2) / s2, -1.0f, 1.0f));
						per = lim > 0.0f ? uint32_t(std::ceil(TWO_PI / lim)) : 24u;
					}
				}
			}
			if (roundUp4(per) * total > 288) per = (288u / total) & ~3u;
			o.rings = std::max(1u, uint32_t(std::ceil(total / float(n))));
			o.perRing = roundUp4(per);
			o.vs = o.rings * o.perRing;
			if (o.perRing < 4 || size_t(o.vs) * n > 288) return false;
			o.phi = n == 1 ? PI / (o.rings + 1) : PI / (o.rings * n + n * 0.5f);
			/* PrecomputeAzimuth (:312-346): half of the packs of a ring (+ the odd one), theta advanced by repeated addition */
			const uint32_t packs = o.perRing >> 2, half = ((packs & ~1u) >> 1) + (packs & 1u);
			const float step = TWO_PI / o.perRing;
			float th[4] = { step * 0.5f, step * 1.5f, step * 2.5f, step * 3.5f };
			o.azSin.resize(half * 4); o.azCos.resize(half * 4);
			for (uint32_t i = 0; i < half; ++i)
				for (int l = 0; l < 4; ++l) { sinCosLane(th[l], o.azSin[4 * i + l], o.azCos[4 * i + l]); th[l] += step * 4.0f; }
		}
		o.channels = n;
		o.weight = 1.0f / o.vs;
		const std::vector<AngleId> S = sortedAngles(src, true);
		const float width = TWO_PI / float(n), off = (src & FRONT_CENTER) ? 0.0f : 0.5f * width;
		struct G { float q[4]; uint32_t id; };
		std::vector<G> g(n);
		for (uint32_t i = 0; i < n && i < S.size(); ++i)
		{
			const float h = 0.5f * (-(off + width * i)); /* Quat::Rotation(up, -angle) (inc/Math/MinimalQuat.h:58-63) */
			const float s = std::sin(h), c = std::cos(h);
			g[i] = { { 0.0f * s, 1.0f * s, 0.0f * s, c }, S[i].id };
		}
		std::stable_sort(g.begin(), g.end(), [](const G& l, const G& r) { return l.id < r.id; });
		for (uint32_t i = 0; i < n; ++i) { std::memcpy(o.quat[i], g[i].q, 16); o.ids[i] = g[i].id; }
		return true;
	}

	/* GenerateSpreadCap: canonical cap of one channel, in the reference's buffer order (pack-major) */
	void spreadCap(const orc_panner& p, const Layout& lay, float focus, std::vector<Vec3>& cap)
	{
		cap.resize(lay.vs);
		const float spread = 1.0f - focus;
		if (!p.is3d)
		{
			/* inc/Panning/VBAPanning2D.h:196-246 */
			const uint32_t packs = lay.vs >> 2, half = (packs & ~1u) >> 1, halfTail = half + (packs & 1u);
			const float phiT = spread * lay.phi, off = (-spread) * lay.offset;
			float ph[4]; for (int l = 0; l < 4; ++l) ph[l] = phiT * (0.5f + float(l)) + off;
			for (uint32_t i = 0; i < halfTail; ++i)
				for (int l = 0; l < 4; ++l) { float s, c; sinCosLane(ph[l], s, c); cap[4 * i + l] = { s, 0.0f, c }; ph[l] += 4.0f * phiT; }
			for (uint32_t i = 0; i < half; ++i)
				for (int l = 0; l < 4; ++l) cap[4 * (halfTail + i) + l] = { 0.0f - cap[4 * i + l].x, 0.0f, cap[4 * i + l].z };
			return;
		}
		/* inc/Panning/VBAPanning3D.h:404-518 */
		std::vector<float> rs(lay.rings), rc(lay.rings);
		const float phiS = spread * lay.phi;
		float ph[4] = { phiS * 1.0f, phiS * 2.0f, phiS * 3.0f, phiS * 4.0f };
		uint32_t i = 0;
		for (; i + 4 <= lay.rings; i += 4)
			for (int l = 0; l < 4; ++l) { sinCosLane(ph[l], rs[i + l], rc[i + l]); ph[l] += 4.0f * phiS; }
		for (uint32_t t = 0; i < lay.rings; ++i, ++t) { rs[i] = std::sin(ph[t]); rc[i] = std::cos(ph[t]); } /* scalar Math::SinCos = libm (inc/Math/Math.h:164-168) */
		const uint32_t packs = lay.perRing >> 2, half = (packs & ~1u) >> 1, tail = packs & 1u;
		uint32_t w = 0;
		for (uint32_t r = 0; r < lay.rings; ++r)
			for (uint32_t k = 0; k < half; ++k)
				for (int l = 0; l < 4; ++l) cap[w++] = { rs[r] * lay.azCos[4 * k + l], rs[r] * lay.azSin[4 * k + l], rc[r] };
		const uint32_t halfCount = w;
		for (uint32_t k = 0; k < halfCount; ++k) cap[w++] = { cap[k].x, 0.0f - cap[k].y, cap[k].z };
		if (tail)
			for (uint32_t r = 0; r < lay.rings; ++r)
				for (int l = 0; l < 4; ++l) cap[w++] = { rs[r] * lay.azCos[4 * half + l], rs[r] * lay.azSin[4 * half + l], rc[r] };
	}

	/* ProcessVBAPDataImpl for one source (inc/Panning/PannerBase.h:520-728) */
	void mdapOne(const orc_panner& p, const Layout& lay, const float* prm, float* out, bool oriented = false)
	{
		const uint32_t L = p.L;
		const Vec3 target{ prm[0], prm[1], prm[2] };
		const float focus = prm[3] == 1.0f ? 0.9999f : prm[3];    /* :589 */
		const float spread = prm[4] == 0.0f ? 0.0001f : prm[4];   /* :646 */
		const Quat pan = oriented ? quatFromUpAndForward(Vec3{ prm[5], prm[6], prm[7] }, Vec3{ prm[8], prm[9], prm[10] })  /* :509 */
		                          : quatLookAt(target);            /* :495 */
		std::vector<Vec3> cap; spreadCap(p, lay, focus, cap);
		std::fill(out, out + size_t(lay.channels) * L, 0.0f);
		float mix[32] = { 0 };
		for (uint32_t ch = 0; ch < lay.channels; ++ch)
		{
			const Basis rot = quatToBasis(quatMul(pan, Quat{ lay.quat[ch][0], lay.quat[ch][1], lay.quat[ch][2], lay.quat[ch][3] })); /* :593-601 */
			float* g = out + size_t(ch) * L;
			for (uint32_t pk = 0; pk < lay.vs; pk += 4) /* one simd pack */
			{
				Vec3 v[4];
				for (int l = 0; l < 4; ++l) v[l] = slerpLane(basisTransformSimd(rot, cap[pk + l]), target, 1.0f - spread); /* :895-919 */
				if (p.is3d)
				{
					/* ProcessVirtualSourcesImplSIMD<true>, 3D (:823-851): 12 (speaker, w*g*g) entries per pack, added in entry order */
					for (int l = 0; l < 4; ++l)
					{
						const uint32_t code = octEncodeSimd(v[l].x, v[l].y, v[l].z);
						for (int k = 0; k < 3; ++k)
						{
							const float gg = p.g3[size_t(code) * 3 + k];
							const uint8_t sp = p.s3[size_t(code) * 3 + k];
							if (sp < L) g[sp] += (lay.weight * gg) * gg;
						}
					}
				}
				else
				{
					/* 2D (:852-875): per speaker (w*g*g).reduce() = (l0+l1)+(l2+l3) (inc/Math/SIMD.h:938-945) */
					uint32_t pos[4]; for (int l = 0; l < 4; ++l) pos[l] = lut2dIndexSimd(v[l].x, v[l].z);
					for (uint32_t s = 0; s < L; ++s)
					{
						float t[4]; for (int l = 0; l < 4; ++l) { const float gg = p.lut2d[size_t(pos[l]) * L + s]; t[l] = (lay.weight * gg) * gg; }
						g[s] += (t[0] + t[1]) + (t[2] + t[3]);
					}
				}
			}
			for (uint32_t s = 0; s < L; ++s) mix[s] += g[s];      /* :683-686 */
		}
		for (uint32_t s = 0; s < L; ++s)                          /* :694-713 */
		{
			const float n = mix[s] > 0.0f ? invSqrtExact(mix[s] * float(lay.channels)) : 0.0f;
			for (uint32_t ch = 0; ch < lay.channels; ++ch) out[size_t(ch) * L + s] *= n;
		}
	}

	/* ---- FourBandCrossover (inc/Auralization/CrossoverFilter.h) */
	struct Svf { float g, d, k; }; /* k = -(R + g), R = sqrt(2) */
	inline void svfTick(const Svf& c, float x, float* s, float& lo, float& bp, float& hi) /* :64-76,154-162 */
	{
		hi = (c.k * s[0] + (x - s[1])) * c.d;
		bp = c.g * hi + s[0];
		s[0] = c.g * hi + bp;
		lo = c.g * bp + s[1];
		s[1] = c.g * bp + lo;
	}
	/* one input sample -> bands in the internal order { b1, b3, b2, b4 } (X2.Process :294-310, X13.ProcessInterleaved :371-401, all-pass :187-193,221-241) */
	void xoverTick(const float* cf, float* st, float x, float band[4])
	{
		const Svf mid{ cf[0], cf[1], cf[2] };
		float l1, b1, h1, l2, b2, h2;
		svfTick(mid, x, st + 0, l1, b1, h1);
		svfTick(mid, l1, st + 2, l2, b2, h2);
		const float low = l2, high = (-SQRT2) * b1 + (l1 + h1 - l2);
		const float in1[4] = { low, high, low, high };
		float L1[4], H1[4], tmp;
		for (int l = 0; l < 2; ++l) svfTick(Svf{ cf[3 + l], cf[7 + l], cf[11 + l] }, in1[l], st + 4 + 2 * l, L1[l], tmp, H1[l]);
		L1[2] = L1[0]; L1[3] = L1[1]; H1[2] = H1[0]; H1[3] = H1[1]; /* lanes 2,3 duplicate 0,1 (:382-387) */
		const float in2[4] = { L1[0], L1[1], H1[2], H1[3] };
		float o2[4];
		for (int l = 0; l < 4; ++l)
		{
			float lo, bp, hi;
			svfTick(Svf{ cf[3 + l], cf[7 + l], cf[11 + l] }, in2[l], st + 8 + 2 * l, lo, bp, hi);
			o2[l] = l < 2 ? lo : hi;
		}
		for (int l = 0; l < 4; ++l)
		{
			float lo, bp, hi;
			svfTick(Svf{ cf[15 + l], cf[19 + l], cf[23 + l] }, o2[l], st + 16 + 2 * l, lo, bp, hi);
			band[l] = (-SQRT2) * bp + (hi + lo);
		}
	}
}

struct orc_dse
{
	float fs; uint32_t C, L;
	float coef[27];
	float alpha, curDelay, tgtDelay;
	float curBand[4], tgtBand[4];
	std::vector<float> curMix, tgtMix;
	struct Ch { std::vector<float> ring; uint32_t w; float y1; float st[24]; };
	std::vector<Ch> ch;
	uint32_t ringSize;
};

extern "C"
{
	void orc_sincos(const float* x, uint32_t n, float* s, float* c) { for (uint32_t i = 0; i < n; ++i) sinCosLane(x[i], s[i], c[i]); }
	void orc_acos(const float* x, uint32_t n, float* out) { for (uint32_t i = 0; i < n; ++i) out[i] = acosLane(x[i]); }
	void orc_tan(const float* x, uint32_t n, float* out) { for (uint32_t i = 0; i < n; ++i) out[i] = tanLane(x[i]); }
	void orc_invsqrt_nr1(const float* x, uint32_t n, float* out) { for (uint32_t i = 0; i < n; ++i) out[i] = invSqrtNR1(x[i]); }
	void orc_invsqrt_nr2(const float* x, uint32_t n, float* out) { for (uint32_t i = 0; i < n; ++i) out[i] = invSqrtNR2(x[i]); }
	void orc_oct_encode_simd(const float* d, uint32_t n, uint32_t* out) { for (uint32_t i = 0; i < n; ++i) out[i] = octEncodeSimd(d[3 * i], d[3 * i + 1], d[3 * i + 2]); }
	void orc_oct_encode_scalar(const float* d, uint32_t n, uint32_t* out) { for (uint32_t i = 0; i < n; ++i) out[i] = octEncodeScalar(d[3 * i], d[3 * i + 1], d[3 * i + 2]); }
	void orc_oct_decode_scalar(const uint32_t* c, uint32_t n, float* out) { for (uint32_t i = 0; i < n; ++i) { const Vec3 v = octDecodeScalar(c[i]); out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z; } }
	void orc_diamond_scalar(const float* xy, uint32_t n, float* out) { for (uint32_t i = 0; i < n; ++i) out[i] = toDiamondScalar(xy[2 * i], xy[2 * i + 1]); }
	void orc_from_diamond(const float* p, uint32_t n, float* o) { for (uint32_t i = 0; i < n; ++i) fromDiamond(p[i], o[2 * i], o[2 * i + 1]); }
	void orc_slerp_simd(float* v0, const float* v1, float t, uint32_t n)
	{
		for (uint32_t i = 0; i < n; ++i)
		{
			const Vec3 r = slerpLane({ v0[3 * i], v0[3 * i + 1], v0[3 * i + 2] }, { v1[0], v1[1], v1[2] }, t); /* one target for all, like SlerpChannelCap */
			v0[3 * i] = r.x; v0[3 * i + 1] = r.y; v0[3 * i + 2] = r.z;
		}
	}
	void orc_quat_lookat(const float* d, uint32_t n, float* out)
	{
		for (uint32_t i = 0; i < n; ++i) { const Quat q = quatLookAt({ d[3 * i], d[3 * i + 1], d[3 * i + 2] }); out[4 * i] = q.x; out[4 * i + 1] = q.y; out[4 * i + 2] = q.z; out[4 * i + 3] = q.w; }
	}

	/* VBAPannerBase::Initialize (inc/Panning/PannerBase.h:455-482) */
	orc_panner* orc_panner_create(uint32_t mask, const uint8_t* tris, uint32_t ntris)
	{
		if (mask == 0 || (mask & ~uint32_t(ALL))) return nullptr;
		orc_panner* p = new orc_panner;
		p->mask = mask; p->L = popcount(mask); p->is3d = mask >= FIRST_TOP;
		bool ok;
		if (p->is3d) ok = p->L >= 3 && build3d(*p, tris, ntris);   /* inc/Panning/VBAPanning3D.h:157-166 */
		else ok = p->L != 1 && build2d(*p);                         /* inc/Panning/VBAPanning2D.h:124-137 */
		if (!ok) { delete p; return nullptr; }
		return p;
	}
	void orc_panner_destroy(orc_panner* p) { delete p; }
	int orc_panner_is3d(const orc_panner* p) { return p->is3d; }
	uint32_t orc_panner_num_channels(const orc_panner* p) { return p->L; }
	float orc_panner_aperture(const orc_panner* p) { return p->aperture; }
	void orc_panner_lut(const orc_panner* p, float* gains, uint8_t* speakers)
	{
		if (p->is3d) { std::memcpy(gains, p->g3.data(), p->g3.size() * 4); if (speakers) std::memcpy(speakers, p->s3.data(), p->s3.size()); }
		else std::memcpy(gains, p->lut2d.data(), p->lut2d.size() * 4);
	}

	/* GetSpeakerGains(direction, outGains) (PannerBase.h:441-445): 3D writes three entries (VBAPLUT3D.h:1001-1009), rows are pre-zeroed here */
	void orc_point_gains(const orc_panner* p, const float* d, uint32_t n, float* out)
	{
		for (uint32_t i = 0; i < n; ++i)
		{
			float* row = out + size_t(i) * p->L;
			std::fill(row, row + p->L, 0.0f);
			if (p->is3d)
			{
				const uint32_t code = octEncodeScalar(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
				for (int k = 0; k < 3; ++k) { const uint8_t s = p->s3[size_t(code) * 3 + k]; if (s < p->L) row[s] = p->g3[size_t(code) * 3 + k]; }
			}
			else
			{
				/* LUTQuery2D::GainsFor (inc/Panning/VBAPLUT2D.h:193-205): zero Y, exact normalise, diamond index, copy the row */
				float x = d[3 * i], z = d[3 * i + 2];
				const float il = invSqrtExact(x * x + 0.0f * 0.0f + z * z);
				x *= il; z *= il;
				std::memcpy(row, &p->lut2d[size_t(lut2dIndexScalar(x, z)) * p->L], p->L * 4);
			}
		}
	}

	int orc_layout_info(const orc_panner* p, uint32_t src, uint32_t info[4], float* phi, float* offset, float* weight, float* quats, uint32_t* ids)
	{
		Layout l;
		if (!makeLayout(*p, src, l)) return 0;
		info[0] = l.channels; info[1] = l.vs; info[2] = l.rings; info[3] = l.perRing;
		if (phi) *phi = l.phi;
		if (offset) *offset = l.offset;
		if (weight) *weight = l.weight;
		for (uint32_t i = 0; i < l.channels; ++i) { if (quats) std::memcpy(quats + 4 * i, l.quat[i], 16); if (ids) ids[i] = l.ids[i]; }
		return 1;
	}

	int orc_mdap(const orc_panner* p, uint32_t src, const float* params, uint32_t n, float* out)
	{
		Layout l;
		if (!makeLayout(*p, src, l)) return 0;
		for (uint32_t i = 0; i < n; ++i) mdapOne(*p, l, params + size_t(i) * 5, out + size_t(i) * l.channels * p->L);
		return 1;
	}

	/* ProcessVirtualSources (PannerBase.h:731-804): weighted linear sum + Algo::NormalizeL2 (inc/Algo/Algorithm.h:98-104); out is accumulated into */
	void orc_process_virtual_sources(const orc_panner* p, const float* vs, uint32_t n, float* out)
	{
		std::vector<float> row(p->L);
		for (uint32_t i = 0; i < n; ++i)
		{
			const float w = vs[4 * i + 3];
			if (p->is3d)
			{
				const uint32_t code = octEncodeScalar(vs[4 * i], vs[4 * i + 1], vs[4 * i + 2]);
				for (int k = 0; k < 3; ++k) { const uint8_t s = p->s3[size_t(code) * 3 + k]; if (s < p->L) out[s] += w * p->g3[size_t(code) * 3 + k]; }
			}
			else
			{
				orc_point_gains(p, vs + 4 * i, 1, row.data());
				for (uint32_t s = 0; s < p->L; ++s) out[s] += w * row[s];
			}
		}
		float e = 0.0f;
		for (uint32_t s = 0; s < p->L; ++s) e = e + out[s] * out[s];
		const float k = invSqrtExact(e);
		for (uint32_t s = 0; s < p->L; ++s) out[s] *= k;
	}

	/* Prepare: X2 scalar (std::tan), X13 and the all-pass through the simd tan (CrossoverFilter.h:50-61,139-150,253-258,343-362) */
	void orc_xover_coeffs(float fs, float f1, float f2, float f3, float out[27])
	{
		const float g = std::tan(PI * (std::clamp(f2, 1.0f, 0.49f * fs) / fs));
		out[0] = g; out[1] = 1.0f / (g * (SQRT2 + g) + 1.0f); out[2] = -(SQRT2 + g);
		const float lanes[4] = { f1, f3, f1, f3 };
		float f[4];
		for (int l = 0; l < 4; ++l) f[l] = laneMax(laneMin(lanes[l], 0.49f * fs), 1.0f);
		for (int set = 0; set < 2; ++set)
			for (int l = 0; l < 4; ++l)
			{
				const float fr = set == 0 ? f[l] : f[3 - l];        /* all-pass lanes reversed: { f3, f1, f3, f1 } */
				const float gl = tanLane(PI * (fr / fs));
				out[3 + 12 * set + l] = gl;
				out[7 + 12 * set + l] = 1.0f / (gl * (SQRT2 + gl) + 1.0f);
				out[11 + 12 * set + l] = 0.0f - (SQRT2 + gl);
			}
	}
	/* ProcessBlock(in, startGains, endGains, out) (:505-535): gains b1..b4, internally interleaved { b1, b3, b2, b4 } (:516-517) */
	void orc_xover_process_ramp(const float coef[27], float state[24], const float* in, uint32_t n, const float g0[4], const float g1[4], float* out)
	{
		float g[4] = { g0[0], g0[2], g0[1], g0[3] };
		const float e[4] = { g1[0], g1[2], g1[1], g1[3] };
		float inc[4]; for (int k = 0; k < 4; ++k) inc[k] = (e[k] - g[k]) / float(n);
		for (uint32_t i = 0; i < n; ++i)
		{
			float b[4]; xoverTick(coef, state, in[i], b);
			out[i] = (b[0] * g[0] + b[1] * g[1]) + (b[2] * g[2] + b[3] * g[3]);
			for (int k = 0; k < 4; ++k) g[k] += inc[k];
		}
	}
	void orc_xover_split(const float coef[27], float state[24], const float* in, uint32_t n, float* bands) /* :462-470 */
	{
		for (uint32_t i = 0; i < n; ++i) { float b[4]; xoverTick(coef, state, in[i], b); bands[4 * i] = b[0]; bands[4 * i + 1] = b[2]; bands[4 * i + 2] = b[1]; bands[4 * i + 3] = b[3]; }
	}
	float orc_smoother_alpha(uint32_t ticks) { return ticks == 0 ? 1.0f : float(1.0 - std::exp(-1.0 / double(ticks))); } /* inc/Utilities/SmoothedValue.h:36-41 */
	void orc_add_gain_ramp(float* dest, const float* src, uint32_t dc, uint32_t sc, uint32_t nd, uint32_t ns, uint32_t frames, float g0, float g1) /* inc/Auralization/ChannelMixing.h:89-107 */
	{
		const float delta = (g1 - g0) / float(frames);
		float g = g0;
		for (uint32_t i = 0; i < frames; ++i, g += delta) dest[size_t(i) * nd + dc] += src[size_t(i) * ns + sc] * g;
	}

	/* DirectSoundEffect ctor (src/Auralization/DirectSound.cpp:34-82) */
	orc_dse* orc_dse_create(float fs, uint32_t nSrc, uint32_t nOut, const float band[4], float delaySeconds, const float* mix)
	{
		orc_dse* e = new orc_dse;
		e->fs = fs; e->C = nSrc; e->L = nOut;
		orc_xover_coeffs(fs, 176.0f, 775.0f, 3408.0f, e->coef);     /* inc/FrequencyBands.h:46 */
		e->alpha = orc_smoother_alpha(480);                           /* cScratchSize ticks (DirectSound.cpp:53, DirectSound.h:90) */
		e->curDelay = e->tgtDelay = delaySeconds * fs;
		std::memcpy(e->curBand, band, 16); std::memcpy(e->tgtBand, band, 16);
		e->curMix.assign(mix, mix + size_t(nSrc) * nOut); e->tgtMix = e->curMix;
		uint32_t size = 1; while (size < uint32_t(fs * 2.0f)) size <<= 1; /* bit_ceil(2 fs) (:79, DelayLine.h:142-146) */
		e->ringSize = size;
		e->ch.resize(nSrc);
		for (auto& c : e->ch) { c.ring.assign(size, 0.0f); c.w = 0; c.y1 = 0.0f; std::fill(c.st, c.st + 24, 0.0f); }
		return e;
	}
	void orc_dse_destroy(orc_dse* e) { delete e; }
	void orc_dse_update(orc_dse* e, const float band[4], float delaySeconds, const float* mix) /* :231-239 */
	{
		std::memcpy(e->tgtBand, band, 16);
		e->tgtDelay = delaySeconds * e->fs;
		std::copy(mix, mix + e->tgtMix.size(), e->tgtMix.begin());
	}
	/* ProcessInterleaved (:95-229). The ring here is indexed by push count (sample pushed k-th lives at k & mask); a read of
	 * integer delay D after P pushes returns push P-1-D, which is what the backward-writing ring + GetReadWindow<2> yield
	 * (inc/Auralization/DelayLine.h:183-213, inc/Utilities/AbstractIndex.h:79-156). */
	void orc_dse_process(orc_dse* e, const float* in, float* out, uint32_t frames)
	{
		const uint32_t C = e->C, L = e->L, mask = e->ringSize - 1;
		uint32_t done = 0, left = frames;
		std::vector<float> delays(480), buf(480);
		while (left > 0)
		{
			const uint32_t n = std::min(left, 480u);
			const float* chunk = in + size_t(done) * C;
			float* o = out + size_t(done) * L;
			/* quirk (:107-128): the chunk span runs to the END of the input, so all remaining frames are pushed */
			const uint32_t remaining = frames - done;
			done += n;
			for (uint32_t c = 0; c < C; ++c)
				for (uint32_t s = 0; s < remaining; ++s) { auto& ch = e->ch[c]; ch.ring[ch.w & mask] = chunk[size_t(s) * C + c]; ++ch.w; }
			for (uint32_t s = 0; s < n; ++s) { e->curDelay = e->alpha * (e->tgtDelay - e->curDelay) + e->curDelay; delays[s] = e->curDelay; } /* :145-150 */
			for (uint32_t c = 0; c < C; ++c)
			{
				auto& ch = e->ch[c];
				for (uint32_t s = 0; s < n; ++s)
				{
					/* DelayTap::SetDelay + Process (DelayLine.h:53-67), Thiran1stInterpolator (FractionalDelay.h:97-136) */
					const float d = float(int(n) - int(s)) + delays[s];
					const uint32_t D = uint32_t(d);
					const float fr = d - float(D);
					const float a = (1.0f - fr) / (1.0f + fr);
					const float x0 = ch.ring[(ch.w - 1u - D) & mask], x1 = ch.ring[(ch.w - 2u - D) & mask];
					const float y = a * (x0 - ch.y1) + x1;
					ch.y1 = y;
					buf[s] = y;
				}
				orc_xover_process_ramp(e->coef, ch.st, buf.data(), n, e->curBand, e->tgtBand, buf.data()); /* :188 */
				for (uint32_t k = 0; k < L; ++k)                                                            /* :200-222 */
				{
					orc_add_gain_ramp(o, buf.data(), k, 0, L, 1, n, e->curMix[size_t(c) * L + k], e->tgtMix[size_t(c) * L + k]);
					e->curMix[size_t(c) * L + k] = e->tgtMix[size_t(c) * L + k];
				}
			}
			std::memcpy(e->curBand, e->tgtBand, 16); /* :225 */
			left -= n;
		}
	}

	double orc_dse_process_batch(orc_dse** fx, uint32_t n, const float* in, uint32_t frames, float* bus, int threads)
	{
		if (n == 0) return 0.0;
		const uint32_t T = uint32_t(std::max(1, threads));
		const uint32_t C = fx[0]->C, L = fx[0]->L;
		std::vector<std::vector<float>> part(T, std::vector<float>(size_t(frames) * L, 0.0f));
		const auto t0 = std::chrono::steady_clock::now();
		std::vector<std::thread> pool;
		for (uint32_t t = 0; t < T; ++t)
			pool.emplace_back([&, t] { for (uint32_t i = uint32_t(uint64_t(n) * t / T); i < uint32_t(uint64_t(n) * (t + 1) / T); ++i) orc_dse_process(fx[i], in + size_t(i) * frames * C, part[t].data(), frames); });
		for (auto& th : pool) th.join();
		for (uint32_t t = 0; t < T; ++t) for (size_t k = 0; k < part[t].size(); ++k) bus[k] += part[t][k];
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	}
	/* ================================================================================================
	 * ReverbBus (src/Auralization/LateReverb.cpp:31-134) + MultiChannelMixedFeedback<16> (inc/Auralization/ReverbDesign.h:573-704) */

	/* JPL::exp(simd), one lane (inc/Math/SIMDMath.h:296-316; SSE2 floor inc/Math/SIMD.h:1861-1879); FMA call sites unfused (SSE2 build) */
	static float expLane(float x)
	{
		x = std::min(std::max(x, -88.3762626647949f), 88.3762626647949f);
		const float v = x * 1.44269504088896341f + 0.5f;
		const float tr = float(int32_t(v));
		const float n = (v == 0.0f) ? v : tr - (tr > v ? 1.0f : 0.0f);
		x = x - (n * 0.693359375f + n * -2.12194440e-4f);
		const float x2 = x * x;
		float y = 1.9875691500e-4f;
		y = y * x + 1.3981999507e-3f; y = y * x + 8.3334519073e-3f; y = y * x + 4.1665795894e-2f;
		y = y * x + 1.6666665459e-1f; y = y * x + 5.0000001201e-1f;
		y = y * x2 + (x + 1.0f);
		const int32_t e = (int32_t(n) + 0x7f) << 23;
		float p; std::memcpy(&p, &e, 4);
		return y * p;
	}
	void orc_db_to_gain(const float* db, uint32_t n, float* out) /* inc/Math/DecibelsAndGain.h:40-56 */
	{
		const float ln10div20 = 2.302585092994046f * 0.05f;
		for (uint32_t i = 0; i < n; ++i) out[i] = expLane(ln10div20 * db[i]);
	}
	/* Math::Sqrt<double, 2> as evaluated at compile time for Hadamard's scalingFactor (inc/Math/Math.h:213-262; ReverbDesign.h:443-446) */
	static float hadamardScale16()
	{
		const double x = 1.0 / 16.0; uint64_t i; std::memcpy(&i, &x, 8);
		i = 0x5fe6eb50c7b537a9ull - (i >> 1);
		double y; std::memcpy(&y, &i, 8);
		for (int k = 0; k < 2; ++k) y = y * (1.5 - 0.5 * x * y * y);
		return float(x * y);
	}
	static const uint32_t kFdnDelays[16] = { 2927, 2593, 2273, 3697, 1877, 3877, 2477, 3461, 1609, 3779, 3541, 4259, 1669, 3539, 3637, 4013 }; /* ReverbDesign.h:578-581 */

	struct orc_reverb
	{
		float fs, invFs, coef[27], slope[4], gains[16][4], st[16][24];
		std::vector<float> ring[16]; uint32_t w[16];
	};
	orc_reverb* orc_reverb_create(float fs, float f1, float f2, float f3) /* Prepare (LateReverb.cpp:42-48) */
	{
		auto* r = new orc_reverb();
		r->fs = fs; r->invFs = 1.0f / fs;
		orc_xover_coeffs(fs, f1 > 0 ? f1 : 176.0f, f2 > 0 ? f2 : 775.0f, f3 > 0 ? f3 : 3408.0f, r->coef);
		for (int c = 0; c < 16; ++c) { r->ring[c].assign(8192, 0.0f); r->w[c] = 0; std::memset(r->st[c], 0, sizeof(r->st[c])); }
		for (int k = 0; k < 4; ++k) r->slope[k] = -60.0f;
		return r;
	}
	void orc_reverb_destroy(orc_reverb* r) { delete r; }
	void orc_reverb_set_rt60(orc_reverb* r, const float rt60[4]) { for (int k = 0; k < 4; ++k) r->slope[k] = -60.0f / std::min(std::max(rt60[k], 0.03f), 10.0f); } /* :31-35 */
	void orc_reverb_set_decay_slope(orc_reverb* r, const float s[4]) { for (int k = 0; k < 4; ++k) r->slope[k] = std::min(std::max(s[k], -60.0f / 0.03f), -60.0f / 10.0f); } /* :37-40 */
	/* mono in [frames]; out [frames][nOut] accumulated into */
	void orc_reverb_process(orc_reverb* r, const float* in, float* out, uint32_t nOut, uint32_t frames)
	{
		for (int c = 0; c < 16; ++c) /* UpdateFilterParameters (ReverbDesign.h:619-628, LateReverb.h:64-68) */
		{
			const float delaySeconds = float(kFdnDelays[c]) * r->invFs;
			float db[4]; for (int k = 0; k < 4; ++k) db[k] = r->slope[k] * delaySeconds;
			orc_db_to_gain(db, 4, r->gains[c]);
		}
		const float hs = hadamardScale16();
		for (uint32_t f = 0; f < frames; ++f)
		{
			const float inj = in[f] * 0.25f; /* InjectNormalized (ReverbDesign.h:455-459) */
			float delayed[16], mixed[16];
			for (int c = 0; c < 16; ++c) delayed[c] = r->ring[c][(r->w[c] - 1u - kFdnDelays[c]) & 8191u]; /* GetReadWindow<1>(D) */
			/* Householder::MixInPlace 16-special (:311-349) */
			for (int c = 0; c < 16; ++c) mixed[c] = delayed[c];
			for (int g = 0; g < 16; g += 4) { const float sum = (mixed[g] + mixed[g + 1]) + (mixed[g + 2] + mixed[g + 3]); for (int k = 0; k < 4; ++k) mixed[g + k] = 2.0f * mixed[g + k] + (-sum); }
			for (int lane = 0; lane < 4; ++lane)
			{
				float v[4] = { mixed[lane], mixed[lane + 4], mixed[lane + 8], mixed[lane + 12] };
				const float sum = (v[0] + v[1]) + (v[2] + v[3]);
				for (int k = 0; k < 4; ++k) mixed[lane + 4 * k] = (2.0f * v[k] + (-sum)) * 0.25f;
			}
			for (int c = 0; c < 16; ++c) /* AttenuationFilter::Process = FourBandCrossover::ProcessSample, then Push (:696-699) */
			{
				const float x = mixed[c] + inj;
				float b[4]; orc_xover_split(r->coef, r->st[c], &x, 1, b);
				const float* g = r->gains[c];
				const float sum = (b[0] * g[0] + b[1] * g[1]) + (b[2] * g[2] + b[3] * g[3]);
				r->ring[c][r->w[c] & 8191u] = sum; ++r->w[c];
			}
			/* FDNOutputMixer::Mix (:480-492): Hadamard::RecursiveUnscaled (:391-437), scale, fold */
			float h[16]; for (int c = 0; c < 16; ++c) h[c] = delayed[c];
			for (int half = 1; half < 16; half <<= 1)
				for (int base = 0; base < 16; base += 2 * half)
					for (int i = 0; i < half; ++i) { const float a = h[base + i], bb = h[base + i + half]; h[base + i] = a + bb; h[base + i + half] = a - bb; }
			for (int i = 0; i < 16; ++i) out[size_t(f) * nOut + (i % nOut)] += h[i] * hs;
		}
	}

	/* ================================================================================================
	 * ERBus (src/Auralization/EarlyReflectionsBus.cpp:83-417) */
	struct orc_er_tap { uint32_t id; int tgtState, rtState; float tgtBand[4], curBand[4], tgtDelay, curDelay, y1, st[24]; std::vector<float> tgtMix, curMix; };
	struct orc_erbus { float fs, alpha, coef[27]; uint32_t L, ringSize, w; std::vector<float> ring; std::vector<orc_er_tap> taps; };
	orc_erbus* orc_erbus_create(float fs, uint32_t L)
	{
		auto* b = new orc_erbus();
		b->fs = fs; b->L = L; b->alpha = orc_smoother_alpha(480);
		orc_xover_coeffs(fs, 176.0f, 775.0f, 3408.0f, b->coef);
		uint32_t sz = 1; while (sz < uint32_t(fs * 2.0f)) sz <<= 1; /* DelayLineType(sampleRate * 2) (:74) */
		b->ringSize = sz; b->ring.assign(sz, 0.0f); b->w = 0;
		return b;
	}
	void orc_erbus_destroy(orc_erbus* b) { delete b; }
	uint32_t orc_erbus_num_taps(const orc_erbus* b) { return uint32_t(b->taps.size()); }
	/* SetTaps (:239-406); states: 0 Stopped, 1 Rendering, 2 FadingOut */
	void orc_erbus_set_taps(orc_erbus* b, uint32_t n, const float* filterGains, const float* delaySeconds, const uint32_t* ids, const float* gains)
	{
		const uint32_t L = b->L;
		/* :246-322: taps whose fade-out has been rendered are dropped; EraseIf (:616-633) is a std::partition + erase of the tail, so it
		 * reorders the survivors the way the library algorithm does */
		b->taps.erase(std::partition(b->taps.begin(), b->taps.end(), [](const orc_er_tap& t) { return t.rtState != 0; }), b->taps.end());
		auto wanted = [&](uint32_t id) { for (uint32_t k = 0; k < n; ++k) if (ids[k] == id) return true; return false; };
		/* PartitionERs (:556-566): std::partition reorders the taps (requested ones first); the order is observable because every tap
		 * adds into the output in vector order (float sums are not associative), so the same library algorithm is used here */
		std::partition(b->taps.begin(), b->taps.end(), [&](const orc_er_tap& t) { return wanted(t.id); });
		for (auto& t : b->taps) /* stale taps fade out (:329-343) */
			if (!wanted(t.id) && t.rtState != 2 && t.tgtState != 2) { std::fill(t.tgtMix.begin(), t.tgtMix.end(), 0.0f); t.tgtState = 2; }
		for (uint32_t k = 0; k < n; ++k)
		{
			orc_er_tap* hit = nullptr;
			for (auto& t : b->taps) if (t.id == ids[k]) { hit = &t; break; }
			if (hit && hit->tgtState == 2) continue; /* dead path in the reference (:366-376) */
			if (!hit)
			{
				orc_er_tap t{}; t.id = ids[k]; t.tgtState = 1; t.rtState = 1;
				t.curMix.assign(L, 0.0f); t.tgtMix.assign(L, 0.0f);
				std::memcpy(t.curBand, filterGains + 4 * size_t(k), 16);
				t.curDelay = delaySeconds[k] * b->fs; t.y1 = 0.0f; std::memset(t.st, 0, sizeof(t.st)); /* AllocateRTData (:421-455) */
				b->taps.push_back(t); hit = &b->taps.back();
			}
			std::memcpy(hit->tgtBand, filterGains + 4 * size_t(k), 16);
			std::copy(gains + size_t(k) * L, gains + size_t(k + 1) * L, hit->tgtMix.begin());
			hit->tgtDelay = delaySeconds[k] * b->fs;
		}
	}
	/* in [frames][nIn]; out [frames][L] accumulated into (:83-225) */
	void orc_erbus_process(orc_erbus* b, const float* in, uint32_t nIn, float* out, uint32_t frames)
	{
		const uint32_t L = b->L, mask = b->ringSize - 1;
		const float norm = 1.0f / float(nIn);
		uint32_t done = 0, left = frames;
		std::vector<float> buf(480);
		while (left > 0)
		{
			const uint32_t n = std::min(left, 480u);
			const float* chunk = in + size_t(done) * nIn;
			float* o = out + size_t(done) * L;
			done += n;
			for (uint32_t s = 0; s < n; ++s) { float smp = 0.0f; for (uint32_t c = 0; c < nIn; ++c) smp += chunk[size_t(s) * nIn + c]; buf[s] = smp; }
			for (uint32_t s = 0; s < n; ++s) buf[s] *= norm;
			for (uint32_t s = 0; s < n; ++s) { b->ring[b->w & mask] = buf[s]; ++b->w; }
			for (auto& t : b->taps)
			{
				if (t.tgtState == 0 && t.rtState == 0) continue;
				if (t.rtState == 0) continue;
				t.rtState = t.tgtState;
				for (uint32_t s = 0; s < n; ++s)
				{
					t.curDelay = b->alpha * (t.tgtDelay - t.curDelay) + t.curDelay;
					const float d = float(int(n) - int(s)) + t.curDelay;
					const uint32_t D = uint32_t(d);
					const float fr = d - float(D);
					const float a = (1.0f - fr) / (1.0f + fr);
					const float x0 = b->ring[(b->w - 1u - D) & mask], x1 = b->ring[(b->w - 2u - D) & mask];
					const float y = a * (x0 - t.y1) + x1;
					t.y1 = y; buf[s] = y;
				}
				orc_xover_process_ramp(b->coef, t.st, buf.data(), n, t.curBand, t.tgtBand, buf.data());
				std::memcpy(t.curBand, t.tgtBand, 16);
				for (uint32_t k = 0; k < L; ++k) { orc_add_gain_ramp(o, buf.data(), k, 0, L, 1, n, t.curMix[k], t.tgtMix[k]); t.curMix[k] = t.tgtMix[k]; }
				if (t.rtState == 2) t.rtState = 0; /* FadingOut -> Stopped after one chunk (:218-220) */
				/* the scratch buffer keeps this tap's output: the next tap overwrites it (buffer[s] = Tap.Process) */
			}
			left -= n;
		}
	}
	/* ================================================================================================
	 * Scene front-end: the per-source body of SpatialManager::AdvanceSimulation (inc/SpatialManager.h:374-420)
	 * and the parameter producers. Positions: 9 floats = location, up, forward. libm calls are the reference's own (this is CPU code). */
	int orc_mdap_oriented(const orc_panner* p, uint32_t src, const float* params, uint32_t n, float* out) /* PannerBase.h:501-515 */
	{
		Layout l;
		if (!makeLayout(*p, src, l)) return 0;
		for (uint32_t i = 0; i < n; ++i) mdapOne(*p, l, params + size_t(i) * 11, out + size_t(i) * l.channels * p->L, true);
		return 1;
	}
	/* DirectPathService::ProcessDirectPath (inc/Services/DirectPathService.h:280-338); out [12] = distance, dot, inverse dot, location, up, forward */
	void orc_direct_path(const float* s, const float* l, float* out)
	{
		const Vec3 sLoc{ s[0], s[1], s[2] }, sUp{ s[3], s[4], s[5] }, sFwd{ s[6], s[7], s[8] };
		const Vec3 lLoc{ l[0], l[1], l[2] }, lY{ l[3], l[4], l[5] }, lZ{ -l[6], -l[7], -l[8] };   /* ToBasisUnsafe (inc/Math/Position.h:58-68) */
		const Vec3 lX = cross(lY, lZ);
		auto inv = [&](const Vec3& v) { return Vec3{ dot(v, lX), dot(v, lY), dot(v, lZ) }; };  /* Basis::InverseTransform (MinimalBasis.h:99-106) */
		Vec3 p = inv(Vec3{ sLoc.x - lLoc.x, sLoc.y - lLoc.y, sLoc.z - lLoc.z });
		if (std::fabs(p.x) <= 1e-6f && std::fabs(p.z) <= 1e-6f) { p.x += 0.0f * 1e-5f; p.y += 0.0f * 1e-5f; p.z += -1.0f * 1e-5f; }  /* :291-294 */
		const float distance = std::sqrt(p.x * p.x + p.y * p.y + p.z * p.z);
		const Vec3 dir{ p.x / distance, p.y / distance, p.z / distance };
		out[0] = distance;
		out[1] = dir.x * 0.0f + dir.y * 0.0f + dir.z * -1.0f;
		out[2] = dot(sFwd, inv(Vec3{ -dir.x, -dir.y, -dir.z }));
		const Vec3 up = inv(sUp), fwd = inv(Vec3{ -sFwd.x, -sFwd.y, -sFwd.z });
		out[3] = dir.x; out[4] = dir.y; out[5] = dir.z; out[6] = up.x; out[7] = up.y; out[8] = up.z; out[9] = fwd.x; out[10] = fwd.y; out[11] = fwd.z;
	}
	float orc_angle_attenuation(float azimuthCos, float inner, float outer) /* ProcessAngleAttenuationImpl (:340-364) */
	{
		const float ci = std::cos(inner * 0.5f), co = std::cos(outer * 0.5f);
		if (azimuthCos > ci) return 0.0f;
		if (azimuthCos > co) return (ci - azimuthCos) / (ci - co);
		return 1.0f;
	}
	float orc_attenuation_model(int model, float mn, float mx, float rolloff, float d) /* AttenuationBaseModel::Evaluate (inc/DistanceAttenuation.h:66-119) */
	{
		const bool valid = mn < mx;
		const float dc = d < mn ? mn : (mx < d ? mx : d);
		switch (model)
		{
		case 1: return valid ? mn / (mn + rolloff * (dc - mn)) : 1.0f;
		case 2: return valid ? 1.0f - rolloff * (dc - mn) / (mx - mn) : 1.0f;
		case 3: return valid ? std::pow(dc / mn, -rolloff) : 1.0f;
		default: return 1.0f;
		}
	}
	static float curveShape(int type, float t) /* Curve::EvaluateCurve (inc/Math/Curves.h:58-85) */
	{
		const float halfPi = 3.14159274f * 0.5f;
		switch (type)
		{
		case 0: return 1.0f;
		case 1: return t;
		case 2: return std::sin(t * halfPi);
		case 3: return 1.0f - std::cos(t * halfPi);
		case 4: return t * t * (3 - 2 * t);
		case 5: return std::log(1.0f + 0.41f * t) * (1.0f / std::log(1.41f));
		case 6: return std::log(1.0f + 2.0f * t) * (1.0f / std::log(3.0f));
		case 7: return (std::pow(1.41f, t) - 1.0f) * (1.0f / 0.41f);
		case 8: return (std::pow(3.0f, t) - 1.0f) * 0.5f;
		default: return t;
		}
	}
	float orc_attenuation_curve(const float* pts, uint32_t k, float d) /* AttenuationCurve::Evaluate (:172-215); pts sorted, [k][3] = distance, value, type */
	{
		if (k == 0) return 1.0f;
		if (d <= pts[0]) return pts[1];
		if (d >= pts[3 * (k - 1)]) return pts[3 * (k - 1) + 1];
		uint32_t i = 0;
		while (i < k && !(d < pts[3 * i])) ++i;                 /* first point with d < Distance (std::upper_bound) */
		const float* a = pts + 3 * (i - 1); const float* b = pts + 3 * i;
		if (int(a[2]) == 0) return a[1];
		float t = (d - a[0]) / (b[0] - a[0]);
		t = t < 0.0f ? 0.0f : (1.0f < t ? 1.0f : t);
		return a[1] + (b[1] - a[1]) * curveShape(int(a[2]), t);
	}
	/* JPL::log(simd) one lane, x > 0 (inc/Math/SIMDMath.h:191-226) */
	static float logLane(float x)
	{
		uint32_t u; std::memcpy(&u, &x, 4);
		if (u < 0x00800000u) u = 0x00800000u;
		float e = float(int32_t(u >> 23) - 0x7f) + 1.0f;
		u = (u & ~0x7f800000u) | 0x3f000000u; std::memcpy(&x, &u, 4);
		const bool lo = x < 0.707106781186547524f;
		if (lo) e -= 1.0f;
		x += (lo ? x : 0.0f) - 1.0f;
		float y = 7.0376836292e-2f;
		const float c[8] = { -1.1514610310e-1f, 1.1676998740e-1f, -1.2420140846e-1f, 1.4249322787e-1f, -1.6668057665e-1f, 2.0000714765e-1f, -2.4999993993e-1f, 3.3333331174e-1f };
		for (float ck : c) y = y * x + ck;
		y *= x;
		const float x2 = x * x;
		y = (e * -2.12194440e-4f + y * x2) - x2 * 0.5f;
		return e * 0.693359375f + (x + y);
	}
	float orc_speed_of_sound(float tC) { return 343.2f * std::sqrt((273.15f + tC) * (1.0f / (20.0f + 273.15f))); } /* inc/SpeedOfSound.h:38-47 */
	/* AirAbsorption::CacheParameters (inc/AirAbsorption.h:179-210, 297-368); out [32] like ref_air_cache, knee [4] */
	void orc_air_cache(const float* bands, float tC, float rh, float pa, float nyquist, float* out, uint32_t* knee)
	{
		const float pr = 101.325f, T = tC + 273.15f, T0 = 293.15f, T01 = 273.16f, tr = T / T0, prel = pa / pr;
		const float psat = pr * ::pow(10.0f, -6.8346f * ::pow(T01 / T, 1.261f) + 4.6151f);
		const float h = rh * (psat / pa);
		const float frO = prel * (24.0f + 4.04e4f * h * ((0.02f + h) / (0.391f + h)));
		const float frN = prel * ::pow(tr, -0.5f) * (9.0f + 280.0f * h * ::exp(-4.170f * (::pow(tr, -1.0f / 3.0f) - 1.0f)));
		const float classical = 1.84e-11f * ::pow(prel, -1.0f) * ::sqrt(tr);
		const float scale = ::pow(tr, -2.5f);
		const float cO = 0.01275f * ::exp(-2239.1f / T), cN = 0.1068f * ::exp(-3352.0f / T);
		const float iO = 1.0f / frO, iN = 1.0f / frN;
		auto alpha = [&](float f) { const float f2 = f * f; return (8.68588963807f * f2) * (classical + scale * (cO / (f2 * iO + frO) + cN / (f2 * iN + frN))); };
		const float f3 = bands[3] * bands[3] / nyquist, f2 = bands[2] * bands[2] / f3, f1 = bands[1] * bands[1] / f2;   /* FrequencyBands.h:52-62 */
		const float lo[4] = { 20.0f, f1, f2, f3 }, hi[4] = { f1, f2, f3, 22050.0f };
		for (int i = 0; i < 4; ++i)
		{
			const float third = lo[i] * expLane((1.0f / 6.0f) * logLane(hi[i] / lo[i]));   /* simd pow = exp(y * log x) (SIMDMath.h:372-379) */
			const float ctr = alpha(bands[i]), low = alpha(third), wtd = alpha(bands[i] >= 1000.0f ? third : bands[i]);
			const float rk = (bands[i] >= 1000.0f ? 14.0f : 6.0f) / ctr;
			out[i] = bands[i]; out[4 + i] = third; out[8 + i] = ctr; out[12 + i] = low; out[16 + i] = wtd;
			out[20 + i] = wtd * orc_speed_of_sound(tC); out[24 + i] = rk; out[28 + i] = rk * (ctr - low);
			knee[i] = bands[i] > 400.0f;
		}
	}
	/* AirAbsorption::ComputeForDistance (:212-234) + dBToGain(simd) of the negated loss; cache = orc_air_cache's out / knee */
	void orc_air_for_distance(const float* cache, const uint32_t* knee, float d, float* db, float* gain)
	{
		for (int i = 0; i < 4; ++i)
		{
			const float lin = d * cache[12 + i] + cache[28 + i], ctr = d * cache[8 + i];
			const float loss = (knee[i] && d >= cache[24 + i]) ? lin : ctr;
			if (db) db[i] = loss;
			if (gain) gain[i] = expLane((2.302585092994046f * 0.05f) * (0.0f - loss));
		}
	}
	void orc_sanitize_source_dir(const float* d, float* out) /* Impl::SanitizeSourceDir (inc/Services/PanningService.h:480-495) */
	{
		/* fallbackDir is constant-initialised: 1 / SqrtConstevalFallback<float, 2>(1 + 0.05^2) (inc/Math/Math.h:213-262) */
		const float x = 0.0f * 0.0f + 1.0f * 1.0f + -5e-2f * -5e-2f;
		uint32_t u; std::memcpy(&u, &x, 4); u = 0x5f3759dfu - (u >> 1);
		float y; std::memcpy(&y, &u, 4);
		for (int i = 0; i < 2; ++i) y = y * (1.5f - (0.5f * x) * y * y);
		float r = x * y; r = 0.5f * (r + x / r);
		const float il = 1.0f / r;
		const bool vertical = std::fabs(std::fabs(d[1]) - 1.0f) <= 1e-6f;
		out[0] = vertical ? 0.0f * il : d[0];
		out[1] = std::copysign(vertical ? 1.0f * il : d[1], d[1]);
		out[2] = vertical ? -5e-2f * il : d[2];
	}
	float orc_spread_from_size(int flavour, float size, float distance) /* SpatialManager.h:617-623 / PannerBase.h:69-74 */
	{
		if (distance <= 0.0f) return 1.0f;
		return flavour == 0 ? std::atan((0.5f * size) / distance) * 0.318309886f : atanf((size * 0.5f) / distance) * (0.318309886f * 2.0f);
	}
}
