// Host-side builders (see jplg_host.hpp). Every function restates, in scalar IEEE-754 float arithmetic with
// one rounding per written operation, what the reference computes once per layout / per filter on the CPU.
// Compile WITHOUT FP contraction (-ffp-contract=off) and without -ffast-math.
#include "jplg_host.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <xmmintrin.h>

namespace jplg
{
	static constexpr float kPi = 3.14159265358979323846f;   // std::numbers::pi_v<float>
	static constexpr float kTwoPi = 2.0f * kPi;             // JPL_TWO_PI (inc/Math/Math.h:36)
	static constexpr float kSqrt2 = 1.41421356237309504880f;

	static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
	static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

	//======================================================================
	// ChannelMap (inc/ChannelMap.h:153-249)
	uint32_t NumChannels(uint32_t mask) { return static_cast<uint32_t>(__builtin_popcount(mask)); }
	bool HasLFE(uint32_t mask) { return (mask & CH_LFE) != 0; }
	bool HasTopChannels(uint32_t mask) { return mask >= CH_TOP_Channels; }
	uint32_t ChannelIndex(uint32_t mask, uint32_t channelBit)
	{
		if (channelBit == 0 || (mask & channelBit) != channelBit) return 0xFFFFFFFFu;
		return static_cast<uint32_t>(__builtin_popcount(mask & (channelBit - 1)));
	}
	uint32_t MaskFromNumChannels(uint32_t n)
	{
		switch (n)
		{
		case 1: return CH_FrontCenter;
		case 2: return CH_FrontLeft | CH_FrontRight;
		case 3: return CH_FrontLeft | CH_FrontRight | CH_FrontCenter;
		case 4: return CH_FrontLeft | CH_FrontRight | CH_BackLeft | CH_BackRight;
		case 5: return CH_FrontLeft | CH_FrontRight | CH_LFE | CH_BackLeft | CH_BackRight;
		case 6: return CH_FrontLeft | CH_FrontRight | CH_FrontCenter | CH_LFE | CH_BackLeft | CH_BackRight;
		case 7: return CH_FrontLeft | CH_FrontRight | CH_FrontCenter | CH_LFE | CH_BackLeft | CH_BackRight | CH_BackCenter;
		case 8: return CH_FrontLeft | CH_FrontRight | CH_FrontCenter | CH_LFE | CH_SideLeft | CH_SideRight | CH_BackLeft | CH_BackRight;
		default: return 0;
		}
	}

	// Speaker direction / angle tables (inc/Panning/PannerBase.h:108-186). Indexed by bit position.
	bool ChannelVector(uint32_t bit, V3& o)
	{
		switch (bit)
		{
		case CH_FrontLeft: o = { -0.7071f, 0.0f, -0.7071f }; return true;
		case CH_FrontRight: o = { 0.7071f, 0.0f, -0.7071f }; return true;
		case CH_FrontCenter: o = { 0.0f, 0.0f, -1.0f }; return true;
		case CH_LFE: o = { 0.0f, 0.0f, -1.0f }; return true;
		case CH_BackLeft: o = { -0.7071f, 0.0f, 0.7071f }; return true;
		case CH_BackRight: o = { 0.7071f, 0.0f, 0.7071f }; return true;
		case CH_FrontLeftCenter: o = { -0.3162f, 0.0f, -0.9487f }; return true;
		case CH_FrontRightCenter: o = { 0.3162f, 0.0f, -0.9487f }; return true;
		case CH_BackCenter: o = { 0.0f, 0.0f, 1.0f }; return true;
		case CH_SideLeft: o = { -1.0f, 0.0f, 0.0f }; return true;
		case CH_SideRight: o = { 1.0f, 0.0f, 0.0f }; return true;
		case CH_TopCenter: o = { 0.0f, 1.0f, 0.0f }; return true;
		case CH_TopFrontLeft: o = { -0.5774f, 0.5774f, -0.5774f }; return true;
		case CH_TopFrontCenter: o = { 0.0f, 0.7071f, -0.7071f }; return true;
		case CH_TopFrontRight: o = { 0.5774f, 0.5774f, -0.5774f }; return true;
		case CH_TopSideLeft: o = { -0.7071f, 0.7071f, 0.0f }; return true;
		case CH_TopSideRight: o = { 0.7071f, 0.7071f, 0.0f }; return true;
		case CH_TopBackLeft: o = { -0.5774f, 0.5774f, 0.5774f }; return true;
		case CH_TopBackCenter: o = { 0.0f, 0.7071f, 0.7071f }; return true;
		case CH_TopBackRight: o = { 0.5774f, 0.5774f, 0.5774f }; return true;
		case CH_WideLeft: o = { -0.866f, 0.0f, -0.5f }; return true;
		case CH_WideRight: o = { 0.866f, 0.0f, -0.5f }; return true;
		default: return false;
		}
	}
	bool ChannelAngle(uint32_t bit, float& o)
	{
		switch (bit)
		{
		case CH_FrontLeft: o = -0.785398f; return true;
		case CH_FrontRight: o = 0.785398f; return true;
		case CH_FrontCenter: o = 0.0f; return true;
		case CH_LFE: o = 0.0f; return true;
		case CH_BackLeft: o = -2.35619f; return true;
		case CH_BackRight: o = 2.35619f; return true;
		case CH_FrontLeftCenter: o = -0.321719f; return true;
		case CH_FrontRightCenter: o = 0.321719f; return true;
		case CH_BackCenter: o = 3.14159f; return true;
		case CH_SideLeft: o = -1.5708f; return true;
		case CH_SideRight: o = 1.5708f; return true;
		case CH_TopCenter: o = 3.14159f; return true;
		case CH_TopFrontLeft: o = -0.785398f; return true;
		case CH_TopFrontCenter: o = 0.0f; return true;
		case CH_TopFrontRight: o = 0.785398f; return true;
		case CH_TopSideLeft: o = -1.5708f; return true;
		case CH_TopSideRight: o = 1.5708f; return true;
		case CH_TopBackLeft: o = -2.35619f; return true;
		case CH_TopBackCenter: o = 3.14159f; return true;
		case CH_TopBackRight: o = 2.35619f; return true;
		case CH_WideLeft: o = -1.0472f; return true;
		case CH_WideRight: o = 1.0472f; return true;
		default: return false;
		}
	}

	//======================================================================
	// small vector helpers with the reference's operation order (inc/Math/MinimalVec3.h)
	static inline float Dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
	static inline V3 Cross(const V3& a, const V3& b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
	static inline V3 Scale(const V3& a, float s) { return { a.x * s, a.y * s, a.z * s }; }
	static inline V3 Add(const V3& a, const V3& b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
	static inline float InvSqrtExact(float x) { return 1.0f / std::sqrt(x); } // Math::InvSqrt runtime path (inc/Math/Math.h:283-289)
	static inline V3 Normalized(V3 v) { const float il = InvSqrtExact(Dot(v, v)); return { v.x * il, v.y * il, v.z * il }; }
	static inline float Sign2Scalar(float v) { return std::copysign(1.0f, v); } // inc/Math/Math.h:131-137
	static inline float AbsScalar(float v) { return v < 0 ? -v : v; }          // Math::Abs (:85-88)

	//======================================================================
	// SIMD-flavoured scalar restatements
	static inline int32_t TruncToInt(float v) { return _mm_cvttss_si32(_mm_set_ss(v)); } // _mm_cvttps_epi32 lane

	void SinCosSimd(float xin, float& outS, float& outC) // inc/Math/SIMDMath.h:488-534
	{
		uint32_t sinSign = f2u(xin) & 0x80000000u;
		float x = u2f(f2u(xin) & 0x7fffffffu);
		uint32_t q = static_cast<uint32_t>(TruncToInt(1.27323954473516f * x + 1.0f)) & ~1u;
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
		const uint32_t cosSign = bit1 ^ bit2;
		outS = u2f(f2u(s) ^ sinSign);
		outC = u2f(f2u(c) ^ cosSign);
	}

	static inline float RoundSimd(float v) // SSE2 path of round() (inc/Math/SIMD.h:1930-1937)
	{
		const uint32_t sign = f2u(v) & 0x80000000u;
		return static_cast<float>(TruncToInt(v + u2f(f2u(0.5f) | sign)));
	}

	float TanSimd(float x) // inc/Math/SIMDMath.h:617-665
	{
		const float nf = RoundSimd(x * 0.63661977236758134308f);
		float r = nf * -1.5703125f + x;
		r = nf * -0.0004837512969970703125f + r;
		r = nf * -7.549789948768648e-8f + r;
		const float x2 = r * r;
		float y = 9.38540185543e-3f * x2 + 3.11992232697e-3f;
		y = y * x2 + 2.44301354525e-2f;
		y = y * x2 + 5.34112807005e-2f;
		y = y * x2 + 1.33387994085e-1f;
		y = y * x2 + 3.33331568548e-1f;
		y = y * (x2 * r) + r;
		const int32_t n = TruncToInt(nf);
		return (n & 1) ? (0.0f - (1.0f / y)) : y;
	}

	V2 FromDiamond(float p) // inc/Math/DirectionEncoding.h:212-227
	{
		const float pSign = Sign2Scalar(p - 0.5f);
		const float pSignN = -pSign;
		V2 v;
		v.x = pSignN * (4.0f * p) + (pSign * 2.0f + 1.0f);
		v.y = pSignN * AbsScalar(v.x) + pSign;
		const float il = InvSqrtExact(v.x * v.x + v.y * v.y);
		v.x *= il; v.y *= il;
		return v;
	}

	V3 OctDecode(uint32_t code) // Octahedron16Bit::Decode<Vec3> scalar (inc/Math/DirectionEncoding.h:389-436)
	{
		constexpr float muInv = 1.0f / 255.0f;
		uint32_t dx = code & 255u, dy = (code >> 8) & 255u;
		dx = std::min(dx, 254u); dy = std::min(dy, 254u);
		const float px = -1.0f + 2.0f * (static_cast<float>(dx) + 0.5f) * muInv;
		const float py = -1.0f + 2.0f * (static_cast<float>(dy) + 0.5f) * muInv;
		float d[3] = { px, py, 1.0f - AbsScalar(px) - AbsScalar(py) };
		const float t = std::max(-d[2], 0.0f);
		d[0] += (d[0] > 0.0f) ? -t : t;
		d[1] += (d[1] > 0.0f) ? -t : t;
		return Normalized({ d[0], d[1], d[2] });
	}

	float ExpAlpha(uint32_t ticks) // inc/Utilities/SmoothedValue.h:36-41
	{
		return ticks == 0 ? 1.0f : static_cast<float>(1.0 - std::exp(-1.0 / static_cast<double>(ticks)));
	}

	//======================================================================
	// RSQRTPS model
	float RsqrtHost(float x) { return _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(x))); }

	static inline float RsqrtModel(const uint32_t* table, float x)
	{
		const uint32_t u = f2u(x);
		const uint32_t E = (u >> 23) & 0xffu;
		const uint32_t par = E & 1u;                     // 1: representative exponent 127 ([1,2)); 0: 128 ([2,4))
		const uint32_t idx = (par << 10) | ((u >> 13) & 0x3ffu);
		const int32_t E0 = par ? 127 : 128;
		const int32_t half = (static_cast<int32_t>(E) - E0) / 2; // exact: E-E0 is even
		return u2f(table[idx] - (static_cast<uint32_t>(half) << 23));
	}

	bool BuildRsqrtTable(uint32_t table[2048])
	{
		for (uint32_t par = 0; par < 2; ++par)
			for (uint32_t m = 0; m < 1024; ++m)
			{
				const uint32_t E0 = par ? 127u : 128u;
				table[(par << 10) | m] = f2u(RsqrtHost(u2f((E0 << 23) | (m << 13))));
			}
		// verify the hypothesis: output depends only on exponent parity + top 10 mantissa bits; exponent scaling exact
		bool ok = true;
		uint64_t s = 0x9E3779B97F4A7C15ull;
		for (uint32_t i = 0; i < (1u << 20) && ok; ++i)
		{
			s += 0x9E3779B97F4A7C15ull;
			uint64_t z = s; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
			const uint32_t mant = static_cast<uint32_t>(z) & 0x7fffffu;
			const uint32_t E = 64u + (static_cast<uint32_t>(z >> 32) % 128u); // exponents 2^-63 .. 2^64
			const float x = u2f((E << 23) | mant);
			if (f2u(RsqrtHost(x)) != f2u(RsqrtModel(table, x))) ok = false;
		}
		// exhaustive over the mantissa for the two base exponents' low bits (every 7th value to bound start-up time)
		for (uint32_t par = 0; par < 2 && ok; ++par)
			for (uint32_t m = 0; m < (1u << 23) && ok; m += 7)
			{
				const float x = u2f(((par ? 127u : 128u) << 23) | m);
				if (f2u(RsqrtHost(x)) != f2u(RsqrtModel(table, x))) ok = false;
			}
		return ok;
	}

	//======================================================================
	// Sorted channel angles (inc/Panning/VBAPEx.h:239-295)
	struct ChAngle { float angle; uint32_t id; };

	static void SortedChannelAngles(uint32_t mask, std::vector<ChAngle>& out, bool skipLFE)
	{
		out.clear();
		const bool hasLFE = HasLFE(mask);
		uint32_t index = 0;
		for (uint32_t v = mask; v != 0; v &= (v - 1), ++index)
		{
			const uint32_t bit = v & (0u - v);
			if (skipLFE && bit == CH_LFE) continue;
			if (bit >= CH_TOP_Channels) continue;
			uint32_t channelIndex = index;
			if (bit > CH_LFE && skipLFE && hasLFE) channelIndex--;
			float a = 0.0f;
			ChannelAngle(bit, a);
			out.push_back({ a < 0.0f ? a + kTwoPi : a, channelIndex });
		}
		// std::ranges::sort on <= 16 elements is a plain insertion sort, i.e. stable (LFE and FrontCenter share angle 0)
		std::stable_sort(out.begin(), out.end(), [](const ChAngle& l, const ChAngle& r) { return l.angle < r.angle; });
	}

	//======================================================================
	// 2D LUT (inc/Panning/VBAPLUT2D.h:536-733)
	namespace
	{
		struct Mat2 { V2 c0, c1; };
		struct ChannelPair { Mat2 invL; uint32_t id1, id2; };

		bool TryInverse2(const Mat2& m, Mat2& out) // inc/Math/MinimalMat.h:49-68
		{
			const float det = m.c0.x * m.c1.y - m.c0.y * m.c1.x;
			if (!(AbsScalar(det) >= 1e-6f)) { out = { {1, 0}, {0, 1} }; return false; }
			const float inv = 1.0f / det;
			const V2 r0{ m.c1.y, -m.c1.x }, r1{ -m.c0.y, m.c0.x };
			const V2 R0{ r0.x * inv, r0.y * inv }, R1{ r1.x * inv, r1.y * inv };
			out.c0 = { R0.x, R1.x };
			out.c1 = { R0.y, R1.y };
			return true;
		}

		// cChannelPlaneRatios (inc/ChannelConversion.h:43-78), indexed by bit position
		const float kPlaneRatios[22][6] = {
			{0.5f, 0.0f, 0.5f, 0.0f, 0.0f, 0.0f}, {0.0f, 0.5f, 0.5f, 0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f}, {0, 0, 0, 0, 0, 0},
			{1.0f, 0, 0, 0, 0, 0}, {0, 1.0f, 0, 0, 0, 0},
			{0.25f, 0.0f, 0.75f, 0, 0, 0}, {0.0f, 0.25f, 0.75f, 0, 0, 0},
			{0.5f, 0.0f, 0.0f, 0.5f, 0, 0}, {0.0f, 0.5f, 0.0f, 0.5f, 0, 0}, {0, 0, 0, 1.0f, 0, 0},
			{1.0f, 0, 0, 0, 0, 0}, {0, 1.0f, 0, 0, 0, 0},
			{0, 0, 0, 0, 0, 1.0f}, {0.33f, 0.0f, 0.33f, 0.0f, 0.0f, 0.34f}, {0, 0, 0.5f, 0, 0, 0.5f}, {0.0f, 0.33f, 0.33f, 0.0f, 0.0f, 0.34f},
			{0.5f, 0, 0, 0, 0, 0.5f}, {0, 0.5f, 0, 0, 0, 0.5f},
			{0.33f, 0.0f, 0.0f, 0.33f, 0.0f, 0.34f}, {0, 0, 0, 0.5f, 0, 0.5f}, {0.0f, 0.33f, 0.0f, 0.33f, 0.0f, 0.34f} };

		inline int ChannelPos(uint32_t bit) { return 31 - __builtin_clz(bit); } // ToChannelPos (:98-101) for single-bit values
		inline float RectWeight(int a, int b) // :84-95
		{
			return kPlaneRatios[a][0] * kPlaneRatios[b][0] + kPlaneRatios[a][1] * kPlaneRatios[b][1] + kPlaneRatios[a][2] * kPlaneRatios[b][2] +
				kPlaneRatios[a][3] * kPlaneRatios[b][3] + kPlaneRatios[a][4] * kPlaneRatios[b][4] + kPlaneRatios[a][5] * kPlaneRatios[b][5];
		}
		inline bool IsSpatial(uint32_t bit) { return bit != CH_LFE && bit != CH_FrontCenter && bit != 0; }

		// ComputeChannelConversionRectangularWeights (:163-236): w[out][in]
		void ConversionWeights(uint32_t maskIn, uint32_t maskOut, std::vector<std::vector<float>>& w)
		{
			const uint32_t nIn = NumChannels(maskIn), nOut = NumChannels(maskOut);
			w.assign(nOut, std::vector<float>(nIn, 0.0f));
			auto forEach = [](uint32_t mask, auto&& fn) { uint32_t i = 0; for (uint32_t v = mask; v; v &= v - 1, ++i) fn(v & (0u - v), i); };
			forEach(maskIn, [&](uint32_t cin, uint32_t iin) { forEach(maskOut, [&](uint32_t cout, uint32_t iout) { if (cin == cout) w[iout][iin] = 1.0f; }); });
			forEach(maskIn, [&](uint32_t cin, uint32_t iin)
			{
				if (!IsSpatial(cin) || (maskOut & cin)) return;
				forEach(maskOut, [&](uint32_t cout, uint32_t iout)
				{
					if (!IsSpatial(cout)) return;
					const float weight = RectWeight(ChannelPos(cin), ChannelPos(cout));
					if (w[iout][iin] == 0.0f) w[iout][iin] = weight;
				});
			});
			forEach(maskOut, [&](uint32_t cout, uint32_t iout)
			{
				if (!IsSpatial(cout) || (maskIn & cout)) return;
				forEach(maskIn, [&](uint32_t cin, uint32_t iin)
				{
					if (!IsSpatial(cin)) return;
					const float weight = RectWeight(ChannelPos(cout), ChannelPos(cin));
					if (w[iout][iin] == 0.0f) w[iout][iin] = weight;
				});
			});
		}

		bool Build2D(uint32_t targetMask, PannerData& out, std::string& error)
		{
			const uint32_t numTarget = NumChannels(targetMask);
			const uint32_t quad = CH_FrontLeft | CH_FrontRight | CH_BackLeft | CH_BackRight;
			const uint32_t internalMask = numTarget < 4 ? quad : targetMask;
			const uint32_t lfeIndex = ChannelIndex(internalMask, CH_LFE);
			const bool convert = internalMask != targetMask;

			std::vector<ChAngle> angles;
			SortedChannelAngles(internalMask, angles, /*skipLFE*/ false);
			if (angles.size() < 2) { error = "2D panner needs at least two ground channels"; return false; }

			// ForEachChannelAnglePair (:674-701), restated including its LFE skipping quirk
			std::vector<std::pair<uint32_t, uint32_t>> pairIdx;
			auto sanitize = [&](uint32_t idx) { return idx + (angles[idx].id == lfeIndex ? 1u : 0u); };
			for (uint32_t ch = 0; ch + 1 < angles.size(); ++ch)
			{
				const uint32_t ch1 = sanitize(ch);
				if (ch1 + 1 >= angles.size()) break;   // the reference would read out of bounds here; unreachable:
				const uint32_t ch2 = sanitize(ch1 + 1); // LFE (angle 0) always sorts to position 0 or 1
				if (ch2 >= angles.size()) break;
				pairIdx.emplace_back(ch1, ch2);
			}
			pairIdx.emplace_back(static_cast<uint32_t>(angles.size()) - 1, sanitize(0));

			// ComputePairMatrices (:645-671)
			std::vector<ChannelPair> pairs;
			for (auto [i1, i2] : pairIdx)
			{
				const ChAngle& a1 = angles[i1]; const ChAngle& a2 = angles[i2];
				const float s1 = std::sin(a1.angle), c1 = std::cos(a1.angle);
				const float s2 = std::sin(a2.angle), c2 = std::cos(a2.angle);
				Mat2 m{ {s1, -c1}, {s2, -c2} }, inv;
				TryInverse2(m, inv);
				pairs.push_back({ inv, a1.id, a2.id });
			}

			// FindShortestAperture (:566-583)
			float shortest = std::numeric_limits<float>::max();
			const uint32_t lastId = angles.back().id;
			for (auto [i1, i2] : pairIdx)
			{
				const ChAngle& a1 = angles[i1]; const ChAngle& a2 = angles[i2];
				if (a1.id == lastId) shortest = std::min(shortest, kTwoPi - a1.angle + a2.angle);
				else shortest = std::min(shortest, a2.angle - a1.angle);
			}
			out.shortestAperture = shortest;

			std::vector<std::vector<float>> weights;
			if (convert) ConversionWeights(internalMask, targetMask, weights);

			// BuildForAllDirections (:619-642) + ComputeCellFor (:586-616)
			out.lut2d.assign(size_t(kLUT2DResolution) * numTarget, 0.0f);
			const float step = 1.0f / static_cast<float>(kLUT2DResolution);
			float diamond = 0.0f;
			for (uint32_t pos = 0; pos < kLUT2DResolution; ++pos, diamond += step)
			{
				const V2 dir = FromDiamond(diamond);
				float* row = &out.lut2d[size_t(pos) * numTarget];
				for (const ChannelPair& p : pairs)
				{
					V2 g{ p.invL.c0.x * dir.x + p.invL.c1.x * dir.y, p.invL.c0.y * dir.x + p.invL.c1.y * dir.y };
					if (g.x < 0.0f || g.y < 0.0f) continue;
					const float il = InvSqrtExact(g.x * g.x + g.y * g.y);
					g.x *= il; g.y *= il;
					if (convert)
					{
						// StoreChannelGainsConverted (:704-719) + ApplyChannelConversion (:721-733) + Algo::NormalizeL2
						float sum2 = 0.0f;
						for (uint32_t o = 0; o < numTarget; ++o)
						{
							row[o] = g.x * weights[o][p.id1] + g.y * weights[o][p.id2];
						}
						for (uint32_t o = 0; o < numTarget; ++o) sum2 += row[o] * row[o];
						const float inv = InvSqrtExact(sum2);
						for (uint32_t o = 0; o < numTarget; ++o) row[o] *= inv;
					}
					else
					{
						row[p.id1] = g.x;
						row[p.id2] = g.y;
					}
					break;
				}
			}
			return true;
		}
	}

	//======================================================================
	// 3D LUT (inc/Panning/VBAPLUT3D.h:354-716)
	void SpeakerVertices(uint32_t mask, std::vector<V3>& verts, uint32_t& numDummies, uint32_t& lfeIndex)
	{
		verts.clear(); numDummies = 0; lfeIndex = 0xFFFFFFFFu;
		uint32_t numTop = 0, index = 0;
		for (uint32_t v = mask; v != 0; v &= (v - 1), ++index)
		{
			const uint32_t bit = v & (0u - v);
			if (bit == CH_LFE) { lfeIndex = index; continue; }
			V3 d{}; ChannelVector(bit, d);
			verts.push_back(d);
			numTop += bit >= CH_TOP_Channels;
		}
		verts.push_back({ 0.0f, -1.0f, 0.0f }); ++numDummies;
		if ((numTop == 6 || numTop == 4) && !(mask & CH_TopCenter)) { verts.push_back({ 0.0f, 1.0f, 0.0f }); ++numDummies; }
	}

	namespace
	{
		struct TriTable { uint32_t mask; uint32_t n; uint8_t tris[64 * 3]; };
		const TriTable kTriTables[] = {
#include "jplg_speaker_tris.inc"
		};

		struct Mat3 { V3 c0, c1, c2; };
		bool TryInverse3(const Mat3& m, Mat3& out) // inc/Math/MinimalMat.h:92-117
		{
			const V3 v1 = Cross(m.c1, m.c2), v2 = Cross(m.c2, m.c0), v3 = Cross(m.c0, m.c1);
			const float det = Dot(m.c0, v1);
			if (!(AbsScalar(det) >= 1e-6f)) { out = { {1, 0, 0}, {0, 1, 0}, {0, 0, 1} }; return false; }
			const float inv = 1.0f / det;
			const V3 R0 = Scale(v1, inv), R1 = Scale(v2, inv), R2 = Scale(v3, inv);
			out.c0 = { R0.x, R1.x, R2.x };
			out.c1 = { R0.y, R1.y, R2.y };
			out.c2 = { R0.z, R1.z, R2.z };
			return true;
		}
		inline V3 Transform3(const Mat3& m, const V3& v) { return Add(Add(Scale(m.c0, v.x), Scale(m.c1, v.y)), Scale(m.c2, v.z)); }

		bool Build3D(uint32_t targetMask, PannerData& out, std::string& error, int lutVariant)
		{
			std::vector<V3> verts; uint32_t numDummies = 0, lfeIndex = 0;
			SpeakerVertices(targetMask, verts, numDummies, lfeIndex);
			const uint32_t numReal = static_cast<uint32_t>(verts.size()) - numDummies;
			auto isDummy = [&](uint32_t idx) { return idx >= numReal; };
			// ELUTSize storage variants (inc/Panning/VBAPLUT3D.h:93-104,186-204). The device serves every variant from the same repacked 16-byte cell
			// table, so a variant is what the reference's own query of that variant RETURNS:
			//   KB_786  gains go through Gain24Bit (inc/Utilities/GainEncoding.h:30-62) on their way into the table and back out of it
			//   KB_851  speakers come from the per-triangle dynamic data (ExtractDynamicData :768-800) through a triplet index. The reference stores
			//           that index as SanitizeSpeakerIndex(triI) (:618-620 -> :516-519), i.e. shifted by one for every triangle index >= the LFE's
			//           CHANNEL index, and the last triangle then reads past the end of the table: with an LFE in the layout the variant has no
			//           defined output, without one the index is untouched and the speakers are the triangle's (dummy replaced), unsanitised —
			//           which is the same as sanitised when there is no LFE to skip.
			if (lutVariant == 1 && (targetMask & uint32_t(CH_LFE)))
			{
				error = "LUT variant KB_851 is undefined for layouts with an LFE: the reference indexes its triangle table with a channel-sanitised triplet index (VBAPLUT3D.h:618-620,516-519) and reads past its end";
				return false;
			}
			auto gain24 = [](float g)
			{
				// Gain24Bit(float): w24 = uint32(g * 2^24); w24 -= w24 >> 24; decode: w24 * (1 / (2^24 - 1)). The float -> uint32 conversion of the
				// reference's x86-64 build is a 64-bit truncating convert whose low 32 bits are kept (negative gains wrap instead of clamping)
				const float scaled = g * 16777216.0f;
				uint32_t w = static_cast<uint32_t>(static_cast<int64_t>(scaled));
				w -= (w >> 24);
				const uint32_t stored = w & 0xffffffu;
				return static_cast<float>(stored) * (1.0f / 16777215.0f);
			};

			out.tris.clear(); out.trisTabulated = false;
			if (verts.size() == 3) out.tris = { 0, 1, 2 };
			else
			{
				for (const TriTable& t : kTriTables)
					if (t.mask == (targetMask & ~uint32_t(CH_LFE))) { out.tris.assign(t.tris, t.tris + 3 * t.n); out.trisTabulated = true; break; }
				if (!out.trisTabulated && !TriangulateGeneric(verts, out.tris)) { error = "Failed to triangulate speaker setup."; return false; }
			}
			const uint32_t numTris = static_cast<uint32_t>(out.tris.size() / 3);

			std::vector<Mat3> inv(numTris);
			for (uint32_t t = 0; t < numTris; ++t)
			{
				const Mat3 Lm{ verts[out.tris[3 * t]], verts[out.tris[3 * t + 1]], verts[out.tris[3 * t + 2]] };
				TryInverse3(Lm, inv[t]);
			}

			// FindShortestAperture (:466-480)
			float maxDot = -std::numeric_limits<float>::max();
			for (uint32_t t = 0; t < numTris; ++t)
			{
				const V3& A = verts[out.tris[3 * t]]; const V3& B = verts[out.tris[3 * t + 1]]; const V3& C = verts[out.tris[3 * t + 2]];
				maxDot = std::max({ maxDot, Dot(A, B), Dot(B, C), Dot(C, A) });
			}
			out.shortestAperture = maxDot;

			auto sanitize = [&](uint32_t idx) { return static_cast<uint8_t>(idx + (idx >= lfeIndex ? 1u : 0u)); };

			out.lut3dGains.assign(size_t(kLUT3DCells) * 3, 0.0f);
			out.lut3dSpeakers.assign(size_t(kLUT3DCells) * 3, 0);
			for (uint32_t dy = 0; dy < 256; ++dy)
				for (uint32_t dx = 0; dx < 256; ++dx)
				{
					if (dx == 255 || dy == 255) continue;
					const uint32_t code = (std::min(dy, 254u) << 8) | std::min(dx, 254u);
					const V3 dir = OctDecode(code);
					bool found = false;
					for (uint32_t t = 0; t < numTris && !found; ++t)
					{
						V3 g = Transform3(inv[t], dir);
						const float eps = -1e-6f;
						if (g.x < eps || g.y < eps || g.z < eps) continue;
						if (AbsScalar(Dot(g, g)) <= 1e-6f) continue;
						uint32_t tri[3] = { out.tris[3 * t], out.tris[3 * t + 1], out.tris[3 * t + 2] };
						int dummy = -1;
						if (isDummy(tri[0])) { dummy = 0; g.x = 0.0f; }
						else if (isDummy(tri[1])) { dummy = 1; g.y = 0.0f; }
						else if (isDummy(tri[2])) { dummy = 2; g.z = 0.0f; }
						if (dummy >= 0)
						{
							uint32_t newI = 0; // FindReaplacementForDummy (:803-815)
							while (tri[0] == newI || tri[1] == newI || tri[2] == newI) ++newI;
							tri[dummy] = newI;
						}
						g = Normalized(g);
						float* gp = &out.lut3dGains[size_t(code) * 3];
						gp[0] = g.x; gp[1] = g.y; gp[2] = g.z;
						if (lutVariant == 2) { gp[0] = gain24(g.x); gp[1] = gain24(g.y); gp[2] = gain24(g.z); }
						if (lutVariant == 1 && dummy >= 0) gp[dummy] = 0.0f;    // GainsFor: gains[dummyIndex] = 0 (:850-854); already zero before the normalisation
						uint8_t* sp = &out.lut3dSpeakers[size_t(code) * 3];
						sp[0] = sanitize(tri[0]); sp[1] = sanitize(tri[1]); sp[2] = sanitize(tri[2]);
						found = true;
					}
					if (!found)
					{
						uint8_t* sp = &out.lut3dSpeakers[size_t(code) * 3];
						sp[0] = sp[1] = sp[2] = 255;
					}
				}
			return true;
		}
	}

	// Generic fallback: brute-force convex hull of <= 32 points with coplanar facets merged and fan-triangulated.
	bool TriangulateGeneric(const std::vector<V3>& verts, std::vector<uint8_t>& outTris)
	{
		outTris.clear();
		const int n = static_cast<int>(verts.size());
		if (n < 4 || n > 64) return false;
		auto sub = [](const V3& a, const V3& b) { return V3{ a.x - b.x, a.y - b.y, a.z - b.z }; };
		std::vector<std::vector<int>> faces; // sorted vertex sets already emitted
		const double eps = 1e-5;
		for (int i = 0; i < n; ++i) for (int j = i + 1; j < n; ++j) for (int k = j + 1; k < n; ++k)
		{
			const V3 e1 = sub(verts[j], verts[i]), e2 = sub(verts[k], verts[i]);
			double nx = double(e1.y) * e2.z - double(e1.z) * e2.y, ny = double(e1.z) * e2.x - double(e1.x) * e2.z, nz = double(e1.x) * e2.y - double(e1.y) * e2.x;
			const double len = std::sqrt(nx * nx + ny * ny + nz * nz);
			if (len < 1e-9) continue;
			nx /= len; ny /= len; nz /= len;
			int pos = 0, neg = 0; std::vector<int> on;
			for (int p = 0; p < n; ++p)
			{
				const V3 d = sub(verts[p], verts[i]);
				const double s = nx * d.x + ny * d.y + nz * d.z;
				if (s > eps) ++pos; else if (s < -eps) ++neg; else on.push_back(p);
			}
			if (pos && neg) continue; // not a hull plane
			if (pos) { nx = -nx; ny = -ny; nz = -nz; } // outward normal
			if (std::find(faces.begin(), faces.end(), on) != faces.end()) continue;
			faces.push_back(on);
			// order polygon counter-clockwise around the outward normal
			double cx = 0, cy = 0, cz = 0; for (int p : on) { cx += verts[p].x; cy += verts[p].y; cz += verts[p].z; }
			cx /= on.size(); cy /= on.size(); cz /= on.size();
			const V3 r0 = sub(verts[on[0]], V3{ float(cx), float(cy), float(cz) });
			double ux = r0.x, uy = r0.y, uz = r0.z; const double ul = std::sqrt(ux * ux + uy * uy + uz * uz); ux /= ul; uy /= ul; uz /= ul;
			const double vx = ny * uz - nz * uy, vy = nz * ux - nx * uz, vz = nx * uy - ny * ux;
			std::vector<std::pair<double, int>> ang;
			for (int p : on)
			{
				const double dx = verts[p].x - cx, dy = verts[p].y - cy, dz = verts[p].z - cz;
				ang.emplace_back(std::atan2(dx * vx + dy * vy + dz * vz, dx * ux + dy * uy + dz * uz), p);
			}
			std::sort(ang.begin(), ang.end());
			for (size_t t = 1; t + 1 < ang.size(); ++t)
			{
				outTris.push_back(static_cast<uint8_t>(ang[0].second));
				outTris.push_back(static_cast<uint8_t>(ang[t].second));
				outTris.push_back(static_cast<uint8_t>(ang[t + 1].second));
			}
		}
		return outTris.size() >= 12 && outTris.size() / 3 <= 64;
	}

	bool BuildPanner(uint32_t targetMask, PannerData& out, std::string& error, int lutVariant)
	{
		out = PannerData{};
		out.mask = targetMask;
		out.numChannels = NumChannels(targetMask);
		if (targetMask == 0 || (targetMask & ~uint32_t(CH_ALL))) { error = "Channel map is invalid."; return false; }
		out.is3D = HasTopChannels(targetMask);
		if (out.is3D)
		{
			// Panning3D::IsValidTargetChannelMap (inc/Panning/VBAPanning3D.h:157-166)
			if (out.numChannels < 3) { error = "Trying to initialize with < 3 channels, 3D panning is not possible for such layout."; return false; }
			return Build3D(targetMask, out, error, lutVariant);
		}
		if (lutVariant != 0) { error = "LUT storage variants exist for 3D panners only (VBAPanning3D.h)."; return false; }
		// Panning2D::IsValidTargetChannelMap (inc/Panning/VBAPanning2D.h:124-137)
		if (out.numChannels == 1) { error = "Trying to initialize with single target channel, panning is not possible."; return false; }
		return Build2D(targetMask, out, error);
	}

	//======================================================================
	// Source layouts
	static inline uint32_t RoundUpBy4(uint32_t n) { return (n + 3u) & ~3u; } // inc/Memory/Bits.h:52

	static uint32_t MinPointsOnCapRing(float angularDiameter, float dotM, uint32_t fallback) // inc/Panning/VBAPanning3D.h:213-235
	{
		const float alpha = 0.5f * angularDiameter;
		const float s = std::sin(alpha), c = std::cos(alpha);
		const float s2 = s * s, c2 = c * c;
		if (!(s2 != 0.0)) return 1u;
		const float rhs = std::clamp((dotM - c2) / s2, -1.0f, 1.0f);
		const float dphiMax = std::acos(rhs);
		if (!(dphiMax > 0.0f)) return fallback;
		return static_cast<uint32_t>(std::ceil(kTwoPi / dphiMax));
	}

	bool BuildSourceLayout(const PannerData& panner, uint32_t sourceMask, LayoutData& out, std::string& error)
	{
		out = LayoutData{};
		out.is3D = panner.is3D; out.sourceMask = sourceMask; out.targetMask = panner.mask;
		// IsValidSourceChannelMap (inc/Panning/PannerBase.h:188-203)
		if (sourceMask == 0 || (sourceMask & ~uint32_t(CH_ALL))) { error = "Failed to initialize SourceLayout, provided invalid channel map."; return false; }
		if (HasTopChannels(sourceMask)) { error = "Failed to initialize SourceLayout for VBAPPanner, provided channel map has top channels that don't participate in panning."; return false; }
		const uint32_t numChannels = NumChannels(sourceMask) - (HasLFE(sourceMask) ? 1u : 0u);
		if (numChannels == 0 || numChannels > kMaxSourceChannels) { error = "unsupported number of source channels"; return false; }
		const float aperture = panner.shortestAperture;

		if (!panner.is3D)
		{
			// VBAPanning2D.h:150-193
			uint32_t totalRings = 4;
			if (aperture < std::numeric_limits<float>::max())
				totalRings = static_cast<uint32_t>(std::ceil(kTwoPi / std::max(aperture, 1e-6f)));
			out.numVS = RoundUpBy4(std::max(totalRings / numChannels, 1u));
			if (out.numVS * numChannels > 32) { error = "source layout needs more than 32 virtual sources (reference buffer capacity)"; return false; }
			out.phiTerm = kTwoPi / (out.numVS * numChannels);
			out.offsetTerm = numChannels > 1 ? out.phiTerm * out.numVS * 0.5f : 0.0f;
		}
		else
		{
			// Dimensions::For (VBAPanning3D.h:534-612)
			uint32_t totalRings = 12, samplesPerRing = 24;
			if (aperture < std::numeric_limits<float>::max())
			{
				const float m = std::acos(aperture);
				totalRings = static_cast<uint32_t>(std::ceil(kPi / std::max(m, 1e-6f)));
				if (numChannels <= 2) samplesPerRing = totalRings << 1;
				else
				{
					const float angularDiameter = kTwoPi / numChannels;
					samplesPerRing = MinPointsOnCapRing(angularDiameter, aperture, 24);
				}
			}
			if (RoundUpBy4(samplesPerRing) * totalRings > 288) samplesPerRing = (288u / totalRings) & 0xFFFFFFFCu;
			out.numRings = std::max(1u, static_cast<uint32_t>(std::ceil(totalRings / static_cast<float>(numChannels))));
			out.samplesPerRing = RoundUpBy4(samplesPerRing);
			out.numVS = out.numRings * out.samplesPerRing;
			if (out.samplesPerRing < 4) { error = "unsupported source layout dimensions"; return false; }
			if (size_t(out.numVS) * numChannels > 288) { error = "source layout needs more than 288 virtual sources (reference buffer capacity)"; return false; }
			// VBAPanning3D.h:267-269
			out.phiTerm = numChannels == 1
				? kPi / (out.numRings + 1)
				: kPi / (out.numRings * numChannels + numChannels * 0.5f);
			// PrecomputeAzimuth (:312-346)
			const uint32_t numSIMDOps = out.samplesPerRing >> 2;
			const uint32_t toMirror = numSIMDOps & ~1u, tail = numSIMDOps & 1u;
			const uint32_t halfAndTail = (toMirror >> 1) + tail;
			const float thetaTerm = kTwoPi / out.samplesPerRing;
			const float ramp[4] = { 0.5f, 1.5f, 2.5f, 3.5f };
			float theta[4]; for (int l = 0; l < 4; ++l) theta[l] = thetaTerm * ramp[l];
			const float thetaDelta = thetaTerm * 4.0f;
			out.azSin.resize(size_t(halfAndTail) * 4); out.azCos.resize(size_t(halfAndTail) * 4);
			for (uint32_t i = 0; i < halfAndTail; ++i)
				for (int l = 0; l < 4; ++l)
				{
					SinCosSimd(theta[l], out.azSin[4 * i + l], out.azCos[4 * i + l]);
					theta[l] += thetaDelta;
				}
		}

		// VBAPLayoutBase::InitializeBase (PannerBase.h:922-985)
		out.numChannels = numChannels;
		out.vsWeight = 1.0f / out.numVS;
		std::vector<ChAngle> sorted;
		SortedChannelAngles(sourceMask, sorted, /*skipLFE*/ true);
		const float channelWidth = kTwoPi / static_cast<float>(numChannels);
		const float channelAngleOffset = !(sourceMask & CH_FrontCenter) ? 0.5f * channelWidth : 0.0f;
		struct Group { float q[4]; uint32_t id; };
		std::vector<Group> groups(numChannels);
		for (uint32_t i = 0; i < sorted.size() && i < numChannels; ++i)
		{
			const float channelAngle = channelAngleOffset + channelWidth * i;
			const float halfAngle = 0.5f * (-channelAngle); // Quat::Rotation (inc/Math/MinimalQuat.h:58-63)
			const float s = std::sin(halfAngle), c = std::cos(halfAngle);
			groups[i] = { { 0.0f * s, 1.0f * s, 0.0f * s, c }, sorted[i].id };
		}
		std::stable_sort(groups.begin(), groups.end(), [](const Group& l, const Group& r) { return l.id < r.id; });
		for (uint32_t i = 0; i < numChannels; ++i) { std::memcpy(out.quats[i], groups[i].q, sizeof(float) * 4); out.channelIds[i] = groups[i].id; }
		return true;
	}

	//======================================================================
	void CrossoverCoefficients(float sampleRate, float f1, float f2, float f3, float out[27])
	{
		// LR4Split::Prepare (inc/Auralization/CrossoverFilter.h:253-258) + StateVariableFilterTPT::Prepare (:50-61)
		{
			const float frequency = std::clamp(f2, 1.0f, 0.49f * sampleRate);
			const float g = std::tan(kPi * (frequency / sampleRate));
			out[0] = g;
			out[1] = 1.0f / (g * (kSqrt2 + g) + 1.0f);
			out[2] = -(kSqrt2 + g);
		}
		// LR4Split2Lanes::Prepare (:343-362) + TPTSVFStageSIMD::Prepare (:139-150)
		const float hi = 0.49f * sampleRate;
		const float fr[4] = { f1, f3, f1, f3 };
		float freq[4];
		for (int l = 0; l < 4; ++l)
		{
			const float mn = fr[l] < hi ? fr[l] : hi;     // _mm_min_ps(value, maxV)
			freq[l] = mn > 1.0f ? mn : 1.0f;              // _mm_max_ps(.., minV)
		}
		auto prep = [&](const float* f, float* g, float* d, float* mrg)
		{
			for (int l = 0; l < 4; ++l)
			{
				g[l] = TanSimd(kPi * (f[l] / sampleRate));
				d[l] = 1.0f / (g[l] * (kSqrt2 + g[l]) + 1.0f);
				mrg[l] = 0.0f - (kSqrt2 + g[l]);
			}
		};
		prep(freq, out + 3, out + 7, out + 11);
		const float rev[4] = { freq[3], freq[2], freq[1], freq[0] };
		prep(rev, out + 15, out + 19, out + 23);
	}
	// State-transition matrix of the whole FourBandCrossover over T samples of zero input (the time-parallel throughput path, jplg_scan_kernels.cu):
	// M[r][k] = state r after T zero-input samples started from the unit state k, state order as ST_FILTER (jplg_kernels.hpp). The network is
	// linear and time-invariant (CrossoverFilter.h:64-76,294-310,371-401), so the state after a tile is M * (state before) + (zero-state response).
	// Evaluated in double from the float coefficients, rounded once.
	void CrossoverTransition(const float coef[27], uint32_t T, float M[24 * 24])
	{
		double cf[27];
		for (int i = 0; i < 27; ++i) cf[i] = coef[i];
		const double minusR = -1.41421356237309504880;
		for (int k = 0; k < 24; ++k)
		{
			double st[24] = {};
			st[k] = 1.0;
			auto svf = [&](int c, int stride, double x, int i, double& yL, double& yB, double& yH)
			{
				const double g = cf[c], d = cf[c + stride], mrg = cf[c + 2 * stride];
				yH = (mrg * st[i] + (x - st[i + 1])) * d;
				yB = g * yH + st[i]; st[i] = g * yH + yB;
				yL = g * yB + st[i + 1]; st[i + 1] = g * yB + yL;
			};
			for (uint32_t n = 0; n < T; ++n)
			{
				double yL, yB, yH, yL2, yB2, yH2;
				svf(0, 1, 0.0, 0, yL, yB, yH);
				svf(0, 1, yL, 2, yL2, yB2, yH2);
				const double low = yL2, high = minusR * yB + (yL + yH - yL2);
				double aL, aB, aH, bL, bB, bH, t0, t1, l0, l1, h2, h3;
				svf(3, 4, low, 4, aL, aB, aH);
				svf(4, 4, high, 6, bL, bB, bH);
				svf(3, 4, aL, 8, l0, t0, t1);
				svf(4, 4, bL, 10, l1, t0, t1);
				svf(5, 4, aH, 12, t0, t1, h2);
				svf(6, 4, bH, 14, t0, t1, h3);
				const double o[4] = { l0, l1, h2, h3 };
				for (int q = 0; q < 4; ++q) { double pL, pB, pH; svf(15 + q, 4, o[q], 16 + 2 * q, pL, pB, pH); }
			}
			for (int r = 0; r < 24; ++r) M[r * 24 + k] = static_cast<float>(st[r]);
		}
	}
	//======================================================================
	// Scene front-end builders
	namespace
	{
		inline uint32_t FBits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
		inline float BitsF(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

		// JPL::exp(simd), one lane, SSE2 semantic: fma = multiply then add (inc/Math/SIMDMath.h:296-316; floor inc/Math/SIMD.h:1861-1879)
		float ExpSimdLane(float x)
		{
			x = std::min(std::max(x, -88.3762626647949f), 88.3762626647949f);
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
			return y * BitsF(static_cast<uint32_t>((static_cast<int32_t>(n) + 0x7f) << 23));
		}
		// JPL::log(simd), one lane, for x > 0 (inc/Math/SIMDMath.h:191-226)
		float LogSimdLane(float x)
		{
			x = std::max(x, BitsF(0x00800000u));
			const int32_t emm0 = static_cast<int32_t>(FBits(x) >> 23) - 0x7f;
			float e = static_cast<float>(emm0) + 1.0f;
			x = BitsF((FBits(x) & ~0x7f800000u) | FBits(0.5f));
			const bool below = x < 0.707106781186547524f;
			e -= below ? 1.0f : 0.0f;
			x += (below ? x : 0.0f) - 1.0f;
			float y = 7.0376836292e-2f;
			y = y * x + -1.1514610310e-1f;
			y = y * x + 1.1676998740e-1f;
			y = y * x + -1.2420140846e-1f;
			y = y * x + 1.4249322787e-1f;
			y = y * x + -1.6668057665e-1f;
			y = y * x + 2.0000714765e-1f;
			y = y * x + -2.4999993993e-1f;
			y = y * x + 3.3333331174e-1f;
			y *= x;
			const float x2 = x * x;
			y = (e * -2.12194440e-4f + y * x2) - x2 * 0.5f;
			return e * 0.693359375f + (x + y);
		}
		// pow(simd, simd) for finite x > 0, x != 1, finite y != 0: exp(y * log(|x|)) (SIMDMath.h:372-379; the rest of the function only patches special values)
		float PowSimdLane(float x, float y) { return ExpSimdLane(y * LogSimdLane(x)); }
		// round(simd), SSE2 path: trunc(x + copysign(0.5, x)) (inc/Math/SIMD.h:1934-1937)
		float RoundSimdLane(float x) { return static_cast<float>(static_cast<int32_t>(x + BitsF(FBits(0.5f) | (FBits(x) & 0x80000000u)))); }

		void LowerThirdCenters(float f1, float f2, float f3, float out[4])
		{
			const float lower[4] = { 20.0f, f1, f2, f3 }, upper[4] = { f1, f2, f3, 22050.0f };
			for (int i = 0; i < 4; ++i) out[i] = lower[i] * PowSimdLane(upper[i] / lower[i], 1.0f / 6.0f);
		}
	}

	void DefaultBandCenters(float out[4])
	{
		const float lower[4] = { 20.0f, 176.0f, 775.0f, 3408.0f }, upper[4] = { 176.0f, 775.0f, 3408.0f, 22050.0f };
		for (int i = 0; i < 4; ++i) out[i] = RoundSimdLane(std::sqrt(lower[i] * upper[i]));
	}

	float SpeedOfSoundForTemperature(float temperatureC)
	{
		constexpr float zeroK = 273.15f;
		constexpr float invRefTemp = 1.0f / (20.0f + zeroK);
		return 343.2f * std::sqrt((zeroK + temperatureC) * invRefTemp);
	}

	void BuildAirAbsorptionCache(const float bands[4], float temperatureC, float humidityPercent, float pressureKPa, float nyquist, AirCacheHost& out)
	{
		// ---- AirAbsorption ctor (ISO 9613-1 terms). The reference calls the C functions ::pow / ::exp / ::sqrt, i.e. the DOUBLE versions, on float
		// arguments; the mixed float / double promotions below follow its expressions term by term.
		const float pRef = 101.325f;
		const float tK = temperatureC + 273.15f;
		const float tRef = 293.15f;
		const float tTriple = 273.16f;
		const float tRel = tK / tRef;
		const float pRel = pressureKPa / pRef;
		const float pSat = pRef * ::pow(10.0f, -6.8346f * ::pow(tTriple / tK, 1.261f) + 4.6151f);
		const float h = humidityPercent * (pSat / pressureKPa);
		const float frO = pRel * (24.0f + 4.04e4f * h * ((0.02f + h) / (0.391f + h)));
		const float frN = pRel * ::pow(tRel, -0.5f) * (9.0f + 280.0f * h * ::exp(-4.170f * (::pow(tRel, -1.0f / 3.0f) - 1.0f)));
		const float classical = 1.84e-11f * ::pow(pRel, -1.0f) * ::sqrt(tRel);
		const float relaxScale = ::pow(tRel, -2.5f);
		const float cO = 0.01275f * ::exp(-2239.1f / tK);
		const float cN = 0.1068f * ::exp(-3352.0f / tK);
		const float frOInv = 1.0f / frO, frNInv = 1.0f / frN;
		auto perMeter = [&](float f) // ComputeAbsorptionPerMeter(simd) lane; Math::FMA = multiply, add in the SSE2 build
		{
			const float f2 = f * f;
			const float o = cO / (f2 * frOInv + frO);
			const float n = cN / (f2 * frNInv + frN);
			const float mol = relaxScale * (o + n);
			const float total = classical + mol;
			return (8.68588963807f * f2) * total;
		};
		// ---- CacheParameters
		float c2[4];
		for (int i = 0; i < 4; ++i) c2[i] = bands[i] * bands[i];
		const float f3 = c2[3] / nyquist, f2 = c2[2] / f3, f1 = c2[1] / f2;   // SplitFrequenciesFromBandCenters
		LowerThirdCenters(f1, f2, f3, out.lowerThirdCenters);
		for (int i = 0; i < 4; ++i)
		{
			out.centers[i] = bands[i];
			const float weighted = bands[i] >= 1000.0f ? out.lowerThirdCenters[i] : bands[i];
			out.centerDb[i] = perMeter(bands[i]);
			out.lowerThirdDb[i] = perMeter(out.lowerThirdCenters[i]);
			out.weightedDb[i] = perMeter(weighted);
			out.hasKnee[i] = bands[i] > 400.0f ? 1u : 0u;
			const float deltaLk = bands[i] >= 1000.0f ? 14.0f : 6.0f;
			out.distanceToKnee[i] = deltaLk / out.centerDb[i];
			out.linFactor[i] = out.distanceToKnee[i] * (out.centerDb[i] - out.lowerThirdDb[i]);
			out.decaySlope[i] = out.weightedDb[i] * SpeedOfSoundForTemperature(temperatureC);
		}
	}

	void SanitizeFallbackDir(float out[3])
	{
		// Normalized(Vec3(0, 1, -5e-2)) with Math::InvSqrt evaluated at compile time: 1 / SqrtConstevalFallback<float, 2>(x)
		const float x = 0.0f * 0.0f + 1.0f * 1.0f + -5e-2f * -5e-2f;
		float y = BitsF(0x5f3759dfu - (FBits(x) >> 1));
		const float halfx = 0.5f * x;
		for (int i = 0; i < 2; ++i) y = y * (1.5f - halfx * y * y);
		float r = x * y;
		r = 0.5f * (r + x / r);
		const float inv = 1.0f / r;
		out[0] = 0.0f * inv; out[1] = 1.0f * inv; out[2] = -5e-2f * inv;
	}
}
