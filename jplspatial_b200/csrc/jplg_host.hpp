// Host-side builders of the B200 spatial renderer: everything the reference computes ONCE per layout /
// per filter on the CPU (LUTs, source-layout constants, crossover coefficients, smoothing alpha) plus the
// RSQRTPS model table the device needs. Plain scalar IEEE float code; compiled without FP contraction.
// Reference paths: inc/ = Spatialization/include/JPLSpatial/.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

namespace jplg
{
	struct V3 { float x, y, z; };
	struct V2 { float x, y; };

	// EChannel bits (inc/ChannelMap.h:38-79)
	enum : uint32_t
	{
		CH_FrontLeft = 1u << 0, CH_FrontRight = 1u << 1, CH_FrontCenter = 1u << 2, CH_LFE = 1u << 3,
		CH_SideLeft = 1u << 4, CH_SideRight = 1u << 5, CH_FrontLeftCenter = 1u << 6, CH_FrontRightCenter = 1u << 7,
		CH_BackLeft = 1u << 8, CH_BackRight = 1u << 9, CH_BackCenter = 1u << 10, CH_WideLeft = 1u << 11, CH_WideRight = 1u << 12,
		CH_TopCenter = 1u << 13, CH_TopFrontLeft = 1u << 14, CH_TopFrontCenter = 1u << 15, CH_TopFrontRight = 1u << 16,
		CH_TopSideLeft = 1u << 17, CH_TopSideRight = 1u << 18, CH_TopBackLeft = 1u << 19, CH_TopBackCenter = 1u << 20, CH_TopBackRight = 1u << 21,
		CH_TOP_Channels = CH_TopCenter, CH_ALL = (1u << 22) - 1
	};

	constexpr uint32_t kMaxChannels = 32;        // VBAPBaseTraits::MAX_CHANNELS (inc/Panning/PannerBase.h:101)
	constexpr uint32_t kMaxSourceChannels = 8;   // MAX_SOURCE_CHANNELS (:103)
	constexpr uint32_t kLUT2DResolution = 1024;  // LUT2D::LUTStats::Resolution (inc/Panning/VBAPLUT2D.h:93)
	constexpr uint32_t kLUT3DCells = 65536;      // Octahedron16Bit::cAxisRange^2 (inc/Panning/VBAPanning3D.h:77-78)

	uint32_t NumChannels(uint32_t mask);
	bool HasLFE(uint32_t mask);
	bool HasTopChannels(uint32_t mask);
	uint32_t ChannelIndex(uint32_t mask, uint32_t channelBit);
	uint32_t MaskFromNumChannels(uint32_t n);
	bool ChannelVector(uint32_t channelBit, V3& out); // VBAPBaseTraits::GetChannelVector (PannerBase.h:150-186)
	bool ChannelAngle(uint32_t channelBit, float& out); // GetChannelAngle (:108-147)

	//======================================================================
	struct PannerData
	{
		uint32_t mask = 0;
		bool is3D = false;
		uint32_t numChannels = 0;
		float shortestAperture = 0.0f;
		// 2D: [1024][L]
		std::vector<float> lut2d;
		// 3D: gains [65536][3], speakers [65536][3]
		std::vector<float> lut3dGains;
		std::vector<uint8_t> lut3dSpeakers;
		std::vector<uint8_t> tris; // [n][3], vertex numbering = non-LFE channel order then dummies
		bool trisTabulated = false;
	};

	// VBAPannerBase::Initialize (PannerBase.h:455-482). Returns false + message like the reference returns false + JPL_ERROR_TAG.
	// lutVariant: 0 KB_983 (the reference's default), 1 KB_851, 2 KB_786 (ELUTSize, inc/Panning/VBAPLUT3D.h:93-104); 3D panners only
	bool BuildPanner(uint32_t targetMask, PannerData& out, std::string& error, int lutVariant = 0);

	// Generic speaker triangulation used for masks that are not in the reference-derived table.
	bool TriangulateGeneric(const std::vector<V3>& verts, std::vector<uint8_t>& outTris);
	// Speaker vertices exactly as LUTBuilder's ctor collects them (VBAPLUT3D.h:354-403): non-LFE channels, dummy (0,-1,0), optional TopCenter dummy
	void SpeakerVertices(uint32_t mask, std::vector<V3>& verts, uint32_t& numDummies, uint32_t& lfeIndex);

	//======================================================================
	struct LayoutData
	{
		bool is3D = false;
		uint32_t sourceMask = 0, targetMask = 0;
		uint32_t numChannels = 0;    // ChannelGroups.size()
		uint32_t numVS = 0;          // virtual sources per channel
		uint32_t numRings = 0, samplesPerRing = 0; // 3D
		float phiTerm = 0.0f, offsetTerm = 0.0f, vsWeight = 0.0f;
		float quats[kMaxSourceChannels][4] = {}; // channel-group rotations {x,y,z,w}, sorted by channel id
		uint32_t channelIds[kMaxSourceChannels] = {};
		std::vector<float> azSin, azCos; // 3D cached azimuth packs (4 floats each)
	};

	// InitializeSourceLayout (VBAPanning2D.h:150-193 / VBAPanning3D.h:242-310 + PannerBase.h:922-985)
	bool BuildSourceLayout(const PannerData& panner, uint32_t sourceMask, LayoutData& out, std::string& error);

	//======================================================================
	// FourBandCrossover::Prepare (inc/Auralization/CrossoverFilter.h:425-431): 27 floats
	// X2{g,d,-(R+g)}, X13{g[4],d[4],mrg[4]} (lanes f1,f3,f1,f3), allpass{g[4],d[4],mrg[4]} (lanes f3,f1,f3,f1)
	void CrossoverCoefficients(float sampleRate, float f1, float f2, float f3, float out[27]);
	void CrossoverTransition(const float coef[27], uint32_t T, float M[24 * 24]);   // zero-input state transition over T samples (time-parallel path)

	// SmoothedValue<float>::GetExpAlpha (inc/Utilities/SmoothedValue.h:36-41)
	float ExpAlpha(uint32_t ticks);

	//======================================================================
	// RSQRTPS model: table[(expParity<<10) | top10MantissaBits] = bits of _mm_rsqrt_ss(x0) for the representative
	// x0 in [1,4); other exponents by exact scaling. Returns true when the model reproduced the host instruction
	// on all probed inputs.
	bool BuildRsqrtTable(uint32_t table[2048]);
	float RsqrtHost(float x); // the host instruction itself

	// SIMD-flavoured math restated in scalar form (used by host-side builders and the CPU checks)
	void SinCosSimd(float x, float& s, float& c);  // inc/Math/SIMDMath.h:488-534
	float TanSimd(float x);                         // :643-665
	V2 FromDiamond(float p);                        // inc/Math/DirectionEncoding.h:212-227
	V3 OctDecode(uint32_t code);                    // :389-436 (scalar Decode<Vec3>)
	//======================================================================
	// Scene front-end builders (host, once per environment, like the reference)
	// AirAbsorption::CacheParameters (inc/AirAbsorption.h:179-210, ctor :297-343, ComputeAbsorptionPerMeter(simd) :358-368,
	// band helpers inc/FrequencyBands.h:40-66 incl. the SIMD pow = exp(y * log(x)) of inc/Math/SIMDMath.h:191-226,296-316,372-379)
	struct AirCacheHost
	{
		float centers[4], lowerThirdCenters[4];
		float centerDb[4], lowerThirdDb[4], weightedDb[4], decaySlope[4];
		float distanceToKnee[4], linFactor[4];
		uint32_t hasKnee[4];
	};
	void BuildAirAbsorptionCache(const float bands[4], float temperatureC, float humidityPercent, float pressureKPa, float nyquist, AirCacheHost& out);
	void DefaultBandCenters(float out[4]);               // cBandCenters = round(ComputeBandCenters(cDefaultFrequencySplits)) (FrequencyBands.h:68)
	float SpeedOfSoundForTemperature(float temperatureC); // SpeedOfSound::ForTemperature (inc/SpeedOfSound.h:38-47)
	// Impl::SanitizeSourceDir's `static const fallbackDir = Normalized(Vec3(0, 1, -5e-2))` (inc/Services/PanningService.h:487) is constant-initialised,
	// i.e. evaluated by the reference's constexpr sqrt fallback (inc/Math/Math.h:213-262,283-289), not by the runtime 1/sqrt: restated here.
	void SanitizeFallbackDir(float out[3]);
}
