// TEST INFRASTRUCTURE — not part of the product.
//
// One translation unit that holds BOTH sides of the drop-in boundary: the unmodified reference (its public headers, plus DirectSound.cpp compiled
// from where it lies) and include/jplg_shim.hpp over libjplg.so. It builds the same JPL::StandardPanner3D::PanUpdateData / DirectSoundEffect
// arguments an engine would, runs them through the reference objects and through the shim, and compares the results bit for bit. This is the
// proof that the adapter a maintainer would add compiles against the reference's real types and returns the reference's values.
//
//   shim_check -1      host-only context: everything that needs no device (LUT / layout construction through the shim) must succeed and every
//                      compute call must fail with JPLG_ERR_NO_DEVICE (no CPU fallback). Run by tests/test_shim_cpu.py.
//   shim_check 0       on the GPU box (prebuilt binary, oracle/_ref/shim_check): full comparison. Run by tests/test_gpu_shim.py.
#include <JPLSpatial/ChannelMap.h>
#include <JPLSpatial/Panning/VBAPanning3D.h>
#include <JPLSpatial/Auralization/DirectSound.h>
#include <JPLSpatial/Auralization/ChannelMixing.h>
#include <JPLSpatial/Math/MinimalVec3.h>

#include "jplg_shim.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

namespace
{
	using Panner = JPL::StandardPanner3D;
	using Update = Panner::PanUpdateData;
	using UpdateOriented = Panner::PanUpdateDataWithOrientation;

	uint32_t gState = 0x1234567u;
	float Uniform() { gState = gState * 1664525u + 1013904223u; return float(gState >> 8) * (1.0f / 16777216.0f); }

	int gFailures = 0;
	void Expect(bool ok, const char* what)
	{
		std::printf("%s %s\n", ok ? "ok  " : "FAIL", what);
		if (!ok) ++gFailures;
	}
	bool SameBits(const float* a, const float* b, size_t n) { return std::memcmp(a, b, n * sizeof(float)) == 0; }
}

int main(int argc, char** argv)
{
	const int device = argc > 1 ? std::atoi(argv[1]) : -1;
	constexpr uint32_t N = 96, C = 2, frames = 512, blocks = 3;
	const auto target = JPL::ChannelMap::FromChannelMask(JPL::ChannelMask::Surround_7_1_4);
	const auto source = JPL::ChannelMap::FromChannelMask(JPL::ChannelMask::Stereo);

	// ---- reference side
	Panner refPanner;
	Expect(refPanner.Initialize(target), "reference panner Initialize(7.1.4)");
	Panner::SourceLayoutType refLayout;
	Expect(refPanner.InitializeSourceLayout(source, refLayout), "reference InitializeSourceLayout(Stereo)");
	const uint32_t L = refPanner.GetNumChannels();

	// ---- shim side
	JPLG::Context ctx(device);
	Expect(ctx.IsValid(), "JPLG::Context");
	JPLG::BatchedPanner<Panner> pan(ctx);
	Expect(pan.Initialize(target), "shim panner Initialize(7.1.4)");
	Expect(pan.InitializeSourceLayout(source), "shim InitializeSourceLayout(Stereo)");
	Expect(pan.GetNumChannels() == L && pan.GetChannelMap().GetChannelMask() == target.GetChannelMask(), "GetNumChannels / GetChannelMap agree");
	Expect(pan.GetNumSourceChannels() == C, "two panned source channels");

	// the arguments an engine builds: the reference's own nested structs
	std::vector<Update> updates(N * (blocks + 1));
	std::vector<UpdateOriented> oriented(N);
	for (auto& u : updates)
	{
		float x = Uniform() * 2.0f - 1.0f, y = (Uniform() * 2.0f - 1.0f) * 0.8f, z = Uniform() * 2.0f - 1.0f;
		const float inv = 1.0f / std::sqrt(x * x + y * y + z * z + 1e-6f);
		u = Update{ .SourceDirection = JPL::MinimalVec3(x * inv, y * inv, z * inv), .Focus = Uniform(), .Spread = Uniform() };
	}
	for (uint32_t i = 0; i < N; ++i)
		oriented[i] = UpdateOriented{ .Pan = updates[i], .Orientation = { .Up = JPL::MinimalVec3(0.0f, 1.0f, 0.0f), .Forward = JPL::MinimalVec3(0.0f, 0.0f, -1.0f) } };

	const size_t mapSize = size_t(C) * L;
	std::vector<float> refMaps(updates.size() * mapSize, 0.0f), shimMaps(updates.size() * mapSize, -1.0f);
	for (size_t i = 0; i < updates.size(); ++i)
		refPanner.ProcessVBAPData(refLayout, updates[i], std::span<float>(&refMaps[i * mapSize], mapSize));
	const jplg_status st = pan.ProcessVBAPData(std::span<const Update>(updates), shimMaps);

	if (device < 0)
	{
		Expect(st == JPLG_ERR_NO_DEVICE, "host-only context: ProcessVBAPData fails with JPLG_ERR_NO_DEVICE (no CPU fallback)");
		std::vector<float> one(mapSize * N);
		Expect(pan.ProcessVBAPData(std::span<const UpdateOriented>(oriented), one) == JPLG_ERR_NO_DEVICE, "host-only context: oriented ProcessVBAPData fails too");
		std::vector<float> gains(N * 4, 1.0f), delays(N, 0.0f);
		JPLG::DirectSoundBatch batch(ctx, 48000.0f, N, C, L, gains, delays, std::span<const float>(refMaps.data(), N * mapSize));
		Expect(batch.Status() == JPLG_ERR_NO_DEVICE, "host-only context: DirectSoundBatch cannot be created");
		float d[4] = { 0, 0, 0, 0 }; const float s[4] = { 1, 2, 3, 4 };
		Expect(JPLG::Mixing::Add(ctx, d, s, 1, 0, 1, 4) == JPLG_ERR_NO_DEVICE, "host-only context: Mixing::Add fails");
		std::printf("shim_check: %d failure(s) (host-only)\n", gFailures);
		return gFailures ? 1 : 0;
	}

	Expect(st == JPLG_OK, "shim ProcessVBAPData status");
	Expect(SameBits(refMaps.data(), shimMaps.data(), refMaps.size()), "ProcessVBAPData(PanUpdateData): shim == reference, bit for bit");
	{
		std::vector<float> a(N * mapSize, 0.0f), b(N * mapSize, -1.0f);
		for (uint32_t i = 0; i < N; ++i) refPanner.ProcessVBAPData(refLayout, oriented[i], std::span<float>(&a[i * mapSize], mapSize));
		Expect(pan.ProcessVBAPData(std::span<const UpdateOriented>(oriented), b) == JPLG_OK && SameBits(a.data(), b.data(), a.size()),
			   "ProcessVBAPData(PanUpdateDataWithOrientation): shim == reference, bit for bit");
	}

	// ---- DirectSoundEffect x N against DirectSoundBatch. Per-source bit equality needs each source on its own bus, so the batch is split into N
	// batches of one source for the exact check, and one batch of all N is checked against the fp32 summation bound.
	std::vector<float> gains((blocks + 1) * N * 4), delays((blocks + 1) * N), audio(size_t(blocks) * N * frames * C);
	for (auto& g : gains) g = 0.1f + 0.9f * Uniform();
	for (auto& d : delays) d = (1.0f + 299.0f * Uniform()) / 343.2f;
	for (auto& a : audio) a = Uniform() * 2.0f - 1.0f;

	std::vector<std::unique_ptr<JPL::DirectSoundEffect>> effects;
	std::vector<std::unique_ptr<JPLG::DirectSoundBatch>> singles;
	for (uint32_t i = 0; i < N; ++i)
	{
		const std::span<const float> map(&refMaps[i * mapSize], mapSize);
		effects.push_back(std::make_unique<JPL::DirectSoundEffect>(48000.0f, C, L, JPL::simd(&gains[i * 4]), delays[i], map));
		if (i < 8) singles.push_back(std::make_unique<JPLG::DirectSoundBatch>(ctx, 48000.0f, 1u, C, L, std::span<const float>(&gains[i * 4], 4), std::span<const float>(&delays[i], 1), map));
	}
	JPLG::DirectSoundBatch batch(ctx, 48000.0f, N, C, L, std::span<const float>(gains.data(), N * 4), std::span<const float>(delays.data(), N),
								 std::span<const float>(refMaps.data(), N * mapSize));
	Expect(batch.Status() == JPLG_OK, "DirectSoundBatch created");

	bool exact = true, bounded = true;
	for (uint32_t b = 0; b < blocks; ++b)
	{
		const float* g = &gains[size_t(b + 1) * N * 4];
		const float* d = &delays[size_t(b + 1) * N];
		const float* m = &refMaps[size_t(b + 1) * N * mapSize];
		const float* in = &audio[size_t(b) * N * frames * C];
		std::vector<double> sum(size_t(frames) * L, 0.0), mag(size_t(frames) * L, 0.0);
		for (uint32_t i = 0; i < N; ++i)
		{
			effects[i]->UpdateParameters(JPL::simd(g + i * 4), d[i], std::span<const float>(m + i * mapSize, mapSize));
			std::vector<float> out(size_t(frames) * L, 0.0f);
			effects[i]->ProcessInterleaved(std::span<const float>(in + size_t(i) * frames * C, size_t(frames) * C), out, frames);
			for (size_t k = 0; k < out.size(); ++k) { sum[k] += out[k]; mag[k] += std::fabs(out[k]); }
			if (i < singles.size())
			{
				// the reference's one-source signature, then Commit()
				singles[i]->UpdateParameters(0u, JPL::simd(g + i * 4), d[i], std::span<const float>(m + i * mapSize, mapSize));
				singles[i]->Commit();
				std::vector<float> one(size_t(frames) * L, 0.0f);
				singles[i]->ProcessInterleaved(std::span<const float>(in + size_t(i) * frames * C, size_t(frames) * C), one, frames);
				exact = exact && SameBits(out.data(), one.data(), out.size());
			}
		}
		batch.UpdateParameters(std::span<const float>(g, N * 4), std::span<const float>(d, N), std::span<const float>(m, N * mapSize));
		std::vector<float> bus(size_t(frames) * L, 0.0f);
		batch.ProcessInterleaved(std::span<const float>(in, size_t(N) * frames * C), bus, frames);
		for (size_t k = 0; k < bus.size(); ++k)
			bounded = bounded && std::fabs(double(bus[k]) - sum[k]) <= 4.0 * 1.1920929e-7 * mag[k] + 1e-12;
	}
	Expect(exact, "DirectSoundEffect::ProcessInterleaved: one-source batches == reference effects, bit for bit, 3 blocks");
	Expect(bounded, "DirectSoundBatch of all sources: bus within the fp32 summation bound of the reference effects' sum");

	// ---- ChannelMixing.h free functions
	{
		std::vector<float> dest(size_t(4) * frames * L), src(size_t(4) * frames * C);
		for (auto& v : dest) v = Uniform();
		for (auto& v : src) v = Uniform();
		auto r = dest, s = dest;
		for (uint32_t i = 0; i < 4; ++i) JPL::Add(&r[size_t(i) * frames * L], &src[size_t(i) * frames * C], 5, 1, L, C, frames);
		JPLG::Mixing::Add(ctx, s.data(), src.data(), 4, 5, 1, L, C, frames);
		Expect(SameBits(r.data(), s.data(), r.size()), "Mixing::Add == JPL::Add");
		std::vector<float> mr(size_t(4) * frames, 0.25f), ms = mr;
		for (uint32_t i = 0; i < 4; ++i) JPL::DownmixToMono(&mr[size_t(i) * frames], &dest[size_t(i) * frames * L], L, frames * L);
		JPLG::Mixing::DownmixToMono(ctx, ms.data(), dest.data(), 4, L, frames);
		Expect(SameBits(mr.data(), ms.data(), mr.size()), "Mixing::DownmixToMono == JPL::DownmixToMono");
	}
	std::printf("shim_check: %d failure(s)\n", gFailures);
	return gFailures ? 1 : 0;
}
