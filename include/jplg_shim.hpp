// jplg_shim.hpp — the reference-side adapter: JPLSpatial's own type names and call shapes over the C ABI of jplg.h.
//
// This is the header a JPLSpatial maintainer adds next to the engine code. It is header-only C++20, includes the
// reference's public headers for the TYPES only (ChannelMap, the panner's nested PanUpdateData, simd) and links against libjplg.so. It is compiled
// by tests/test_shim_cpu.py against /root/reference/Spatialization/include, so a change of the reference's signatures or of jplg.h breaks the CPU
// suite, not an engine build.
//
// What replaces what:
//   JPLG::BatchedPanner<P>    P::Initialize / InitializeSourceLayout / ProcessVBAPData      inc/Panning/PannerBase.h:315,324,369,375
//   JPLG::DirectSoundBatch    a container of JPL::DirectSoundEffect and the loop over it      inc/Auralization/DirectSound.h:44-65
//   JPLG::Mixing              the free functions of inc/Auralization/ChannelMixing.h:32-136, batched
#pragma once

#include <JPLSpatial/ChannelMap.h>
#include <JPLSpatial/Math/SIMD.h>
#include <JPLSpatial/Math/Vec3Traits.h>

#include "jplg.h"

#include <cstdint>
#include <span>
#include <vector>

namespace JPLG
{
	/// One per device. Not copyable; everything created from it must be destroyed first.
	class Context
	{
	public:
		explicit Context(int device = 0, uint32_t flags = 0) { if (jplg_context_create_ex(device, flags, &h) != JPLG_OK) h = nullptr; }
		~Context() { if (h) jplg_context_destroy(h); }
		Context(const Context&) = delete;
		Context& operator=(const Context&) = delete;

		[[nodiscard]] bool IsValid() const noexcept { return h != nullptr; }
		[[nodiscard]] const char* LastError() const noexcept { return jplg_last_error(h); }
		void Synchronize() const { jplg_context_synchronize(h); }

		jplg_context h = nullptr;
	};

	//==========================================================================
	/// Batched counterpart of a reference panner type (JPL::VBAPanner2D<...> / VBAPanner3D<...>): same method names, spans of per-source
	/// update data instead of one. `ReferencePanner` only supplies the nested parameter types, so engine code keeps building its
	/// PanUpdateData exactly as before.
	template<class ReferencePanner>
	class BatchedPanner
	{
	public:
		using PanUpdateData = typename ReferencePanner::PanUpdateData;
		using PanUpdateDataWithOrientation = typename ReferencePanner::PanUpdateDataWithOrientation;
		using Vec3Type = typename ReferencePanner::Vec3Type;

		explicit BatchedPanner(Context& context) : mContext(context) {}
		~BatchedPanner()
		{
			if (mLayout) jplg_source_layout_destroy(mLayout);
			if (mPanner) jplg_panner_destroy(mPanner);
		}
		BatchedPanner(const BatchedPanner&) = delete;
		BatchedPanner& operator=(const BatchedPanner&) = delete;

		/// VBAPannerBase::Initialize (PannerBase.h:315)
		bool Initialize(JPL::ChannelMap channelMap)
		{
			return jplg_panner_create(mContext.h, channelMap.GetChannelMask(), &mPanner) == JPLG_OK;
		}
		/// VBAPannerBase::InitializeSourceLayout (PannerBase.h:324); the layout is kept by the shim (one source layout per batch)
		bool InitializeSourceLayout(JPL::ChannelMap channelMap)
		{
			if (mLayout) { jplg_source_layout_destroy(mLayout); mLayout = nullptr; }
			return jplg_source_layout_create(mPanner, channelMap.GetChannelMask(), &mLayout) == JPLG_OK;
		}
		[[nodiscard]] uint32_t GetNumChannels() const { return jplg_panner_num_channels(mPanner); }                       // :346
		[[nodiscard]] JPL::ChannelMap GetChannelMap() const { return JPL::ChannelMap::FromChannelMask(jplg_panner_channel_mask(mPanner)); }   // :349
		[[nodiscard]] uint32_t GetNumSourceChannels() const { return jplg_source_layout_num_channels(mLayout); }

		/// ProcessVBAPData(sourceLayout, updateData, outChannelMixMap) (PannerBase.h:369) for updates.size() sources:
		/// outChannelMixMaps is [n][source channels * target channels]
		jplg_status ProcessVBAPData(std::span<const PanUpdateData> updates, std::span<float> outChannelMixMaps)
		{
			mParams.resize(updates.size() * 5);
			for (size_t i = 0; i < updates.size(); ++i) Pack(updates[i], &mParams[5 * i]);
			return jplg_panner_mdap(mPanner, mLayout, mParams.data(), static_cast<uint32_t>(updates.size()), outChannelMixMaps.data(), JPLG_MEM_HOST);
		}
		/// ProcessVBAPData(sourceLayout, PanUpdateDataWithOrientation, outChannelMixMap) (PannerBase.h:375)
		jplg_status ProcessVBAPData(std::span<const PanUpdateDataWithOrientation> updates, std::span<float> outChannelMixMaps)
		{
			mParams.resize(updates.size() * 11);
			for (size_t i = 0; i < updates.size(); ++i)
			{
				float* q = &mParams[11 * i];
				Pack(updates[i].Pan, q);
				PackVec(updates[i].Orientation.Up, q + 5);
				PackVec(updates[i].Orientation.Forward, q + 8);
			}
			return jplg_panner_mdap_oriented(mPanner, mLayout, mParams.data(), static_cast<uint32_t>(updates.size()), outChannelMixMaps.data(), JPLG_MEM_HOST);
		}

		[[nodiscard]] jplg_panner Handle() const noexcept { return mPanner; }
		[[nodiscard]] jplg_source_layout LayoutHandle() const noexcept { return mLayout; }

		/// [n][5] rows of { dir.x, dir.y, dir.z, focus, spread }: what the device-resident entry points (jplg_renderer_update_panned) consume
		static void Pack(const PanUpdateData& u, float* q)
		{
			PackVec(u.SourceDirection, q);
			q[3] = u.Focus;
			q[4] = u.Spread;
		}

	private:
		static void PackVec(const Vec3Type& v, float* q)
		{
			q[0] = static_cast<float>(JPL::GetX(v));
			q[1] = static_cast<float>(JPL::GetY(v));
			q[2] = static_cast<float>(JPL::GetZ(v));
		}

		Context& mContext;
		jplg_panner mPanner = nullptr;
		jplg_source_layout mLayout = nullptr;
		std::vector<float> mParams;
	};

	//==========================================================================
	/// Replaces a container of JPL::DirectSoundEffect and the audio callback's loop over it (src/Auralization/DirectSound.cpp:95).
	/// All sources of one batch share the source and output channel counts, like the effects of one source layout do.
	class DirectSoundBatch
	{
	public:
		/// DirectSoundEffect ctor (DirectSound.h:44-49) for numSources effects:
		/// initialFilterGains [N][4], initialDelayTimesInSeconds [N], initialMixMaps [N][numSourceChannels * numOutputChannels]
		DirectSoundBatch(Context& context, float sampleRate, uint32_t numSources, uint32_t numSourceChannels, uint32_t numOutputChannels,
						 std::span<const float> initialFilterGains, std::span<const float> initialDelayTimesInSeconds, std::span<const float> initialMixMaps,
						 uint32_t maxBlockFrames = 512)
		{
			jplg_renderer_config cfg{};
			cfg.sample_rate = sampleRate;
			cfg.num_sources = numSources;
			cfg.source_channels = numSourceChannels;
			cfg.output_channels = numOutputChannels;
			cfg.ring_size = 0;                          // the reference's bit_ceil(2 * sampleRate) (DirectSound.cpp:79)
			cfg.max_block_frames = maxBlockFrames;
			cfg.mode = JPLG_MODE_PARITY;
			mStatus = jplg_renderer_create(context.h, &cfg, initialFilterGains.data(), initialDelayTimesInSeconds.data(), initialMixMaps.data(), &mRenderer);
		}
		~DirectSoundBatch() { if (mRenderer) jplg_renderer_destroy(mRenderer); }
		DirectSoundBatch(const DirectSoundBatch&) = delete;
		DirectSoundBatch& operator=(const DirectSoundBatch&) = delete;

		[[nodiscard]] jplg_status Status() const noexcept { return mStatus; }

		/// DirectSoundEffect::ProcessInterleaved (DirectSound.h:61), audio thread: input [N][numFrames][C] (every source's block interleaved as the
		/// reference's `input` span), output [numFrames][L], accumulated into
		jplg_status ProcessInterleaved(std::span<const float> input, std::span<float> output, uint32_t numFrames)
		{
			return jplg_renderer_process_block(mRenderer, input.data(), output.data(), numFrames, JPLG_MEM_HOST);
		}
		/// DirectSoundEffect::UpdateParameters (DirectSound.h:65), non-audio thread, all sources at once
		jplg_status UpdateParameters(std::span<const float> filterGains, std::span<const float> delayTimesSeconds, std::span<const float> channelMixMaps)
		{
			return jplg_renderer_update(mRenderer, filterGains.data(), delayTimesSeconds.data(), channelMixMaps.data(), JPLG_MEM_HOST);
		}
		/// Same signature as the reference for ONE source of the batch, for call sites that are converted one at a time: gathers into the
		/// staging arrays; Commit() hands them over in one call.
		void UpdateParameters(uint32_t source, const JPL::simd& filterGains, float delayTimeSeconds, std::span<const float> channelMixMap)
		{
			if (mGains.size() < size_t(source + 1) * 4) { mGains.resize(size_t(source + 1) * 4, 1.0f); mDelays.resize(source + 1, 0.0f); }
			if (mMixMaps.size() < size_t(source + 1) * channelMixMap.size()) mMixMaps.resize(size_t(source + 1) * channelMixMap.size(), 0.0f);
			filterGains.store(&mGains[size_t(source) * 4]);
			mDelays[source] = delayTimeSeconds;
			for (size_t i = 0; i < channelMixMap.size(); ++i) mMixMaps[size_t(source) * channelMixMap.size() + i] = channelMixMap[i];
		}
		jplg_status Commit() { return UpdateParameters(mGains, mDelays, mMixMaps); }

		/// Panning and parameter update fused on the device (PannerBase.h:369 -> DirectSound.cpp:231-239 per source): device pointers
		template<class ReferencePanner>
		jplg_status UpdatePanned(const BatchedPanner<ReferencePanner>& panner, const float* devicePanParams, const float* deviceFilterGains, const float* deviceDelays)
		{
			return jplg_renderer_update_panned(mRenderer, panner.Handle(), panner.LayoutHandle(), devicePanParams, deviceFilterGains, deviceDelays, JPLG_MEM_DEVICE);
		}

		[[nodiscard]] jplg_renderer Handle() const noexcept { return mRenderer; }

	private:
		jplg_renderer mRenderer = nullptr;
		jplg_status mStatus = JPLG_OK;
		std::vector<float> mGains, mDelays, mMixMaps;
	};

	//==========================================================================
	/// inc/Auralization/ChannelMixing.h free functions over `n` (dest, source) pairs laid out back to back (host memory)
	namespace Mixing
	{
		inline jplg_status Add(Context& c, float* dest, const float* source, uint32_t n, uint32_t destChannel, uint32_t sourceChannel,
							   uint32_t destNumChannels, uint32_t sourceNumChannels, uint32_t numFrames)                                   // :32-39
		{
			return jplg_mix_add(c.h, dest, source, n, destChannel, sourceChannel, destNumChannels, sourceNumChannels, numFrames, JPLG_MEM_HOST);
		}
		inline jplg_status Add(Context& c, float* dest, const float* source, uint32_t n, uint32_t destChannel, uint32_t destNumChannels, uint32_t numFrames)   // :41-48
		{
			return jplg_mix_add(c.h, dest, source, n, destChannel, 0, destNumChannels, 1, numFrames, JPLG_MEM_HOST);
		}
		inline jplg_status ApplyGain(Context& c, float* dest, uint32_t n, uint32_t numFrames, const float* gains)                          // :50-63
		{
			return jplg_mix_apply_gain(c.h, dest, n, numFrames, gains, JPLG_MEM_HOST);
		}
		inline jplg_status AddAndApplyGainRamp(Context& c, float* dest, const float* source, uint32_t n, uint32_t destChannel, uint32_t sourceChannel,
											   uint32_t destNumChannels, uint32_t sourceNumChannels, uint32_t numFrames,
											   const float* gainStart, const float* gainEnd)                                               // :89-107
		{
			return jplg_mix_add_and_apply_gain_ramp(c.h, dest, source, n, destChannel, sourceChannel, destNumChannels, sourceNumChannels, numFrames, gainStart, gainEnd, JPLG_MEM_HOST);
		}
		inline jplg_status DownmixToMono(Context& c, float* dest, const float* source, uint32_t n, uint32_t numChannels, uint32_t numFrames) // :129-136
		{
			return jplg_mix_downmix_to_mono(c.h, dest, source, n, numChannels, numFrames, JPLG_MEM_HOST);
		}
	}
} // namespace JPLG
