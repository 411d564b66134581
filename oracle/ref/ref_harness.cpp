// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// C-ABI wrapper around the UNMODIFIED reference (JPL Spatial) compiled from the
// sources where they lie under /root/reference (see oracle/Makefile). Nothing of
// the reference is copied: this file only #includes its public headers and calls
// its public API. The resulting shared object lands in oracle/_ref/ and is used
// by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference arm as
// (a) the checker the oracle restatement (oracle/port) is pinned against and
// (b) the timed CPU baseline. The product (jplspatial_b200/) never links it.
//
// Built with -fno-access-control so a few private constants (source-layout
// phi term, cached azimuths) can be read for finer-grained pinning.

#include <JPLSpatial/ChannelMap.h>
#include <JPLSpatial/Panning/VBAPanning2D.h>
#include <JPLSpatial/Panning/VBAPanning3D.h>
#include <JPLSpatial/Auralization/DirectSound.h>
#include <JPLSpatial/Auralization/LateReverb.h>
#include <JPLSpatial/Auralization/EarlyReflectionsBus.h>
#include <JPLSpatial/Math/DecibelsAndGain.h>
#include <JPLSpatial/Auralization/CrossoverFilter.h>
#include <JPLSpatial/Auralization/DelayLine.h>
#include <JPLSpatial/Auralization/FractionalDelay.h>
#include <JPLSpatial/Auralization/ChannelMixing.h>
#include <JPLSpatial/Math/DirectionEncoding.h>
#include <JPLSpatial/Math/SIMD.h>
#include <JPLSpatial/Math/SIMDMath.h>
#include <JPLSpatial/Math/Vec3Math.h>
#include <JPLSpatial/Utilities/SmoothedValue.h>
#include <JPLSpatial/Math/MinimalVec3.h>
#include <JPLSpatial/Services/DirectPathService.h>
#include <JPLSpatial/Services/PanningService.h>
#include <JPLSpatial/SpatialManager.h>
#include <JPLSpatial/DistanceAttenuation.h>
#include <JPLSpatial/AirAbsorption.h>
#include <JPLSpatial/SpeedOfSound.h>

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <functional>
#include <atomic>
#include <vector>

using namespace JPL;

namespace
{
	using Panner2D = StandardPanner2D;
	using Panner3D = StandardPanner3D;
	using Layout2D = typename Panner2D::SourceLayoutType;
	using Layout3D = typename Panner3D::SourceLayoutType;

	struct RefPanner
	{
		bool is3D = false;
		bool ok = false;
		Panner2D p2;
		Panner3D p3;
	};

	struct RefLayout
	{
		RefPanner* panner = nullptr;
		bool ok = false;
		Layout2D l2;
		Layout3D l3;
	};

	struct RefDSE
	{
		std::unique_ptr<DirectSoundEffect> fx;
		uint32 nSrc = 0, nOut = 0;
	};
}


//==========================================================================
// Persistent worker pool for the TIMED batch helpers (bench.py's CPU arm): threads are created once per thread count and parked on a
// condition variable, so a timed call pays a wake-up, not std::thread creation + join.
namespace
{
	class WorkerPool
	{
	public:
		explicit WorkerPool(int n) : mN(n)
		{
			for (int t = 1; t < n; ++t) mThreads.emplace_back([this, t] { Loop(t); });
		}
		~WorkerPool()
		{
			{ std::lock_guard<std::mutex> lk(mMtx); mStop = true; ++mEpoch; }
			mWake.notify_all();
			for (auto& th : mThreads) th.join();
		}
		// runs fn(0..n-1), fn(0) on the calling thread; returns when all are done
		void Run(const std::function<void(int)>& fn)
		{
			if (mN == 1) { fn(0); return; }
			{ std::lock_guard<std::mutex> lk(mMtx); mFn = &fn; mPending = mN - 1; ++mEpoch; }
			mWake.notify_all();
			fn(0);
			std::unique_lock<std::mutex> lk(mMtx);
			mDone.wait(lk, [this] { return mPending == 0; });
			mFn = nullptr;
		}
	private:
		void Loop(int t)
		{
			uint64_t seen = 0;
			for (;;)
			{
				const std::function<void(int)>* fn = nullptr;
				{
					std::unique_lock<std::mutex> lk(mMtx);
					mWake.wait(lk, [&] { return mEpoch != seen; });
					seen = mEpoch;
					if (mStop) return;
					fn = mFn;
				}
				(*fn)(t);
				{ std::lock_guard<std::mutex> lk(mMtx); if (--mPending == 0) mDone.notify_one(); }
			}
		}
		int mN;
		std::vector<std::thread> mThreads;
		std::mutex mMtx;
		std::condition_variable mWake, mDone;
		const std::function<void(int)>* mFn = nullptr;
		int mPending = 0;
		uint64_t mEpoch = 0;
		bool mStop = false;
	};
	void RunParallel(int threads, const std::function<void(int)>& fn)
	{
		static std::mutex sMtx;
		static std::vector<std::unique_ptr<WorkerPool>> sPools;   // one pool per thread count, kept for the life of the process
		static std::vector<int> sCounts;
		if (threads < 1) threads = 1;
		WorkerPool* pool = nullptr;
		{
			std::lock_guard<std::mutex> lk(sMtx);
			for (size_t i = 0; i < sCounts.size(); ++i) if (sCounts[i] == threads) pool = sPools[i].get();
			if (!pool) { sPools.emplace_back(new WorkerPool(threads)); sCounts.push_back(threads); pool = sPools.back().get(); }
		}
		pool->Run(fn);
	}
}

extern "C"
{
	//==========================================================================
	// build info
	int ref_build_flags()
	{
		int f = 0;
#if defined(JPL_USE_SSE4_1)
		f |= 1;
#endif
#if defined(JPL_USE_AVX2)
		f |= 2;
#endif
#if defined(JPL_USE_FMADD)
		f |= 4;
#endif
		return f;
	}

	//==========================================================================
	// Panners (inc/Panning/PannerBase.h:315-389)
	void* ref_panner_create(uint32_t targetMask)
	{
		auto* p = new RefPanner();
		const ChannelMap map = ChannelMap::FromChannelMask(targetMask);
		p->is3D = map.HasTopChannels();
		p->ok = p->is3D ? p->p3.Initialize(map) : p->p2.Initialize(map);
		return p;
	}
	void ref_panner_destroy(void* h) { delete static_cast<RefPanner*>(h); }
	int ref_panner_ok(void* h) { return static_cast<RefPanner*>(h)->ok; }
	int ref_panner_is3d(void* h) { return static_cast<RefPanner*>(h)->is3D; }
	int ref_panner_initialized(void* h)
	{
		auto* p = static_cast<RefPanner*>(h);
		return p->is3D ? p->p3.IsInitialized() : p->p2.IsInitialized();
	}
	uint32_t ref_panner_num_channels(void* h)
	{
		auto* p = static_cast<RefPanner*>(h);
		return p->is3D ? p->p3.GetNumChannels() : p->p2.GetNumChannels();
	}
	float ref_panner_aperture(void* h)
	{
		auto* p = static_cast<RefPanner*>(h);
		return p->is3D ? p->p3.GetShortestSpeakerAperture() : p->p2.GetShortestSpeakerAperture();
	}
	// 3D LUT: gains[65536*3] floats, speakers[65536*3] bytes (inc/Panning/VBAPLUT3D.h:163-171)
	void ref_panner_lut3d(void* h, float* outGains, uint8_t* outSpeakers)
	{
		auto* p = static_cast<RefPanner*>(h);
		const auto* lut = p->p3.GetLUT();
		std::memcpy(outGains, lut->Gains.data(), sizeof(lut->Gains));
		std::memcpy(outSpeakers, lut->Speakers.data(), sizeof(lut->Speakers));
	}
	// 2D LUT: 1024 * numChannels floats (inc/Panning/VBAPLUT2D.h:91-99)
	uint32_t ref_panner_lut2d(void* h, float* out)
	{
		auto* p = static_cast<RefPanner*>(h);
		const auto* lut = p->p2.GetLUT();
		const size_t n = lut->GetLUTSize();
		if (out)
			for (size_t i = 0; i < n; ++i) out[i] = lut->GetLUTValue(static_cast<uint32>(i));
		return static_cast<uint32_t>(n);
	}

	// Speaker triangulation (inc/Panning/VBAPEx.h:177-234): vertices incl. dummies and triangle list
	int ref_triangulate(uint32_t targetMask, float* outVerts /*[32*3]*/, int* outNumVerts, uint8_t* outTris /*[64*3]*/, int* outNumTris)
	{
		std::pmr::vector<MinimalVec3> verts;
		std::pmr::vector<SpeakerTriangulation::Vec3i> tris;
		const bool ok = SpeakerTriangulation::TriangulateSpeakerLayout<&VBAPStandardTraits::GetChannelVector, MinimalVec3>(
			ChannelMap::FromChannelMask(targetMask), verts, tris);
		if (!ok) return 0;
		*outNumVerts = static_cast<int>(verts.size());
		*outNumTris = static_cast<int>(tris.size());
		for (size_t i = 0; i < verts.size(); ++i) { outVerts[3 * i] = verts[i].X; outVerts[3 * i + 1] = verts[i].Y; outVerts[3 * i + 2] = verts[i].Z; }
		for (size_t i = 0; i < tris.size(); ++i) { outTris[3 * i] = tris[i][0]; outTris[3 * i + 1] = tris[i][1]; outTris[3 * i + 2] = tris[i][2]; }
		return 1;
	}

	// GetSpeakerGains(dir, span) scalar (PannerBase.h:441-445). `out` [n][L] is zero-filled first
	// because the 3D query writes only three entries (VBAPLUT3D.h:1001-1009).
	void ref_point_gains(void* h, const float* dirs, uint32_t n, float* out)
	{
		auto* p = static_cast<RefPanner*>(h);
		const uint32 L = ref_panner_num_channels(h);
		std::memset(out, 0, sizeof(float) * size_t(n) * L);
		for (uint32_t i = 0; i < n; ++i)
		{
			const MinimalVec3 d(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
			std::span<float> g(out + size_t(i) * L, L);
			if (p->is3D) p->p3.GetSpeakerGains(d, g); else p->p2.GetSpeakerGains(d, g);
		}
	}

	// GetSpeakerGains SIMD overload (PannerBase.h:448-452); n multiple of 4; 2D only gives full rows.
	void ref_point_gains_simd2d(void* h, const float* dirs, uint32_t n, float* out)
	{
		auto* p = static_cast<RefPanner*>(h);
		const uint32 L = p->p2.GetNumChannels();
		std::vector<simd> g(L);
		for (uint32_t i = 0; i + 3 < n; i += 4)
		{
			simd x(dirs[3 * i], dirs[3 * (i + 1)], dirs[3 * (i + 2)], dirs[3 * (i + 3)]);
			simd y(dirs[3 * i + 1], dirs[3 * (i + 1) + 1], dirs[3 * (i + 2) + 1], dirs[3 * (i + 3) + 1]);
			simd z(dirs[3 * i + 2], dirs[3 * (i + 1) + 2], dirs[3 * (i + 2) + 2], dirs[3 * (i + 3) + 2]);
			p->p2.GetSpeakerGains(x, y, z, std::span<simd>(g));
			for (uint32 c = 0; c < L; ++c)
			{
				float v[4]; g[c].store(v);
				for (int l = 0; l < 4; ++l) out[size_t(i + l) * L + c] = v[l];
			}
		}
	}

	//==========================================================================
	// Source layouts (PannerBase.h:324; VBAPanning2D.h:150-193; VBAPanning3D.h:242-310)
	void* ref_layout_create(void* hPanner, uint32_t sourceMask)
	{
		auto* p = static_cast<RefPanner*>(hPanner);
		auto* l = new RefLayout();
		l->panner = p;
		const ChannelMap map = ChannelMap::FromChannelMask(sourceMask);
		l->ok = p->is3D ? p->p3.InitializeSourceLayout(map, l->l3) : p->p2.InitializeSourceLayout(map, l->l2);
		return l;
	}
	void ref_layout_destroy(void* h) { delete static_cast<RefLayout*>(h); }
	int ref_layout_ok(void* h) { return static_cast<RefLayout*>(h)->ok; }
	// info[0]=numSourceChannels (non-LFE), [1]=numVS per channel, [2]=rings (3D) or 0, [3]=samples per ring (3D) or 0
	void ref_layout_info(void* h, uint32_t* info, float* phiTerm, float* offsetTerm, float* vsWeight)
	{
		auto* l = static_cast<RefLayout*>(h);
		if (l->panner->is3D)
		{
			info[0] = static_cast<uint32_t>(l->l3.ChannelGroups.size());
			info[1] = static_cast<uint32_t>(l->l3.GetNumVirtualSources());
			info[2] = l->l3.GetDimensions().NumRings;
			info[3] = l->l3.GetDimensions().NumSamplesPerRing;
			*phiTerm = l->l3.mPhiTerm;
			*offsetTerm = 0.0f;
			*vsWeight = l->l3.mDefaultVSWeight;
		}
		else
		{
			info[0] = static_cast<uint32_t>(l->l2.ChannelGroups.size());
			info[1] = static_cast<uint32_t>(l->l2.GetNumVirtualSources());
			info[2] = 0; info[3] = 0;
			*phiTerm = l->l2.mPhiTerm;
			*offsetTerm = l->l2.mOffsetTerm;
			*vsWeight = l->l2.mDefaultVSWeight;
		}
	}
	// channel-group quaternions [nCh][4] = {x,y,z,w} and channel ids
	void ref_layout_groups(void* h, float* outQuats, uint32_t* outChannelIds)
	{
		auto* l = static_cast<RefLayout*>(h);
		auto copy = [&](const auto& groups)
		{
			for (size_t i = 0; i < groups.size(); ++i)
			{
				outQuats[4 * i + 0] = groups[i].Rotation.V.X;
				outQuats[4 * i + 1] = groups[i].Rotation.V.Y;
				outQuats[4 * i + 2] = groups[i].Rotation.V.Z;
				outQuats[4 * i + 3] = groups[i].Rotation.W;
				outChannelIds[i] = groups[i].Channel;
			}
		};
		if (l->panner->is3D) copy(l->l3.ChannelGroups); else copy(l->l2.ChannelGroups);
	}
	// cached azimuth sin/cos of the 3D layout (VBAPanning3D.h:312-346); returns number of simd packs
	uint32_t ref_layout_azimuth(void* h, float* outSin, float* outCos)
	{
		auto* l = static_cast<RefLayout*>(h);
		const size_t n = l->l3.mSamplesSin.size();
		for (size_t i = 0; i < n; ++i)
		{
			l->l3.mSamplesSin[i].store(outSin + 4 * i);
			l->l3.mSamplesCos[i].store(outCos + 4 * i);
		}
		return static_cast<uint32_t>(n);
	}

	// ProcessVBAPData (PannerBase.h:486-500). params [n][5] = {dx,dy,dz,focus,spread}; out [n][nSrcCh*L]
	void ref_mdap(void* hPanner, void* hLayout, const float* params, uint32_t n, float* out)
	{
		auto* p = static_cast<RefPanner*>(hPanner);
		auto* l = static_cast<RefLayout*>(hLayout);
		const uint32 L = ref_panner_num_channels(hPanner);
		if (p->is3D)
		{
			const size_t stride = l->l3.ChannelGroups.size() * L;
			for (uint32_t i = 0; i < n; ++i)
			{
				const float* q = params + 5 * size_t(i);
				typename Panner3D::PanUpdateData d{ .SourceDirection = MinimalVec3(q[0], q[1], q[2]), .Focus = q[3], .Spread = q[4] };
				p->p3.ProcessVBAPData(l->l3, d, std::span<float>(out + stride * i, stride));
			}
		}
		else
		{
			const size_t stride = l->l2.ChannelGroups.size() * L;
			for (uint32_t i = 0; i < n; ++i)
			{
				const float* q = params + 5 * size_t(i);
				typename Panner2D::PanUpdateData d{ .SourceDirection = MinimalVec3(q[0], q[1], q[2]), .Focus = q[3], .Spread = q[4] };
				p->p2.ProcessVBAPData(l->l2, d, std::span<float>(out + stride * i, stride));
			}
		}
	}

	// multi-threaded variant for timing; returns seconds
	double ref_mdap_timed(void* hPanner, void* hLayout, const float* params, uint32_t n, float* out, int threads)
	{
		auto* l = static_cast<RefLayout*>(hLayout);
		const uint32 L = ref_panner_num_channels(hPanner);
		const size_t nCh = static_cast<RefPanner*>(hPanner)->is3D ? l->l3.ChannelGroups.size() : l->l2.ChannelGroups.size();
		const size_t stride = nCh * L;
		const auto t0 = std::chrono::steady_clock::now();
		RunParallel(threads, [&](int t)
		{
			const uint32_t b = uint32_t(uint64_t(n) * t / threads), e = uint32_t(uint64_t(n) * (t + 1) / threads);
			([=] { ref_mdap(hPanner, hLayout, params + 5 * size_t(b), e - b, out + stride * b); })();
		});
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	}

	// ProcessVirtualSources (PannerBase.h:731-742). vs [n][4] = {dx,dy,dz,weight}; out[L] must be zeroed by caller
	void ref_process_virtual_sources(void* hPanner, const float* vs, uint32_t n, float* out)
	{
		auto* p = static_cast<RefPanner*>(hPanner);
		const uint32 L = ref_panner_num_channels(hPanner);
		if (p->is3D)
		{
			std::vector<typename Panner3D::VirtualSource> v(n);
			for (uint32_t i = 0; i < n; ++i) v[i] = { MinimalVec3(vs[4 * i], vs[4 * i + 1], vs[4 * i + 2]), vs[4 * i + 3] };
			p->p3.ProcessVirtualSources(v, std::span<float>(out, L));
		}
		else
		{
			std::vector<typename Panner2D::VirtualSource> v(n);
			for (uint32_t i = 0; i < n; ++i) v[i] = { MinimalVec3(vs[4 * i], vs[4 * i + 1], vs[4 * i + 2]), vs[4 * i + 3] };
			p->p2.ProcessVirtualSources(v, std::span<float>(out, L));
		}
	}
	// NOTE: VBAPannerBase::NormalizeWeights (PannerBase.h:879-892) cannot be instantiated: it calls an
	// unqualified `Accumulate` that only exists as JPL::Algo::Accumulate (GCC 13: "not declared in this scope"),
	// so no reference output exists for it; the restatement follows the source text.

	//==========================================================================
	// Math primitives used on the path (for pinning the restatement op by op)
	void ref_sincos(const float* x, uint32_t n, float* s, float* c) // SIMDMath.h:488-534, n multiple of 4
	{
		for (uint32_t i = 0; i + 3 < n; i += 4) { simd ss, cc; Math::SinCos(simd(x + i), ss, cc); ss.store(s + i); cc.store(c + i); }
	}
	void ref_acos(const float* x, uint32_t n, float* out) // SIMDMath.h:710-714
	{
		for (uint32_t i = 0; i + 3 < n; i += 4) JPL::acos(simd(x + i)).store(out + i);
	}
	void ref_tan(const float* x, uint32_t n, float* out) // SIMDMath.h:643-665
	{
		for (uint32_t i = 0; i + 3 < n; i += 4) JPL::tan(simd(x + i)).store(out + i);
	}
	void ref_invsqrt(const float* x, uint32_t n, float* outEst, float* outNR1, float* outNR2) // SIMD.h:1700-1790
	{
		for (uint32_t i = 0; i + 3 < n; i += 4)
		{
			const simd v(x + i);
			simd(_mm_rsqrt_ps(v)).store(outEst + i);
			Math::InvSqrtFast(v).store(outNR1 + i);
			Math::InvSqrt(v).store(outNR2 + i);
		}
	}
	void ref_oct_encode_simd(const float* dirs, uint32_t n, uint32_t* out) // DirectionEncoding.h:346-386
	{
		for (uint32_t i = 0; i + 3 < n; i += 4)
		{
			simd x(dirs[3 * i], dirs[3 * (i + 1)], dirs[3 * (i + 2)], dirs[3 * (i + 3)]);
			simd y(dirs[3 * i + 1], dirs[3 * (i + 1) + 1], dirs[3 * (i + 2) + 1], dirs[3 * (i + 3) + 1]);
			simd z(dirs[3 * i + 2], dirs[3 * (i + 1) + 2], dirs[3 * (i + 2) + 2], dirs[3 * (i + 3) + 2]);
			Octahedron16Bit::Encode(x, y, z).store(out + i);
		}
	}
	void ref_oct_encode_scalar(const float* dirs, uint32_t n, uint32_t* out) // DirectionEncoding.h:292-343
	{
		for (uint32_t i = 0; i < n; ++i) out[i] = Octahedron16Bit::Encode(MinimalVec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
	}
	void ref_oct_decode_scalar(const uint32_t* codes, uint32_t n, float* out) // DirectionEncoding.h:389-436
	{
		for (uint32_t i = 0; i < n; ++i)
		{
			const MinimalVec3 v = Octahedron16Bit::Decode<MinimalVec3>(static_cast<uint16>(codes[i]));
			out[3 * i] = v.X; out[3 * i + 1] = v.Y; out[3 * i + 2] = v.Z;
		}
	}
	void ref_diamond_scalar(const float* xy, uint32_t n, float* out) // DirectionEncoding.h:197-209
	{
		for (uint32_t i = 0; i < n; ++i) out[i] = ToDiamond(Vec2(xy[2 * i], xy[2 * i + 1]));
	}
	void ref_from_diamond(const float* p, uint32_t n, float* outXY) // DirectionEncoding.h:212-227
	{
		for (uint32_t i = 0; i < n; ++i) { const Vec2 v = FromDiamond(p[i]); outXY[2 * i] = v.X; outXY[2 * i + 1] = v.Y; }
	}
	// SIMD Slerp (Vec3Math.h:206-307): v0 [n][3] in/out, v1 single target, t scalar; n multiple of 4
	void ref_slerp_simd(float* v0, const float* v1, float t, uint32_t n)
	{
		const Vec3Pack target(MinimalVec3(v1[0], v1[1], v1[2]));
		for (uint32_t i = 0; i + 3 < n; i += 4)
		{
			Vec3Pack v(simd(v0[3 * i], v0[3 * (i + 1)], v0[3 * (i + 2)], v0[3 * (i + 3)]),
					   simd(v0[3 * i + 1], v0[3 * (i + 1) + 1], v0[3 * (i + 2) + 1], v0[3 * (i + 3) + 1]),
					   simd(v0[3 * i + 2], v0[3 * (i + 1) + 2], v0[3 * (i + 2) + 2], v0[3 * (i + 3) + 2]));
			Math::Slerp(v, target, simd(t));
			float xs[4], ys[4], zs[4]; v.X.store(xs); v.Y.store(ys); v.Z.store(zs);
			for (int l = 0; l < 4; ++l) { v0[3 * (i + l)] = xs[l]; v0[3 * (i + l) + 1] = ys[l]; v0[3 * (i + l) + 2] = zs[l]; }
		}
	}
	// QuatLookAt(dir, worldUp) (MinimalQuat.h:141-144) -> {x,y,z,w}
	void ref_quat_lookat(const float* dirs, uint32_t n, float* out)
	{
		const MinimalVec3 up(0, 1, 0);
		for (uint32_t i = 0; i < n; ++i)
		{
			const auto q = Math::QuatLookAt(MinimalVec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]), up);
			out[4 * i] = q.V.X; out[4 * i + 1] = q.V.Y; out[4 * i + 2] = q.V.Z; out[4 * i + 3] = q.W;
		}
	}

	//==========================================================================
	// DSP components
	// FourBandCrossover (CrossoverFilter.h:416-535)
	void* ref_xover_create(float fs) { auto* f = new FourBandCrossover(); f->Prepare(fs, {}); return f; }
	void ref_xover_destroy(void* h) { delete static_cast<FourBandCrossover*>(h); }
	void ref_xover_reset(void* h) { static_cast<FourBandCrossover*>(h)->Reset(); }
	// coefficient dump: 3 scalar (X2) + 3x4 (X13) + 3x4 (all-pass) = 27 floats
	void ref_xover_coeffs(void* h, float* out)
	{
		auto* f = static_cast<FourBandCrossover*>(h);
		out[0] = f->X2.Params.g; out[1] = f->X2.Params.d; out[2] = f->X2.Params.MinusRg;
		f->X13.Params.g.store(out + 3); f->X13.Params.d.store(out + 7); f->X13.Params.MinusRg.store(out + 11);
		f->X13.AllpassEqualizer.Params.g.store(out + 15); f->X13.AllpassEqualizer.Params.d.store(out + 19); f->X13.AllpassEqualizer.Params.MinusRg.store(out + 23);
	}
	// ProcessBlock(in, startGains, endGains, out) (CrossoverFilter.h:505-535)
	void ref_xover_process_ramp(void* h, const float* in, uint32_t n, const float* g0, const float* g1, float* out)
	{
		static_cast<FourBandCrossover*>(h)->ProcessBlock(std::span<const float>(in, n), simd(g0), simd(g1), std::span<float>(out, n));
	}
	// ProcessBlock(in, gains, out) (:473-502)
	void ref_xover_process_gains(void* h, const float* in, uint32_t n, const float* g, float* out)
	{
		static_cast<FourBandCrossover*>(h)->ProcessBlock(std::span<const float>(in, n), simd(g), std::span<float>(out, n));
	}
	// ProcessBlock(in, outBands) (:462-470): out [n][4] in band order b1..b4
	void ref_xover_split(void* h, const float* in, uint32_t n, float* outBands)
	{
		std::vector<simd> b(n);
		static_cast<FourBandCrossover*>(h)->ProcessBlock(std::span<const float>(in, n), std::span<simd>(b));
		for (uint32_t i = 0; i < n; ++i) b[i].store(outBands + 4 * i);
	}

	// DelayLine<482> + DelayTap<Thiran> exactly as DirectSoundEffect's ChannelRTData uses them (DirectSound.h:88-91)
	struct RefDelay
	{
		using Line = DelayLine<482>;
		Line line;
		DelayTap<Thiran1stInterpolator> thiran;
		DelayTap<LinearInterpolator> linear;
		DelayTapXFade xfade;
		explicit RefDelay(uint32 minSize) : line(minSize) {}
	};
	void* ref_delay_create(uint32_t minSize) { return new RefDelay(minSize); }
	void ref_delay_destroy(void* h) { delete static_cast<RefDelay*>(h); }
	uint32_t ref_delay_size(void* h) { return static_cast<RefDelay*>(h)->line.GetSize(); }
	void ref_delay_push(void* h, const float* x, uint32_t n) { auto* d = static_cast<RefDelay*>(h); for (uint32_t i = 0; i < n; ++i) d->line.Push(x[i]); }
	float ref_delay_read_int(void* h, uint32_t delay) { return static_cast<RefDelay*>(h)->line.GetReadWindow<1>(delay); }
	float ref_delay_tap_thiran(void* h, float delay) { auto* d = static_cast<RefDelay*>(h); d->thiran.SetDelay(delay); return d->thiran.Process(d->line); }
	float ref_delay_tap_linear(void* h, float delay) { auto* d = static_cast<RefDelay*>(h); d->linear.SetDelay(delay); return d->linear.Process(d->line); }

	float ref_delay_tap_xfade(void* h, float delay, float alpha) { auto* d = static_cast<RefDelay*>(h); d->xfade.SetCrossfade(alpha); d->xfade.SetDelay(delay); return d->xfade.Process(d->line); }

	// SmoothedValue exp smoothing (SmoothedValue.h:36-41,73-77)
	float ref_smoother_alpha(uint32_t ticks) { return SmoothedValue<float>::GetExpAlpha(ticks); }
	void ref_smoother_run(float current, float target, uint32_t ticks, uint32_t n, float* out)
	{
		auto s = SmoothedValue<float>::CreateExpSmoothing(current, target, ticks);
		for (uint32_t i = 0; i < n; ++i) out[i] = s.GetNext();
	}

	// ChannelMixing.h:89-107 strided AddAndApplyGainRamp
	void ref_add_gain_ramp(float* dest, const float* src, uint32_t destCh, uint32_t srcCh, uint32_t nDest, uint32_t nSrc, uint32_t frames, float g0, float g1)
	{
		AddAndApplyGainRamp(dest, src, destCh, srcCh, nDest, nSrc, frames, g0, g1);
	}

	// ChannelMixing.h:32-48 Add (nSrc == 0 selects the mono-source overload) and :129-136 DownmixToMono
	void ref_mix_add(float* dest, const float* src, uint32_t destCh, uint32_t srcCh, uint32_t nDest, uint32_t nSrc, uint32_t frames)
	{
		if (nSrc == 0) Add(dest, src, destCh, nDest, frames);
		else Add(dest, src, destCh, srcCh, nDest, nSrc, frames);
	}
	void ref_downmix_to_mono(float* dest, const float* src, uint32_t numChannels, uint32_t numSamples)
	{
		DownmixToMono(dest, src, numChannels, numSamples);
	}

	//==========================================================================
	// DirectSoundEffect (DirectSound.h:44-65)
	void* ref_dse_create(float fs, uint32_t nSrc, uint32_t nOut, const float* bandGains, float delaySeconds, const float* mixMap)
	{
		auto* d = new RefDSE();
		d->nSrc = nSrc; d->nOut = nOut;
		d->fx = std::make_unique<DirectSoundEffect>(fs, nSrc, nOut, simd(bandGains), delaySeconds, std::span<const float>(mixMap, size_t(nSrc) * nOut));
		return d;
	}
	void ref_dse_destroy(void* h) { delete static_cast<RefDSE*>(h); }
	void ref_dse_update(void* h, const float* bandGains, float delaySeconds, const float* mixMap)
	{
		auto* d = static_cast<RefDSE*>(h);
		d->fx->UpdateParameters(simd(bandGains), delaySeconds, std::span<const float>(mixMap, size_t(d->nSrc) * d->nOut));
	}
	// in: interleaved [frames][nSrc]; out: interleaved [frames][nOut], ACCUMULATED INTO (DirectSound.cpp:200-222)
	void ref_dse_process(void* h, const float* in, float* out, uint32_t frames)
	{
		auto* d = static_cast<RefDSE*>(h);
		d->fx->ProcessInterleaved(std::span<const float>(in, size_t(frames) * d->nSrc), std::span<float>(out, size_t(frames) * d->nOut), frames);
	}

	// Batch helper: N mono-or-multichannel effects, one block each, sources partitioned over threads.
	// in: planar per source [N][frames*nSrc] interleaved per source; per-thread private buses are summed
	// in thread order into `bus` (accumulated). Returns seconds spent (for the CPU baseline).
	double ref_dse_process_batch(void** effects, uint32_t n, const float* in, uint32_t frames, float* bus, int threads)
	{
		if (n == 0) return 0.0;
		auto* d0 = static_cast<RefDSE*>(effects[0]);
		const uint32 nSrc = d0->nSrc, nOut = d0->nOut;
		const size_t inStride = size_t(frames) * nSrc, busSize = size_t(frames) * nOut;
		if (threads < 1) threads = 1;
		std::vector<std::vector<float>> partial(threads);
		const auto t0 = std::chrono::steady_clock::now();
		if (threads == 1)
		{
			for (uint32_t i = 0; i < n; ++i) ref_dse_process(effects[i], in + inStride * i, bus, frames);
			return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
		}
		RunParallel(threads, [&](int t)
		{
			partial[t].assign(busSize, 0.0f);
			const uint32_t b = uint32_t(uint64_t(n) * t / threads), e = uint32_t(uint64_t(n) * (t + 1) / threads);
			([=, &partial] { for (uint32_t i = b; i < e; ++i) ref_dse_process(effects[i], in + inStride * i, partial[t].data(), frames); })();
		});
		for (int t = 0; t < threads; ++t) for (size_t i = 0; i < busSize; ++i) bus[i] += partial[t][i];
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	}
	void ref_dse_update_batch(void** effects, uint32_t n, const float* bandGains /*[n][4]*/, const float* delaySeconds /*[n]*/, const float* mixMaps /*[n][nSrc*nOut]*/)
	{
		if (n == 0) return;
		auto* d0 = static_cast<RefDSE*>(effects[0]);
		const size_t mm = size_t(d0->nSrc) * d0->nOut;
		for (uint32_t i = 0; i < n; ++i) ref_dse_update(effects[i], bandGains + 4 * size_t(i), delaySeconds[i], mixMaps + mm * i);
	}
	// UpdateParameters for every effect, sources partitioned over the pool's threads (each effect has its own RealtimeObject: no sharing)
	void ref_dse_update_batch_mt(void** effects, uint32_t n, const float* bandGains, const float* delaySeconds, const float* mixMaps, int threads)
	{
		if (n == 0) return;
		auto* d0 = static_cast<RefDSE*>(effects[0]);
		const size_t mm = size_t(d0->nSrc) * d0->nOut;
		RunParallel(threads, [&](int t)
		{
			const uint32_t b = uint32_t(uint64_t(n) * t / threads), e = uint32_t(uint64_t(n) * (t + 1) / threads);
			for (uint32_t i = b; i < e; ++i) ref_dse_update(effects[i], bandGains + 4 * size_t(i), delaySeconds[i], mixMaps + mm * i);
		});
	}
	// build flags of this library (bench.py reports them with the CPU baseline)
	const char* ref_build_info()
	{
		return
#if defined(JPL_USE_AVX2)
			"AVX2+FMA (-mavx2 -mfma, JPL_USE_AVX2=1 JPL_USE_FMADD=1)"
#else
			"SSE2 baseline (JPL_SIMD_CONFIGURED=1)"
#endif
#if defined(__OPTIMIZE__) && defined(JPLREF_O3)
			", -O3 -DNDEBUG"
#else
			", -O2 -DNDEBUG"
#endif
			;
	}
	//==========================================================================
	// ReverbBus (LateReverb.h:84-108)
	void* ref_reverb_create(float fs, float f1, float f2, float f3)
	{
		auto* r = new ReverbBus();
		SplitFrequencies sp{}; if (f1 > 0.0f) { sp.F1 = f1; sp.F2 = f2; sp.F3 = f3; }
		r->Prepare(fs, sp);
		return r;
	}
	void ref_reverb_destroy(void* h) { delete static_cast<ReverbBus*>(h); }
	void ref_reverb_set_rt60(void* h, const float* rt60) { static_cast<ReverbBus*>(h)->SetRT60(simd(rt60)); }
	void ref_reverb_set_decay_slope(void* h, const float* slope) { static_cast<ReverbBus*>(h)->SetDecaySlope(simd(slope)); }
	// in: interleaved [frames][nIn]; out: interleaved [frames][nOut], ACCUMULATED INTO (ReverbDesign.h:480-492)
	void ref_reverb_process(void* h, const float* in, uint32_t nIn, float* out, uint32_t nOut, uint32_t frames)
	{
		static_cast<ReverbBus*>(h)->ProcessInterleaved(std::span<const float>(in, size_t(frames) * nIn), std::span<float>(out, size_t(frames) * nOut), frames);
	}
	// dBToGain on 4 lanes (DecibelsAndGain.h:40-56): the decay gains of one FDN channel = dBToGain(slope * delaySeconds)
	void ref_db_to_gain(const float* db, float* out) { dBToGain(simd(db)).store(out); }
	double ref_reverb_process_batch(void** buses, uint32_t n, const float* in /*[n][frames]*/, float* out /*[n][frames][nOut]*/, uint32_t nOut, uint32_t frames, int threads)
	{
		if (threads < 1) threads = 1;
		const auto t0 = std::chrono::steady_clock::now();
		RunParallel(threads, [&](int t)
		{
			const uint32_t b = uint32_t(uint64_t(n) * t / threads), e = uint32_t(uint64_t(n) * (t + 1) / threads);
			([=] { for (uint32_t i = b; i < e; ++i) ref_reverb_process(buses[i], in + size_t(i) * frames, 1, out + size_t(i) * frames * nOut, nOut, frames); })();
		});
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	}

	//==========================================================================
	// ERBus (EarlyReflectionsBus.h:60-88)
	void* ref_erbus_create(float fs, uint32_t numChannels) { return new ERBus(fs, numChannels); }
	void ref_erbus_destroy(void* h) { delete static_cast<ERBus*>(h); }
	uint32_t ref_erbus_num_taps(void* h) { return static_cast<ERBus*>(h)->GetNumTaps(); }
	// taps: n x { filterGains[4], delaySeconds, id (as float bits), gains[numChannels] }
	void ref_erbus_set_taps(void* h, uint32_t n, const float* filterGains /*[n][4]*/, const float* delaySeconds /*[n]*/, const uint32_t* ids /*[n]*/, const float* gains /*[n][L]*/)
	{
		auto* b = static_cast<ERBus*>(h);
		const uint32 L = b->GetNumChannels();
		std::vector<ERBus::ERUpdateData> upd(n);
		for (uint32_t i = 0; i < n; ++i)
		{
			upd[i].FilterGains = simd(filterGains + 4 * size_t(i));
			upd[i].Gains.assign(gains + size_t(i) * L, gains + size_t(i + 1) * L);
			upd[i].Delay = delaySeconds[i];
			upd[i].Id = ids[i];
		}
		b->SetTaps(upd);
	}
	// in: interleaved [frames][nIn]; out: interleaved [frames][L], ACCUMULATED INTO
	void ref_erbus_process(void* h, const float* in, uint32_t nIn, float* out, uint32_t frames)
	{
		auto* b = static_cast<ERBus*>(h);
		b->ProcessInterleaved(std::span<const float>(in, size_t(frames) * nIn), std::span<float>(out, size_t(frames) * b->GetNumChannels()), nIn, frames);
	}
	double ref_erbus_process_batch(void** buses, uint32_t n, const float* in /*[n][frames]*/, float* out /*[n][frames][L]*/, uint32_t frames, int threads)
	{
		if (threads < 1) threads = 1;
		const auto t0 = std::chrono::steady_clock::now();
		RunParallel(threads, [&](int t)
		{
			const uint32_t b = uint32_t(uint64_t(n) * t / threads), e = uint32_t(uint64_t(n) * (t + 1) / threads);
			([=] { for (uint32_t i = b; i < e; ++i) { const uint32 L = static_cast<ERBus*>(buses[i])->GetNumChannels(); ref_erbus_process(buses[i], in + size_t(i) * frames, 1, out + size_t(i) * frames * L, frames); } })();
		});
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	}
	//==========================================================================
	// Scene front-end: DirectPathService, SanitizeSourceDir, spread-from-size,
	// distance attenuation, air absorption, speed of sound. Positions are 9 floats: location, up, forward.
	static Position<MinimalVec3> ToPos(const float* p)
	{
		return Position<MinimalVec3>{ .Location = MinimalVec3(p[0], p[1], p[2]),
			.Orientation = { .Up = MinimalVec3(p[3], p[4], p[5]), .Forward = MinimalVec3(p[6], p[7], p[8]) } };
	}

	// out [n][12] = distance, directionDot, invDirectionDot, location[3], up[3], forward[3]
	void ref_direct_path(const float* sources, const float* listeners, const uint32_t* listenerIndex, uint32_t n, float* out)
	{
		for (uint32_t i = 0; i < n; ++i)
		{
			const auto r = DirectPathService::ProcessDirectPath(ToPos(sources + size_t(i) * 9), ToPos(listeners + size_t(listenerIndex ? listenerIndex[i] : 0) * 9));
			float* o = out + size_t(i) * 12;
			o[0] = r.Distance; o[1] = r.DirectionDot; o[2] = r.InvDirectionDot;
			o[3] = r.Position.Location.X; o[4] = r.Position.Location.Y; o[5] = r.Position.Location.Z;
			o[6] = r.Position.Orientation.Up.X; o[7] = r.Position.Orientation.Up.Y; o[8] = r.Position.Orientation.Up.Z;
			o[9] = r.Position.Orientation.Forward.X; o[10] = r.Position.Orientation.Forward.Y; o[11] = r.Position.Orientation.Forward.Z;
		}
	}
	// kind 0: ProcessAngleAttenuationImpl(dot, cone) (what EvaluateDirection caches); 1: azimuth overload; 2: position overload (x = [n][3], refPoint [9])
	void ref_angle_attenuation(int kind, const float* x, const float* refPoint, const float* cones, uint32_t n, float* out)
	{
		for (uint32_t i = 0; i < n; ++i)
		{
			const AttenuationCone cone{ cones[2 * i], cones[2 * i + 1] };
			if (kind == 0) out[i] = DirectPathService::ProcessAngleAttenuationImpl(x[i], cone);
			else if (kind == 1) out[i] = DirectPathService::ProcessAngleAttenuation(x[i], cone);
			else out[i] = DirectPathService::ProcessAngleAttenuation(MinimalVec3(x[3 * i], x[3 * i + 1], x[3 * i + 2]), ToPos(refPoint), cone);
		}
	}
	void ref_attenuation_model(int model, float minDistance, float maxDistance, float rolloff, const float* distance, uint32_t n, float* out)
	{
		AttenuationBaseModel m;
		m.Model = static_cast<AttenuationModel>(model);
		m.Parameters = { minDistance, maxDistance, rolloff };
		const AttenuationFunction& f = m;
		for (uint32_t i = 0; i < n; ++i) out[i] = f.Evaluate(distance[i]);
	}
	// points [k][3] = distance, value, Curve::EType (as float)
	void ref_attenuation_curve(const float* points, uint32_t k, int sortPoints, const float* distance, uint32_t n, float* out)
	{
		AttenuationCurve c;
		for (uint32_t j = 0; j < k; ++j)
		{
			AttenuationCurve::Point p;
			p.Distance = points[3 * j]; p.Value = points[3 * j + 1]; p.FunctionType = static_cast<Curve::EType>(int(points[3 * j + 2]));
			c.Points.push_back(p);
		}
		if (sortPoints) c.SortPoints();
		const AttenuationFunction& f = c;
		for (uint32_t i = 0; i < n; ++i) out[i] = f.Evaluate(distance[i]);
	}
	// out [32]: centres, lowerThirdCentres, centre dB, lowerThird dB, weighted dB, decay slope, distanceToKnee, linearised factor; knee mask [4]
	void ref_air_cache(const float* bands, float tempC, float humidity, float pressureKPa, float nyquist, float* out, uint32_t* kneeMask)
	{
		const AirAbsorptionCache c = AirAbsorption::CacheParameters(simd(bands[0], bands[1], bands[2], bands[3]), AirAbsorptionParams{ tempC, humidity, pressureKPa }, nyquist);
		c.FrequencyBandCenters.store(out); c.FrequencyBandLowerThirdCenters.store(out + 4);
		c.CenterFrequencyAbsorption_dB.store(out + 8); c.LowerThirdBandAbsorption_dB.store(out + 12);
		c.HighFreqWeightedAbsorption_dB.store(out + 16); c.FrequencyDecaySlope.store(out + 20);
		c.LinAprox.DistanceToKnee.store(out + 24); c.LinAprox.LinearizedHFLossFactor.store(out + 28);
		float m[4]; simd::select(c.LinAprox.HasKnee, simd(1.0f), simd(0.0f)).store(m);
		for (int i = 0; i < 4; ++i) kneeMask[i] = m[i] != 0.0f;
	}
	void ref_default_bands(float* out) { JPL::cBandCenters.store(out); }
	// ComputeForDistance -> dB loss [n][4]; gains [n][4] = dBToGain(-loss) (simd flavour) when non-null
	void ref_air_for_distance(const float* bands, float tempC, float humidity, float pressureKPa, float nyquist, const float* distance, uint32_t n, float* outDb, float* outGain)
	{
		const AirAbsorptionCache c = AirAbsorption::CacheParameters(simd(bands[0], bands[1], bands[2], bands[3]), AirAbsorptionParams{ tempC, humidity, pressureKPa }, nyquist);
		for (uint32_t i = 0; i < n; ++i)
		{
			const simd loss = AirAbsorption::ComputeForDistance(distance[i], c);
			loss.store(outDb + size_t(i) * 4);
			if (outGain) dBToGain(simd(0.0f) - loss).store(outGain + size_t(i) * 4);
		}
	}
	float ref_air_for_frequency(float f, float tempC, float humidity, float pressureKPa) { return AirAbsorption::ComputeForFrequency(f, AirAbsorptionParams{ tempC, humidity, pressureKPa }); }
	float ref_speed_of_sound(float tempC) { return SpeedOfSound::ForTemperature(tempC); }
	float ref_speed_of_sound_reference(int inverse) { return inverse ? SpeedOfSound::cInvReference : SpeedOfSound::cReference; }
	void ref_sanitize_source_dir(const float* dirs, uint32_t n, float* out)
	{
		for (uint32_t i = 0; i < n; ++i)
		{
			const MinimalVec3 d = Impl::SanitizeSourceDir(MinimalVec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
			out[3 * i] = d.X; out[3 * i + 1] = d.Y; out[3 * i + 2] = d.Z;
		}
	}
	// flavour 0: SpatialManager::GetSpreadFromSourceSize (atan * 1/pi, what AdvanceSimulation calls); 1: the free function of PannerBase.h:69-74 (atanf * 2/pi)
	void ref_spread_from_size(int flavour, const float* size, const float* distance, uint32_t n, float* out)
	{
		for (uint32_t i = 0; i < n; ++i)
			out[i] = flavour == 0 ? Spatial::SpatialManager<MinimalVec3>::GetSpreadFromSourceSize(size[i], distance[i]) : JPL::GetSpreadFromSourceSize(size[i], distance[i]);
	}
	// The per-source body of SpatialManager::AdvanceSimulation (inc/SpatialManager.h:374-420) with the engine-side glue spelled out:
	// direct path -> curve -> cone -> spread from size -> sanitised direction; air absorption and delay from the distance.
	// Timed (single thread x `threads`), returns seconds; outputs like the GPU front-end.
	double ref_frontend_timed(const float* sources, const float* listeners, uint32_t n, const float* cones, const float* sizes, const float* model /* [4] */,
							  float* outDirect /* [n][4] */, float* outPan /* [n][5] */, float* outBand /* [n][4] */, float* outDelay, float* outAtt, float focus, int threads)
	{
		const AirAbsorptionCache& c = cDefaultAirAbsCache;
		AttenuationBaseModel m; m.Model = static_cast<AttenuationModel>(int(model[0])); m.Parameters = { model[1], model[2], model[3] };
		const auto t0 = std::chrono::steady_clock::now();
		RunParallel(threads, [&](int t)
		{
			const uint32_t b = uint32_t(uint64_t(n) * t / threads), e = uint32_t(uint64_t(n) * (t + 1) / threads);
			([=, &c, &m] {
				const Position<MinimalVec3> lis = ToPos(listeners);
				for (uint32_t i = b; i < e; ++i)
				{
					const auto r = DirectPathService::ProcessDirectPath(ToPos(sources + size_t(i) * 9), lis);
					outDirect[4 * i] = r.Distance; outDirect[4 * i + 1] = r.DirectionDot; outDirect[4 * i + 2] = r.InvDirectionDot;
					outDirect[4 * i + 3] = DirectPathService::ProcessAngleAttenuationImpl(r.InvDirectionDot, AttenuationCone{ cones[2 * i], cones[2 * i + 1] });
					outAtt[i] = static_cast<const AttenuationFunction&>(m).Evaluate(r.Distance);
					const float spread = Spatial::SpatialManager<MinimalVec3>::GetSpreadFromSourceSize(sizes[i], r.Distance);
					const MinimalVec3 d = Impl::SanitizeSourceDir(r.Position.Location);
					outPan[5 * i] = d.X; outPan[5 * i + 1] = d.Y; outPan[5 * i + 2] = d.Z; outPan[5 * i + 3] = focus; outPan[5 * i + 4] = spread;
					dBToGain(simd(0.0f) - AirAbsorption::ComputeForDistance(r.Distance, c)).store(outBand + size_t(i) * 4);
					outDelay[i] = r.Distance * SpeedOfSound::cInvReference;
				}
			})();
		});
		return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	}

	// ProcessVBAPData(sourceLayout, PanUpdateDataWithOrientation, out) (PannerBase.h:375,501-515); params [n][11] = dir, focus, spread, up, forward
	void ref_mdap_oriented(void* panner, void* layout, const float* params, uint32_t n, float* out)
	{
		auto* p = static_cast<RefPanner*>(panner);
		auto* l = static_cast<RefLayout*>(layout);
		if (!p->is3D) return;
		const uint32_t row = static_cast<uint32_t>(l->l3.ChannelGroups.size()) * p->p3.GetNumChannels();
		for (uint32_t i = 0; i < n; ++i)
		{
			const float* q = params + size_t(i) * 11;
			typename Panner3D::PanUpdateDataWithOrientation u{
				.Pan = { .SourceDirection = MinimalVec3(q[0], q[1], q[2]), .Focus = q[3], .Spread = q[4] },
				.Orientation = { .Up = MinimalVec3(q[5], q[6], q[7]), .Forward = MinimalVec3(q[8], q[9], q[10]) } };
			p->p3.ProcessVBAPData(l->l3, u, std::span<float>(out + size_t(i) * row, row));
		}
	}
	//==========================================================================
	// 3D panners with the other LUT storage variants (inc/Panning/VBAPLUT3D.h:93-104): kind 0 KB_983, 1 KB_851, 2 KB_786, 3 KB_655, 4 KB_65.
	// One-shot: builds the panner + source layout, runs ProcessVBAPData for n rows and GetSpeakerGains for m directions.
	} // extern "C" (templates need C++ linkage)
	template<VBAP::ELUTSize K>
	static int RunLutVariant(uint32_t targetMask, uint32_t sourceMask, const float* params, uint32_t n, float* outMix, const float* dirs, uint32_t m, float* outPoint)
	{
		using P = VBAPanner3D<VBAPStandardTraits, K>;
		auto p = std::make_unique<P>();
		if (!p->Initialize(ChannelMap::FromChannelMask(targetMask))) return 0;
		typename P::SourceLayoutType lay;
		if (!p->InitializeSourceLayout(ChannelMap::FromChannelMask(sourceMask), lay)) return 0;
		const uint32 L = p->GetNumChannels();
		const size_t stride = lay.ChannelGroups.size() * L;
		for (uint32_t i = 0; i < n; ++i)
		{
			const float* q = params + 5 * size_t(i);
			typename P::PanUpdateData d{ .SourceDirection = MinimalVec3(q[0], q[1], q[2]), .Focus = q[3], .Spread = q[4] };
			p->ProcessVBAPData(lay, d, std::span<float>(outMix + stride * i, stride));
		}
		for (uint32_t i = 0; i < m; ++i)
		{
			float* o = outPoint + size_t(i) * L;
			for (uint32 c = 0; c < L; ++c) o[c] = 0.0f;
			p->GetSpeakerGains(MinimalVec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]), std::span<float>(o, L));
		}
		return static_cast<int>(stride);
	}
	extern "C" {
	int ref_lut_variant(int kind, uint32_t targetMask, uint32_t sourceMask, const float* params, uint32_t n, float* outMix, const float* dirs, uint32_t m, float* outPoint)
	{
		switch (kind)
		{
		case 0: return RunLutVariant<VBAP::ELUTSize::KB_983>(targetMask, sourceMask, params, n, outMix, dirs, m, outPoint);
		case 1: return RunLutVariant<VBAP::ELUTSize::KB_851>(targetMask, sourceMask, params, n, outMix, dirs, m, outPoint);
		case 2: return RunLutVariant<VBAP::ELUTSize::KB_786>(targetMask, sourceMask, params, n, outMix, dirs, m, outPoint);
		case 3: return RunLutVariant<VBAP::ELUTSize::KB_655>(targetMask, sourceMask, params, n, outMix, dirs, m, outPoint);
		case 4: return RunLutVariant<VBAP::ELUTSize::KB_65>(targetMask, sourceMask, params, n, outMix, dirs, m, outPoint);
		default: return 0;
		}
	}
} // extern "C"
