// C-ABI of the scene front-end (include/jplg.h "Scene front-end"). Part of jplg_api.cu's translation unit (included at its end, inside
// extern "C") because it uses the context internals (launch accounting, error state, stream). Host work here is table upload, staging
// and the once-per-environment builders of jplg_host.cpp; the per-source arithmetic is k_scene_frontend (jplg_frontend_kernels.cu).

struct jplg_frontend_t
{
	jplg_context ctx = nullptr;
	AttFunctionDev* dFuncs = nullptr;
	float* dPoints = nullptr;
	uint32_t numFuncs = 0;
	AirCacheDev air{};
	float invSpeed = 0.0f;
	float fallback[3] = {};
	int32_t spreadFlavour = 0;
	// staging arena for JPLG_MEM_HOST calls
	unsigned char* arena = nullptr;
	size_t arenaBytes = 0;
};

void jplg_air_absorption_cache_build(const float* bands, float temperature_c, float humidity_percent, float pressure_kpa, float nyquist,
									 jplg_air_absorption_cache* out)
{
	if (!out) return;
	float def[4];
	if (!bands) { DefaultBandCenters(def); bands = def; }
	AirCacheHost h;
	BuildAirAbsorptionCache(bands, temperature_c, humidity_percent, pressure_kpa, nyquist, h);
	std::memcpy(out->band_centers, h.centers, 16); std::memcpy(out->lower_third_centers, h.lowerThirdCenters, 16);
	std::memcpy(out->center_db, h.centerDb, 16); std::memcpy(out->lower_third_db, h.lowerThirdDb, 16);
	std::memcpy(out->weighted_db, h.weightedDb, 16); std::memcpy(out->decay_slope, h.decaySlope, 16);
	std::memcpy(out->distance_to_knee, h.distanceToKnee, 16); std::memcpy(out->linearized_loss_factor, h.linFactor, 16);
	std::memcpy(out->has_knee, h.hasKnee, 16);
}
void jplg_default_band_centers(float out[4]) { if (out) DefaultBandCenters(out); }
float jplg_speed_of_sound_for_temperature(float temperature_c) { return SpeedOfSoundForTemperature(temperature_c); }
void jplg_sanitize_fallback_dir(float out[3]) { if (out) SanitizeFallbackDir(out); }

jplg_status jplg_frontend_create(jplg_context ctx, const jplg_frontend_config* cfg, jplg_frontend* out)
{
	if (!out || !cfg) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	JPLG_NEED_DEVICE(ctx);
	if ((cfg->num_functions && !cfg->functions) || (cfg->num_points && !cfg->points)) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_frontend_create: null table");
	std::vector<AttFunctionDev> funcs(cfg->num_functions);
	std::vector<float> points(cfg->points, cfg->points + size_t(cfg->num_points) * 3);
	for (uint32_t i = 0; i < cfg->num_functions; ++i)
	{
		const jplg_attenuation_function& f = cfg->functions[i];
		if (f.kind == 1 && (uint64_t(f.first_point) + f.num_points > cfg->num_points))
			return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_frontend_create: curve point range outside the point table");
		funcs[i] = { f.kind, f.model, f.min_distance, f.max_distance, f.rolloff, f.first_point, f.num_points };
		if (f.kind == 1 && cfg->sort_points)
		{
			// AttenuationCurve::SortPoints (inc/DistanceAttenuation.h:149-155): std::sort by Distance
			struct P { float d, v, t; };
			P* b = reinterpret_cast<P*>(points.data()) + f.first_point;
			std::sort(b, b + f.num_points, [](const P& a, const P& c) { return a.d < c.d; });
		}
	}
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	auto* f = new jplg_frontend_t();
	f->ctx = ctx;
	f->numFuncs = cfg->num_functions;
	jplg_air_absorption_cache def;
	const jplg_air_absorption_cache* air = cfg->air;
	if (!air) { jplg_air_absorption_cache_build(nullptr, 20.0f, 50.0f, 101.325f, 22050.0f, &def); air = &def; }
	std::memcpy(f->air.centerDb, air->center_db, 16); std::memcpy(f->air.lowerThirdDb, air->lower_third_db, 16);
	std::memcpy(f->air.distanceToKnee, air->distance_to_knee, 16); std::memcpy(f->air.linFactor, air->linearized_loss_factor, 16);
	std::memcpy(f->air.hasKnee, air->has_knee, 16);
	f->invSpeed = cfg->speed_of_sound > 0.0f ? 1.0f / cfg->speed_of_sound : 1.0f / 343.2f;   // SpeedOfSound::cInvReference
	SanitizeFallbackDir(f->fallback);
	f->spreadFlavour = cfg->spread_flavour;
	cudaError_t e = cudaSuccess;
	if (!funcs.empty())
	{
		e = cudaMalloc(&f->dFuncs, funcs.size() * sizeof(AttFunctionDev));
		if (e == cudaSuccess) e = cudaMemcpy(f->dFuncs, funcs.data(), funcs.size() * sizeof(AttFunctionDev), cudaMemcpyHostToDevice);
	}
	if (e == cudaSuccess && !points.empty())
	{
		e = cudaMalloc(&f->dPoints, points.size() * 4);
		if (e == cudaSuccess) e = cudaMemcpy(f->dPoints, points.data(), points.size() * 4, cudaMemcpyHostToDevice);
	}
	if (e != cudaSuccess)
	{
		cudaGetLastError();
		if (f->dFuncs) cudaFree(f->dFuncs);
		if (f->dPoints) cudaFree(f->dPoints);
		delete f;
		return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, std::string("jplg_frontend_create: ") + cudaGetErrorString(e));
	}
	*out = f;
	return JPLG_OK;
}

void jplg_frontend_destroy(jplg_frontend f)
{
	if (!f) return;
	if (f->ctx && f->ctx->device >= 0)
	{
		cudaSetDevice(f->ctx->device);
		cudaStreamSynchronize(f->ctx->stream);
		if (f->dFuncs) cudaFree(f->dFuncs);
		if (f->dPoints) cudaFree(f->dPoints);
		if (f->arena) cudaFree(f->arena);
	}
	delete f;
}

jplg_status jplg_frontend_process(jplg_frontend f, const jplg_frontend_io* io, int32_t mem)
{
	if (mem != JPLG_MEM_HOST && mem != JPLG_MEM_DEVICE) return JPLG_ERR_INVALID_ARGUMENT;   // JPLG_MEM_HOST_ASYNC: mdap / update / process_block only
	if (!f || !io) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = f->ctx;
	JPLG_NEED_DEVICE(ctx);
	const uint32_t n = io->n;
	if (n == 0) return JPLG_OK;
	if (io->sources && (!io->listeners || io->num_listeners == 0)) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_frontend_process: sources without listeners");
	if (io->pan_params && io->pan_stride != 5 && io->pan_stride != 11) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_frontend_process: pan_stride must be 5 or 11");
	if (io->pan_params && io->pan_stride == 11 && !io->sources) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_frontend_process: oriented pan rows need source positions");
	if (io->cones && !io->direct) return Fail(ctx, JPLG_ERR_INVALID_ARGUMENT, "jplg_frontend_process: cones given but no `direct` output for the cone factor");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);

	FrontendArgs fa{};
	fa.n = n;
	fa.focusConst = io->focus; fa.spreadConst = io->spread;
	fa.att = f->dFuncs; fa.numAtt = f->numFuncs; fa.attPoints = f->dPoints;
	fa.air = f->air; fa.invSpeedOfSound = f->invSpeed;
	fa.fallbackDir[0] = f->fallback[0]; fa.fallbackDir[1] = f->fallback[1]; fa.fallbackDir[2] = f->fallback[2];
	fa.spreadFlavour = f->spreadFlavour;
	fa.panStride = io->pan_params ? io->pan_stride : 0;

	struct In { const void* host; size_t bytes; const void** dev; };
	struct Out { void* host; size_t bytes; void** dev; };
	const size_t N = n;
	In ins[] = {
		{ io->sources, N * 36, reinterpret_cast<const void**>(&fa.sources) },
		{ io->sources ? io->listeners : nullptr, size_t(io->num_listeners) * 36, reinterpret_cast<const void**>(&fa.listeners) },
		{ io->sources ? io->listener_index : nullptr, N * 4, reinterpret_cast<const void**>(&fa.listenerIndex) },
		{ io->sources ? nullptr : io->distance_in, N * 4, reinterpret_cast<const void**>(&fa.distanceIn) },
		{ io->sources ? nullptr : io->dot_in, N * 4, reinterpret_cast<const void**>(&fa.dotIn) },
		{ io->cones, N * 8, reinterpret_cast<const void**>(&fa.cones) },
		{ io->object_sizes, N * 4, reinterpret_cast<const void**>(&fa.sizes) },
		{ io->focus_spread, N * 8, reinterpret_cast<const void**>(&fa.focusSpread) },
		{ io->function_index, N * 4, reinterpret_cast<const void**>(&fa.attIndex) },
	};
	Out outs[] = {
		{ io->direct, N * 16, reinterpret_cast<void**>(&fa.direct) },
		{ io->relative, N * 36, reinterpret_cast<void**>(&fa.rel) },
		{ io->pan_params, N * 4 * fa.panStride, reinterpret_cast<void**>(&fa.pan) },
		{ io->distance_attenuation, N * 4, reinterpret_cast<void**>(&fa.distAtt) },
		{ io->air_db, N * 16, reinterpret_cast<void**>(&fa.airDb) },
		{ io->band_gains, N * 16, reinterpret_cast<void**>(&fa.bandGains) },
		{ io->delay_seconds, N * 4, reinterpret_cast<void**>(&fa.delay) },
	};
	if (mem == JPLG_MEM_HOST)
	{
		size_t total = 0;
		auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
		for (auto& i : ins) if (i.host) total += pad(i.bytes);
		for (auto& o : outs) if (o.host) total += pad(o.bytes);
		if (total > f->arenaBytes)
		{
			if (f->arena) { JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(f->arena); f->arena = nullptr; f->arenaBytes = 0; }
			if (cudaMalloc(&f->arena, total + total / 4) != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_OUT_OF_MEMORY, "front-end staging allocation failed"); }
			f->arenaBytes = total + total / 4;
		}
		size_t off = 0;
		for (auto& i : ins) if (i.host)
		{
			JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(f->arena + off, i.host, i.bytes, cudaMemcpyHostToDevice, ctx->stream));
			*i.dev = f->arena + off; off += pad(i.bytes);
		}
		for (auto& o : outs) if (o.host) { *o.dev = f->arena + off; off += pad(o.bytes); }
	}
	else
	{
		for (auto& i : ins) *i.dev = i.host;
		for (auto& o : outs) *o.dev = o.host;
	}
	JPLG_LAUNCH(ctx, JPLG_K_FRONTEND, LaunchFrontend(fa, ctx->numSMs, ctx->stream));
	if (mem == JPLG_MEM_HOST)
	{
		for (auto& o : outs) if (o.host) JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(o.host, *o.dev, o.bytes, cudaMemcpyDeviceToHost, ctx->stream));
		JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return JPLG_OK;
}
