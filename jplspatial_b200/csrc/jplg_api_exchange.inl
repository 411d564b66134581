// C-ABI of the peer-memory bus exchange (include/jplg.h "Multi-GPU"). Part of jplg_api.cu's translation unit (context internals).

struct jplg_bus_exchange_t
{
	jplg_context ctx = nullptr;
	uint32_t world = 0, rank = 0, maxFloats = 0;
	bool root = false;
	unsigned char* base = nullptr;   // root: cudaMalloc; peer: cudaIpcOpenMemHandle mapping of the root's allocation
	// the peer traffic runs on a side stream from a private copy of the bus, so the context stream goes on with the next block while the
	// root waits for the other ranks' stores: staging[seq & 1], evReady = copy done (context stream), evDone = send / gather done (side stream)
	cudaStream_t side = nullptr;
	float* staging = nullptr;        // [2][maxFloats]
	cudaEvent_t evReady[2] = {}, evDone[2] = {};
	uint64_t submitted = 0;
	long long spinLimitCycles = 0;   // device-side wait time-out (jplg_bus_exchange_set_timeout_ms; default 2 s)
	uint32_t* errHost = nullptr;     // page-locked, device-mapped word the kernels raise on a time-out: readable without touching the device
	uint32_t* errDev = nullptr;      // its device alias
	int clockKHz = 0;
};

static jplg_status ExchangeSideInit(jplg_bus_exchange x)
{
	jplg_context ctx = x->ctx;
	{
		// highest priority: the exchange kernels are tiny and latency-bound, they should take the first free slot
		int prioLeast = 0, prioGreatest = 0;
		JPLG_CUDA_TRY(ctx, cudaDeviceGetStreamPriorityRange(&prioLeast, &prioGreatest));
		JPLG_CUDA_TRY(ctx, cudaStreamCreateWithPriority(&x->side, cudaStreamNonBlocking, prioGreatest));
	}
	JPLG_CUDA_TRY(ctx, cudaMalloc(reinterpret_cast<void**>(&x->staging), size_t(2) * x->maxFloats * sizeof(float)));
	JPLG_CUDA_TRY(ctx, cudaHostAlloc(reinterpret_cast<void**>(&x->errHost), sizeof(uint32_t), cudaHostAllocMapped));
	*x->errHost = 0;
	JPLG_CUDA_TRY(ctx, cudaHostGetDevicePointer(reinterpret_cast<void**>(&x->errDev), x->errHost, 0));
	JPLG_CUDA_TRY(ctx, cudaDeviceGetAttribute(&x->clockKHz, cudaDevAttrClockRate, ctx->device));
	if (x->spinLimitCycles == 0) x->spinLimitCycles = 2000ll * x->clockKHz;   // 2 s
	for (int i = 0; i < 2; ++i)
	{
		JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&x->evReady[i], cudaEventDisableTiming));
		JPLG_CUDA_TRY(ctx, cudaEventCreateWithFlags(&x->evDone[i], cudaEventDisableTiming));
	}
	return JPLG_OK;
}

static size_t ExchangeBytes(uint32_t world, uint32_t maxFloats) { return kExchangeDataOffset + size_t(2) * world * maxFloats * sizeof(float); }

jplg_status jplg_bus_exchange_create_root(jplg_context ctx, uint32_t world, uint32_t max_floats, jplg_bus_exchange* out, uint8_t handle_out[JPLG_IPC_HANDLE_BYTES])
{
	if (!out || !handle_out || world == 0 || world > kMaxWorld || max_floats == 0) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	JPLG_NEED_DEVICE(ctx);
	static_assert(sizeof(cudaIpcMemHandle_t) == JPLG_IPC_HANDLE_BYTES, "IPC handle size");
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	max_floats = (max_floats + 63u) & ~63u;   // slots stay 256-byte aligned
	auto* x = new jplg_bus_exchange_t{ ctx, world, 0, max_floats, true, nullptr };
	const size_t bytes = ExchangeBytes(world, max_floats);
	cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&x->base), bytes);
	if (e == cudaSuccess) e = cudaMemset(x->base, 0, bytes);
	cudaIpcMemHandle_t h;
	if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, x->base);
	if (e != cudaSuccess)
	{
		cudaGetLastError();
		if (x->base) cudaFree(x->base);
		delete x;
		return Fail(ctx, JPLG_ERR_CUDA, std::string("jplg_bus_exchange_create_root: ") + cudaGetErrorString(e));
	}
	std::memcpy(handle_out, &h, JPLG_IPC_HANDLE_BYTES);
	*out = x;
	return JPLG_OK;
}

jplg_status jplg_bus_exchange_open_peer(jplg_context ctx, uint32_t world, uint32_t rank, uint32_t max_floats, const uint8_t handle[JPLG_IPC_HANDLE_BYTES], jplg_bus_exchange* out)
{
	if (!out || !handle || world == 0 || world > kMaxWorld || rank == 0 || rank >= world || max_floats == 0) return JPLG_ERR_INVALID_ARGUMENT;
	*out = nullptr;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	max_floats = (max_floats + 63u) & ~63u;
	cudaIpcMemHandle_t h;
	std::memcpy(&h, handle, JPLG_IPC_HANDLE_BYTES);
	void* p = nullptr;
	const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
	if (e != cudaSuccess) { cudaGetLastError(); return Fail(ctx, JPLG_ERR_CUDA, std::string("jplg_bus_exchange_open_peer: ") + cudaGetErrorString(e)); }
	*out = new jplg_bus_exchange_t{ ctx, world, rank, max_floats, false, static_cast<unsigned char*>(p) };
	return JPLG_OK;
}

void jplg_bus_exchange_destroy(jplg_bus_exchange x)
{
	if (!x) return;
	if (x->ctx && x->ctx->device >= 0)
	{
		cudaSetDevice(x->ctx->device);
		cudaStreamSynchronize(x->ctx->stream);
		if (x->side)
		{
			cudaStreamSynchronize(x->side);
			cudaStreamDestroy(x->side);
			if (x->staging) cudaFree(x->staging);
			if (x->errHost) cudaFreeHost(x->errHost);
			for (int i = 0; i < 2; ++i) { cudaEventDestroy(x->evReady[i]); cudaEventDestroy(x->evDone[i]); }
		}
		if (x->base) { if (x->root) cudaFree(x->base); else cudaIpcCloseMemHandle(x->base); }
		cudaGetLastError();
	}
	delete x;
}

jplg_status jplg_bus_exchange_submit(jplg_bus_exchange x, const float* device_bus, uint32_t n, uint64_t seq, float* device_out, int32_t accumulate)
{
	if (!x || !device_bus || n == 0 || n > x->maxFloats || (x->root && !device_out)) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = x->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	if (!x->side) { const jplg_status st = ExchangeSideInit(x); if (st != JPLG_OK) return st; }
	auto* hdr = reinterpret_cast<ExchangeHeader*>(x->base);
	float* data = reinterpret_cast<float*>(x->base + kExchangeDataOffset);
	const int slot = static_cast<int>(seq & 1ull);
	float* stage = x->staging + size_t(slot) * x->maxFloats;
	if (x->submitted >= 2) JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, x->evDone[slot], 0));   // the block that last used this staging slot has left
	JPLG_CUDA_TRY(ctx, cudaMemcpyAsync(stage, device_bus, size_t(n) * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaEventRecord(x->evReady[slot], ctx->stream));
	JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(x->side, x->evReady[slot], 0));
	if (x->root)
		JPLG_LAUNCH_ON(ctx, JPLG_K_EXCHANGE, x->side, LaunchBusGather(hdr, data, x->world, x->maxFloats, stage, n, seq, device_out, accumulate ? 1 : 0, x->spinLimitCycles, x->errDev, x->side));
	else
		JPLG_LAUNCH_ON(ctx, JPLG_K_EXCHANGE, x->side, LaunchBusSend(hdr, data + (size_t(slot) * x->world + x->rank) * x->maxFloats, stage, n, x->rank, seq, x->spinLimitCycles, x->errDev, x->side));
	JPLG_CUDA_TRY(ctx, cudaEventRecord(x->evDone[slot], x->side));
	x->submitted++;
	return JPLG_OK;
}

jplg_status jplg_bus_exchange_wait(jplg_bus_exchange x)
{
	if (!x) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = x->ctx;
	JPLG_NEED_DEVICE(ctx);
	if (!x->side || x->submitted == 0) return JPLG_OK;
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	// the side stream is in order: waiting for the most recent block covers the earlier one
	JPLG_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, x->evDone[(x->submitted - 1) & 1ull], 0));
	return JPLG_OK;
}

jplg_status jplg_bus_exchange_set_timeout_ms(jplg_bus_exchange x, uint32_t milliseconds)
{
	if (!x || milliseconds == 0) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = x->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	int khz = 0;
	JPLG_CUDA_TRY(ctx, cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device));
	x->spinLimitCycles = static_cast<long long>(milliseconds) * khz;
	return JPLG_OK;
}

int32_t jplg_bus_exchange_poll_error(jplg_bus_exchange x)
{
	if (!x) return -1;
	return x->errHost ? static_cast<int32_t>(*reinterpret_cast<volatile uint32_t*>(x->errHost)) : 0;
}

jplg_status jplg_bus_exchange_clear_error(jplg_bus_exchange x)
{
	if (!x) return JPLG_ERR_INVALID_ARGUMENT;
	jplg_context ctx = x->ctx;
	JPLG_NEED_DEVICE(ctx);
	std::lock_guard<std::mutex> lk(ctx->mtx);
	cudaSetDevice(ctx->device);
	JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	if (x->side) JPLG_CUDA_TRY(ctx, cudaStreamSynchronize(x->side));
	if (x->errHost) *x->errHost = 0;
	if (x->root)
	{
		// the root owns the header: sequence flags and the acknowledgement restart from zero, every rank restarts its sequence numbers at 0
		JPLG_CUDA_TRY(ctx, cudaMemset(x->base, 0, kExchangeDataOffset));
	}
	x->submitted = 0;
	return JPLG_OK;
}

int32_t jplg_bus_exchange_error(jplg_bus_exchange x)
{
	if (!x || !x->base || x->ctx->device < 0) return -1;
	if (x->errHost && *reinterpret_cast<volatile uint32_t*>(x->errHost) != 0) return static_cast<int32_t>(*reinterpret_cast<volatile uint32_t*>(x->errHost));
	cudaSetDevice(x->ctx->device);
	uint32_t e = 0;
	if (cudaMemcpy(&e, x->base + offsetof(ExchangeHeader, error), 4, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return -1; }
	return static_cast<int32_t>(e);
}
