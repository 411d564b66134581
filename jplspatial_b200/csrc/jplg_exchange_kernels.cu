// Peer-memory combine of the per-GPU partial speaker buses (the path's only exchange, [frames][L] floats per rank and block).
//
// The reference is a single process; here sources are sharded across the GPUs of one box and the buses must be summed. Instead of handing the
// 24-33 KB to a collective library, every non-root rank stores its bus straight into a slot of the ROOT GPU's gather buffer through an NVLink /
// NVSwitch peer mapping (CUDA IPC) and publishes a sequence flag; the root's gather kernel waits for the flags and adds the slots in rank order
// (fixed order: run-to-run deterministic). Two slots per rank: block k+1 may be stored while block k is being summed; a rank does not overwrite a
// slot before the root has acknowledged the block that used it (the acknowledgement lives in the root's memory and is polled over the peer mapping).
//
// Layout of the root allocation (ExchangeHeader, then data):
//   ack        uint64   number of blocks the root has consumed
//   done       uint32   CTA completion counter of the running gather
//   error      uint32   != 0 after a wait timed out (a rank died or fell behind by more than the exchange's time-out, 2 s by default); the kernel that
//                       timed out neither stores nor sums anything (fail closed), every rank also raises its own host-visible error word, nothing hangs
//   flags      uint64 [2][kMaxWorld]   (seq + 1) of the block a rank last published into that slot
//   data       float  [2][world][maxFloats]
#include "jplg_kernels.hpp"

namespace jplg
{
	__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
	{
		unsigned long long v;
		asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
		return v;
	}
	__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
	{
		asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
	}

	// non-root rank: one small CTA (128 threads x 32 registers: it has to fit beside the CTAs of the next block's panning / push kernels, which fill
	// whatever the render grid leaves free — a 1,024-thread CTA needed half an SM's register file and waited a whole block for it at 8 GPUs).
	// Wait until the root has consumed block seq-2 (the previous user of this slot), store the bus, publish.
	__global__ void __launch_bounds__(128) k_bus_send(ExchangeHeader* __restrict__ hdr, float* __restrict__ slotData, const float* __restrict__ bus,
													   uint32_t n, uint32_t rank, unsigned long long seq, long long spinLimit, uint32_t* __restrict__ localError)
	{
		__shared__ int ok;
		if (threadIdx.x == 0)
		{
			ok = 1;
			const long long t0 = clock64();
			while (ld_acquire_sys(&hdr->ack) + 2ull <= seq)
				if (clock64() - t0 > spinLimit) { ok = 0; atomicExch(&hdr->error, 1u); *localError = 1u; __threadfence_system(); break; }
		}
		__syncthreads();
		if (!ok) return;   // fail closed: the slot still belongs to a block the root has not consumed; nothing is stored or published
		const float4* src = reinterpret_cast<const float4*>(bus);
		float4* dst = reinterpret_cast<float4*>(slotData);
		for (uint32_t i = threadIdx.x; i < n / 4; i += blockDim.x) dst[i] = src[i];
		for (uint32_t i = (n & ~3u) + threadIdx.x; i < n; i += blockDim.x) slotData[i] = bus[i];
		__threadfence_system();
		__syncthreads();
		if (threadIdx.x == 0) st_release_sys(&hdr->flags[seq & 1ull][rank], seq + 1ull);
	}

	// root: wait for every peer's block `seq`, out (+)= own + slots in rank order, then acknowledge (last CTA).
	__global__ void __launch_bounds__(128) k_bus_gather(ExchangeHeader* __restrict__ hdr, const float* __restrict__ data, uint32_t world, uint32_t maxFloats,
														const float* __restrict__ own, uint32_t n, unsigned long long seq, float* __restrict__ out, int accumulate,
														long long spinLimit, uint32_t* __restrict__ localError)
	{
		__shared__ int bad;
		if (threadIdx.x == 0) bad = 0;
		__syncthreads();
		if (threadIdx.x > 0 && threadIdx.x < world)
		{
			const long long t0 = clock64();
			while (ld_acquire_sys(&hdr->flags[seq & 1ull][threadIdx.x]) != seq + 1ull)
				if (clock64() - t0 > spinLimit) { bad = 1; atomicExch(&hdr->error, 2u); *localError = 2u; __threadfence_system(); break; }
		}
		__syncthreads();
		// fail closed: a missing peer's slot holds an older block, so nothing is summed or written; the block is still acknowledged, otherwise every
		// sender would run into its own time-out for each of the following blocks
		const bool skip = bad != 0 || *reinterpret_cast<volatile uint32_t*>(&hdr->error) != 0;
		const float* slots = data + size_t(seq & 1ull) * world * maxFloats;
		// All of a thread's slot loads are issued before the first add: they bypass the caches (written by peer GPUs), and as a load-add-load-add
		// chain their latencies added up to a gather as long as the block itself at 8 GPUs (7 x 8 dependent loads per thread, ~3 us each beside a
		// running render), which made the root the slowest rank. The sum itself stays in rank order.
		const uint32_t n4 = ((reinterpret_cast<uintptr_t>(own) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(slots)) & 15u) == 0 && (maxFloats & 3u) == 0 ? n / 4 : 0;
		for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < (skip ? 0u : n4); i += gridDim.x * blockDim.x)
		{
			float4 v[kMaxWorld];
			#pragma unroll
			for (uint32_t r = 1; r < kMaxWorld; ++r)
				if (r < world) v[r] = __ldcv(reinterpret_cast<const float4*>(slots + size_t(r) * maxFloats) + i);
			float4 acc = reinterpret_cast<const float4*>(own)[i];
			#pragma unroll
			for (uint32_t r = 1; r < kMaxWorld; ++r)
				if (r < world) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
			float4* o4 = reinterpret_cast<float4*>(out) + i;
			if (accumulate) { const float4 o = *o4; acc.x = o.x + acc.x; acc.y = o.y + acc.y; acc.z = o.z + acc.z; acc.w = o.w + acc.w; }
			*o4 = acc;
		}
		for (uint32_t i = n4 * 4 + blockIdx.x * blockDim.x + threadIdx.x; i < (skip ? 0u : n); i += gridDim.x * blockDim.x)
		{
			float acc = own[i];
			for (uint32_t r = 1; r < world; ++r) acc += __ldcv(slots + size_t(r) * maxFloats + i);
			out[i] = accumulate ? out[i] + acc : acc;
		}
		__threadfence();
		__syncthreads();
		if (threadIdx.x == 0)
		{
			if (atomicAdd(&hdr->done, 1u) + 1u == gridDim.x)
			{
				hdr->done = 0;
				__threadfence_system();
				st_release_sys(&hdr->ack, seq + 1ull);
			}
		}
	}

	cudaError_t LaunchBusSend(ExchangeHeader* hdr, float* slotData, const float* bus, uint32_t n, uint32_t rank, uint64_t seq, long long spinLimitCycles,
							  uint32_t* localError, cudaStream_t st)
	{
		k_bus_send<<<1, 128, 0, st>>>(hdr, slotData, bus, n, rank, seq, spinLimitCycles, localError);
		return cudaGetLastError();
	}
	cudaError_t LaunchBusGather(ExchangeHeader* hdr, const float* data, uint32_t world, uint32_t maxFloats, const float* own, uint32_t n, uint64_t seq,
								float* out, int accumulate, long long spinLimitCycles, uint32_t* localError, cudaStream_t st)
	{
		const uint32_t grid = (n + 511) / 512 ? (n + 511) / 512 : 1;      // one float4 per thread
		k_bus_gather<<<grid, 128, 0, st>>>(hdr, data, world, maxFloats, own, n, seq, out, accumulate, spinLimitCycles, localError);
		return cudaGetLastError();
	}
}
