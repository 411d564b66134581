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
//   error      uint32   != 0 after a wait timed out (a rank died or fell > ~2 s behind); results are then undefined, nothing hangs
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
	constexpr long long kSpinLimitCycles = 4000000000ll;   // ~2 s at 1.965 GHz

	// non-root rank: one CTA. Wait until the root has consumed block seq-2 (the previous user of this slot), store the bus, publish.
	__global__ void __launch_bounds__(1024) k_bus_send(ExchangeHeader* __restrict__ hdr, float* __restrict__ slotData, const float* __restrict__ bus,
													   uint32_t n, uint32_t rank, unsigned long long seq)
	{
		__shared__ int ok;
		if (threadIdx.x == 0)
		{
			ok = 1;
			const long long t0 = clock64();
			while (ld_acquire_sys(&hdr->ack) + 2ull <= seq)
				if (clock64() - t0 > kSpinLimitCycles) { ok = 0; atomicExch(&hdr->error, 1u); break; }
		}
		__syncthreads();
		const float4* src = reinterpret_cast<const float4*>(bus);
		float4* dst = reinterpret_cast<float4*>(slotData);
		for (uint32_t i = threadIdx.x; i < n / 4; i += blockDim.x) dst[i] = src[i];
		for (uint32_t i = (n & ~3u) + threadIdx.x; i < n; i += blockDim.x) slotData[i] = bus[i];
		__threadfence_system();
		__syncthreads();
		if (threadIdx.x == 0) st_release_sys(&hdr->flags[seq & 1ull][rank], seq + 1ull);
		(void)ok;
	}

	// root: wait for every peer's block `seq`, out (+)= own + slots in rank order, then acknowledge (last CTA).
	__global__ void __launch_bounds__(256) k_bus_gather(ExchangeHeader* __restrict__ hdr, const float* __restrict__ data, uint32_t world, uint32_t maxFloats,
														const float* __restrict__ own, uint32_t n, unsigned long long seq, float* __restrict__ out, int accumulate)
	{
		if (threadIdx.x > 0 && threadIdx.x < world)
		{
			const long long t0 = clock64();
			while (ld_acquire_sys(&hdr->flags[seq & 1ull][threadIdx.x]) != seq + 1ull)
				if (clock64() - t0 > kSpinLimitCycles) { atomicExch(&hdr->error, 2u); break; }
		}
		__syncthreads();
		const float* slots = data + size_t(seq & 1ull) * world * maxFloats;
		for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
		{
			float acc = own[i];
			for (uint32_t r = 1; r < world; ++r) acc += __ldcv(slots + size_t(r) * maxFloats + i);   // written by a peer GPU: bypass stale lines
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

	cudaError_t LaunchBusSend(ExchangeHeader* hdr, float* slotData, const float* bus, uint32_t n, uint32_t rank, uint64_t seq, cudaStream_t st)
	{
		k_bus_send<<<1, 1024, 0, st>>>(hdr, slotData, bus, n, rank, seq);
		return cudaGetLastError();
	}
	cudaError_t LaunchBusGather(ExchangeHeader* hdr, const float* data, uint32_t world, uint32_t maxFloats, const float* own, uint32_t n, uint64_t seq,
								float* out, int accumulate, cudaStream_t st)
	{
		const uint32_t grid = (n + 1023) / 1024 ? (n + 1023) / 1024 : 1;
		k_bus_gather<<<grid, 256, 0, st>>>(hdr, data, world, maxFloats, own, n, seq, out, accumulate);
		return cudaGetLastError();
	}
}
