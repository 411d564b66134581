// Microbenchmark of k_scan_front's P3 loop (Thiran recurrence, lanes 0-3 of one warp, operands as one float4 per sample in shared memory):
// cycles per sample for several operand / result layouts. nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o thiran_loop_bench thiran_loop.cu
// Results: profiles/r02zz_scan_width_ab.txt (16.8 / 14.3 / 15.6 / 15.9 / 15.3 cycles per sample for V0..V4).
#include <cstdio>
#include <cuda_runtime.h>
constexpr int kT = 16, nT = 32, kRow = nT * (kT + 1) + 2;
// V0: the kernel's loop (16 LDS.128 per tile, STS.32 of y into .w per sample)   V1: no store   V2: y to a separate array, one STS.128 per 4 samples
// V3: V2 + operands as structure of arrays (3 LDS.128 per 4 samples)             V4: V2 with the next tile's operands loaded before this tile's chain
template<int V> __global__ void k(float* out, int reps, unsigned long long* stamps)
{
	extern __shared__ __align__(16) float4 sm[];
	float4* tp = sm + threadIdx.x * kRow;
	float4* ys = sm + 4 * kRow + threadIdx.x * (nT * kT / 4 + 1);
	float* soa = reinterpret_cast<float*>(sm + 4 * kRow + 4 * (nT * kT / 4 + 1)) + threadIdx.x * (3 * nT * kT + 4);
	for (int i = 0; i < kRow; ++i) tp[i] = make_float4(0.37f + 1e-3f * i, 0.01f * (i % 37), 0.02f * (i % 11), 0.f);
	for (int i = 0; i < 3 * nT * kT; ++i) soa[i] = 0.3f + 1e-4f * i;
	__syncthreads();
	float y1 = out[threadIdx.x];
	const unsigned long long c0 = clock64();
	for (int r = 0; r < reps; ++r)
	{
		if (V <= 2)
		{
			for (int t = 0; t < nT; ++t)
			{
				const int n0 = t * (kT + 1);
				float4 v[kT];
				#pragma unroll
				for (int i = 0; i < kT; ++i) v[i] = tp[n0 + i];
				float y[kT];
				#pragma unroll
				for (int i = 0; i < kT; ++i)
				{
					y1 = v[i].x * (v[i].y - y1) + v[i].z; y[i] = y1;
					if (V == 0) tp[n0 + i].w = y1;
					if (V == 2 && (i & 3) == 3) ys[t * 4 + (i >> 2)] = make_float4(y[i - 3], y[i - 2], y[i - 1], y[i]);
				}
			}
		}
		else if (V == 3)
		{
			const float4* a4 = reinterpret_cast<const float4*>(soa); const float4* x4 = a4 + nT * kT / 4; const float4* z4 = x4 + nT * kT / 4;
			for (int t = 0; t < nT; ++t)
			{
				float4 a[4], x[4], z[4];
				#pragma unroll
				for (int i = 0; i < 4; ++i) { a[i] = a4[t * 4 + i]; x[i] = x4[t * 4 + i]; z[i] = z4[t * 4 + i]; }
				#pragma unroll
				for (int i = 0; i < 4; ++i)
				{
					float4 y;
					y1 = a[i].x * (x[i].x - y1) + z[i].x; y.x = y1;
					y1 = a[i].y * (x[i].y - y1) + z[i].y; y.y = y1;
					y1 = a[i].z * (x[i].z - y1) + z[i].z; y.z = y1;
					y1 = a[i].w * (x[i].w - y1) + z[i].w; y.w = y1;
					ys[t * 4 + i] = y;
				}
			}
		}
		else
		{
			float4 v[kT], w[kT];
			#pragma unroll
			for (int i = 0; i < kT; ++i) v[i] = tp[i];
			for (int t = 0; t < nT; ++t)
			{
				const int n1 = min(t + 1, nT - 1) * (kT + 1);
				#pragma unroll
				for (int i = 0; i < kT; ++i) w[i] = tp[n1 + i];
				float y[kT];
				#pragma unroll
				for (int i = 0; i < kT; ++i)
				{
					y1 = v[i].x * (v[i].y - y1) + v[i].z; y[i] = y1;
					if ((i & 3) == 3) ys[t * 4 + (i >> 2)] = make_float4(y[i - 3], y[i - 2], y[i - 1], y[i]);
				}
				#pragma unroll
				for (int i = 0; i < kT; ++i) v[i] = w[i];
			}
		}
	}
	const unsigned long long c1 = clock64();
	out[blockIdx.x * blockDim.x + threadIdx.x] = y1 + tp[5].w + ys[3].x;
	if (threadIdx.x == 0 && blockIdx.x == 0) stamps[0] = c1 - c0;
}
int main()
{
	float* o; cudaMalloc(&o, 1 << 20); cudaMemset(o, 0, 1 << 20);
	unsigned long long* st; cudaMallocManaged(&st, 16);
	const int reps = 8; const size_t smem = 100 * 1024;
	for (int blocks : {1, 148, 296})
	{
		#define RUN(V, name) { cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); k<V><<<blocks, 4, smem>>>(o, reps, st); cudaDeviceSynchronize(); k<V><<<blocks, 4, smem>>>(o, reps, st); cudaDeviceSynchronize(); \
			printf("%-52s blocks %3d: %6.2f cycles per sample (%s)\n", name, blocks, double(st[0]) / (reps * nT * kT), cudaGetErrorString(cudaGetLastError())); }
		RUN(0, "V0 kernel loop: LDS.128 + STS.32 per sample");
		RUN(1, "V1 no store");
		RUN(2, "V2 y to its own array, STS.128 per 4 samples");
		RUN(3, "V3 SoA operands (3 LDS.128 per 4) + STS.128 per 4");
		RUN(4, "V4 V2 + next tile's operands prefetched");
	}
	return 0;
}
