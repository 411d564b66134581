// Microbenchmark: latency of dependent FP32 operations on sm_100a (one warp per SM), and the SM clock seen through clock64 vs globaltimer.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
// variant 0: alpha as kernel parameter (uniform register operand); 1: alpha forced into a register; 2: FFMA chain; 3: same as 0 with a shared-memory store per step
template<int V> __global__ void k(float* out, float alpha, float tgt, int n, unsigned long long* stamps)
{
	__shared__ float sbuf[1024];
	float cur = out[threadIdx.x];
	float a = alpha;
	if (V == 1) asm volatile("mov.f32 %0, %1;" : "=f"(a) : "f"(alpha));
	const unsigned long long c0 = clock64(), g0 = gtime();
	#pragma unroll 16
	for (int i = 0; i < n; ++i)
	{
		if (V == 2) cur = __fmaf_rn(a, cur, tgt);
		else cur = a * (tgt - cur) + cur;
		if (V == 3) sbuf[(i & 1023)] = cur;
	}
	const unsigned long long c1 = clock64(), g1 = gtime();
	out[blockIdx.x * blockDim.x + threadIdx.x] = cur + (V == 3 ? sbuf[5] : 0.0f);
	if (threadIdx.x == 0 && blockIdx.x == 0) { stamps[0] = c1 - c0; stamps[1] = g1 - g0; }
}
int main()
{
	float* o; cudaMalloc(&o, 1 << 20); cudaMemset(o, 0, 1 << 20);
	unsigned long long* st; cudaMallocManaged(&st, 16);
	const int n = 1 << 16;
	for (int threads : {32, 4}) for (int blocks : {1, 148, 296})
	{
		#define RUN(V, ops, name) { k<V><<<blocks, threads>>>(o, 0.0005f, 3.0f, n, st); cudaDeviceSynchronize(); k<V><<<blocks, threads>>>(o, 0.0005f, 3.0f, n, st); cudaDeviceSynchronize(); \
			printf("%-28s blocks %3d threads %2d: %6.2f cycles per dependent op, SM clock %.0f MHz (clock64 / globaltimer)\n", name, blocks, threads, double(st[0]) / n / ops, double(st[0]) / double(st[1]) * 1e3); }
		RUN(0, 3, "sub,mul,add (uniform alpha)"); RUN(1, 3, "sub,mul,add (register alpha)"); RUN(2, 1, "ffma chain"); RUN(3, 3, "sub,mul,add + STS");
	}
	return 0;
}
