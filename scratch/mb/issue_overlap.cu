// Microbenchmark: can a scheduler issue non-FP32 instructions in the second pipe cycle of a packed FFMA2 (sm_100a)?
// Per scheduler: WP warps run independent FFMA2 chains, WI warps run independent integer (LOP3/IADD3) chains or LDS loads, same instruction count.
// If the FP32 pipe's two-cycle occupancy leaves the issue slot free, time(mixed) ~ max(time(fp only), time(int only)); if not, ~ their sum.
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
// role: 0 = packed FP32, 1 = integer ALU, 2 = shared-memory loads, 3 = scalar FP32
__global__ void k(int nPackedWarps, int otherRole, int nOtherWarps, u64 c, int iters, long long* cyc, float* sink)
{
  __shared__ float sm[4096];
  const int warp = threadIdx.x >> 5, sched = warp & 3, slot = warp >> 2;   // slot-th warp on its scheduler
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = (float)i;
  __syncthreads();
  long long t0 = clock64();
  if (slot < nPackedWarps)
  {
    u64 x[8]; for (int k2 = 0; k2 < 8; ++k2) x[k2] = c + k2 + threadIdx.x;
    for (int i = 0; i < iters; ++i) {
      #pragma unroll
      for (int k2 = 0; k2 < 8; ++k2) x[k2] = fma2(x[k2], c, c);
    }
    u64 a = 0; for (int k2 = 0; k2 < 8; ++k2) a ^= x[k2];
    if (a == 12345) sink[0] = 1.0f;
  }
  else if (slot < nPackedWarps + nOtherWarps)
  {
    if (otherRole == 1)
    {
      unsigned y[8]; for (int k2 = 0; k2 < 8; ++k2) y[k2] = threadIdx.x * 7 + k2;
      for (int i = 0; i < iters; ++i) {
        #pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[k2]) : "r"(y[(k2 + 1) & 7]), "r"((unsigned)c));
      }
      unsigned a = 0; for (int k2 = 0; k2 < 8; ++k2) a ^= y[k2];
      if (a == 12345) sink[1] = 1.0f;
    }
    else if (otherRole == 2)
    {
      float acc[8] = {0,0,0,0,0,0,0,0}; unsigned idx = threadIdx.x;
      for (int i = 0; i < iters; ++i) {
        #pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)(__cvta_generic_to_shared(sm) + (((idx + k2 * 32) & 4095) << 2)))); acc[k2] = v; }
        idx += 1;
      }
      float a = 0; for (int k2 = 0; k2 < 8; ++k2) a += acc[k2];
      if (a == 12345.f) sink[2] = a;
    }
    else
    {
      float y[8]; for (int k2 = 0; k2 < 8; ++k2) y[k2] = threadIdx.x * 0.5f + k2;
      const float cc = __uint_as_float((unsigned)c);
      for (int i = 0; i < iters; ++i) {
        #pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(y[k2]) : "f"(cc));
      }
      float a = 0; for (int k2 = 0; k2 < 8; ++k2) a += y[k2];
      if (a == 12345.f) sink[3] = a;
    }
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main()
{
  long long* cyc; cudaMallocManaged(&cyc, 8); float* sink; cudaMalloc(&sink, 64);
  const float one = 1.0001f; unsigned b; memcpy(&b, &one, 4); u64 C = ((u64)b << 32) | b;
  const int iters = 20000;
  const char* names[] = {"", "int-ALU", "LDS", "scalar-FP32"};
  printf("# cycles per iteration (8 instructions per warp and iteration), per scheduler; 148 CTAs\n");
  for (int role = 1; role <= 3; ++role)
    for (int np = 0; np <= 2; ++np)
      for (int no = 0; no <= 2; ++no)
      {
        if (np + no == 0) continue;
        const int threads = (np + no) * 4 * 32;
        k<<<148, threads>>>(np, role, no, C, iters, cyc, sink); cudaDeviceSynchronize();
        k<<<148, threads>>>(np, role, no, C, iters, cyc, sink); cudaError_t e = cudaDeviceSynchronize();
        printf("packed warps/sched=%d  %s warps/sched=%d : %8.2f cycles/iter  (%s)\n", np, names[role], no, double(*cyc) / iters, e ? cudaGetErrorString(e) : "ok");
      }
  return 0;
}
