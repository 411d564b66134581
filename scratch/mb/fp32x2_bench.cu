// Microbenchmark: scalar FMUL/FADD SVF recurrences vs packed FMUL2/FFMA2 (f32x2) on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2(u64 a, u64 b){ u64 c; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(c) : "l"(a), "l"(b)); return c; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b){ u64 c; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(c) : "l"(a), "l"(b)); return c; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b){ u64 c; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(c) : "l"(a), "l"(b)); return c; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 pk(float x, float y){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(x), "f"(y)); return r; }
__device__ __forceinline__ float lo(u64 r){ float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(r)); return x; }

template<int ILP> __global__ void k_scalar(float* o, float g, float d, float mrg, int n, long long* cyc){
  float s1[ILP], s2[ILP], x[ILP];
  for (int k=0;k<ILP;k++){ s1[k]=o[threadIdx.x+k]; s2[k]=o[threadIdx.x+32+k]; x[k]=o[threadIdx.x+64+k]; }
  long long t0 = clock64();
  for (int i=0;i<n;i++){
    #pragma unroll
    for (int k=0;k<ILP;k++){
      float yH = (mrg*s1[k] + (x[k]-s2[k]))*d; float p = g*yH; float yB = p + s1[k]; s1[k] = p + yB; float q = g*yB; float yL = q + s2[k]; s2[k] = q + yL; x[k] = yL*0.5f + yH;
    }
  }
  long long t1 = clock64();
  float acc=0; for (int k=0;k<ILP;k++) acc += s1[k]+s2[k]+x[k];
  o[blockIdx.x*blockDim.x+threadIdx.x]=acc; if (threadIdx.x==0 && blockIdx.x==0) *cyc = t1-t0; }

template<int ILP> __global__ void k_packed(float* o, u64 g, u64 d, u64 mrg, u64 one, u64 half, int n, long long* cyc){
  u64 s1[ILP], s2[ILP], x[ILP];
  for (int k=0;k<ILP;k++){ s1[k]=pk(o[threadIdx.x+k],o[threadIdx.x+k+1]); s2[k]=pk(o[threadIdx.x+32+k],o[threadIdx.x+33+k]); x[k]=pk(o[threadIdx.x+64+k],o[threadIdx.x+65+k]); }
  long long t0 = clock64();
  for (int i=0;i<n;i++){
    #pragma unroll
    for (int k=0;k<ILP;k++){
      u64 t = sub2(x[k], s2[k]); u64 yH = mul2(fma2(mul2(mrg, s1[k]), one, t), d); u64 p = mul2(g, yH); u64 yB = fma2(p, one, s1[k]); s1[k] = fma2(p, one, yB);
      u64 q = mul2(g, yB); u64 yL = fma2(q, one, s2[k]); s2[k] = fma2(q, one, yL); x[k] = fma2(mul2(yL, half), one, yH);
    }
  }
  long long t1 = clock64();
  float acc=0; for (int k=0;k<ILP;k++) acc += lo(s1[k])+lo(s2[k])+lo(x[k]);
  o[blockIdx.x*blockDim.x+threadIdx.x]=acc; if (threadIdx.x==0 && blockIdx.x==0) *cyc = t1-t0; }

int main(){
  float* o; cudaMalloc(&o, 1<<24); cudaMemset(o, 0, 1<<24); long long* cyc; cudaMallocManaged(&cyc, 8);
  const int n = 20000; const float one1=1.0f; u64 ONE; { unsigned int b; memcpy(&b,&one1,4); ONE = (u64(b)<<32)|b; }
  auto P=[&](float v){ unsigned int b; memcpy(&b,&v,4); return (u64(b)<<32)|b; };
  printf("# instr per iter: scalar ILP*12, packed ILP*12 (each packed instr = 2 lanes)\n");
  for (int warps : {1,2,4,8,16}) {
    int threads = warps*32;   // one CTA per SM: warps/4 per scheduler (warps<4: fewer schedulers busy)
    #define RUN(K, ILPV, ...) { K<ILPV><<<148, threads>>>(__VA_ARGS__); cudaDeviceSynchronize(); K<ILPV><<<148, threads>>>(__VA_ARGS__); cudaError_t e=cudaDeviceSynchronize(); \
       printf("%-9s ILP=%d warps/SM=%2d  cycles/iter=%8.2f  cyc per warp-instr per scheduler=%6.3f %s\n", #K, ILPV, warps, double(*cyc)/n, double(*cyc)/n/(ILPV*12.0)/ (warps<4?1.0:warps/4.0), e?cudaGetErrorString(e):""); }
    RUN(k_scalar, 1, o, 0.01f, 0.9f, -1.4f, n, cyc); RUN(k_scalar, 2, o, 0.01f, 0.9f, -1.4f, n, cyc); RUN(k_scalar, 4, o, 0.01f, 0.9f, -1.4f, n, cyc); RUN(k_scalar, 8, o, 0.01f, 0.9f, -1.4f, n, cyc);
    RUN(k_packed, 1, o, P(0.01f), P(0.9f), P(-1.4f), ONE, P(0.5f), n, cyc); RUN(k_packed, 2, o, P(0.01f), P(0.9f), P(-1.4f), ONE, P(0.5f), n, cyc); RUN(k_packed, 4, o, P(0.01f), P(0.9f), P(-1.4f), ONE, P(0.5f), n, cyc); RUN(k_packed, 8, o, P(0.01f), P(0.9f), P(-1.4f), ONE, P(0.5f), n, cyc);
  }
  return 0; }
