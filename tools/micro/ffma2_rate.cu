#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 pk(float a, float b) { u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ float lo(u64 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return x + y; }
// V0: scalar FFMA acc=fma(c,x,acc), c from LDS.128, 16 chains (8 x regs x 2 "nodes")
// V1: FFMA2 acc2=fma2(bc(c),x2,acc2), 8 packed chains
// V2: FFMA2 all-packed operands (no bcast)
// V3: scalar FFMA, all register operands, no LDS
template <int V>
__global__ void __launch_bounds__(512) k(const float* xin, float* y, int iters, long long* cyc) {
  __shared__ float4 cs[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) cs[i] = make_float4(xin[i & 63], xin[(i + 1) & 63], xin[(i + 2) & 63], xin[(i + 3) & 63]);
  __syncthreads();
  float xa[8], xb[8], aa[8], ab[8]; u64 x2[8], a2[8];
  for (int i = 0; i < 8; ++i) { xa[i] = xin[threadIdx.x % 32 + i]; xb[i] = xin[threadIdx.x % 32 + 8 + i]; aa[i] = 0; ab[i] = 0; x2[i] = pk(xa[i], xb[i]); a2[i] = pk(0.f, 0.f); }
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float4 c = cs[(it * 8 + q) & 1023];
      float cc[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int r = (j * 2 + i) & 7;
          if (V == 0) { aa[r] = fmaf(cc[j], xa[(r + q) & 7], aa[r]); ab[r] = fmaf(cc[j], xb[(r + q) & 7], ab[r]); }
          if (V == 1) { a2[r] = ffma2(pk(cc[j], cc[j]), x2[(r + q) & 7], a2[r]); }
          if (V == 2) { a2[r] = ffma2(x2[(r + 3) & 7], x2[(r + q) & 7], a2[r]); }
          if (V == 3) { aa[r] = fmaf(xa[(r + 3) & 7], xa[(r + q) & 7], aa[r]); ab[r] = fmaf(xb[(r + 3) & 7], xb[(r + q) & 7], ab[r]); }
        }
      }
    }
  }
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; ++i) s += aa[i] + ab[i] + lo(a2[i]);
  y[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int V> void run(const char* name, float* x, float* y, long long* cyc, int threads) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int iters = 20000; int blocks = 148;
  k<V><<<blocks, threads>>>(x, y, 100, cyc); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<V><<<blocks, threads>>>(x, y, iters, cyc); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  // scalar-equivalent FMAs per thread per iter: 8 q * 4 j * 2 i * 2 nodes = 128
  double fma_thread = 128.0 * iters;
  double warps_per_smsp = threads / 32 / 4.0;
  double lane_fma_per_clk_smsp = fma_thread * 32 * warps_per_smsp / c;
  printf("%-28s thr=%4d: %.3f ms  %.1f TFLOP/s  cycles=%lld  => %.1f scalar-FMA lanes/clk/SMSP  (clk %.0f MHz) %s\n", name, threads, ms,
         2 * fma_thread * blocks * threads / ms * 1e-9, c, lane_fma_per_clk_smsp, c / ms * 1e-3, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  float *x, *y; long long* cyc; cudaMalloc(&x, 4096 * 4); cudaMalloc(&y, 148 * 512 * 4); cudaMalloc(&cyc, 8); cudaMemset(x, 0, 4096 * 4);
  for (int thr : {128, 256, 512}) {
    run<0>("FFMA c=LDS.128 bcast", x, y, cyc, thr); run<1>("FFMA2 bcast c=LDS.128", x, y, cyc, thr);
    run<2>("FFMA2 packed regs", x, y, cyc, thr); run<3>("FFMA regs", x, y, cyc, thr);
  }
  return 0;
}
