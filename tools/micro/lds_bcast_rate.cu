#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 pk(float a, float b) { u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ float lo(u64 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return x + y; }
// R = FFMA2 per LDS.128 (broadcast). MODE 0: broadcast float4 (all lanes same address); 1: per-lane distinct float4 (conflict-free 512B);
// 2: broadcast float (LDS.32) x4; 3: no LDS (coef regs)
template <int R, int MODE>
__global__ void __launch_bounds__(512) k(const float* xin, float* y, int iters, long long* cyc) {
  __shared__ float4 cs[2048];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) cs[i] = make_float4(xin[i & 63], xin[(i + 1) & 63], xin[(i + 2) & 63], xin[(i + 3) & 63]);
  __syncthreads();
  u64 x2[8], a2[8];
  for (int i = 0; i < 8; ++i) { x2[i] = pk(xin[threadIdx.x % 32 + i], xin[threadIdx.x % 32 + 8 + i]); a2[i] = pk(0.f, 0.f); }
  const float* cf = reinterpret_cast<const float*>(cs);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float cc[4];
      if (MODE == 0) { float4 c = cs[(it * 8 + q) & 1023]; cc[0] = c.x; cc[1] = c.y; cc[2] = c.z; cc[3] = c.w; }
      if (MODE == 1) { float4 c = cs[((it * 8 + q) * 32 & 1023) + (threadIdx.x & 31)]; cc[0] = c.x; cc[1] = c.y; cc[2] = c.z; cc[3] = c.w; }
      if (MODE == 2) { int b = ((it * 8 + q) & 1023) * 4; cc[0] = cf[b]; cc[1] = cf[b + 1]; cc[2] = cf[b + 2]; cc[3] = cf[b + 3]; }
      if (MODE == 3) { cc[0] = lo(x2[q]); cc[1] = cc[0]; cc[2] = cc[0]; cc[3] = cc[0]; }
#pragma unroll
      for (int r = 0; r < R; ++r) a2[r & 7] = ffma2(pk(cc[r & 3], cc[r & 3]), x2[(r + q) & 7], a2[r & 7]);
    }
  }
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; ++i) s += lo(a2[i]);
  y[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int R, int MODE> void run(float* x, float* y, long long* cyc, int threads) {
  int iters = 5000;
  k<R, MODE><<<148, threads>>>(x, y, 100, cyc); cudaDeviceSynchronize();
  k<R, MODE><<<148, threads>>>(x, y, iters, cyc); cudaDeviceSynchronize();
  long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  double lds_per_sm = 8.0 * iters * threads / 32;     // LDS(.128-equivalent) warp-instrs per SM
  double ffma2_per_smsp = 8.0 * iters * R * threads / 32 / 4;
  printf("R=%2d mode=%d thr=%3d: %.2f clk per LDS(SM-wide)  %.2f clk per FFMA2 (per SMSP)  %s\n", R, MODE, threads, c / lds_per_sm, c / ffma2_per_smsp,
         cudaGetErrorString(cudaGetLastError()));
}
int main() {
  float *x, *y; long long* cyc; cudaMalloc(&x, 4096 * 4); cudaMalloc(&y, 148 * 512 * 4); cudaMalloc(&cyc, 8); cudaMemset(x, 0, 4096 * 4);
  run<1, 0>(x, y, cyc, 512); run<1, 1>(x, y, cyc, 512); run<1, 2>(x, y, cyc, 512);
  run<6, 0>(x, y, cyc, 512); run<6, 1>(x, y, cyc, 512); run<6, 2>(x, y, cyc, 512); run<6, 3>(x, y, cyc, 512);
  run<12, 0>(x, y, cyc, 512); run<12, 2>(x, y, cyc, 512); run<12, 3>(x, y, cyc, 512);
  run<24, 0>(x, y, cyc, 512); run<24, 3>(x, y, cyc, 512);
  run<6, 0>(x, y, cyc, 256); run<12, 0>(x, y, cyc, 256);
  return 0;
}
