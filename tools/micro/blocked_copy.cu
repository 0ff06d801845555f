// Does the per-irrep-block access order limit the K = 32 linears?  Copies a [N, 800] fp32 tensor (lmax 4, 32 channels)
//   A: block-major (what the block linears do): one pass per l over all nodes, 128-byte rows, (2l+1)*128 B contiguous per node
//   B: node-major: the same bytes, but a warp walks whole nodes (3200 contiguous bytes)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a blocked_copy.cu -o blocked_copy.bin
#include <cstdio>
#include <cuda_runtime.h>
__global__ void copy_block_major(const float* __restrict__ in, float* __restrict__ out, int N) {
  const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  for (int l = 0; l <= 4; ++l) {
    const int rows = 2 * l + 1, off = l * l * 32;
    const long total = (long)N * rows;
    const long per = (total + nw - 1) / nw;
    for (long r = warp * per; r < (warp + 1) * per && r < total; ++r) {
      const long n = r / rows, m = r % rows;
      out[n * 800 + off + m * 32 + lane] = in[n * 800 + off + m * 32 + lane] * 2.f;
    }
  }
}
__global__ void copy_node_major(const float* __restrict__ in, float* __restrict__ out, int N) {
  const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  const long per = ((long)N + nw - 1) / nw;
  for (long n = warp * per; n < (warp + 1) * per && n < N; ++n)
#pragma unroll 5
    for (int r = 0; r < 25; ++r) out[n * 800 + r * 32 + lane] = in[n * 800 + r * 32 + lane] * 2.f;
}
int main() {
  const int N = 30720;
  float *a, *b;
  cudaMalloc(&a, (size_t)N * 800 * 4 * 2);
  cudaMalloc(&b, (size_t)N * 800 * 4 * 2);
  cudaMemset(a, 0, (size_t)N * 800 * 4 * 2);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode)
    for (int ctas = 148 * 2; ctas <= 148 * 16; ctas *= 2) {
      float best = 1e9f;
      for (int it = 0; it < 6; ++it) {
        const float* src = a + (size_t)(it & 1) * N * 800;   // alternate two tensors: 400 MB working set > L2
        float* dst = b + (size_t)(it & 1) * N * 800;
        cudaEventRecord(e0);
        if (mode == 0) copy_block_major<<<ctas, 256>>>(src, dst, N); else copy_node_major<<<ctas, 256>>>(src, dst, N);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (it >= 2 && ms < best) best = ms;
      }
      printf("%s ctas=%5d: %.1f us  %.0f GB/s (read+write)\n", mode ? "node-major " : "block-major", ctas, best * 1e3, 2.0 * N * 800 * 4 / best / 1e6);
    }
  return 0;
}
