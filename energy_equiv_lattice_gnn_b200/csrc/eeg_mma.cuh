// Warp-level tensor-core helpers shared by the K = 32 chain kernels (eeg_readout_mma.cu, eeg_linpair_mma.cu):
// mma.sync.m16n8k8 TF32 with fp32 accumulation, 3xTF32 operand split, and the accumulator -> A fragment re-layout.
#pragma once
#include <stdint.h>

constexpr int CM_FRAG = 4 * 32;   // floats per (8-column tile, k-step) of a weight image: [b0 hi | b0 lo | b1 hi | b1 lo][lane]

__device__ __forceinline__ uint32_t cm_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void cm_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// acc[j] += A (k-step s, fp32 values split hi/lo) x W fragments (tile j, k-step s) for j < NT
template <int NT>
__device__ __forceinline__ void cm_kstep(float (&acc)[4][4], const float (&av)[4], const float* frag, int s) {
  uint32_t ah[4], al[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    ah[q] = cm_tf32(av[q]);
    al[q] = __float_as_uint(av[q] - __uint_as_float(ah[q]));
  }
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const float* f = frag + (j * 4 + s) * CM_FRAG;
    const uint32_t b0h = __float_as_uint(f[0]), b0l = __float_as_uint(f[32]), b1h = __float_as_uint(f[64]), b1l = __float_as_uint(f[96]);
    cm_mma(acc[j], al, b0h, b1h);
    cm_mma(acc[j], ah, b0l, b1l);
    cm_mma(acc[j], ah, b0h, b1h);
  }
}
// accumulator fragment (columns 8 s .. 8 s + 7 of a 16-row tile) -> A fragment of k-step s
__device__ __forceinline__ void cm_c2a(const float (&c)[4], float (&a)[4], int lane) {
  const int t = lane & 3, src1 = (lane & ~3) | (t >> 1), src2 = src1 + 2;
  const bool odd = t & 1;
  const float p0 = __shfl_sync(0xffffffffu, c[0], src1), p1 = __shfl_sync(0xffffffffu, c[1], src1);
  const float p2 = __shfl_sync(0xffffffffu, c[2], src1), p3 = __shfl_sync(0xffffffffu, c[3], src1);
  const float q0 = __shfl_sync(0xffffffffu, c[0], src2), q1 = __shfl_sync(0xffffffffu, c[1], src2);
  const float q2 = __shfl_sync(0xffffffffu, c[2], src2), q3 = __shfl_sync(0xffffffffu, c[3], src2);
  a[0] = odd ? p1 : p0;   // (row g,     col t)
  a[1] = odd ? p3 : p2;   // (row g + 8, col t)
  a[2] = odd ? q1 : q0;   // (row g,     col t + 4)
  a[3] = odd ? q3 : q2;   // (row g + 8, col t + 4)
}

