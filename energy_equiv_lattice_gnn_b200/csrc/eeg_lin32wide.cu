// Per-irrep channel mixing from multiplicity 32 to WIDE outputs (o3.Linear backward of the path-stacked interaction.linear,
// gnn/blocks.py:552-559: dM = dmsg W^T, K = 32 -> 160..320 columns per irrep block) on warp-level tensor-core MMAs.
// eeg_lin32mma.cu binds one warp to one (block, 32-column group) and stops at 16 groups; this shape has 42.  Here a CTA is
// bound to ONE irrep block: its weights (all column groups, <= 11 x 8 KB of hi / lo TF32 B fragments) sit in shared memory, a
// warp loads the A fragments of a 16-row tile ONCE (straight from global memory, next tile prefetched) and sweeps all column
// groups of the block with them: 3 x mma.sync.m16n8k8 (3xTF32) per (k-step, 8-column tile).  Accumulation chains are 12 MMAs
// long (K = 32), so the tensor core's truncating fp32 accumulation is not an issue here.  The CUDA-core GEMM ran this at
// 22 TFLOP/s (0.33 ms at 512 lattices), bound by shared-memory loads per FMA.
#include "eeg_common.cuh"

namespace {

constexpr int LW_WARPS = 8, LW_MAXG = 11, LW_MAXB = 8;

struct LwBlock {
  int rows, nout, groups, in_off, out_off;
  unsigned inv_rows;     // ceil(2^32 / rows) for rows > 1
  int64_t w_off;
  float scale;
  int cta_begin, n_ctas;
  int64_t tiles;         // 16-row tiles of the block
};
struct LwParams {
  LwBlock b[LW_MAXB];
  int n_blocks, in_stride, out_stride, transB;
};

__device__ __forceinline__ uint32_t lw_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void lw_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(LW_WARPS * 32, 2) k_lin32_wide(const float* __restrict__ in, const float* __restrict__ W,
                                                                float* __restrict__ out, int64_t N, const LwParams P) {
  extern __shared__ __align__(16) float lw_smem[];   // per group: B fragments [8-column tile j][k-step s][b0|b1][hi|lo][lane]
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3, warp = threadIdx.x >> 5;
  int bi = 0;
#pragma unroll
  for (int q = 1; q < LW_MAXB; ++q)
    if (q < P.n_blocks && (int)blockIdx.x >= P.b[q].cta_begin) bi = q;
  const LwBlock d = P.b[bi];
  // ---- weights of this block -> B fragments: b0 = W[k = 8 s + t][n], b1 = W[k = 8 s + t + 4][n], n = 32 grp + 8 j + g
  {
    const float* Wb = W + d.w_off;
    for (int grp = warp; grp < d.groups; grp += LW_WARPS) {
      float* sw = lw_smem + grp * 2048 + lane;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = 32 * grp + 8 * j + g;
#pragma unroll
        for (int s = 0; s < 4; ++s)
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int k = 8 * s + t + 4 * hh;
            const float w = d.scale * __ldg(P.transB ? Wb + (int64_t)n * 32 + k : Wb + (int64_t)k * d.nout + n);
            const float hi = __uint_as_float(lw_tf32(w));
            sw[(((j * 4 + s) * 2 + hh) * 2 + 0) * 32] = hi;
            sw[(((j * 4 + s) * 2 + hh) * 2 + 1) * 32] = w - hi;
          }
      }
    }
  }
  __syncthreads();

  // this CTA's tiles of the block: contiguous range, warps round robin
  const int c = blockIdx.x - d.cta_begin;
  const int64_t tile_lo = d.tiles * c / d.n_ctas, tile_hi = d.tiles * (c + 1) / d.n_ctas;
  const int64_t total_rows = N * d.rows;
  auto load = [&](int64_t tile, float (&x)[16], int64_t (&oo)[2], bool (&ok)[2]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int64_t r = tile * 16 + g + 8 * h;
      ok[h] = r < total_rows;
      const unsigned r32 = ok[h] ? (unsigned)r : 0u;
      const unsigned n = d.rows == 1 ? r32 : __umulhi(r32, d.inv_rows), m = r32 - n * d.rows;
      const float* p = in + (int64_t)n * P.in_stride + d.in_off + m * 32 + t;
      oo[h] = (int64_t)n * P.out_stride + d.out_off + (int64_t)m * d.nout;
#pragma unroll
      for (int s = 0; s < 4; ++s) {   // k-step s: columns 8 s + t and 8 s + t + 4
        x[(s * 2 + 0) * 2 + h] = ok[h] ? __ldg(p + 8 * s) : 0.f;
        x[(s * 2 + 1) * 2 + h] = ok[h] ? __ldg(p + 8 * s + 4) : 0.f;
      }
    }
  };
  float xn[16];
  int64_t on[2];
  bool okn[2];
  int64_t tile = tile_lo + warp;
  if (tile < tile_hi) load(tile, xn, on, okn);
  for (; tile < tile_hi; tile += LW_WARPS) {
    uint32_t ah[4][4], al[4][4];   // [k-step][a0: (row g, k t), a1: (row g+8, k t), a2: (row g, k t+4), a3: (row g+8, k t+4)]
    int64_t oo[2];
    bool ok[2];
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float v = xn[(s * 2 + (q >> 1)) * 2 + (q & 1)];
        ah[s][q] = lw_tf32(v);
        al[s][q] = __float_as_uint(v - __uint_as_float(ah[s][q]));
      }
#pragma unroll
    for (int h = 0; h < 2; ++h) { oo[h] = on[h]; ok[h] = okn[h]; }
    if (tile + LW_WARPS < tile_hi) load(tile + LW_WARPS, xn, on, okn);   // next tile's loads fly during the sweep below
#pragma unroll 1
    for (int grp = 0; grp < d.groups; ++grp) {
      const float* sw = lw_smem + grp * 2048 + lane;
      float acc[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[j][q] = 0.f;
#pragma unroll
      for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float* fr = sw + ((j * 4 + s) * 4) * 32;
          const uint32_t b0h = __float_as_uint(fr[0]), b0l = __float_as_uint(fr[32]), b1h = __float_as_uint(fr[64]), b1l = __float_as_uint(fr[96]);
          lw_mma(acc[j], al[s], b0h, b1h);
          lw_mma(acc[j], ah[s], b0l, b1l);
          lw_mma(acc[j], ah[s], b0h, b1h);
        }
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (ok[h]) {
          float* o = out + oo[h] + 32 * grp + 2 * t;
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<float2*>(o + 8 * j) = make_float2(acc[j][2 * h], acc[j][2 * h + 1]);
        }
    }
  }
}

}  // namespace

// 0 = launched, -1 = shape outside this kernel: every non-empty block must have input multiplicity 32 and an output
// multiplicity that is a multiple of 32 (<= 352), at least one of them wider than 32; no bias / residual / activation.
int eeg_lin32_wide_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                          int transB, float* out, const EegLayout* lout, int64_t N, cudaStream_t stream) {
  LwParams P;
  P.in_stride = lin->stride;
  P.out_stride = lout->stride;
  P.transB = transB;
  if ((lin->stride | lout->stride) & 1) return -1;
  int nb = 0, max_groups = 0;
  double cost[LW_MAXB], total_cost = 0.0;
  for (int b = 0; b < lin->n_blocks; ++b) {
    const int rows = lin->rows[b], nout = lout->mul[b];
    if (lin->mul[b] == 0 || nout == 0) continue;
    if (lin->mul[b] != 32 || rows != lout->rows[b] || rows <= 0 || nout % 32 || nout > 32 * LW_MAXG || (lout->off[b] & 1) || nb >= LW_MAXB) return -1;
    if (N * rows >= (1LL << 28)) return -1;
    LwBlock& d = P.b[nb];
    d.rows = rows;
    d.nout = nout;
    d.groups = nout / 32;
    d.in_off = lin->off[b];
    d.out_off = lout->off[b];
    d.inv_rows = rows > 1 ? (unsigned)(((1ULL << 32) + rows - 1) / rows) : 0u;
    d.w_off = host_w_off[b];
    d.scale = host_scale[b];
    d.tiles = (N * rows + 15) / 16;
    cost[nb] = (double)d.tiles * (d.groups + 0.5);
    total_cost += cost[nb];
    if (d.groups > max_groups) max_groups = d.groups;
    ++nb;
  }
  if (nb == 0 || max_groups < 2) return -1;
  P.n_blocks = nb;
  // CTAs per block in proportion to cost, two resident CTAs per SM
  const int target = 2 * EEG_NUM_SMS;
  int ctas = 0;
  for (int q = 0; q < nb; ++q) {
    int n = (int)(target * cost[q] / total_cost);
    if (n < 1) n = 1;
    if (n > P.b[q].tiles) n = (int)P.b[q].tiles;
    P.b[q].cta_begin = ctas;
    P.b[q].n_ctas = n;
    ctas += n;
  }
  const size_t smem = (size_t)max_groups * 2048 * sizeof(float);
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_lin32_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, LW_MAXG * 2048 * (int)sizeof(float));
    if (e != cudaSuccess) {
      eeg_set_error("eeg_lin32_wide: cannot opt in to shared memory: %s", cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  k_lin32_wide<<<(unsigned)ctas, LW_WARPS * 32, smem, stream>>>(in, W, out, N, P);
  EEG_CHECK_LAUNCH("eeg_lin32_wide");
  return 0;
}
