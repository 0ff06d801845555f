// Per-irrep channel mixing for multiplicity-32 inputs (o3.Linear with mul_in = 32: gnn/blocks.py:516-521,552-559,471-476,
// 276-279) on warp-level tensor-core MMAs, node-major.  K = 32 is too short for the tcgen05 pipeline (128-row tiles, TMEM
// round trip: 140-250 us measured) and the CUDA-core versions are bound by the per-block sweep (2.9 TB/s); here a CTA walks
// chunks of whole nodes, each warp is bound to one (irrep block, 32-column group) - in proportion to the rows per node - and
// handles 16-row tiles: A fragments of mma.sync.m16n8k8 are loaded straight from global memory (the next tile's loads fly
// during the current MMAs), the group's weights sit in shared memory as hi/lo TF32 B fragments, 3 MMAs per (k-step,
// 8-column tile) give fp32-grade accuracy (3xTF32).
#include "eeg_common.cuh"

namespace {

constexpr int LM_WARPS = 16, LM_NB = 32;   // warps per CTA; nodes per chunk
constexpr int LM_MAXG = 16;                // (block, 32-column group) pairs

struct LmBlock {
  int rows, nout, in_off, out_off, n0;   // n0: first output column of this (block, column group)
  unsigned inv_rows;                     // ceil(2^32 / rows) for rows > 1
  int64_t w_off;
  float scale;
  int ntile8;                            // number of 8-column tiles in this group (<= 4)
  int has_bias;
};
struct LmParams {
  LmBlock b[LM_MAXG];
  int n_groups, in_stride, out_stride, transB;
  unsigned char warp_group[LM_WARPS], warp_rank[LM_WARPS], group_warps[LM_MAXG];
};

__device__ __forceinline__ uint32_t lm_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void lm_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(LM_WARPS * 32, 1) k_lin32_mma(const float* __restrict__ in, const float* __restrict__ W,
                                                               const float* __restrict__ bias, const float* __restrict__ residual,
                                                               float* __restrict__ out, int64_t N, LmParams P, int n_chunks) {
  extern __shared__ __align__(16) float lm_smem[];   // per group: B fragments [8-column tile][k-step][b0|b1][hi|lo][lane]
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3, warp = threadIdx.x >> 5;
  const int gi = P.warp_group[warp], rank = P.warp_rank[warp], nw = P.group_warps[gi];
  const LmBlock d = P.b[gi];
  float* sw = lm_smem + gi * 2048 + lane;
  if (rank == 0) {   // b0 = W[k = 8 s + t][n], b1 = W[k = 8 s + t + 4][n], n = n0 + 8 j + g; scale folded in; hi / lo split
    const float* Wb = W + d.w_off;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = d.n0 + 8 * j + g;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const int k = 8 * s + t + 4 * hh;
          const float w = j < d.ntile8 ? d.scale * __ldg(P.transB ? Wb + (int64_t)n * 32 + k : Wb + (int64_t)k * d.nout + n) : 0.f;
          const float hi = __uint_as_float(lm_tf32(w));
          sw[(((j * 4 + s) * 2 + hh) * 2 + 0) * 32] = hi;
          sw[(((j * 4 + s) * 2 + hh) * 2 + 1) * 32] = w - hi;
        }
      }
    }
  }
  float bv[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool okb = j < d.ntile8 && d.has_bias;
    bv[j][0] = okb ? __ldg(bias + d.n0 + 8 * j + 2 * t) : 0.f;
    bv[j][1] = okb ? __ldg(bias + d.n0 + 8 * j + 2 * t + 1) : 0.f;
  }
  __syncthreads();

  // this warp's tiles, flattened over the CTA's chunks: r0 advances by nw tiles inside a chunk, then to the next chunk
  int chunk = blockIdx.x;
  unsigned r0 = rank * 16;
  auto rows_in = [&](int c) {
    const int64_t left = N - (int64_t)c * LM_NB;
    return (unsigned)((left < LM_NB ? left : LM_NB) * d.rows);
  };
  int shift = 0;   // tiles of this group seen so far (mod nw): the group's warps take tiles round robin ACROSS chunks, so the
                   // leftover tiles of a chunk do not always land on the same warps
  auto settle = [&]() {   // move to the first existing tile at or after (chunk, r0)
    while (chunk < n_chunks && r0 >= rows_in(chunk)) {
      shift = (shift + (int)((rows_in(chunk) + 15) / 16)) % nw;
      chunk += gridDim.x;
      r0 = (unsigned)((rank + nw - shift) % nw) * 16;
    }
  };
  auto load = [&](float (&x)[16], float2 (&rr)[4][2], int64_t (&oo)[2], bool (&ok)[2]) {
    const unsigned rows_here = rows_in(chunk), r_base = (unsigned)((int64_t)chunk * LM_NB * d.rows);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const unsigned rl = r0 + g + 8 * h, r = r_base + rl;
      ok[h] = rl < rows_here;
      const unsigned n = d.rows == 1 ? r : __umulhi(r, d.inv_rows), m = r - n * d.rows;
      const int64_t oi = ok[h] ? (int64_t)n * P.in_stride + d.in_off + m * 32 : 0;
      oo[h] = ok[h] ? (int64_t)n * P.out_stride + d.out_off + m * d.nout + d.n0 : 0;
      const float* p = in + oi + t;
#pragma unroll
      for (int s = 0; s < 4; ++s) {   // k-step s: columns 8 s + t and 8 s + t + 4
        x[(s * 2 + 0) * 2 + h] = ok[h] ? __ldg(p + 8 * s) : 0.f;
        x[(s * 2 + 1) * 2 + h] = ok[h] ? __ldg(p + 8 * s + 4) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        rr[j][h] = (residual != nullptr && ok[h] && j < d.ntile8) ? __ldg(reinterpret_cast<const float2*>(residual + oo[h] + 8 * j + 2 * t))
                                                                  : make_float2(0.f, 0.f);
    }
  };

  settle();
  float xn[16];
  float2 rn[4][2];
  int64_t on[2];
  bool okn[2];
  if (chunk < n_chunks) load(xn, rn, on, okn);
  while (chunk < n_chunks) {
    float x[16];
    int64_t oo[2];
    bool ok[2];
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = xn[i];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      oo[h] = on[h];
      ok[h] = okn[h];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[j][2 * h + 0] = bv[j][0] + rn[j][h].x;
        acc[j][2 * h + 1] = bv[j][1] + rn[j][h].y;
      }
    }
    r0 += nw * 16;
    settle();
    if (chunk < n_chunks) load(xn, rn, on, okn);   // next tile's loads fly during the MMAs below
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      uint32_t ah[4], al[4];   // a0: (row g, k t), a1: (row g+8, k t), a2: (row g, k t+4), a3: (row g+8, k t+4)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float v = x[(s * 2 + (q >> 1)) * 2 + (q & 1)];
        ah[q] = lm_tf32(v);
        al[q] = __float_as_uint(v - __uint_as_float(ah[q]));
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j < d.ntile8) {
          const float* f = sw + ((j * 4 + s) * 4) * 32;
          const uint32_t b0h = __float_as_uint(f[0]), b0l = __float_as_uint(f[32]), b1h = __float_as_uint(f[64]), b1l = __float_as_uint(f[96]);
          lm_mma(acc[j], al, b0h, b1h);
          lm_mma(acc[j], ah, b0l, b1l);
          lm_mma(acc[j], ah, b0h, b1h);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (j < d.ntile8) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (ok[h]) *reinterpret_cast<float2*>(out + oo[h] + 8 * j + 2 * t) = make_float2(acc[j][2 * h], acc[j][2 * h + 1]);
      }
    }
  }
}

}  // namespace

// 0 = launched, -1 = shape outside this kernel (input multiplicity 32, output multiplicities multiples of 8, <= 16 groups)
int eeg_lin32_mma_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                         const float* bias, int transB, const float* residual, float* out, const EegLayout* lout, int64_t N,
                         cudaStream_t stream) {
  LmParams P;
  P.in_stride = lin->stride;
  P.out_stride = lout->stride;
  P.transB = transB;
  if ((lin->stride | lout->stride) & 1) return -1;
  int ng = 0;
  for (int b = 0; b < lin->n_blocks; ++b) {
    const int rows = lin->rows[b], nout = lout->mul[b];
    if (lin->mul[b] != 32 || rows != lout->rows[b] || rows <= 0 || nout <= 0 || nout % 8 || (lout->off[b] & 1)) return -1;
    if (N * rows >= (1LL << 28)) return -1;
    for (int n0 = 0; n0 < nout; n0 += 32) {
      if (ng >= LM_MAXG) return -1;
      LmBlock& d = P.b[ng++];
      d.rows = rows;
      d.nout = nout;
      d.in_off = lin->off[b];
      d.out_off = lout->off[b];
      d.n0 = n0;
      d.inv_rows = rows > 1 ? (unsigned)(((1ULL << 32) + rows - 1) / rows) : 0u;
      d.w_off = host_w_off[b];
      d.scale = host_scale[b];
      d.ntile8 = (nout - n0 >= 32 ? 32 : nout - n0) / 8;
      d.has_bias = bias != nullptr && b == 0;
    }
  }
  if (ng == 0 || ng > LM_WARPS) return -1;
  P.n_groups = ng;
  // warps per group in proportion to the rows per node: one each, then greedily to the most loaded group
  int wcount[LM_MAXG];
  for (int q = 0; q < ng; ++q) wcount[q] = 1;
  for (int w = ng; w < LM_WARPS; ++w) {
    int best = 0;
    for (int q = 1; q < ng; ++q)
      if ((int64_t)P.b[q].rows * wcount[best] > (int64_t)P.b[best].rows * wcount[q]) best = q;
    ++wcount[best];
  }
  for (int q = 0, w = 0; q < ng; ++q) {
    P.group_warps[q] = (unsigned char)wcount[q];
    for (int r = 0; r < wcount[q]; ++r, ++w) {
      P.warp_group[w] = (unsigned char)q;
      P.warp_rank[w] = (unsigned char)r;
    }
  }
  const size_t smem = (size_t)ng * 2048 * sizeof(float);
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_lin32_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, LM_MAXG * 2048 * (int)sizeof(float));
    if (e != cudaSuccess) {
      eeg_set_error("eeg_lin32_mma: cannot opt in to shared memory: %s", cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  const int64_t chunks = eeg_div_up(N, LM_NB);
  if (chunks >= (1LL << 31)) return -1;
  const int64_t ctas = chunks < EEG_NUM_SMS ? chunks : EEG_NUM_SMS;
  k_lin32_mma<<<(unsigned)ctas, LM_WARPS * 32, smem, stream>>>(in, W, bias, residual, out, N, P, (int)chunks);
  EEG_CHECK_LAUNCH("eeg_lin32_mma");
  return 0;
}
