// Explicit backward kernels of the readout / tail / linear stages (the reference relies on autograd over e3nn ops):
//   eeg_block_linear_wgrad : dW_b[u,w] += scale_b sum_{n,m} x[n,b,m,u] dy[n,b,m,w]  (+ bias gradient on the 0e block)
//   eeg_colsum             : out[c] += sum_r a[r] x[r,c]   (bias gradients of the radial MLP, embedding gradients)
//   eeg_silu_fwd / _bwd    : SiLU of the radial MLP (gnn/blocks.py:538-549), recomputed in the backward pass
//   eeg_gate_bwd           : e3nn.nn.Gate backward (gnn/blocks.py:268-282)
//   eeg_pool_tail_bwd      : PositiveLayer, Cart_4_to_Mandel, Spherical_to_Cartesian, final o3.Linear and graph pooling
//                            (gnn/model.py:100-111, gnn/blocks.py:185-229,392-442) + eeg_tail_wgrad for its weights
// The input-gradient of a linear is eeg_block_linear_fwd with transB = 1 (dX = dY W^T).
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "eeg_common.cuh"
#include "eeg_eig6.cuh"

namespace {

// ------------------------------------------------------------------------------------------------ linear weight gradient
constexpr int WG_ROWS_MAX = 4096;   // rows per CTA (upper bound; see wg_rows_per_cta)

struct WgBlock {
  int rows, k, nout, in_off, in_ld, out_off;
  int64_t w_off;
  float scale;
  int cta_begin, ktiles, ntiles;
};
struct WgParams {
  WgBlock b[EEG_MAX_L];
  int n_blocks, in_stride, out_stride, transB;
  int wg_rows;   // rows per CTA
};

// Rows per CTA: enough row slices that the grid holds >= ~4 CTAs per SM (the K = 32 -> 32 linears have ONE 32 x 32 tile per
// irrep block: with 4096 rows per CTA their weight gradients ran on 94 CTAs at 0.8 TB/s), multiple of 32, at most WG_ROWS_MAX.
// K tile of a launch: 64 when every non-empty block has K >= 64, else 32 (the five 32 -> 32 mixes)
static int wg_tile_k(const EegLayout* lin, const EegLayout* lout, const int32_t* k_override) {
  int tk = 64;
  for (int b = 0; b < lin->n_blocks; ++b) {
    const int k = k_override ? k_override[b] : lin->mul[b], n = lout ? lout->mul[b] : 32;
    if (k > 0 && n > 0 && k < 64) tk = 32;
  }
  return tk;
}
static int wg_rows_per_cta(const EegLayout* lin, const EegLayout* lout, const int32_t* k_override, int64_t N) {
  const int TK = wg_tile_k(lin, lout, k_override);
  int64_t rows_tiles = 0;   // sum over blocks of rows * tiles
  for (int b = 0; b < lin->n_blocks; ++b) {
    const int k = k_override ? k_override[b] : lin->mul[b], n = lout ? lout->mul[b] : 32;
    if (k <= 0 || n <= 0) continue;
    rows_tiles += N * lin->rows[b] * (int64_t)((k + TK - 1) / TK) * ((n + 31) / 32);
  }
  int64_t r = rows_tiles / (4 * EEG_NUM_SMS);
  r = (r + 31) / 32 * 32;
  if (r < 256) r = 256;
  if (r > WG_ROWS_MAX) r = WG_ROWS_MAX;
  return (int)r;
}

// ws != nullptr: deterministic mode - the row slice `cta / tiles` of a tile writes its partial sums to
// ws[slice * ws_stride + element] (bias partials behind the weights at n_weights + column) and k_wgrad_reduce adds the
// slices in order; ws == nullptr: one atomicAdd per element and slice (order depends on scheduling).
// KPT = k values per thread: CTA tile = (16 * KPT) x 32 outputs, thread tile KPT x 4.  KPT = 4 (64 x 32 tile, used when every
// block has K >= 64) needs two LDS.128 per 16 FMAs; the original KPT = 2 needed two loads per 8 FMAs and ran the K = 1344
// radial-MLP and K = 160..320 interaction weight gradients at 17 TFLOP/s.  Rows are accumulated in the same order for both.
template <int KPT>
__global__ void __launch_bounds__(128) k_linear_wgrad(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dW,
                                                      float* __restrict__ dbias, int64_t N, WgParams P, float* __restrict__ ws,
                                                      int64_t ws_stride, int64_t n_weights) {
  constexpr int TK = 16 * KPT, XLD = TK + 4;
  __shared__ __align__(16) float Xs[32][XLD];
  __shared__ __align__(16) float Ys[32][36];
  int cta = blockIdx.x, bi = 0;
#pragma unroll
  for (int q = 1; q < EEG_MAX_L; ++q)
    if (q < P.n_blocks && cta >= P.b[q].cta_begin) bi = q;
  const WgBlock d = P.b[bi];
  cta -= d.cta_begin;
  const int tiles = d.ktiles * d.ntiles;
  const int kt = (cta % tiles) / d.ntiles, nt = (cta % tiles) % d.ntiles;
  const int64_t r_begin = (int64_t)(cta / tiles) * P.wg_rows, total = N * d.rows;
  const int64_t r_end = r_begin + P.wg_rows < total ? r_begin + P.wg_rows : total;
  const int t = threadIdx.x, tk = t >> 3, tn = t & 7;   // thread: k = KPT tk + {0..KPT-1}, n = 4 tn + {0..3}
  const int lr = t >> 2, lq = t & 3;                     // loader: row lr, float4 columns lq, lq + 4, ...
  float acc[KPT][4];
#pragma unroll
  for (int i = 0; i < KPT; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};
  const bool want_bias = dbias != nullptr && bi == 0 && kt == 0 && tk == 0;
  // the next 32-row chunk is loaded into registers while the current one is multiplied (the loop was bound by the exposed
  // load latency: two barriers and one global round trip per 32 rows)
  float4 xv[TK / 16], yv[2];
  auto fetch = [&](int64_t r0) {
    const int64_t r = r0 + lr;
    const bool ok = r < r_end;
    const int64_t n = ok ? r / d.rows : 0;
    const int m = ok ? (int)(r - n * d.rows) : 0;
    const float* xr = x + n * P.in_stride + d.in_off + (int64_t)m * d.in_ld + kt * TK;
    const float* yr = dy + n * P.out_stride + d.out_off + (int64_t)m * d.nout + nt * 32;
#pragma unroll
    for (int h = 0; h < TK / 16; ++h) {
      const int c4 = (lq + 4 * h) * 4;
      xv[h] = (ok && kt * TK + c4 < d.k) ? __ldg(reinterpret_cast<const float4*>(xr + c4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int c4 = (lq + 4 * h) * 4;
      yv[h] = (ok && nt * 32 + c4 < d.nout) ? __ldg(reinterpret_cast<const float4*>(yr + c4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  if (r_begin < r_end) fetch(r_begin);
  for (int64_t r0 = r_begin; r0 < r_end; r0 += 32) {
#pragma unroll
    for (int h = 0; h < TK / 16; ++h) *reinterpret_cast<float4*>(&Xs[lr][(lq + 4 * h) * 4]) = xv[h];
#pragma unroll
    for (int h = 0; h < 2; ++h) *reinterpret_cast<float4*>(&Ys[lr][(lq + 4 * h) * 4]) = yv[h];
    __syncthreads();
    if (r0 + 32 < r_end) fetch(r0 + 32);
#pragma unroll 8
    for (int rr = 0; rr < 32; ++rr) {
      float a[KPT];
      if constexpr (KPT == 4) {
        const float4 a4 = *reinterpret_cast<const float4*>(&Xs[rr][4 * tk]);
        a[0] = a4.x; a[1] = a4.y; a[2] = a4.z; a[3] = a4.w;
      } else {
        const float2 a2 = *reinterpret_cast<const float2*>(&Xs[rr][2 * tk]);
        a[0] = a2.x; a[1] = a2.y;
      }
      const float4 b = *reinterpret_cast<const float4*>(&Ys[rr][4 * tn]);
#pragma unroll
      for (int i = 0; i < KPT; ++i) {
        acc[i][0] = fmaf(a[i], b.x, acc[i][0]); acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
        acc[i][2] = fmaf(a[i], b.z, acc[i][2]); acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
      }
      if (want_bias) { bsum[0] += b.x; bsum[1] += b.y; bsum[2] += b.z; bsum[3] += b.w; }
    }
    __syncthreads();
  }
  float* const wsl = ws ? ws + (int64_t)(cta / tiles) * ws_stride : nullptr;
  float* Wb = (ws ? wsl : dW) + d.w_off;
#pragma unroll
  for (int i = 0; i < KPT; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = kt * TK + KPT * tk + i, n = nt * 32 + 4 * tn + j;
      if (k < d.k && n < d.nout) {
        float* dst = Wb + (P.transB ? (int64_t)n * d.k + k : (int64_t)k * d.nout + n);
        if (ws) *dst = acc[i][j] * d.scale;
        else atomicAdd(dst, acc[i][j] * d.scale);
      }
    }
  if (want_bias)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (nt * 32 + 4 * tn + j < d.nout) {
        if (ws) wsl[n_weights + nt * 32 + 4 * tn + j] = bsum[j];
        else atomicAdd(dbias + nt * 32 + 4 * tn + j, bsum[j]);
      }
}

// dW[i] += sum over the row slices of i's block in a fixed order (bitwise repeatable).  Block = 32 elements x 32 partial lanes:
// lane y adds the slices y, y + 32, ..., then the 32 lane sums are added in lane order.  (One thread per element chained up to
// ~200 dependent loads: 15-25 us of the 60 us the K = 32 weight gradients took.)
__global__ void __launch_bounds__(1024) k_wgrad_reduce(const float* __restrict__ ws, int64_t ws_stride, int64_t n_weights, int bias_cols,
                                                       float* __restrict__ dW, float* __restrict__ dbias, int64_t N, WgParams P) {
  __shared__ float part[32][33];
  const int64_t i = blockIdx.x * 32LL + threadIdx.x;
  const int y = threadIdx.y;
  int b = -1;
  if (i < n_weights + bias_cols) {
    if (i >= n_weights) b = 0;
    else
      for (int q = 0; q < P.n_blocks; ++q)
        if (P.b[q].k > 0 && P.b[q].nout > 0 && i >= P.b[q].w_off && i < P.b[q].w_off + (int64_t)P.b[q].k * P.b[q].nout) b = q;
  }
  float s = 0.f;
  if (b >= 0) {
    const int64_t slices = (N * P.b[b].rows + P.wg_rows - 1) / P.wg_rows;
    for (int64_t sl = y; sl < slices; sl += 32) s += ws[sl * ws_stride + i];
  }
  part[y][threadIdx.x] = s;
  __syncthreads();
  if (y == 0 && b >= 0) {
    float r = 0.f;
#pragma unroll
    for (int q = 0; q < 32; ++q) r += part[q][threadIdx.x];
    if (i >= n_weights) dbias[i - n_weights] += r;
    else dW[i] += r;
  }
}

// ------------------------------------------------------------------------------------------------ column sums
// block = 32 columns x 8 row lanes; a CTA sums rows_per_cta rows (fixed order inside the CTA), one atomicAdd per column
__global__ void __launch_bounds__(256) k_colsum(const float* __restrict__ x, int64_t R, int ld, int cols, const float* __restrict__ a,
                                                float* __restrict__ out, int rows_per_cta, float* __restrict__ ws) {
  __shared__ float part[8][33];
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = blockIdx.y * 32 + cl;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta, r1 = r0 + rows_per_cta < R ? r0 + rows_per_cta : R;
  float s = 0.f;
  if (c < cols)
    for (int64_t r = r0 + rl; r < r1; r += 8) s = fmaf(a ? __ldg(a + r) : 1.0f, __ldg(x + r * ld + c), s);
  part[rl][cl] = s;
  __syncthreads();
  if (rl == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += part[q][cl];
    if (ws) ws[(int64_t)blockIdx.x * cols + c] = t;   // deterministic mode: k_colsum_reduce adds the CTA partials in order
    else atomicAdd(out + c, t);
  }
}
__global__ void k_colsum_reduce(const float* __restrict__ ws, int64_t parts, int cols, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (int64_t p = 0; p < parts; ++p) s += ws[p * cols + c];
  out[c] += s;
}

__global__ void k_silu_fwd(const float* __restrict__ a, float* __restrict__ s, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) s[i] = eeg_silu(a[i]);
}
__device__ __forceinline__ float silu_grad(float x) {
  const float sg = 1.0f / (1.0f + __expf(-x));
  return sg * (1.0f + x * (1.0f - sg));
}
__global__ void k_silu_bwd(const float* __restrict__ a, const float* __restrict__ ds, float* __restrict__ da, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) da[i] = ds[i] * silu_grad(a[i]);
}

// ------------------------------------------------------------------------------------------------ gate backward
struct GateBwdParams {
  EegLayout lz, lg;
  int gate_off[EEG_MAX_L];
  int n_scalars;
  float c;
};
__global__ void k_gate_bwd(const float* __restrict__ z, const float* __restrict__ dg, float* __restrict__ dz, GateBwdParams P) {
  const int64_t n = blockIdx.x;
  const float* zr = z + n * P.lz.stride;
  const float* gr = dg + n * P.lg.stride;
  float* dr = dz + n * P.lz.stride;
  for (int u = threadIdx.x; u < P.n_scalars; u += blockDim.x) dr[u] = gr[u] * P.c * silu_grad(zr[u]);
  // one warp per gated block (lane = channel): the blocks of a node run side by side instead of one after the other with a
  // quarter of the CTA active
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int b = 1 + warp; b < P.lz.n_blocks; b += nwarps) {
    const int mul = P.lz.mul[b], rows = P.lz.rows[b];
    for (int u = lane; u < mul; u += 32) {
      const float zg = zr[P.gate_off[b] + u];
      const float gv = P.c * eeg_silu(zg);
      float s = 0.f;
      for (int m = 0; m < rows; ++m) {
        const float d = gr[P.lg.off[b] + m * mul + u];
        s = fmaf(d, zr[P.lz.off[b] + m * mul + u], s);
        dr[P.lz.off[b] + m * mul + u] = d * gv;
      }
      dr[P.gate_off[b] + u] = s * P.c * silu_grad(zg);
    }
  }
}

// ------------------------------------------------------------------------------------------------ tail backward
struct TailBwdParams {
  EegLayout lo;
  int qmul[EEG_MAX_L], qoff[EEG_MAX_L], woff[EEG_MAX_L], poff[EEG_MAX_L];
  float qscale[EEG_MAX_L];
  int nq, npool, positive, reduce;
};
__device__ __forceinline__ void mat6_sq_bwd(const float* A, const float* dC, float* dA, int t) {  // C = A A: dA = dC A^T + A^T dC
  if (t < 36) {
    const int r = t / 6, c = t % 6;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) s += dC[r * 6 + k] * A[c * 6 + k] + A[k * 6 + r] * dC[k * 6 + c];
    dA[t] = s;
  }
}
__device__ __forceinline__ void mat6_mul_b(const float* A, const float* B, float* Cm, int t) {
  if (t < 36) {
    const int r = t / 6, c = t % 6;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) s = fmaf(A[r * 6 + k], B[k * 6 + c], s);
    Cm[t] = s;
  }
}
__global__ void __launch_bounds__(128) k_pool_tail_bwd(const float* __restrict__ o, const int32_t* __restrict__ gp,
                                                       const float* __restrict__ Wq, const float* __restrict__ Pm,
                                                       const float* __restrict__ A_in, const float* __restrict__ dC,
                                                       float* __restrict__ d_o, float* __restrict__ dq_out,
                                                       float* __restrict__ pooled_out, TailBwdParams P) {
  extern __shared__ float sm[];
  float* dpool = sm;            // [npool]
  float* dq = dpool + P.npool;  // [nq]
  float* A = dq + P.nq;         // 5 x [36]
  float* B = A + 36;
  float* dCs = B + 36;
  float* dB = dCs + 36;
  float* dA = dB + 36;
  const int g = blockIdx.x, t = threadIdx.x;
  const int nb = gp[g], ne = gp[g + 1];
  const float inv = (P.reduce == 0) ? 1.0f / (float)max(ne - nb, 1) : 1.0f;
  if (t < 36) {
    A[t] = A_in[(int64_t)g * 36 + t];
    dCs[t] = dC[(int64_t)g * 36 + t];
  }
  __syncthreads();
  if (P.positive == 0) {
    if (t < 36) dA[t] = dCs[t];
  } else if (P.positive == 5) {
    // adjoint of exp at symmetric A = V diag(ev) V^T:  dA = V [(V^T dC V) o Phi] V^T,  Phi_ij = (e^ev_i - e^ev_j) / (ev_i - ev_j)
    float* ev = dA + 36;
    if (t == 0) eeg_eigh6(A, ev, B);   // B = V
    __syncthreads();
    float s = 0.f;
    if (t < 36) {   // dB = V^T dC
#pragma unroll
      for (int k = 0; k < 6; ++k) s = fmaf(B[k * 6 + t / 6], dCs[k * 6 + t % 6], s);
      dB[t] = s;
    }
    __syncthreads();
    if (t < 36) {   // A <- (dB V) o Phi
      s = 0.f;
#pragma unroll
      for (int k = 0; k < 6; ++k) s = fmaf(dB[(t / 6) * 6 + k], B[k * 6 + t % 6], s);
      A[t] = s * eeg_exp_divdiff(ev[t / 6], ev[t % 6]);
    }
    __syncthreads();
    if (t < 36) {   // dB = V A
      s = 0.f;
#pragma unroll
      for (int k = 0; k < 6; ++k) s = fmaf(B[(t / 6) * 6 + k], A[k * 6 + t % 6], s);
      dB[t] = s;
    }
    __syncthreads();
    if (t < 36) {   // dA = dB V^T
      s = 0.f;
#pragma unroll
      for (int k = 0; k < 6; ++k) s = fmaf(dB[(t / 6) * 6 + k], B[(t % 6) * 6 + k], s);
      dA[t] = s;
    }
  } else if (P.positive == 1) {
    mat6_sq_bwd(A, dCs, dA, t);
  } else {
    float post = 1.f;
    if (P.positive == 3 || P.positive == 4) {
      const float kk = (P.positive == 3) ? 2.f : 4.f;
      if (t < 36) A[t] = A[t] / kk + ((t / 6 == t % 6) ? 1.f : 0.f);
      post = 1.f / kk;
      __syncthreads();
    }
    if (P.positive == 3) {
      mat6_sq_bwd(A, dCs, dA, t);
    } else {  // fourth power: B = A^2, C = B^2
      mat6_mul_b(A, A, B, t);
      __syncthreads();
      mat6_sq_bwd(B, dCs, dB, t);
      __syncthreads();
      mat6_sq_bwd(A, dB, dA, t);
    }
    __syncthreads();
    if (t < 36) dA[t] *= post;
  }
  __syncthreads();
  for (int k = t; k < P.nq; k += blockDim.x) {
    float s = 0.f;
    for (int e = 0; e < 36; ++e) s = fmaf(__ldg(Pm + k * 36 + e), dA[e], s);
    dq[k] = s;
    dq_out[(int64_t)g * P.nq + k] = s;
  }
  __syncthreads();
  // d pooled_b[m][u] = scale_b sum_w dq[b][w][m] W_b[u][w];  pooled itself is re-reduced for the weight gradient
  for (int i = t; i < P.npool; i += blockDim.x) {
    int b = -1;
    for (int k = 0; k < P.lo.n_blocks; ++k)
      if (P.qmul[k] > 0 && i >= P.poff[k]) b = k;
    const int rows = P.lo.rows[b], mul = P.lo.mul[b], r = i - P.poff[b];
    const int m = r / mul, u = r % mul;
    float s = 0.f;
    for (int w = 0; w < P.qmul[b]; ++w) s = fmaf(dq[P.qoff[b] + w * rows + m], __ldg(Wq + P.woff[b] + u * P.qmul[b] + w), s);
    dpool[i] = s * P.qscale[b] * inv;
    const float* src = o + (int64_t)nb * P.lo.stride + P.lo.off[b] + r;
    float ps = 0.f;
    for (int n = nb; n < ne; ++n, src += P.lo.stride) ps += __ldg(src);
    pooled_out[(int64_t)g * P.npool + i] = ps * inv;
  }
  __syncthreads();
  // d o[n] = d pooled / count for the pooled blocks, zero for the blocks the final linear ignores
  for (int i = t; i < P.lo.stride; i += blockDim.x) {
    int b = 0;
    for (int k = 1; k < P.lo.n_blocks; ++k)
      if (i >= P.lo.off[k]) b = k;
    const float v = P.qmul[b] > 0 ? dpool[P.poff[b] + i - P.lo.off[b]] : 0.f;
    for (int n = nb; n < ne; ++n) d_o[(int64_t)n * P.lo.stride + i] = v;
  }
}

// dWq_b[u][w] = scale_b sum_g sum_m pooled[g][b][m][u] dq[g][b][w][m];  dbias[w] = sum_g dq[g][0][w][0]
__global__ void __launch_bounds__(128) k_tail_wgrad(const float* __restrict__ pooled, const float* __restrict__ dq,
                                                     float* __restrict__ dWq, float* __restrict__ dbias, int64_t G, TailBwdParams P,
                                                     int n_w) {
  // one CTA per output element; threads stride over the graphs, fixed-order tree reduction (deterministic)
  const int i = blockIdx.x;
  float s = 0.f;
  float scale = 1.f;
  if (i < n_w) {
    int b = -1;
    for (int k = 0; k < P.lo.n_blocks; ++k)
      if (P.qmul[k] > 0 && i >= P.woff[k]) b = k;
    const int rows = P.lo.rows[b], mul = P.lo.mul[b], r = i - P.woff[b];
    const int u = r / P.qmul[b], w = r % P.qmul[b];
    const float* pp = pooled + P.poff[b] + u;
    const float* qq = dq + P.qoff[b] + w * rows;
    for (int64_t g = threadIdx.x; g < G; g += blockDim.x)
      for (int m = 0; m < rows; ++m) s = fmaf(pp[g * P.npool + m * mul], qq[g * P.nq + m], s);
    scale = P.qscale[b];
  } else {
    const int w = i - n_w;
    for (int64_t g = threadIdx.x; g < G; g += blockDim.x) s += dq[g * P.nq + P.qoff[0] + w * P.lo.rows[0]];
  }
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ float ws[4];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    const float t = (ws[0] + ws[1]) + (ws[2] + ws[3]);
    if (i < n_w) dWq[i] = t * scale;
    else dbias[i - n_w] = t;
  }
}

void fill_tail(TailBwdParams& T, const EegLayout* lo, const int32_t* host_qmul, const float* host_qscale, int positive, int reduce,
               int* n_w) {
  T.lo = *lo;
  int nq = 0, npool = 0, woff = 0;
  for (int b = 0; b < EEG_MAX_L; ++b) { T.qmul[b] = 0; T.qoff[b] = 0; T.woff[b] = 0; T.poff[b] = 0; T.qscale[b] = 0.f; }
  for (int b = 0; b < lo->n_blocks; ++b) {
    T.qmul[b] = host_qmul[b];
    T.qscale[b] = host_qscale[b];
    T.qoff[b] = nq;
    T.woff[b] = woff;
    T.poff[b] = npool;
    if (host_qmul[b] > 0) {
      nq += host_qmul[b] * lo->rows[b];
      woff += host_qmul[b] * lo->mul[b];
      npool += lo->rows[b] * lo->mul[b];
    }
  }
  T.nq = nq; T.npool = npool; T.positive = positive; T.reduce = reduce;
  *n_w = woff;
}

}  // namespace

// tensor-core path for the wide weight matrices (eeg_wgrad_umma.cu)
int64_t eeg_wgrad_umma_ws_floats(const EegLayout* lin, const EegLayout* lout, int64_t N);
int eeg_wgrad_umma_launch(const float* x, const EegLayout* lin, const float* dy, const EegLayout* lout, const int64_t* host_w_off,
                          const float* host_scale, int transB, float* dW, int64_t N, float* ws, int64_t ws_floats, cudaStream_t stream);
extern "C" int eeg_colsum(const float* x, int64_t R, int ld, int cols, const float* row_scale, float* out, float* ws, int64_t ws_floats,
                          void* stream);
static bool wg_use_tc() {
  static const char* env = getenv("EEG_WGRAD");   // "cuda": keep every weight gradient on the CUDA-core kernel (A/B timing)
  return !(env && !strcmp(env, "cuda"));
}
// warp-MMA path for 32 output channels (eeg_wgrad_mma.cu)
int64_t eeg_wgrad_mma_ws_floats(const EegLayout* lin, const EegLayout* lout, int64_t N);
int eeg_wgrad_mma_launch(const float* x, const EegLayout* lin, const float* dy, const EegLayout* lout, const int64_t* host_w_off,
                         const float* host_scale, int transB, float* dW, int64_t N, float* ws, int64_t ws_floats, cudaStream_t stream);
// workspace of a tensor-core path (*which = 1: tcgen05, 2: warp MMA) incl. the bias partials behind it; -1 when the call cannot
// take either
static int64_t wg_umma_need(const EegLayout* lin, const EegLayout* lout, int bias_cols, int64_t N, int64_t* bias_need, int* which) {
  *bias_need = 0;
  *which = 0;
  if (!wg_use_tc()) return -1;
  int64_t un = eeg_wgrad_umma_ws_floats(lin, lout, N);
  *which = 1;
  if (un < 0) {
    un = eeg_wgrad_mma_ws_floats(lin, lout, N);
    *which = 2;
  }
  if (un < 0) return -1;
  if (bias_cols > 0) {
    if (lin->rows[0] != 1 || lout->mul[0] != bias_cols) return -1;
    *bias_need = (N + 255) / 256 * bias_cols;
  }
  return un;
}

extern "C" int eeg_block_linear_wgrad(const float* x, const EegLayout* lin, const float* dy, const EegLayout* lout,
                                      const int64_t* host_w_off, const float* host_scale, int transB, const int32_t* in_ld,
                                      const int32_t* k_override, float* dW, float* dbias, int64_t N, int64_t n_weights, float* ws,
                                      int64_t ws_floats, void* stream) {
  EEG_REQUIRE(lin && lout && lin->n_blocks == lout->n_blocks && lin->n_blocks <= EEG_MAX_L, "eeg_block_linear_wgrad: layout mismatch");
  if (N == 0) return 0;
  if (ws != nullptr && in_ld == nullptr && k_override == nullptr) {
    int64_t bias_need = 0;
    int which = 0;
    const int64_t un = wg_umma_need(lin, lout, dbias ? lout->mul[0] : 0, N, &bias_need, &which);
    if (un >= 0 && ws_floats >= un + bias_need) {
      const int rc = which == 1 ? eeg_wgrad_umma_launch(x, lin, dy, lout, host_w_off, host_scale, transB, dW, N, ws, un, (cudaStream_t)stream)
                                : eeg_wgrad_mma_launch(x, lin, dy, lout, host_w_off, host_scale, transB, dW, N, ws, un, (cudaStream_t)stream);
      if (rc > 0) return rc;
      if (rc == 0) return dbias ? eeg_colsum(dy + lout->off[0], N, lout->stride, lout->mul[0], nullptr, dbias, ws + un, bias_need, stream) : 0;
    }
  }
  WgParams P;
  P.n_blocks = lin->n_blocks;
  P.in_stride = lin->stride;
  P.out_stride = lout->stride;
  P.transB = transB;
  P.wg_rows = wg_rows_per_cta(lin, lout, k_override, N);
  const int TK = wg_tile_k(lin, lout, k_override);
  int64_t ctas = 0;
  for (int b = 0; b < P.n_blocks; ++b) {
    EEG_REQUIRE(lin->rows[b] == lout->rows[b], "eeg_block_linear_wgrad: block %d rows differ", b);
    WgBlock& d = P.b[b];
    d.rows = lin->rows[b];
    d.k = k_override ? k_override[b] : lin->mul[b];
    d.in_ld = in_ld ? in_ld[b] : lin->mul[b];
    d.nout = lout->mul[b];
    d.in_off = lin->off[b];
    d.out_off = lout->off[b];
    d.w_off = host_w_off[b];
    d.scale = host_scale[b];
    EEG_REQUIRE(d.k % 4 == 0 && d.nout % 4 == 0 && d.in_off % 4 == 0 && d.out_off % 4 == 0 && d.in_ld % 4 == 0 &&
                    lin->stride % 4 == 0 && lout->stride % 4 == 0,
                "eeg_block_linear_wgrad: block %d not float4-aligned", b);
    d.cta_begin = (int)ctas;
    d.ktiles = (d.k + TK - 1) / TK;
    d.ntiles = (d.nout + 31) / 32;
    const int64_t chunks = (N * d.rows + P.wg_rows - 1) / P.wg_rows;
    ctas += (d.k > 0 && d.nout > 0) ? chunks * d.ktiles * d.ntiles : 0;
    if (d.k == 0 || d.nout == 0) { d.ktiles = 1; d.ntiles = 1; }
    EEG_REQUIRE(ctas < (1LL << 31), "eeg_block_linear_wgrad: too many tiles");
  }
  if (ctas == 0) return 0;
  const int bias_cols = dbias ? P.b[0].nout : 0;
  const int64_t stride = n_weights + bias_cols;
  if (ws != nullptr) {
    int64_t slices = 0;
    for (int b = 0; b < P.n_blocks; ++b) {
      const int64_t sl = (N * P.b[b].rows + P.wg_rows - 1) / P.wg_rows;
      if (P.b[b].k > 0 && P.b[b].nout > 0 && sl > slices) slices = sl;
      EEG_REQUIRE(P.b[b].k == 0 || P.b[b].nout == 0 || P.b[b].w_off + (int64_t)P.b[b].k * P.b[b].nout <= n_weights,
                  "eeg_block_linear_wgrad: block %d exceeds n_weights", b);
    }
    EEG_REQUIRE(ws_floats >= slices * stride, "eeg_block_linear_wgrad: workspace too small (%lld < %lld floats)", (long long)ws_floats,
                (long long)(slices * stride));
  }
  if (TK == 64) k_linear_wgrad<4><<<(unsigned)ctas, 128, 0, (cudaStream_t)stream>>>(x, dy, dW, dbias, N, P, ws, stride, n_weights);
  else k_linear_wgrad<2><<<(unsigned)ctas, 128, 0, (cudaStream_t)stream>>>(x, dy, dW, dbias, N, P, ws, stride, n_weights);
  EEG_CHECK_LAUNCH("eeg_block_linear_wgrad");
  if (ws != nullptr) {
    k_wgrad_reduce<<<(unsigned)((stride + 31) / 32), dim3(32, 32), 0, (cudaStream_t)stream>>>(ws, stride, n_weights, bias_cols, dW, dbias, N, P);
    EEG_CHECK_LAUNCH("eeg_block_linear_wgrad (reduce)");
  }
  return 0;
}

// floats of workspace eeg_block_linear_wgrad needs in deterministic mode
extern "C" int64_t eeg_block_linear_wgrad_ws_floats(const EegLayout* lin, const EegLayout* lout, int64_t n_weights, int bias_cols,
                                                    int64_t N) {
  int64_t slices = 0;
  const int wg_rows = lin ? wg_rows_per_cta(lin, lout, nullptr, N) : WG_ROWS_MAX;
  for (int b = 0; lin && b < lin->n_blocks; ++b) {
    const int64_t sl = (N * lin->rows[b] + wg_rows - 1) / wg_rows;
    if (sl > slices) slices = sl;
  }
  int64_t need = slices * (n_weights + bias_cols), bias_need = 0;
  int which = 0;
  const int64_t un = lin && lout ? wg_umma_need(lin, lout, bias_cols, N, &bias_need, &which) : -1;
  if (un >= 0 && un + bias_need > need) need = un + bias_need;
  return need;
}

extern "C" int eeg_colsum(const float* x, int64_t R, int ld, int cols, const float* row_scale, float* out, float* ws, int64_t ws_floats,
                          void* stream) {
  if (R == 0 || cols == 0) return 0;
  const int rows_per_cta = 256;
  dim3 grid((unsigned)((R + rows_per_cta - 1) / rows_per_cta), (unsigned)((cols + 31) / 32));
  EEG_REQUIRE(ws == nullptr || ws_floats >= (int64_t)grid.x * cols, "eeg_colsum: workspace too small (need ceil(R/256) * cols floats)");
  k_colsum<<<grid, 256, 0, (cudaStream_t)stream>>>(x, R, ld, cols, row_scale, out, rows_per_cta, ws);
  EEG_CHECK_LAUNCH("eeg_colsum");
  if (ws != nullptr) {
    k_colsum_reduce<<<(unsigned)((cols + 127) / 128), 128, 0, (cudaStream_t)stream>>>(ws, grid.x, cols, out);
    EEG_CHECK_LAUNCH("eeg_colsum (reduce)");
  }
  return 0;
}

// ---- eeg_colsum_static: the deterministic column sum WITHOUT a caller workspace and without any allocation ----------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The CTA partials live in a file-scope
// device array of EEG_COLSUM_STATIC_MAX_FLOATS floats (4 MiB), held by every GPU that loads the library until exit.
// In HOST code `eeg_colsum_static_scratch` is only a stand-in for the symbol, not a device address: the launcher resolves it
// once per device with cudaGetSymbolAddress (that device current) and passes the resolved pointer to the kernels.
// Nothing in the scratch has to start at zero: every partial that k_colsum_reduce reads is written by k_colsum of the SAME
// call first (all (row slice, column < cols) entries), so the contents left by an earlier call are never observed.
// All calls on a device share the array: two calls that overlap on the GPU (different streams) race.
#define EEG_COLSUM_STATIC_MAX_FLOATS (1 << 20)
static __device__ __align__(256) float eeg_colsum_static_scratch[EEG_COLSUM_STATIC_MAX_FLOATS];

extern "C" int64_t eeg_colsum_static_max_floats(void) { return EEG_COLSUM_STATIC_MAX_FLOATS; }

extern "C" int eeg_colsum_static(const float* x, int64_t R, int ld, int cols, const float* row_scale, float* out, void* stream) {
  EEG_REQUIRE(R >= 0 && cols >= 0 && ld >= cols, "eeg_colsum_static: bad shape (R=%lld, ld=%d, cols=%d)", (long long)R, ld, cols);
  if (R == 0 || cols == 0) return 0;
  const int rows_per_cta = 256;
  const int64_t parts = (R + rows_per_cta - 1) / rows_per_cta;
  EEG_REQUIRE(parts * cols <= (int64_t)EEG_COLSUM_STATIC_MAX_FLOATS,
              "eeg_colsum_static: ceil(R/256) * cols = %lld floats of scratch exceed the built-in maximum of %d; use eeg_colsum with a "
              "caller workspace", (long long)(parts * cols), EEG_COLSUM_STATIC_MAX_FLOATS);
  static float* scratch_of_device[64] = {};   // resolved device address of the array, per device ordinal
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    eeg_set_error("eeg_colsum_static: cudaGetDevice failed or device ordinal %d out of range: %s", dev, cudaGetErrorString(e));
    return 2;
  }
  if (scratch_of_device[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, eeg_colsum_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      eeg_set_error("eeg_colsum_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return 2;
    }
    scratch_of_device[dev] = p;
  }
  float* ws = scratch_of_device[dev];
  dim3 grid((unsigned)parts, (unsigned)((cols + 31) / 32));
  k_colsum<<<grid, 256, 0, (cudaStream_t)stream>>>(x, R, ld, cols, row_scale, out, rows_per_cta, ws);
  EEG_CHECK_LAUNCH("eeg_colsum_static");
  k_colsum_reduce<<<(unsigned)((cols + 127) / 128), 128, 0, (cudaStream_t)stream>>>(ws, parts, cols, out);
  EEG_CHECK_LAUNCH("eeg_colsum_static (reduce)");
  return 0;
}

extern "C" int eeg_silu_fwd(const float* a, float* s, int64_t n, void* stream) {
  if (n == 0) return 0;
  k_silu_fwd<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(a, s, n);
  EEG_CHECK_LAUNCH("eeg_silu_fwd");
  return 0;
}
extern "C" int eeg_silu_bwd(const float* a, const float* ds, float* da, int64_t n, void* stream) {
  if (n == 0) return 0;
  k_silu_bwd<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(a, ds, da, n);
  EEG_CHECK_LAUNCH("eeg_silu_bwd");
  return 0;
}

extern "C" int eeg_gate_bwd(const float* z, const EegLayout* lz, const float* dg, const EegLayout* lg, int n_scalars, float act_const,
                            float* dz, int64_t N, void* stream) {
  EEG_REQUIRE(lz && lg && lz->n_blocks == lg->n_blocks, "eeg_gate_bwd: layout mismatch");
  if (N == 0) return 0;
  GateBwdParams P;
  P.lz = *lz;
  P.lg = *lg;
  P.n_scalars = n_scalars;
  P.c = act_const;
  int go = n_scalars;
  for (int b = 0; b < EEG_MAX_L; ++b) P.gate_off[b] = 0;
  for (int b = 1; b < lz->n_blocks; ++b) {
    P.gate_off[b] = lz->off[0] + go;
    go += lz->mul[b];
  }
  k_gate_bwd<<<(unsigned)N, 128, 0, (cudaStream_t)stream>>>(z, dg, dz, P);
  EEG_CHECK_LAUNCH("eeg_gate_bwd");
  return 0;
}

extern "C" int eeg_pool_tail_bwd(const float* o, const EegLayout* lo, const int32_t* graph_ptr, const float* Wq,
                                 const int32_t* host_qmul, const float* host_qscale, const float* P, int positive, int reduce,
                                 const float* A, const float* dC, float* d_o, float* dq_ws /*[G,nq]*/, float* pooled_ws /*[G,npool]*/,
                                 float* dWq, float* dbias, int64_t G, void* stream) {
  EEG_REQUIRE(lo && lo->n_blocks <= EEG_MAX_L && positive >= 0 && positive <= 5, "eeg_pool_tail_bwd: bad arguments");
  if (G == 0) return 0;
  TailBwdParams T;
  int n_w = 0;
  fill_tail(T, lo, host_qmul, host_qscale, positive, reduce, &n_w);
  const size_t smem = (size_t)(T.npool + T.nq + 36 * 5 + 8) * sizeof(float);
  EEG_REQUIRE(smem <= 48 * 1024, "eeg_pool_tail_bwd: pooled row too large");
  k_pool_tail_bwd<<<(unsigned)G, 128, smem, (cudaStream_t)stream>>>(o, graph_ptr, Wq, P, A, dC, d_o, dq_ws, pooled_ws, T);
  EEG_CHECK_LAUNCH("eeg_pool_tail_bwd");
  const int tot = n_w + (dbias ? T.qmul[0] : 0);
  k_tail_wgrad<<<tot, 128, 0, (cudaStream_t)stream>>>(pooled_ws, dq_ws, dWq, dbias, G, T, n_w);
  EEG_CHECK_LAUNCH("eeg_tail_wgrad");
  return 0;
}

// ------------------------------------------------------------------------------------------------ optimizer
// torch.optim.AdamW(amsgrad) + Lightning gradient_clip_val (scripts/train_utils.py:38-44, scripts/train_main.py:89-92) on
// one flat fp32 buffer: g <- grad_scale * g (e.g. 1/world after an all-reduce sum), clipped by max_norm / (|g| + 1e-6).
namespace {
// ONE CTA, fixed summation order (the flat gradient is < 1 MB: a multi-CTA atomicAdd made the clip factor, and with it
// every parameter, depend on the scheduling order)
__global__ void __launch_bounds__(1024) k_sumsq(const float* __restrict__ x, int64_t n, float* __restrict__ out) {
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s = fmaf(x[i], x[i], s);
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ float ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += ws[w];
    out[0] += t;
  }
}
__global__ void k_adamw(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                        float* __restrict__ vmax, int64_t n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2,
                        const float* __restrict__ gnorm_sq, float max_norm, float grad_scale) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float clip = grad_scale;
  if (gnorm_sq != nullptr && max_norm > 0.f) {
    const float nrm = sqrtf(gnorm_sq[0]) * grad_scale;
    clip *= fminf(max_norm / (nrm + 1e-6f), 1.0f);
  }
  const float gi = g[i] * clip;
  float pi = p[i] * (1.0f - lr * wd);                    // decoupled weight decay
  const float mi = b1 * m[i] + (1.0f - b1) * gi;
  const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  float vh = vi;
  if (vmax != nullptr) {
    vh = fmaxf(vmax[i], vi);
    vmax[i] = vh;
  }
  p[i] = pi - (lr / bc1) * mi / (sqrtf(vh) / sqrtf(bc2) + eps);
}
}  // namespace

extern "C" int eeg_sumsq(const float* x, int64_t n, float* out, void* stream) {
  if (n == 0) return 0;
  k_sumsq<<<1, 1024, 0, (cudaStream_t)stream>>>(x, n, out);
  EEG_CHECK_LAUNCH("eeg_sumsq");
  return 0;
}

extern "C" int eeg_adamw_step(float* p, const float* g, float* m, float* v, float* vmax, int64_t n, float lr, float beta1, float beta2,
                              float eps, float weight_decay, int step, const float* gnorm_sq, float max_norm, float grad_scale,
                              void* stream) {
  EEG_REQUIRE(step >= 1, "eeg_adamw_step: step counts from 1");
  if (n == 0) return 0;
  const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
  k_adamw<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, vmax, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                                                        gnorm_sq, max_norm, grad_scale);
  EEG_CHECK_LAUNCH("eeg_adamw_step");
  return 0;
}
