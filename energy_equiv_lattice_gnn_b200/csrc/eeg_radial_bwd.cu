// Backward of the two hidden layers of the radial MLP (conv_tp_weights, gnn/blocks.py:538-549) in ONE pass over the edges:
//   a1 = W0 f + b0, s1 = silu(a1), a2 = W1 s1 + b1, s2 = silu(a2)            (recomputed, never stored: s2 optionally written)
//   da2 = ds2 * silu'(a2), dW1 += da2^T s1, db1 += sum da2, ds1 = da2 W1, da1 = ds1 * silu'(a1), dW0 += da1^T f, db0 += sum da1
// for hidden width 64 and F <= 32 edge features.  It replaces thirteen launches per layer of the unfused chain (two recompute
// GEMMs + two SiLU, two SiLU backward, two weight gradients, two column sums, one input-gradient GEMM: 0.25 ms of kernels that
// each stream [E x 64] tensors) by one kernel that reads f and ds2 once.
//
// CTA = 256 threads, tile = 64 edges, 2 CTAs per SM.  Activations of a tile live in shared memory feature-major ([feature][edge],
// row pitch 68 floats).  Thread (ty, tx) of the 16 x 16 grid owns edges 4 ty .. 4 ty + 3 and the INTERLEAVED features tx + 16 c:
// with that ownership the transposed float4 stores, the float4 loads along the edge axis of the reductions over edges and the
// weight loads (weights are kept column-permuted so that a thread's four columns are contiguous) are all bank-conflict free.
// Weight-gradient partial sums stay in registers over the CTA's tiles, go to a workspace at the end and are added over the CTAs
// in fixed order (bitwise repeatable, no atomics).
#include "eeg_common.cuh"

namespace {

constexpr int RB_H = 64, RB_TE = 64, RB_LD = 68, RB_FMAX = 32, RB_THREADS = 256;

__device__ __forceinline__ float rb_sigmoid(float x) { return 1.0f / (1.0f + __expf(-x)); }

struct RbSmem {
  float W0P[RB_FMAX][RB_H];    // W0P[i][4 tx + c] = W0[tx + 16 c][i]
  float W1P[RB_H][RB_H];       // W1P[k][4 tx + c] = W1[tx + 16 c][k]     (a2 = s1 W1^T)
  float W1N[RB_H][RB_H];       // W1N[j][4 tx + c] = W1[j][tx + 16 c]     (ds1 = da2 W1)
  float B0[RB_H], B1[RB_H];
  float FT[RB_FMAX][RB_LD];    // f tile, feature-major
  float S1T[RB_H][RB_LD];      // s1 tile, later the da1 tile
  float DA2T[RB_H][RB_LD];
};

__global__ void __launch_bounds__(RB_THREADS, 2)
k_radial_hidden_bwd(const float* __restrict__ f, int F, int f_stride, const float* __restrict__ W0, const float* __restrict__ b0,
                    const float* __restrict__ W1, const float* __restrict__ b1, const float* __restrict__ ds2, int64_t E,
                    float* __restrict__ s2_out, float* __restrict__ ws) {
  extern __shared__ __align__(16) uint8_t rb_raw[];
  RbSmem& S = *reinterpret_cast<RbSmem*>(rb_raw);
  const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
  for (int idx = t; idx < F * RB_H; idx += RB_THREADS) {
    const int i = idx / RB_H, s = idx % RB_H, j = (s >> 2) + 16 * (s & 3);
    S.W0P[i][s] = __ldg(W0 + j * F + i);
  }
  for (int idx = t; idx < RB_H * RB_H; idx += RB_THREADS) {
    const int k = idx / RB_H, s = idx % RB_H, j = (s >> 2) + 16 * (s & 3);
    S.W1P[k][s] = __ldg(W1 + j * RB_H + k);
    S.W1N[k][s] = __ldg(W1 + k * RB_H + j);   // row k of W1, permuted columns
  }
  if (t < RB_H) { S.B0[t] = __ldg(b0 + t); S.B1[t] = __ldg(b1 + t); }

  float accW1[4][4], accW0[RB_FMAX / 4], accb1 = 0.f, accb0 = 0.f;
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int d = 0; d < 4; ++d) accW1[c][d] = 0.f;
#pragma unroll
  for (int q = 0; q < RB_FMAX / 4; ++q) accW0[q] = 0.f;
  const int fq = F >> 2;                       // features per P5 thread
  const int pj = t & 63, pig = t >> 6;         // P5 ownership: dW0[pj][pig * fq + q]

  const int64_t n_tiles = (E + RB_TE - 1) / RB_TE;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t e0 = tile * RB_TE;
    __syncthreads();   // previous tile's readers are done (also orders the weight staging before the first tile)
    {  // f tile -> FT[i][e]
      const int e = t & 63, part = t >> 6;
      const bool ok = e0 + e < E;
      const float* fr = f + (e0 + e) * (int64_t)f_stride;
      for (int q = 0; q < fq; ++q) {
        const int i = part * fq + q;
        S.FT[i][e] = ok ? __ldg(fr + i) : 0.f;
      }
    }
    __syncthreads();
    // ---- P1: a1 = f W0^T + b0 for (edges 4 ty + r, features tx + 16 c)
    float v[4][4], g1[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) v[r][c] = S.B0[tx + 16 * c];
    for (int i = 0; i < F; ++i) {
      const float4 a = *reinterpret_cast<const float4*>(&S.FT[i][4 * ty]);
      const float4 b = *reinterpret_cast<const float4*>(&S.W0P[i][4 * tx]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) v[r][c] = fmaf(av[r], bv[c], v[r][c]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float s1[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float x = v[r][c], sg = rb_sigmoid(x);
        s1[r] = x * sg;
        g1[r][c] = sg * (1.0f + x * (1.0f - sg));
      }
      *reinterpret_cast<float4*>(&S.S1T[tx + 16 * c][4 * ty]) = make_float4(s1[0], s1[1], s1[2], s1[3]);
    }
    __syncthreads();
    // ---- P2: a2 = s1 W1^T + b1, da2 = ds2 * silu'(a2)
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) v[r][c] = S.B1[tx + 16 * c];
#pragma unroll 8
    for (int k = 0; k < RB_H; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&S.S1T[k][4 * ty]);
      const float4 b = *reinterpret_cast<const float4*>(&S.W1P[k][4 * tx]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) v[r][c] = fmaf(av[r], bv[c], v[r][c]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float da[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int64_t e = e0 + 4 * ty + r;
        const float x = v[r][c], sg = rb_sigmoid(x);
        const float d2 = e < E ? __ldg(ds2 + e * RB_H + tx + 16 * c) : 0.f;
        da[r] = d2 * sg * (1.0f + x * (1.0f - sg));
        if (s2_out != nullptr && e < E) s2_out[e * RB_H + tx + 16 * c] = x * sg;
      }
      *reinterpret_cast<float4*>(&S.DA2T[tx + 16 * c][4 * ty]) = make_float4(da[0], da[1], da[2], da[3]);
    }
    __syncthreads();
    // ---- P3: dW1[ty + 16 c][tx + 16 d] += sum_e da2[e][.] s1[e][.];  db1
#pragma unroll 4
    for (int e4 = 0; e4 < RB_TE / 4; ++e4) {
      float4 A[4], B[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) A[c] = *reinterpret_cast<const float4*>(&S.DA2T[ty + 16 * c][4 * e4]);
#pragma unroll
      for (int d = 0; d < 4; ++d) B[d] = *reinterpret_cast<const float4*>(&S.S1T[tx + 16 * d][4 * e4]);
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int d = 0; d < 4; ++d)
          accW1[c][d] = fmaf(A[c].x, B[d].x, fmaf(A[c].y, B[d].y, fmaf(A[c].z, B[d].z, fmaf(A[c].w, B[d].w, accW1[c][d]))));
    }
    if (t < RB_H) {
#pragma unroll 4
      for (int e4 = 0; e4 < RB_TE / 4; ++e4) {
        const float4 a = *reinterpret_cast<const float4*>(&S.DA2T[t][4 * e4]);
        accb1 += (a.x + a.y) + (a.z + a.w);
      }
    }
    __syncthreads();   // S1T is overwritten by da1 below
    // ---- P4: ds1 = da2 W1, da1 = ds1 * silu'(a1)
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) v[r][c] = 0.f;
#pragma unroll 8
    for (int j = 0; j < RB_H; ++j) {
      const float4 a = *reinterpret_cast<const float4*>(&S.DA2T[j][4 * ty]);
      const float4 b = *reinterpret_cast<const float4*>(&S.W1N[j][4 * tx]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) v[r][c] = fmaf(av[r], bv[c], v[r][c]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c)
      *reinterpret_cast<float4*>(&S.S1T[tx + 16 * c][4 * ty]) =
          make_float4(v[0][c] * g1[0][c], v[1][c] * g1[1][c], v[2][c] * g1[2][c], v[3][c] * g1[3][c]);
    __syncthreads();
    // ---- P5: dW0[pj][pig * fq + q] += sum_e da1[e][pj] f[e][i];  db0
#pragma unroll 4
    for (int e4 = 0; e4 < RB_TE / 4; ++e4) {
      const float4 a = *reinterpret_cast<const float4*>(&S.S1T[pj][4 * e4]);
#pragma unroll
      for (int q = 0; q < RB_FMAX / 4; ++q) {
        if (q < fq) {
          const float4 b = *reinterpret_cast<const float4*>(&S.FT[pig * fq + q][4 * e4]);
          accW0[q] = fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, fmaf(a.w, b.w, accW0[q]))));
        }
      }
      if (pig == 0) accb0 += (a.x + a.y) + (a.z + a.w);
    }
  }
  // ---- partial sums of this CTA: [dW1 64 x 64][db1 64][dW0 64 x F][db0 64]
  float* w = ws + (int64_t)blockIdx.x * (RB_H * RB_H + RB_H + RB_H * F + RB_H);
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int d = 0; d < 4; ++d) w[(ty + 16 * c) * RB_H + tx + 16 * d] = accW1[c][d];
  if (t < RB_H) w[RB_H * RB_H + t] = accb1;
#pragma unroll
  for (int q = 0; q < RB_FMAX / 4; ++q)
    if (q < fq) w[RB_H * RB_H + RB_H + pj * F + pig * fq + q] = accW0[q];
  if (pig == 0) w[RB_H * RB_H + RB_H + RB_H * F + pj] = accb0;
}

// out[element] += sum over the CTA partials, fixed order: block = 32 elements x 32 partial lanes (as eeg_wgrad_umma.cu)
__global__ void __launch_bounds__(1024) k_radial_hidden_reduce(const float* __restrict__ ws, int n_parts, int F, float* __restrict__ dW0,
                                                               float* __restrict__ db0, float* __restrict__ dW1, float* __restrict__ db1) {
  __shared__ float part[32][33];
  const int total = RB_H * RB_H + RB_H + RB_H * F + RB_H;
  const int i = blockIdx.x * 32 + threadIdx.x, y = threadIdx.y;
  float s = 0.f;
  if (i < total)
    for (int p = y; p < n_parts; p += 32) s += ws[(int64_t)p * total + i];
  part[y][threadIdx.x] = s;
  __syncthreads();
  if (y == 0 && i < total) {
    float r = 0.f;
#pragma unroll
    for (int q = 0; q < 32; ++q) r += part[q][threadIdx.x];
    if (i < RB_H * RB_H) dW1[i] += r;
    else if (i < RB_H * RB_H + RB_H) db1[i - RB_H * RB_H] += r;
    else if (i < RB_H * RB_H + RB_H + RB_H * F) dW0[i - RB_H * RB_H - RB_H] += r;
    else db0[i - RB_H * RB_H - RB_H - RB_H * F] += r;
  }
}

int rb_grid(int64_t E) {
  const int64_t tiles = (E + RB_TE - 1) / RB_TE;
  return (int)(tiles < 2 * EEG_NUM_SMS ? tiles : 2 * EEG_NUM_SMS);
}

}  // namespace

extern "C" int64_t eeg_radial_hidden_bwd_ws_floats(int F, int64_t E) {
  return (int64_t)rb_grid(E) * (RB_H * RB_H + RB_H + RB_H * F + RB_H);
}

extern "C" int eeg_radial_hidden_bwd(const float* f, int F, int f_stride, const float* W0, const float* b0, const float* W1, const float* b1,
                                     const float* ds2, int64_t E, float* s2_out, float* dW0, float* db0, float* dW1, float* db1, float* ws,
                                     int64_t ws_floats, void* stream) {
  EEG_REQUIRE(F > 0 && F % 4 == 0 && F <= RB_FMAX && f_stride >= F, "eeg_radial_hidden_bwd: F must be a multiple of 4, at most %d (got %d)",
              RB_FMAX, F);
  if (E == 0) return 0;
  EEG_REQUIRE(ws != nullptr && ws_floats >= eeg_radial_hidden_bwd_ws_floats(F, E), "eeg_radial_hidden_bwd: workspace too small");
  static EegDevFlags attr_dev; bool& attr_set = eeg_dev_flag(attr_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_radial_hidden_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RbSmem));
    if (e != cudaSuccess) {
      eeg_set_error("eeg_radial_hidden_bwd: cannot opt in to %zu B shared memory: %s", sizeof(RbSmem), cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int grid = rb_grid(E);
  k_radial_hidden_bwd<<<grid, RB_THREADS, sizeof(RbSmem), (cudaStream_t)stream>>>(f, F, f_stride, W0, b0, W1, b1, ds2, E, s2_out, ws);
  EEG_CHECK_LAUNCH("eeg_radial_hidden_bwd");
  const int total = RB_H * RB_H + RB_H + RB_H * F + RB_H;
  k_radial_hidden_reduce<<<(total + 31) / 32, dim3(32, 32), 0, (cudaStream_t)stream>>>(ws, grid, F, dW0, db0, dW1, db1);
  EEG_CHECK_LAUNCH("eeg_radial_hidden_bwd (reduce)");
  return 0;
}
