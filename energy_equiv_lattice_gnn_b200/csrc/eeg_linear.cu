// Generic fp32 block GEMM over blocked (m-major) irreps layouts.
// Replaces e3nn o3.Linear (einsum 'zui,uw->zwi' per irrep -> cuBLAS bmm; gnn/blocks.py:516-521,552-559,471-476,276-279)
// and the torch.nn.Linear layers of the radial MLP (gnn/blocks.py:538-549).
// CUDA-core path: 128x32x16 tiles, 256 threads, 4x4 register micro-tile, float4 smem reads.
#include <stdlib.h>
#include <string.h>
#include "eeg_common.cuh"

namespace {

constexpr int BM = 128, BN = 32, BK = 16, APAD = 4;

struct BlockDesc {
  int rows;      // 2l+1
  int k;         // mul_in
  int nout;      // mul_out
  int in_off, out_off;
  int64_t w_off;
  float scale;
  int tile_begin;  // first tile index of this block
  int n_col_tiles;
};

struct LinearParams {
  BlockDesc b[EEG_MAX_L];
  int n_blocks;
  int in_stride, out_stride;
  int transB, act;
};

// PREFETCH: the next K-step is loaded into registers during the current one's FMAs (K = 1344: 0.376 -> 0.349 ms); with only two
// K-steps (K = 32 -> 160..320) the extra registers cost more than the overlap gains (0.348 -> 0.368 ms), so K < 64 loads in place.
template <bool PREFETCH>
__global__ void __launch_bounds__(256) k_block_linear(const float* __restrict__ in, const float* __restrict__ W,
                                                      const float* __restrict__ bias, const float* __restrict__ residual,
                                                      float* __restrict__ out, int64_t N, LinearParams P) {
  __shared__ __align__(16) float As[BK][BM + APAD];
  __shared__ __align__(16) float Bs[BK][BN];
  int tile = blockIdx.x;
  int bi = 0;
#pragma unroll
  for (int q = 1; q < EEG_MAX_L; ++q)
    if (q < P.n_blocks && tile >= P.b[q].tile_begin) bi = q;
  const BlockDesc d = P.b[bi];
  tile -= d.tile_begin;
  const int col_tile = tile % d.n_col_tiles;
  const int64_t row0 = (int64_t)(tile / d.n_col_tiles) * BM;
  const int n0 = col_tile * BN;
  const int64_t total_rows = N * d.rows;
  const int t = threadIdx.x;
  const int tx = t & 7, ty = t >> 3;

  // A-load mapping: 4 lanes per row (4 float4 = 16 k), 64 rows per iteration, 2 iterations
  const int a_kq = t & 3, a_r = t >> 2;
  const float* a_ptr[2];
  bool a_ok[2];
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    int64_t r = row0 + a_r + 64 * it;
    a_ok[it] = r < total_rows;
    int64_t n = a_ok[it] ? r / d.rows : 0;
    int m = a_ok[it] ? (int)(r - n * d.rows) : 0;
    a_ptr[it] = in + n * P.in_stride + d.in_off + (int64_t)m * d.k + a_kq * 4;
  }
  const float* Wb = W + d.w_off;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // the next K-step's A / B pieces are loaded into registers while the current one is multiplied
  float4 av4[2], bv4;
  auto fetch = [&](int k0) {
#pragma unroll
    for (int it = 0; it < 2; ++it)
      av4[it] = (a_ok[it] && k0 + a_kq * 4 < d.k) ? __ldg(reinterpret_cast<const float4*>(a_ptr[it] + k0)) : make_float4(0.f, 0.f, 0.f, 0.f);
    bv4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < 128) {
      if (!P.transB) {   // W[k][n]: 16 rows x 32 cols = 128 float4
        const int kk = t >> 3, c4 = (t & 7) * 4;
        if (k0 + kk < d.k && n0 + c4 < d.nout) bv4 = __ldg(reinterpret_cast<const float4*>(Wb + (int64_t)(k0 + kk) * d.nout + n0 + c4));
      } else {           // W[n][k]: 32 cols x 16 k = 128 float4 along k
        const int c = t >> 2, kq = t & 3;
        if (n0 + c < d.nout && k0 + kq * 4 < d.k) bv4 = __ldg(reinterpret_cast<const float4*>(Wb + (int64_t)(n0 + c) * d.k + k0 + kq * 4));
      }
    }
  };
  if (PREFETCH) fetch(0);
  for (int k0 = 0; k0 < d.k; k0 += BK) {
    if (!PREFETCH) fetch(k0);
#pragma unroll
    for (int it = 0; it < 2; ++it) {   // A tile, transposed into smem
      const int r = a_r + 64 * it;
      As[a_kq * 4 + 0][r] = av4[it].x;
      As[a_kq * 4 + 1][r] = av4[it].y;
      As[a_kq * 4 + 2][r] = av4[it].z;
      As[a_kq * 4 + 3][r] = av4[it].w;
    }
    if (t < 128) {
      if (!P.transB) {
        *reinterpret_cast<float4*>(&Bs[t >> 3][(t & 7) * 4]) = bv4;
      } else {
        const int c = t >> 2, kq = t & 3;
        Bs[kq * 4 + 0][c] = bv4.x;
        Bs[kq * 4 + 1][c] = bv4.y;
        Bs[kq * 4 + 2][c] = bv4.z;
        Bs[kq * 4 + 3][c] = bv4.w;
      }
    }
    __syncthreads();
    if (PREFETCH && k0 + BK < d.k) fetch(k0 + BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }

  const int col = n0 + tx * 4;
  if (col >= d.nout) return;
  float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
  if (bias != nullptr && bi == 0) bv = __ldg(reinterpret_cast<const float4*>(bias + col));
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t r = row0 + ty * 4 + i;
    if (r >= total_rows) continue;
    int64_t n = r / d.rows;
    int m = (int)(r - n * d.rows);
    int64_t o = n * P.out_stride + d.out_off + (int64_t)m * d.nout + col;
    float4 v = make_float4(fmaf(acc[i][0], d.scale, bv.x), fmaf(acc[i][1], d.scale, bv.y), fmaf(acc[i][2], d.scale, bv.z),
                           fmaf(acc[i][3], d.scale, bv.w));
    if (P.act == 1) {
      v.x = eeg_silu(v.x); v.y = eeg_silu(v.y); v.z = eeg_silu(v.z); v.w = eeg_silu(v.w);
    }
    if (residual != nullptr) {
      float4 rr = __ldg(reinterpret_cast<const float4*>(residual + o));
      v.x += rr.x; v.y += rr.y; v.z += rr.z; v.w += rr.w;
    }
    *reinterpret_cast<float4*>(out + o) = v;
  }
}

}  // namespace

// tcgen05 path (eeg_tclinear.cu): 0 = launched, -1 = shape outside the tensor-core path, > 0 = error
int eeg_tc_linear_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                         const float* bias, int transB, const float* residual, float* out, const EegLayout* lout, int64_t N,
                         int gate, int gate_off, float act_const, const int32_t* in_ld, const int32_t* k_override, cudaStream_t stream);

// multiplicity-32 streaming path (eeg_lin32.cu): 0 = launched, -1 = shape outside it
int eeg_lin32_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                     const float* bias, int transB, const float* residual, float* out, const EegLayout* lout, int64_t N,
                     int gate, int n_scalars, float act_const, const int32_t* in_ld, cudaStream_t stream);

// multiplicity-32 warp-MMA path (eeg_lin32mma.cu): 0 = launched, -1 = shape outside it
int eeg_lin32_mma_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                         const float* bias, int transB, const float* residual, float* out, const EegLayout* lout, int64_t N,
                         cudaStream_t stream);

// long-K, 64-output input gradient on tcgen05 with TMA operands (eeg_dx_umma.cu): 0 = launched, -1 = shape outside it
int eeg_dx_umma_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                       int transB, float* out, const EegLayout* lout, int64_t N, cudaStream_t stream);

// multiplicity 32 -> wide outputs on warp MMAs (eeg_lin32wide.cu): 0 = launched, -1 = shape outside it
int eeg_lin32_wide_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                          int transB, float* out, const EegLayout* lout, int64_t N, cudaStream_t stream);

static bool eeg_linear_use_tc() {
  static const char* env = getenv("EEG_LINEAR");  // "cuda": force the CUDA-core GEMM (A/B timing, debugging)
  return !(env && !strcmp(env, "cuda"));
}

extern "C" int eeg_gate_linear_fwd(const float* z, const EegLayout* lz, int n_scalars, float act_const, const float* W,
                                   const int64_t* host_w_off, const float* host_scale, float* out, const EegLayout* lout,
                                   int64_t N, void* stream) {
  EEG_REQUIRE(lz && lout && lz->n_blocks == lout->n_blocks && lz->n_blocks <= EEG_MAX_L, "eeg_gate_linear_fwd: layout mismatch");
  if (N == 0) return 0;
  int32_t in_ld[EEG_MAX_L], kk[EEG_MAX_L];
  int gates = 0;
  for (int b = 0; b < lz->n_blocks; ++b) {
    EEG_REQUIRE(lz->rows[b] == lout->rows[b], "eeg_gate_linear_fwd: block %d rows differ", b);
    in_ld[b] = lz->mul[b];
    kk[b] = b == 0 ? n_scalars : lz->mul[b];
    if (b > 0) {
      EEG_REQUIRE(lz->mul[b] == lz->mul[1], "eeg_gate_linear_fwd: gated blocks must share one multiplicity");
      gates += lz->mul[b];
    }
  }
  EEG_REQUIRE(lz->mul[0] == n_scalars + gates, "eeg_gate_linear_fwd: block 0 must hold %d scalars + %d gates", n_scalars, gates);
  static const char* env = getenv("EEG_LINEAR");
  if (!(env && !strcmp(env, "tc"))) {
    const int r32 = eeg_lin32_launch(z, lz, W, host_w_off, host_scale, nullptr, 0, nullptr, out, lout, N, 1, n_scalars, act_const, in_ld,
                                     (cudaStream_t)stream);
    if (r32 != -1) return r32;
  }
  const int rc = eeg_tc_linear_launch(z, lz, W, host_w_off, host_scale, nullptr, 0, nullptr, out, lout, N, 1, n_scalars, act_const,
                                      in_ld, kk, (cudaStream_t)stream);
  if (rc == -1) {
    eeg_set_error("eeg_gate_linear_fwd: shape outside the tensor-core path (multiplicities must be multiples of 32 / 16)");
    return 3;
  }
  return rc;
}

extern "C" int eeg_block_linear_fwd(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off,
                                    const float* host_scale, const float* bias, int transB, int act, const float* residual,
                                    float* out, const EegLayout* lout, int64_t N, void* stream) {
  EEG_REQUIRE(lin && lout && lin->n_blocks == lout->n_blocks && lin->n_blocks <= EEG_MAX_L, "eeg_block_linear_fwd: layout mismatch");
  if (N == 0) return 0;
  static const char* env32 = getenv("EEG_LINEAR");   // "lin32" / "tc" / "cuda": route the K = 32 shapes to those kernels (A/B timing)
  if (act == 0 && env32 == nullptr) {
    // multiplicity-32 inputs: node-major warp-level tensor-core kernel (eeg_lin32mma.cu)
    const int rm = eeg_lin32_mma_launch(in, lin, W, host_w_off, host_scale, bias, transB, residual, out, lout, N, (cudaStream_t)stream);
    if (rm != -1) return rm;
    if (bias == nullptr && residual == nullptr) {
      const int rw = eeg_lin32_wide_launch(in, lin, W, host_w_off, host_scale, transB, out, lout, N, (cudaStream_t)stream);
      if (rw != -1) return rw;
    }
  }
  if (act == 0 && env32 && !strcmp(env32, "lin32")) {
    // measured on B200 (N = 30720): 107 / 102 us (product+residual / linear_1) vs 101 / 96 us on the tiled GEMM below, so
    // the streaming kernel only serves the gate-fused linear_2 (131 us vs 160 us on the tensor-core kernel) by default
    const int r32 = eeg_lin32_launch(in, lin, W, host_w_off, host_scale, bias, transB, residual, out, lout, N, 0, 0, 0.f, nullptr,
                                     (cudaStream_t)stream);
    if (r32 != -1) return r32;
  }
  if (act == 0 && eeg_linear_use_tc()) {
    // Measured on B200 (N = 30720): the tensor-core kernel wins for the path-stacked interaction.linear (K = 160..320:
    // 393 us vs 478 us) but not for the K = 32 mixes (140-250 us vs 80-100 us on the tiled GEMM: its LDG/STS-staged
    // producers and the epilogue share one congested memory-instruction queue); those go to eeg_lin32 above.
    bool same_rows = true;
    int max_k = 0;
    for (int b = 0; b < lin->n_blocks; ++b) {
      same_rows &= lin->rows[b] == lout->rows[b];
      if (lin->mul[b] > max_k) max_k = lin->mul[b];
    }
    static const char* env = getenv("EEG_LINEAR");
    if (same_rows && (max_k >= 128 || (env && !strcmp(env, "tc")))) {
      const int rc = eeg_tc_linear_launch(in, lin, W, host_w_off, host_scale, bias, transB, residual, out, lout, N, 0, 0, 0.f,
                                          nullptr, nullptr, (cudaStream_t)stream);
      if (rc != -1) return rc;
    }
  }
  if (act == 0 && bias == nullptr && residual == nullptr && eeg_linear_use_tc()) {
    const int rd = eeg_dx_umma_launch(in, lin, W, host_w_off, host_scale, transB, out, lout, N, (cudaStream_t)stream);
    if (rd != -1) return rd;
  }
  LinearParams P;
  P.n_blocks = lin->n_blocks;
  P.in_stride = lin->stride;
  P.out_stride = lout->stride;
  P.transB = transB;
  P.act = act;
  EEG_REQUIRE(lin->stride % 4 == 0 && lout->stride % 4 == 0, "eeg_block_linear_fwd: strides must be multiples of 4");
  int64_t tiles = 0;
  for (int b = 0; b < P.n_blocks; ++b) {
    EEG_REQUIRE(lin->rows[b] == lout->rows[b], "eeg_block_linear_fwd: block %d rows differ", b);
    EEG_REQUIRE(lin->mul[b] % 4 == 0 && lout->mul[b] % 4 == 0 && lin->off[b] % 4 == 0 && lout->off[b] % 4 == 0 &&
                    host_w_off[b] % 4 == 0,
                "eeg_block_linear_fwd: block %d not float4-aligned", b);
    BlockDesc& d = P.b[b];
    d.rows = lin->rows[b];
    d.k = lin->mul[b];
    d.nout = lout->mul[b];
    d.in_off = lin->off[b];
    d.out_off = lout->off[b];
    d.w_off = host_w_off[b];
    d.scale = host_scale[b];
    d.tile_begin = (int)tiles;
    d.n_col_tiles = (d.nout + BN - 1) / BN;
    int64_t row_tiles = (N * d.rows + BM - 1) / BM;
    tiles += (d.nout > 0 && d.k > 0) ? row_tiles * d.n_col_tiles : 0;
    if (d.nout == 0 || d.k == 0) d.n_col_tiles = 1;
    EEG_REQUIRE(tiles < (1LL << 31), "eeg_block_linear_fwd: too many tiles");
  }
  if (tiles == 0) return 0;
  int max_k = 0;
  for (int b = 0; b < P.n_blocks; ++b)
    if (P.b[b].nout > 0 && P.b[b].k > max_k) max_k = P.b[b].k;
  if (max_k >= 64) k_block_linear<true><<<(unsigned)tiles, 256, 0, (cudaStream_t)stream>>>(in, W, bias, residual, out, N, P);
  else k_block_linear<false><<<(unsigned)tiles, 256, 0, (cudaStream_t)stream>>>(in, W, bias, residual, out, N, P);
  EEG_CHECK_LAUNCH("eeg_block_linear_fwd");
  return 0;
}
