// Inference tail of the node pipeline as ONE streaming pass per irrep block (multiplicity-32 models):
//   h   = product.linear(B) + h_in           EquivariantProductBlock.linear + residual (gnn/blocks.py:471-476,488-490; model.py:95-96)
//   z   = linear_1(h)                         GeneralNonLinearReadoutBlock (gnn/blocks.py:250-283): 0e -> 32 scalars + 32 gates per l >= 1
//   g   = Gate(z)                             c * silu on the scalars, gated rows * c * silu(gate)   (e3nn.nn.Gate)
//   o   = linear_2(g)                         per-l 32 -> mul_out(l) (<= 32)
// The three K = 32 mixes of a row are chained inside one warp: the row tile is staged once (cp.async ring, as in
// eeg_lin32.cu), each stage's result goes back to the warp's shared-memory slot as the broadcast operand of the next one,
// and only o [N, D, mul_out] (plus the [N, 32 (L-1)] gate activations, written by the 0e pass) reaches HBM: h and z are
// never materialised.  Two launches: the 0e block first (it produces the gates), then the gated blocks.
#include <stdlib.h>
#include <string.h>
#include "eeg_common.cuh"

namespace {

constexpr int RC_WARPS = 8, RC_CHUNKS_PER_WARP = 8, RC_STAGES = 3, RC_MAXR = 10;
constexpr int RC_RING = RC_WARPS * RC_STAGES * 2 * RC_MAXR * 32;   // floats: per warp and stage [x | residual][rows][32]
constexpr int RC_W2 = 32 * 32;                                    // floats: linear_2 block, [u][column] padded to 32 columns
constexpr int RC_W1S = 32 * 32 * EEG_MAX_L;                       // floats: 0e linear_1 block, up to EEG_MAX_L column groups
constexpr int RC_SMEM = (RC_RING + RC_W2 + RC_W1S) * 4;

__host__ __device__ constexpr int rc_npc(int rows) { return rows == 1 ? 8 : rows == 3 ? 3 : rows == 5 ? 2 : 1; }

__device__ __forceinline__ void rc_cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}

struct RcBlock {
  int rows, nout, in_off, out_off;
  int64_t wp_off, w1_off, w2_off;
  float sp, s1, s2;
  int tile_begin, n_rt;
};
struct RcParams {
  RcBlock b[EEG_MAX_L];
  int n_blocks, b_first, in_stride, out_stride, n_gate_blocks;
  float act_const;
};

// acc[i] += sum_u slot[i][u] * w[u]   (slot rows read as broadcast float4)
template <int R>
__device__ __forceinline__ void rc_mix(const float* slot, const float (&w)[32], float (&acc)[R]) {
  const float4* xv = reinterpret_cast<const float4*>(slot);
#pragma unroll
  for (int i0 = 0; i0 < R; i0 += 5) {   // five rows at a time: enough independent chains, few live operands
#pragma unroll
    for (int u4 = 0; u4 < 8; ++u4) {
#pragma unroll
      for (int i = i0; i < i0 + 5 && i < R; ++i) {
        const float4 v = xv[i * 8 + u4];
        acc[i] = fmaf(v.x, w[4 * u4 + 0], acc[i]);
        acc[i] = fmaf(v.y, w[4 * u4 + 1], acc[i]);
        acc[i] = fmaf(v.z, w[4 * u4 + 2], acc[i]);
        acc[i] = fmaf(v.w, w[4 * u4 + 3], acc[i]);
      }
    }
  }
}
// same with the weight column read from shared memory ws[u * 32 + lane] (conflict free)
template <int R>
__device__ __forceinline__ void rc_mix_smem(const float* slot, const float* ws, float (&acc)[R]) {
  const float4* xv = reinterpret_cast<const float4*>(slot);
#pragma unroll 2
  for (int u4 = 0; u4 < 8; ++u4) {
    const float w0 = ws[(4 * u4 + 0) * 32], w1 = ws[(4 * u4 + 1) * 32], w2 = ws[(4 * u4 + 2) * 32], w3 = ws[(4 * u4 + 3) * 32];
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const float4 v = xv[i * 8 + u4];
      acc[i] = fmaf(v.x, w0, acc[i]);
      acc[i] = fmaf(v.y, w1, acc[i]);
      acc[i] = fmaf(v.z, w2, acc[i]);
      acc[i] = fmaf(v.w, w3, acc[i]);
    }
  }
}

template <int ROWS>
__device__ __forceinline__ void rc_tile(const float* __restrict__ Bx, const float* __restrict__ res, float* __restrict__ gates,
                                        float* __restrict__ out, int64_t N, const RcParams& P, const RcBlock& d, int bi, int rt,
                                        const float (&wp)[32], const float (&w1)[32], const float* w1s, const float* w2s,
                                        float* ring, int lane, int warp) {
  constexpr int NPC = rc_npc(ROWS), R = NPC * ROWS;
  const int64_t node0 = (int64_t)rt * (RC_WARPS * RC_CHUNKS_PER_WARP * NPC);
  auto chunk_node = [&](int c) { return node0 + (int64_t)(c * RC_WARPS + warp) * NPC; };
  const int gstride = P.n_gate_blocks * 32;

  auto issue = [&](int c) {
    const int64_t n0 = chunk_node(c);
    if (c < RC_CHUNKS_PER_WARP && n0 < N) {
      float* xs = ring + (c % RC_STAGES) * (2 * RC_MAXR * 32) + lane;
#pragma unroll
      for (int j = 0; j < NPC; ++j) {
        if (n0 + j < N) {
          const int64_t o = (n0 + j) * P.in_stride + d.in_off + lane;
#pragma unroll
          for (int m = 0; m < ROWS; ++m) rc_cp_async4(xs + (j * ROWS + m) * 32, Bx + o + m * 32);
          if (res != nullptr) {
#pragma unroll
            for (int m = 0; m < ROWS; ++m) rc_cp_async4(xs + (RC_MAXR + j * ROWS + m) * 32, res + o + m * 32);
          }
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int c = 0; c < RC_STAGES - 1; ++c) issue(c);

#pragma unroll 1
  for (int c = 0; c < RC_CHUNKS_PER_WARP; ++c) {
    const int64_t n0 = chunk_node(c);
    if (n0 >= N) break;
    issue(c + RC_STAGES - 1);
    asm volatile("cp.async.wait_group %0;" ::"n"(RC_STAGES - 1) : "memory");
    __syncwarp();
    float* slot = ring + (c % RC_STAGES) * (2 * RC_MAXR * 32);
    // rows of nodes past the end were not copied: zero them so that nothing non-finite enters the chain
#pragma unroll
    for (int j = 0; j < NPC; ++j)
      if (n0 + j >= N) {
#pragma unroll
        for (int m = 0; m < ROWS; ++m) slot[(j * ROWS + m) * 32 + lane] = 0.f, slot[(RC_MAXR + j * ROWS + m) * 32 + lane] = 0.f;
      }
    __syncwarp();
    // ---- h = product.linear(B) + h_in
    float a[R];
#pragma unroll
    for (int i = 0; i < R; ++i) a[i] = res != nullptr ? slot[(RC_MAXR + i) * 32 + lane] : 0.f;
    rc_mix<R>(slot, wp, a);
    __syncwarp();
#pragma unroll
    for (int i = 0; i < R; ++i) slot[i * 32 + lane] = a[i];
    __syncwarp();
    // ---- z = linear_1(h); Gate
    float g[R];
    if (ROWS == 1) {
      // 0e: column group 0 = the 32 scalars, groups 1.. = gates of the blocks l >= 1 (stored for the second launch)
#pragma unroll
      for (int i = 0; i < R; ++i) g[i] = 0.f;
      rc_mix<R>(slot, w1, g);
#pragma unroll
      for (int i = 0; i < R; ++i) g[i] = P.act_const * eeg_silu(g[i]);
      for (int cg = 1; cg <= P.n_gate_blocks; ++cg) {
        float z[R];
#pragma unroll
        for (int i = 0; i < R; ++i) z[i] = 0.f;
        rc_mix_smem<R>(slot, w1s + cg * 1024 + lane, z);
#pragma unroll
        for (int i = 0; i < R; ++i)
          if (n0 + i < N) gates[(n0 + i) * gstride + (cg - 1) * 32 + lane] = P.act_const * eeg_silu(z[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < R; ++i) g[i] = 0.f;
      rc_mix<R>(slot, w1, g);
#pragma unroll
      for (int j = 0; j < NPC; ++j) {
        const float gt = n0 + j < N ? __ldg(gates + (n0 + j) * gstride + (bi - 1) * 32 + lane) : 0.f;
#pragma unroll
        for (int m = 0; m < ROWS; ++m) g[j * ROWS + m] *= gt;
      }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < R; ++i) slot[i * 32 + lane] = g[i];
    __syncwarp();
    // ---- o = linear_2(g)
    float o[R];
#pragma unroll
    for (int i = 0; i < R; ++i) o[i] = 0.f;
    rc_mix_smem<R>(slot, w2s + lane, o);
    if (lane < d.nout) {
#pragma unroll
      for (int j = 0; j < NPC; ++j) {
        if (n0 + j < N) {
          float* po = out + (n0 + j) * P.out_stride + d.out_off + lane;
#pragma unroll
          for (int m = 0; m < ROWS; ++m) po[m * d.nout] = o[j * ROWS + m];
        }
      }
    }
    __syncwarp();   // the stage is refilled by the next iteration's issue
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

__global__ void __launch_bounds__(RC_WARPS * 32, 2) k_readout_chain(const float* __restrict__ Bx, const float* __restrict__ res,
                                                                    const float* __restrict__ Wp, const float* __restrict__ W1,
                                                                    const float* __restrict__ W2, float* __restrict__ gates,
                                                                    float* __restrict__ out, int64_t N, RcParams P, int n_tiles) {
  extern __shared__ __align__(16) float rc_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* ring = rc_smem + warp * (RC_STAGES * 2 * RC_MAXR * 32);
  float* w2s = rc_smem + RC_RING;    // [32 u][32 columns] (columns >= nout: 0), scale folded in
  float* w1s = w2s + RC_W2;          // 0e block only: [column group][32 u][32]
  float wp[32], w1[32];
  int cur_bi = -1;
  const int per = (n_tiles + gridDim.x - 1) / gridDim.x;
  const int t_end = min(n_tiles, ((int)blockIdx.x + 1) * per);
  for (int tile = blockIdx.x * per; tile < t_end; ++tile) {
    int bi = P.b_first;
    for (int q = P.b_first + 1; q < P.n_blocks; ++q)
      if (tile >= P.b[q].tile_begin) bi = q;
    const RcBlock d = P.b[bi];
    if (bi != cur_bi) {   // CTA-uniform: the weights of this block
      __syncthreads();
      for (int i = threadIdx.x; i < 32 * 32; i += RC_WARPS * 32) {
        const int u = i >> 5, col = i & 31;
        w2s[i] = col < d.nout ? d.s2 * __ldg(W2 + d.w2_off + (int64_t)u * d.nout + col) : 0.f;
      }
      if (d.rows == 1) {
        const int n1 = 32 * (1 + P.n_gate_blocks);
        for (int i = threadIdx.x; i < 32 * n1; i += RC_WARPS * 32) {
          const int u = i / n1, col = i % n1;
          w1s[(col >> 5) * 1024 + u * 32 + (col & 31)] = d.s1 * __ldg(W1 + d.w1_off + (int64_t)u * n1 + col);
        }
      }
      const int n1 = d.rows == 1 ? 32 * (1 + P.n_gate_blocks) : 32;
#pragma unroll
      for (int u = 0; u < 32; ++u) {
        wp[u] = d.sp * __ldg(Wp + d.wp_off + u * 32 + lane);
        w1[u] = d.s1 * __ldg(W1 + d.w1_off + (int64_t)u * n1 + lane);
      }
      __syncthreads();
      cur_bi = bi;
    }
    const int rt = tile - d.tile_begin;
    switch (d.rows) {
      case 1: rc_tile<1>(Bx, res, gates, out, N, P, d, bi, rt, wp, w1, w1s, w2s, ring, lane, warp); break;
      case 3: rc_tile<3>(Bx, res, gates, out, N, P, d, bi, rt, wp, w1, w1s, w2s, ring, lane, warp); break;
      case 5: rc_tile<5>(Bx, res, gates, out, N, P, d, bi, rt, wp, w1, w1s, w2s, ring, lane, warp); break;
      case 7: rc_tile<7>(Bx, res, gates, out, N, P, d, bi, rt, wp, w1, w1s, w2s, ring, lane, warp); break;
      default: rc_tile<9>(Bx, res, gates, out, N, P, d, bi, rt, wp, w1, w1s, w2s, ring, lane, warp); break;
    }
  }
}

}  // namespace

// warp-level tensor-core version (eeg_readout_mma.cu): 0 = launched, -1 = shape outside it
int eeg_readout_chain_mma_launch(const float* B, const float* residual, const EegLayout* lh, const float* Wp,
                                 const int64_t* host_wp_off, const float* host_sp, const float* W1, const int64_t* host_w1_off,
                                 const float* host_s1, const float* W2, const int64_t* host_w2_off, const float* host_s2,
                                 float act_const, float* gates, float* out, const EegLayout* lout, int64_t N, cudaStream_t stream);

extern "C" int eeg_readout_chain_fwd(const float* B, const float* residual, const EegLayout* lh, const float* Wp,
                                     const int64_t* host_wp_off, const float* host_sp, const float* W1, const int64_t* host_w1_off,
                                     const float* host_s1, const float* W2, const int64_t* host_w2_off, const float* host_s2,
                                     float act_const, float* gates, float* out, const EegLayout* lout, int64_t N, void* stream) {
  EEG_REQUIRE(lh && lout && lh->n_blocks == lout->n_blocks && lh->n_blocks >= 1 && lh->n_blocks <= EEG_MAX_L,
              "eeg_readout_chain_fwd: layout mismatch");
  if (N == 0) return 0;
  {
    static const char* env = getenv("EEG_CHAIN");   // "cuda": force the CUDA-core chain below (A/B timing)
    if (!(env && !strcmp(env, "cuda"))) {
      const int rm = eeg_readout_chain_mma_launch(B, residual, lh, Wp, host_wp_off, host_sp, W1, host_w1_off, host_s1, W2, host_w2_off,
                                                  host_s2, act_const, gates, out, lout, N, (cudaStream_t)stream);
      if (rm != -1) return rm;
    }
  }
  RcParams P;
  P.n_blocks = lh->n_blocks;
  P.in_stride = lh->stride;
  P.out_stride = lout->stride;
  P.n_gate_blocks = lh->n_blocks - 1;
  P.act_const = act_const;
  for (int b = 0; b < P.n_blocks; ++b) {
    const int rows = lh->rows[b];
    EEG_REQUIRE(lh->mul[b] == 32, "eeg_readout_chain_fwd: block %d: input multiplicity must be 32 (got %d)", b, lh->mul[b]);
    EEG_REQUIRE(rows == lout->rows[b] && rows == 2 * b + 1 && rows <= 9, "eeg_readout_chain_fwd: block %d must have 2l+1 rows, l <= 4", b);
    EEG_REQUIRE(lout->mul[b] >= 1 && lout->mul[b] <= 32, "eeg_readout_chain_fwd: block %d: output multiplicity must be 1..32", b);
    RcBlock& d = P.b[b];
    d.rows = rows;
    d.nout = lout->mul[b];
    d.in_off = lh->off[b];
    d.out_off = lout->off[b];
    d.wp_off = host_wp_off[b];
    d.w1_off = host_w1_off[b];
    d.w2_off = host_w2_off[b];
    d.sp = host_sp[b];
    d.s1 = host_s1[b];
    d.s2 = host_s2[b];
    d.n_rt = eeg_div_up(N, RC_WARPS * RC_CHUNKS_PER_WARP * rc_npc(rows));
  }
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_readout_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, RC_SMEM);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_readout_chain_fwd: cannot opt in to %d B shared memory: %s", RC_SMEM, cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  // launch 1: the 0e block (writes the gate activations); launch 2: the gated blocks
  for (int pass = 0; pass < 2; ++pass) {
    const int b0 = pass == 0 ? 0 : 1, b1 = pass == 0 ? 1 : P.n_blocks;
    int64_t tiles = 0;
    for (int b = b0; b < b1; ++b) {
      P.b[b].tile_begin = (int)tiles;
      tiles += P.b[b].n_rt;
    }
    if (tiles == 0) continue;
    EEG_REQUIRE(tiles < (1LL << 31), "eeg_readout_chain_fwd: too many tiles");
    RcParams Q = P;
    Q.b_first = b0;
    Q.n_blocks = b1;
    const unsigned grid = (unsigned)(tiles < 2 * EEG_NUM_SMS ? tiles : 2 * EEG_NUM_SMS);
    k_readout_chain<<<grid, RC_WARPS * 32, RC_SMEM, (cudaStream_t)stream>>>(B, residual, Wp, W1, W2, gates, out, N, Q, (int)tiles);
    EEG_CHECK_LAUNCH("eeg_readout_chain_fwd");
  }
  return 0;
}
