// Per-irrep channel mixing for multiplicity-32 inputs (o3.Linear with mul_in = 32: gnn/blocks.py:516-521,552-559,471-476,
// 276-279; optionally with the e3nn Gate of gnn/blocks.py:268-282 applied to the input rows).  These GEMMs have K = 32, so
// they are pure streaming: the kernel keeps one weight column (32 values, scale folded in) in the registers of each lane
// (lane = output channel) and streams rows (and the residual) through a per-warp 4-stage cp.async ring in shared memory
// (coalesced 128-byte rows in, float4 broadcast reads out: 8 LDS.128 + 32 FFMA per row), 3 chunks of whole nodes (8-10
// rows, contiguous in memory, compile-time offsets) in flight per warp; persistent CTAs (2 per SM) walk the (block, column
// group, node tile) list.
#include "eeg_common.cuh"

namespace {

constexpr int L32_WARPS = 8, L32_CHUNKS_PER_WARP = 8, L32_STAGES = 4, L32_MAXR = 10;   // <= 10 rows per chunk
constexpr int L32_SMEM = L32_WARPS * L32_STAGES * 2 * L32_MAXR * 32 * 4;   // per warp and stage: [x | residual][rows][32]

// nodes per chunk for a block of `rows` rows per node: a chunk is whole nodes, so its rows are contiguous in memory
__host__ __device__ constexpr int l32_npc(int rows) { return rows == 1 ? 8 : rows == 3 ? 3 : rows == 5 ? 2 : 1; }

__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}

struct L32Block {
  int rows, nout, in_off, out_off;
  int64_t w_off;
  float scale;
  int tile_begin, n_cg, n_rt;
};

struct L32Params {
  L32Block b[EEG_MAX_L];
  int n_blocks, in_stride, out_stride, transB, n_scalars;
  float act_const;
};

// One tile = 64 chunks (8 per warp, round robin) of NPC nodes x ROWS rows of one block and one 32-column group.
template <int ROWS, bool GATE>
__device__ __forceinline__ void l32_tile(const float* __restrict__ in, const float* __restrict__ residual, float* __restrict__ out,
                                         int64_t N, const L32Params& P, const L32Block& d, int bi, int rt, int col, bool col_ok,
                                         const float (&w)[32], float bv, float* ring, int lane, int warp) {
  constexpr int NPC = l32_npc(ROWS), R = NPC * ROWS;
  const int64_t node0 = (int64_t)rt * (L32_WARPS * L32_CHUNKS_PER_WARP * NPC);
  auto chunk_node = [&](int c) { return node0 + (int64_t)(c * L32_WARPS + warp) * NPC; };
  const float* in_l = in + d.in_off + lane;
  const int64_t out_l = d.out_off + col;

  auto issue = [&](int c) {
    const int64_t n0 = chunk_node(c);
    if (c < L32_CHUNKS_PER_WARP && n0 < N) {
      float* xs = ring + (c % L32_STAGES) * (2 * L32_MAXR * 32) + lane;
#pragma unroll
      for (int j = 0; j < NPC; ++j) {
        if (n0 + j < N) {
          const float* src = in_l + (n0 + j) * P.in_stride;
#pragma unroll
          for (int m = 0; m < ROWS; ++m) cp_async4(xs + (j * ROWS + m) * 32, src + m * 32);
          if (residual != nullptr && col_ok) {
            const float* rs = residual + (n0 + j) * P.out_stride + out_l;
#pragma unroll
            for (int m = 0; m < ROWS; ++m) cp_async4(xs + (L32_MAXR + j * ROWS + m) * 32, rs + m * d.nout);
          }
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int c = 0; c < L32_STAGES - 1; ++c) issue(c);

#pragma unroll 1
  for (int c = 0; c < L32_CHUNKS_PER_WARP; ++c) {
    const int64_t n0 = chunk_node(c);
    if (n0 >= N) break;
    issue(c + L32_STAGES - 1);
    asm volatile("cp.async.wait_group %0;" ::"n"(L32_STAGES - 1) : "memory");
    __syncwarp();
    float* slot = ring + (c % L32_STAGES) * (2 * L32_MAXR * 32);
    if (GATE) {   // e3nn Gate on the input rows: scalars -> act_const * silu, gated rows * act_const * silu(gate of the node)
#pragma unroll
      for (int j = 0; j < NPC; ++j) {
        if (n0 + j < N) {
          float g = 1.f;
          if (ROWS > 1) g = P.act_const * eeg_silu(__ldg(in + (n0 + j) * P.in_stride + P.n_scalars + (bi - 1) * 32 + lane));
#pragma unroll
          for (int m = 0; m < ROWS; ++m) {
            float* e = slot + (j * ROWS + m) * 32 + lane;
            *e = ROWS == 1 ? P.act_const * eeg_silu(*e) : *e * g;
          }
        }
      }
      __syncwarp();
    }
    float acc[R];
#pragma unroll
    for (int i = 0; i < R; ++i) acc[i] = bv;
    const float4* xv = reinterpret_cast<const float4*>(slot);
#pragma unroll
    for (int u4 = 0; u4 < 8; ++u4) {
#pragma unroll
      for (int i = 0; i < R; ++i) {
        const float4 v = xv[i * 8 + u4];
        acc[i] = fmaf(v.x, w[4 * u4 + 0], acc[i]);
        acc[i] = fmaf(v.y, w[4 * u4 + 1], acc[i]);
        acc[i] = fmaf(v.z, w[4 * u4 + 2], acc[i]);
        acc[i] = fmaf(v.w, w[4 * u4 + 3], acc[i]);
      }
    }
    if (col_ok) {
#pragma unroll
      for (int j = 0; j < NPC; ++j) {
        if (n0 + j < N) {
          float* po = out + (n0 + j) * P.out_stride + out_l;
#pragma unroll
          for (int m = 0; m < ROWS; ++m) {
            float v = acc[j * ROWS + m];
            if (residual != nullptr) v += slot[(L32_MAXR + j * ROWS + m) * 32 + lane];
            po[m * d.nout] = v;
          }
        }
      }
    }
    __syncwarp();   // the stage is refilled by the next iteration's issue
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <bool GATE>
__global__ void __launch_bounds__(L32_WARPS * 32, 2) k_lin32(const float* __restrict__ in, const float* __restrict__ W,
                                                             const float* __restrict__ bias, const float* __restrict__ residual,
                                                             float* __restrict__ out, int64_t N, L32Params P, int n_tiles) {
  extern __shared__ __align__(16) float l32_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* ring = l32_smem + warp * (L32_STAGES * 2 * L32_MAXR * 32);   // this warp's stages
  float w[32];
  float bv = 0.f;
  int cur_bi = -1, cur_cg = -1;

  // contiguous tile range per CTA: it stays inside one (block, column group) most of the time, so the weights are rarely reloaded
  const int per = (n_tiles + gridDim.x - 1) / gridDim.x;
  const int t_end = min(n_tiles, ((int)blockIdx.x + 1) * per);
  for (int tile = blockIdx.x * per; tile < t_end; ++tile) {
    int bi = 0;
#pragma unroll
    for (int q = 1; q < EEG_MAX_L; ++q)
      if (q < P.n_blocks && tile >= P.b[q].tile_begin) bi = q;
    const L32Block d = P.b[bi];
    const int lt = tile - d.tile_begin;
    const int n_rt = d.n_rt;
    const int cg = lt / n_rt, rt = lt - cg * n_rt;
    const int col = cg * 32 + lane;
    const bool col_ok = col < d.nout;
    if (bi != cur_bi || cg != cur_cg) {   // weight column of this lane (scale folded in): once per (block, column group)
      const float* Wb = W + d.w_off;
#pragma unroll
      for (int u = 0; u < 32; ++u)
        w[u] = col_ok ? d.scale * __ldg(P.transB ? Wb + (int64_t)col * 32 + u : Wb + (int64_t)u * d.nout + col) : 0.f;
      bv = (bias != nullptr && bi == 0 && col_ok) ? __ldg(bias + col) : 0.f;
      cur_bi = bi;
      cur_cg = cg;
    }
    switch (d.rows) {
      case 1: l32_tile<1, GATE>(in, residual, out, N, P, d, bi, rt, col, col_ok, w, bv, ring, lane, warp); break;
      case 3: l32_tile<3, GATE>(in, residual, out, N, P, d, bi, rt, col, col_ok, w, bv, ring, lane, warp); break;
      case 5: l32_tile<5, GATE>(in, residual, out, N, P, d, bi, rt, col, col_ok, w, bv, ring, lane, warp); break;
      case 7: l32_tile<7, GATE>(in, residual, out, N, P, d, bi, rt, col, col_ok, w, bv, ring, lane, warp); break;
      default: l32_tile<9, GATE>(in, residual, out, N, P, d, bi, rt, col, col_ok, w, bv, ring, lane, warp); break;
    }
  }
}

}  // namespace

// 0 = launched, -1 = shape outside this kernel (every input block: multiplicity 32, 2l+1 rows with l <= 4)
int eeg_lin32_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                     const float* bias, int transB, const float* residual, float* out, const EegLayout* lout, int64_t N,
                     int gate, int n_scalars, float act_const, const int32_t* in_ld, cudaStream_t stream) {
  L32Params P;
  P.n_blocks = lin->n_blocks;
  P.in_stride = lin->stride;
  P.out_stride = lout->stride;
  P.transB = transB;
  P.n_scalars = n_scalars;
  P.act_const = act_const;
  int64_t tiles = 0;
  for (int b = 0; b < P.n_blocks; ++b) {
    const int rows = lin->rows[b];
    const int k = (gate && b == 0) ? n_scalars : lin->mul[b];   // gate: block 0 of z = [scalars | gates], only the scalars feed the GEMM
    if (k != 32 || rows != lout->rows[b] || lout->mul[b] <= 0) return -1;
    if (rows != 1 && rows != 3 && rows != 5 && rows != 7 && rows != 9) return -1;
    if (rows > 1 && in_ld && in_ld[b] != 32) return -1;
    L32Block& d = P.b[b];
    d.rows = rows;
    d.nout = lout->mul[b];
    d.in_off = lin->off[b];
    d.out_off = lout->off[b];
    d.w_off = host_w_off[b];
    d.scale = host_scale[b];
    d.tile_begin = (int)tiles;
    d.n_cg = (d.nout + 31) / 32;
    d.n_rt = eeg_div_up(N, L32_WARPS * L32_CHUNKS_PER_WARP * l32_npc(rows));
    tiles += (int64_t)d.n_rt * d.n_cg;
    if (tiles >= (1LL << 31)) return -1;
  }
  if (tiles == 0) return 0;
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_lin32<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, L32_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_lin32<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, L32_SMEM);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_lin32: cannot opt in to %d B shared memory: %s", L32_SMEM, cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  const unsigned grid = (unsigned)(tiles < 2 * EEG_NUM_SMS ? tiles : 2 * EEG_NUM_SMS);   // persistent: 2 CTAs per SM
  if (gate)
    k_lin32<true><<<grid, L32_WARPS * 32, L32_SMEM, stream>>>(in, W, bias, residual, out, N, P, (int)tiles);
  else
    k_lin32<false><<<grid, L32_WARPS * 32, L32_SMEM, stream>>>(in, W, bias, residual, out, N, P, (int)tiles);
  EEG_CHECK_LAUNCH("eeg_lin32");
  return 0;
}
