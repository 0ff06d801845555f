// Per-irrep channel mixing (e3nn o3.Linear) on the 5th-gen tensor cores.
// Replaces einsum 'zui,uw->zwi' per irrep (gnn/blocks.py:516-521 linear_up, :552-559 interaction.linear, :471-476
// product linear, :276-279 readout linears); optionally applies e3nn.nn.Gate (gnn/blocks.py:268-282) to the input
// rows while they are staged, so the gated tensor is never written to HBM.
//
// One persistent CTA per SM.  A tile = 128 rows (row = (node, m) of one irrep block) x nout outputs; K = mul_in is
// consumed in chunks of 32 floats (one 128-byte line per row).
//   warps 0-3  (epilogue)  : tcgen05.ld accumulator rows -> per-warp shared staging (transpose) -> scale (+bias on 0e,
//                            +residual, prefetched before the accumulator wait) -> full-line coalesced stores
//   warps 4-19 (producers) : 4 groups of 128 threads taking chunks round-robin (~64 KB of loads in flight per SM);
//                            8 lanes read one row's 128-byte line (4 full lines per LDG.128), every value is split into
//                            tf32 hi + lo (3xTF32, eeg_umma.cuh) and both terms are stored in the K-major UMMA layout
//                            whose K-chunk stride is padded to 2064 B, which makes those stores bank-conflict free
//   warp 20    (one lane)  : tcgen05.mma kind::tf32.  B = [W_hi | W_lo] stacked along N, so one K-step is
//                            A_hi x [W_hi|W_lo] (N = 2 nout) + A_lo x W_hi (N = nout): the A tile, which dominates the
//                            shared-memory traffic of these skinny GEMMs, is read twice instead of three times; the two
//                            accumulator halves are added in the epilogue.  (nout > 128: three N = nout MMAs.)
// Weights are split into a resident hi/lo B operand in shared memory: all blocks at once when they fit (no pipeline
// drain between irrep blocks), else once per (CTA, block).
// Work is partitioned statically by cost (K-chunks): CTA c owns the tiles whose first unit falls in its unit range.
#include <stdlib.h>
#include "eeg_common.cuh"
#include "eeg_umma.cuh"

namespace {
using namespace eeg_umma;

constexpr int TL_ROWS = 128;
constexpr int TL_KC = 32;                                  // floats per K-chunk
constexpr uint32_t TL_A_LBO = TL_ROWS * 16 + 16;           // 2064 B between 4-float K-chunks (padded: see above)
constexpr uint32_t TL_TERM_BYTES = (TL_KC / 4) * TL_A_LBO; // 16512 B: one term (hi or lo) of an A chunk
constexpr uint32_t TL_STAGE_BYTES = 2 * TL_TERM_BYTES;     // 33024 B
constexpr int TL_MAX_STAGES = 4;
constexpr int TL_GROUPS = 4;                               // producer groups (128 threads each)
constexpr int TL_THREADS = (4 + 4 * TL_GROUPS + 1) * 32;   // 672
constexpr int TL_MMA_WARP = 4 + 4 * TL_GROUPS;
constexpr int TL_ACC_COLS = 256;                           // TMEM columns per accumulator stage
constexpr int TL_EPI_LD = 36;                              // floats per staged output row (32 + 4 pad: conflict free)
constexpr uint32_t TL_EPI_BYTES = 4 * 32 * TL_EPI_LD * 4;  // 18432 B: one 32 x 32 staging tile per epilogue warp
constexpr uint32_t TL_SMEM_BUDGET = 226 * 1024;

struct TcBlock {
  int rows, k, nout, in_off, in_ld, out_off;
  int64_t w_off;
  float scale;
  int tiles, kc;      // row tiles, K-chunks per tile
  int w_sm;           // float offset of this block's B operand inside the weight image
  int64_t unit0;      // first work unit (tile-major, then K-chunk) of this block
};
struct TcParams {
  TcBlock b[EEG_MAX_L];
  int n_blocks, in_stride, out_stride, transB, stages;
  int gate, gate_off;   // gate != 0: input rows are gated on the fly, gates of block b >= 1 at in_off[0] + gate_off + mul*(b-1)
  float act_const;
  int64_t units;
  uint32_t w_bytes;     // bytes reserved for the weight image
  int resident;         // all blocks' weights are staged once at kernel start
};

// Bounded pipeline wait: a protocol bug traps (launch error on the host) instead of hanging the GPU.
// SLEEP_NS > 0 backs off between polls (producers spend most of their time blocked on a free stage).
template <int SLEEP_NS>
__device__ __forceinline__ void tl_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
    if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
    if (spins > (1u << 22)) __trap();
  }
}

__device__ __forceinline__ float silu_tc(float x) { return x / (1.0f + __expf(-x)); }
// round-to-nearest (ties away) to tf32, as cvt.rna.tf32.f32 does for finite values, in two integer operations
__device__ __forceinline__ float rna_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

// B operand of one block: [K/4 chunks][2*nout rows][4 floats]; rows [0, nout) = tf32-hi of W^T, rows [nout, 2 nout) = lo.
__device__ __forceinline__ void stage_weights(const float* __restrict__ Wb, float* img, int k_dim, int nout, int transB, int tid) {
  const int total = k_dim * nout;
  for (int i = tid; i < total; i += TL_THREADS) {
    int n, k;
    if (transB) { n = i / k_dim; k = i % k_dim; } else { k = i / nout; n = i % nout; }
    const float x = __ldg(Wb + i);
    const float hi = rna_tf32(x);
    const int o = (k >> 2) * (2 * nout * 4) + n * 4 + (k & 3);
    img[o] = hi;
    img[o + nout * 4] = x - hi;
  }
}

// (n, m) of global row `row` of a block with `urows` rows per node, advanced in steps of 4 rows
struct RowIt {
  uint32_t n, m;
  __device__ __forceinline__ void init(uint32_t row, uint32_t urows) { n = row / urows; m = row - n * urows; }
  __device__ __forceinline__ void step4(uint32_t urows) {
    m += 4;
    while (m >= urows) { m -= urows; ++n; }
  }
};

__global__ void __launch_bounds__(TL_THREADS, 1)
k_tc_linear(const float* __restrict__ in, const float* __restrict__ W, const float* __restrict__ bias,
            const float* __restrict__ residual, float* __restrict__ out, int64_t N, const TcParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a_st = smem;                                               // [stages][2 terms][8 chunks (2064 B)][128 rows][4]
  uint8_t* w_sm = smem + (size_t)P.stages * TL_STAGE_BYTES;           // weight image(s)
  float* epi = reinterpret_cast<float*>(w_sm + P.w_bytes);            // [4 warps][32][TL_EPI_LD]
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_sm + P.w_bytes + TL_EPI_BYTES);
  uint64_t* a_full = bars;                          // [stages]  producers (128 arrivals) -> MMA
  uint64_t* a_empty = a_full + TL_MAX_STAGES;       // [stages]  MMA (commit) -> producers
  uint64_t* acc_full = a_empty + TL_MAX_STAGES;     // [2]       MMA -> epilogue
  uint64_t* acc_empty = acc_full + 2;               // [2]       epilogue (128 arrivals) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int S = P.stages;

  if (tid == 0) {
    for (int s = 0; s < TL_MAX_STAGES; ++s) { mbar_init(&a_full[s], 128); mbar_init(&a_empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 128); }
    mbar_fence_init();
  }
  if (warp == TL_MMA_WARP) tmem_alloc(tmem_slot, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  float* const w_img = reinterpret_cast<float*>(w_sm);
  if (P.resident) {
    for (int bi = 0; bi < P.n_blocks; ++bi)
      if (P.b[bi].tiles) stage_weights(W + P.b[bi].w_off, w_img + P.b[bi].w_sm, P.b[bi].k, P.b[bi].nout, P.transB, tid);
    fence_proxy_async_smem();
    __syncthreads();
  }

  // this CTA's share of the work units
  const int64_t u_lo = P.units * blockIdx.x / gridDim.x, u_hi = P.units * (blockIdx.x + 1) / gridDim.x;
  uint32_t chunk_it = 0, tile_it = 0;   // running counters (identical in every role) -> ring stage / phase
  const int r4 = lane >> 3, q = lane & 7;   // coalesced row mapping: 8 lanes per 128-byte row segment, 4 rows per instruction

  for (int bi = 0; bi < P.n_blocks; ++bi) {
    const TcBlock d = P.b[bi];
    if (d.tiles == 0) continue;
    int64_t t0 = (u_lo - d.unit0 + d.kc - 1) / d.kc, t1 = (u_hi - d.unit0 + d.kc - 1) / d.kc;
    if (u_lo <= d.unit0) t0 = 0;
    if (t0 < 0) t0 = 0;
    if (t1 > d.tiles) t1 = d.tiles;
    if (t0 >= t1) continue;   // uniform over the CTA
    if (!P.resident) {
      __syncthreads();        // previous block fully drained (epilogue consumed its last accumulator)
      stage_weights(W + d.w_off, w_img, d.k, d.nout, P.transB, tid);
      fence_proxy_async_smem();
      __syncthreads();
    }
    const uint32_t total_rows = (uint32_t)(N * d.rows), urows = (uint32_t)d.rows;
    const uint32_t n_tiles = (uint32_t)(t1 - t0), n_chunks = n_tiles * (uint32_t)d.kc;
    const bool cat = d.nout <= 128;   // [W_hi | W_lo] stacked along N

    if (warp < 4) {
      // ===================================================== epilogue
      const uint32_t lane_base = ((uint32_t)(warp * 32)) << 16;
      float* stg = epi + warp * (32 * TL_EPI_LD);
      for (uint32_t tl = 0; tl < n_tiles; ++tl) {
        const uint32_t it = tile_it + tl, a = it & 1;
        const uint32_t row0 = (uint32_t)(t0 + tl) * TL_ROWS + warp * 32;   // this warp's 32 rows
        for (int c0 = 0; c0 < d.nout; c0 += 32) {
          const int nq = d.nout - c0 < 32 ? (d.nout - c0) / 4 : 8;   // float4 columns of this 32-column piece that exist
          // residual for the coalesced mapping (row = row0 + 4 i + r4, float4 column q), issued before the accumulator wait
          float4 rr[8];
          if (residual != nullptr) {
            RowIt ri;
            ri.init(row0 + r4 < total_rows ? row0 + r4 : 0, urows);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const bool okr = row0 + 4 * i + r4 < total_rows && q < nq;
              rr[i] = okr ? __ldg(reinterpret_cast<const float4*>(residual + (int64_t)ri.n * P.out_stride + d.out_off + ri.m * d.nout + c0) + q)
                          : make_float4(0.f, 0.f, 0.f, 0.f);
              ri.step4(urows);
            }
          }
          if (c0 == 0) {
            tl_wait<32>(&acc_full[a], (it >> 1) & 1);
            tc_fence_after_sync();
          }
          uint32_t v[32];
          tmem_ld_32x32b_x32(tmem_base + lane_base + a * TL_ACC_COLS + c0, v);
          tmem_ld_wait();
          if (cat) {
            uint32_t w2[32];
            tmem_ld_32x32b_x32(tmem_base + lane_base + a * TL_ACC_COLS + d.nout + c0, w2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w2[j]));
          }
          if (c0 + 32 >= d.nout) {   // last piece read: the accumulator stage may be overwritten
            tc_fence_before_sync();
            mbar_arrive(&acc_empty[a]);
          }
          // transpose through the warp's staging tile: thread = row  ->  8 lanes per row
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(stg + lane * TL_EPI_LD + 4 * j) =
                make_float4(__uint_as_float(v[4 * j]) * d.scale, __uint_as_float(v[4 * j + 1]) * d.scale,
                            __uint_as_float(v[4 * j + 2]) * d.scale, __uint_as_float(v[4 * j + 3]) * d.scale);
          __syncwarp();
          float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
          if (bias != nullptr && bi == 0 && q < nq) bb = __ldg(reinterpret_cast<const float4*>(bias + c0) + q);
          RowIt ro;
          ro.init(row0 + r4 < total_rows ? row0 + r4 : 0, urows);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (row0 + 4 * i + r4 < total_rows && q < nq) {
              float4 y = *reinterpret_cast<const float4*>(stg + (4 * i + r4) * TL_EPI_LD + 4 * q);
              y.x += bb.x; y.y += bb.y; y.z += bb.z; y.w += bb.w;
              if (residual != nullptr) { y.x += rr[i].x; y.y += rr[i].y; y.z += rr[i].z; y.w += rr[i].w; }
              *(reinterpret_cast<float4*>(out + (int64_t)ro.n * P.out_stride + d.out_off + ro.m * d.nout + c0) + q) = y;
            }
            ro.step4(urows);
          }
          __syncwarp();
        }
      }
    } else if (warp < 4 + 4 * TL_GROUPS) {
      // ===================================================== producers: group g takes chunks it = g (mod G)
      // Only min(TL_GROUPS, S) groups run: a group may start the wait for a stage's next use only when that stage's
      // previous use is at most one phase behind (mbarrier waits are by parity), i.e. active groups <= stages.
      const int g = (warp - 4) >> 2, wq = (tid >> 5) & 3;
      const int G = TL_GROUPS < S ? TL_GROUPS : S;
      uint32_t c = g < G ? (uint32_t)((g + G - (int)(chunk_it % G)) % G) : n_chunks;
      uint32_t tl = c / d.kc, kc = c % d.kc;
      for (; c < n_chunks; c += G) {
        const uint32_t it = chunk_it + c;
        const uint32_t s = it % S;
        const uint32_t row0 = (uint32_t)(t0 + tl) * TL_ROWS + wq * 32;   // this warp's 32 rows; lane: rows row0 + 4 i + r4
        float4 x[8];
        RowIt ri;
        ri.init(row0 + r4 < total_rows ? row0 + r4 : 0, urows);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const bool ok = row0 + 4 * i + r4 < total_rows;
          const float* nb = in + (int64_t)ri.n * P.in_stride;
          x[i] = ok ? __ldg(reinterpret_cast<const float4*>(nb + d.in_off + ri.m * d.in_ld + kc * TL_KC) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
          if (P.gate) {  // e3nn Gate on the fly: scalars -> c*silu, gated rows x c*silu(gate)
            if (bi == 0) {
              x[i].x = P.act_const * silu_tc(x[i].x); x[i].y = P.act_const * silu_tc(x[i].y);
              x[i].z = P.act_const * silu_tc(x[i].z); x[i].w = P.act_const * silu_tc(x[i].w);
            } else {
              const float4 gq = ok ? __ldg(reinterpret_cast<const float4*>(nb + P.b[0].in_off + P.gate_off + d.k * (bi - 1) + kc * TL_KC) + q)
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
              x[i].x *= P.act_const * silu_tc(gq.x); x[i].y *= P.act_const * silu_tc(gq.y);
              x[i].z *= P.act_const * silu_tc(gq.z); x[i].w *= P.act_const * silu_tc(gq.w);
            }
          }
          ri.step4(urows);
        }
        tl_wait<200>(&a_empty[s], ((it / S) & 1) ^ 1);
        uint8_t* hi_dst = a_st + (size_t)s * TL_STAGE_BYTES + q * TL_A_LBO + (wq * 32 + r4) * 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 v = x[i];
          float4 hi, lo;
          hi.x = rna_tf32(v.x); hi.y = rna_tf32(v.y); hi.z = rna_tf32(v.z); hi.w = rna_tf32(v.w);
          lo.x = v.x - hi.x; lo.y = v.y - hi.y; lo.z = v.z - hi.z; lo.w = v.w - hi.w;
          *reinterpret_cast<float4*>(hi_dst + i * 64) = hi;
          *reinterpret_cast<float4*>(hi_dst + i * 64 + TL_TERM_BYTES) = lo;
        }
        fence_proxy_async_smem();
        mbar_arrive(&a_full[s]);
        kc += G;   // advance (tile, K-chunk) by G chunks
        while (kc >= (uint32_t)d.kc) { kc -= d.kc; ++tl; }
      }
    } else if (lane == 0) {
      // ===================================================== MMA issuer
      // Descriptor low word = (addr >> 4) | (LBO >> 4) << 16, high word = (SBO >> 4) | version 1 << 14 (constant): only
      // the address field of the low word moves, in 32-bit arithmetic.
      const uint32_t idesc_n = make_idesc_tf32(TL_ROWS, d.nout), idesc_2n = make_idesc_tf32(TL_ROWS, 2 * d.nout);
      const uint32_t b_lbo = (uint32_t)d.nout * 2 * 16;
      const uint32_t desc_hi = (128u >> 4) | (1u << 14);
      const uint32_t a_lo0 = ((smem_u32(a_st) >> 4) & 0x3FFFu) | ((TL_A_LBO >> 4) << 16);
      const uint32_t b_lo0 = (((smem_u32(w_sm) + (P.resident ? (uint32_t)d.w_sm * 4u : 0u)) >> 4) & 0x3FFFu) | ((b_lbo >> 4) << 16);
      const uint32_t lo_a = TL_TERM_BYTES >> 4, lo_b = ((uint32_t)d.nout * 16) >> 4;
      const uint32_t step_a = (2 * TL_A_LBO) >> 4, step_b = (2 * b_lbo) >> 4;   // one K-step (8 tf32) = 2 K-chunks of 4
      auto desc = [desc_hi](uint32_t lo) {
        uint64_t dd;
        asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(desc_hi));
        return dd;
      };
      uint32_t c = 0;
      for (uint32_t tl = 0; tl < n_tiles; ++tl) {
        const uint32_t it = tile_it + tl, a = it & 1;
        tl_wait<0>(&acc_empty[a], ((it >> 1) & 1) ^ 1);
        tc_fence_after_sync();
        const uint32_t dcol = tmem_base + a * TL_ACC_COLS;
        uint32_t bd = b_lo0;
        for (int kc = 0; kc < d.kc; ++kc, ++c) {
          const uint32_t cit = chunk_it + c, s = cit % S;
          tl_wait<0>(&a_full[s], (cit / S) & 1);
          tc_fence_after_sync();
          uint32_t ad = a_lo0 + s * (TL_STAGE_BYTES >> 4);
#pragma unroll
          for (int j = 0; j < TL_KC / 8; ++j) {
            if (cat) {
              mma_tf32(dcol, desc(ad), desc(bd), idesc_2n, (kc | j) ? 1u : 0u);   // A_hi x [W_hi | W_lo]
              mma_tf32(dcol, desc(ad + lo_a), desc(bd), idesc_n, 1u);             // A_lo x W_hi -> first half
            } else {
              mma_tf32(dcol, desc(ad + lo_a), desc(bd), idesc_n, (kc | j) ? 1u : 0u);
              mma_tf32(dcol, desc(ad), desc(bd + lo_b), idesc_n, 1u);
              mma_tf32(dcol, desc(ad), desc(bd), idesc_n, 1u);
            }
            ad += step_a;
            bd += step_b;
          }
          mma_commit(&a_empty[s]);
        }
        mma_commit(&acc_full[a]);
      }
    }
    chunk_it += n_chunks;
    tile_it += n_tiles;
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (warp == TL_MMA_WARP) tmem_dealloc(tmem_base, 512);
}

}  // namespace

// Returns 0 when launched, -1 when the shape is outside the tensor-core path (caller falls back to the CUDA-core GEMM).
int eeg_tc_linear_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                         const float* bias, int transB, const float* residual, float* out, const EegLayout* lout, int64_t N,
                         int gate, int gate_off, float act_const, const int32_t* in_ld, const int32_t* k_override, cudaStream_t stream) {
  TcParams P;
  P.n_blocks = lin->n_blocks;
  P.in_stride = lin->stride;
  P.out_stride = lout->stride;
  P.transB = transB;
  P.gate = gate;
  P.gate_off = gate_off;
  P.act_const = act_const;
  int64_t units = 0;
  int max_kn = 0, sum_kn = 0;
  for (int b = 0; b < P.n_blocks; ++b) {
    TcBlock& d = P.b[b];
    d.rows = lin->rows[b];
    d.k = k_override ? k_override[b] : lin->mul[b];
    d.in_ld = in_ld ? in_ld[b] : lin->mul[b];
    d.nout = lout->mul[b];
    d.in_off = lin->off[b];
    d.out_off = lout->off[b];
    d.w_off = host_w_off[b];
    d.scale = host_scale[b];
    d.unit0 = units;
    d.w_sm = 0;
    if (d.k == 0 || d.nout == 0) {
      d.tiles = 0;
      d.kc = 1;
      continue;
    }
    if (d.k % TL_KC != 0 || d.nout % 16 != 0 || d.nout > 256 || d.in_ld % 4 != 0 || d.in_off % 4 != 0 || d.out_off % 4 != 0 ||
        d.w_off % 4 != 0 || lin->stride % 4 != 0 || lout->stride % 4 != 0)
      return -1;
    d.kc = d.k / TL_KC;
    const int64_t tiles = (N * d.rows + TL_ROWS - 1) / TL_ROWS;
    if (tiles >= (1 << 24)) return -1;   // 32-bit row arithmetic in the kernel
    d.tiles = (int)tiles;
    units += tiles * d.kc;
    if (d.k * d.nout > max_kn) max_kn = d.k * d.nout;
    d.w_sm = 2 * sum_kn;
    sum_kn += d.k * d.nout;
  }
  if (units == 0) return 0;
  P.units = units;
  if (N * (int64_t)lin->stride >= (1LL << 31) || N * (int64_t)lout->stride >= (1LL << 31)) return -1;
  const uint32_t fixed = TL_EPI_BYTES + 1024;
  P.resident = 2 * 4 * sum_kn + fixed + TL_MAX_STAGES * TL_STAGE_BYTES <= TL_SMEM_BUDGET;   // all blocks + 4 A stages fit
  P.w_bytes = (uint32_t)((2 * 4 * (P.resident ? sum_kn : max_kn) + 1023) / 1024 * 1024);
  if (P.w_bytes + fixed + 2 * TL_STAGE_BYTES > TL_SMEM_BUDGET) return -1;
  const int stages = (int)((TL_SMEM_BUDGET - fixed - P.w_bytes) / TL_STAGE_BYTES);
  P.stages = stages > TL_MAX_STAGES ? TL_MAX_STAGES : stages;
  static const int cap = getenv("EEG_TC_STAGES") ? atoi(getenv("EEG_TC_STAGES")) : TL_MAX_STAGES;  // debugging knob
  if (P.stages > cap && cap >= 1) P.stages = cap;
  const size_t smem = (size_t)P.stages * TL_STAGE_BYTES + P.w_bytes + fixed;
  static EegDevFlags attr_set_dev; bool& attr_set = eeg_dev_flag(attr_set_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_tc_linear, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL_SMEM_BUDGET);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_block_linear_fwd (tcgen05): cannot opt in to %u B shared memory: %s", TL_SMEM_BUDGET, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const unsigned grid = (unsigned)(units < EEG_NUM_SMS ? units : EEG_NUM_SMS);
  k_tc_linear<<<grid, TL_THREADS, smem, stream>>>(in, W, bias, residual, out, N, P);
  EEG_CHECK_LAUNCH("eeg_block_linear_fwd (tcgen05)");
  return 0;
}
