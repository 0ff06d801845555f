// Weight gradients of the wide linears of the backward pass on the 5th-gen tensor cores:
//   dW_b[u, w] = scale_b * sum over rows r = (node, m) of block b of  x[r, u] * dy[r, w]
// for the shapes where the reduction is long and the weight matrix is wide: the last radial-MLP layer
// (gnn/blocks.py:538-549: x = dw [E x 1344], dy = act [E x 64]) and the path-stacked interaction.linear
// (gnn/blocks.py:552-559: x = M [N x (2l+1) x 160..320], dy = dmsg [N x (2l+1) x 32]).  The CUDA-core kernel
// (eeg_bwd.cu k_linear_wgrad) ran these at 22 TFLOP/s; here they are bound by reading x once from HBM.
//
// GEMM orientation: D[u (M = 128 per tile), w (N = nout)] += A[u, r] * B[w, r], K = rows.  Both operands are stored
// row-major in HBM with the REDUCTION index r as the slow index, i.e. "M-major"; the producers transpose while they
// stage: thread = one column u, four consecutive rows -> four conflict-free 4-byte shared-memory loads from the raw rows,
// 3xTF32 split (eeg_umma.cuh), one 16-byte store per term into the canonical K-major UMMA layout (lanes hit consecutive
// 16-byte slots: conflict free).
//
// Work item = (irrep block, group of <= 4 weight-row tiles, row slice); one CTA per item, ~148 items sized by cost.  The
// accumulators of a group stay in TMEM ([hi | lo] halves: A_hi x [B_hi | B_lo] + A_lo x B_hi, two MMAs per tile and K-step)
// for one PASS of 512 rows and are then drained into the workspace; k_wu_reduce adds the passes in fixed order (bitwise
// repeatable), applies the scale and accumulates into dW.  Why passes: the tensor core's fp32 accumulation truncates, so an
// accumulator that lives for n K-steps comes out ~(2/3) n 2^-24 too small (measured: 6.5e-6 after 142 steps, 1.7e-5 after
// 308); 64 steps per pass bound that at 2.5e-6 for any batch size, for 13 % more HBM traffic (partials out and back in).
//   warps 0-15 : cp.async (16 bytes per thread and copy, zero fill past the ends) of the stage 5 ahead into a raw row-major
//                ring (90 KB in flight per SM); transform the current stage's raw rows -> K-major hi / lo operand tiles
//                (4 LDS.32, 8 ALU, 2 STS.128 per task); drain the accumulators at pass ends
//   warp 16    : one lane issues tcgen05.mma kind::tf32 and commits operand stages back to the transform warps
// Measured alternatives (radial-MLP shape, 512 lattices; the CUDA-core kernel takes 0.475 ms):
//   per-thread LDG.32 straight into registers: 2.6 warp instructions per element of address arithmetic, 0.23 ms;
//   one cp.async.bulk per row (2 KB / 256 B) from a loader warp: ~110 clk per bulk copy whatever its size, 0.19-0.22 ms.
#include <stdlib.h>
#include <string.h>
#include "eeg_common.cuh"
#include "eeg_umma.cuh"
#include "eeg_tma.cuh"

namespace {
using namespace eeg_umma;

constexpr int WU_ROWS = 8;                        // reduction rows per stage (one K = 8 MMA step)
constexpr int WU_CH = WU_ROWS / 4;                // 4-row K-chunks per stage
constexpr int WU_RS = 6, WU_OS = 3;               // raw-row stages (5 x 18 KB of cp.async in flight), operand stages
constexpr int WU_PASS_ROWS = 512;                 // most rows accumulated in TMEM before the accumulators are drained (see above)
constexpr int WU_MAXT = 4;                        // 128-row weight tiles per work item
constexpr int WU_MAXN = 64;                       // output columns (nout = 32 or 64)
constexpr int WU_XFORM = 512;                     // transform threads (16 warps)
constexpr int WU_MMA_WARP = WU_XFORM / 32, WU_TMA_WARP = WU_MMA_WARP + 1;
constexpr int WU_THREADS = WU_XFORM + 64;
constexpr int WU_BOXC = 256;                      // columns of one TMA box of x (the hardware limit per box dimension)
constexpr int WU_MAXG = 16;                       // (block, tile group) pairs per launch
constexpr uint32_t WU_LBO = 128 * 16;             // bytes between 4-row K-chunks of an A tile
constexpr uint32_t WU_TERM = WU_CH * WU_LBO;      // 4096 B: one term (hi or lo) of one tile
constexpr uint32_t WU_TILE = 2 * WU_TERM;         // 8192 B
constexpr uint32_t WU_A_BYTES = WU_MAXT * WU_TILE;     // 32768 B
constexpr uint32_t WU_B_BYTES = WU_CH * 2 * WU_MAXN * 16;   // 4096 B: [chunks][hi rows | lo rows][16 B]
constexpr uint32_t WU_OP = WU_A_BYTES + WU_B_BYTES;    // 36864 B: one operand stage
constexpr int WU_TPT = WU_CH * WU_MAXT * 128 / 512;    // A tasks per transform thread and stage (2)
constexpr int WU_CPT = WU_ROWS * WU_MAXT * 128 / 4 / 512;   // 16-byte copy chunks of x per thread and stage (2)
constexpr int WU_XLD = WU_MAXT * 128;             // floats per raw x row
constexpr uint32_t WU_RAWX = WU_ROWS * WU_XLD * 4;     // 16384 B
constexpr uint32_t WU_RAW = WU_RAWX + WU_ROWS * WU_MAXN * 4;   // 18432 B: one raw stage (x rows, then dy rows)
constexpr uint32_t WU_BAR_OFF = WU_OS * WU_OP + WU_RS * WU_RAW;   // multiple of 1024: raw stages stay 128-byte aligned for TMA
constexpr uint32_t WU_SMEM = WU_BAR_OFF + 256;

struct WuGroup {
  int rows, nout, in_off, in_ld, out_off;   // block geometry: rows = 2l+1, x block at in_off (+ m * in_ld), dy at out_off (+ m * nout)
  int k;                                     // weight rows (input multiplicity) of the block
  int col0, ncols, tiles;                    // this group's first weight row, number of valid rows, 128-row tiles
  int cta_begin, n_pass, pps, spp, elem_begin;   // first CTA, passes in total and per CTA, stages (8 rows) per pass, first element
                                             // of the group in k_wu_reduce's index space
  unsigned inv_rows;                         // ceil(2^32 / rows) for rows > 1
  int64_t total_rows, ws_off, w_off;
  float scale;
};
struct WuParams {
  WuGroup g[WU_MAXG];
  int n_groups, in_stride, out_stride, transB, prof;
};

template <int SLEEP_NS>
__device__ __forceinline__ void wu_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
    if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
    if (spins > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ void wu_cp16(void* dst, const void* src, uint32_t src_bytes) {   // src_bytes = 0: zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void wu_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void wu_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void wu_bar_xform() { asm volatile("bar.sync 1, %0;" ::"n"(WU_XFORM) : "memory"); }
__device__ __forceinline__ float wu_hi(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

// Stall accounting of CTA 0 (tools/lin_bwd_time.py --prof): cycles of transform warp 0 in [0] waiting for copies + the stage
// barrier, [1] waiting for a free operand stage, [2] transforming, [3] draining; of the MMA lane in [4] waiting for operands,
// [5] issuing, [6] waiting for the drain; [7] whole kernel.
__device__ unsigned long long wu_prof[8];

// TMA = true (one irrep block with one row per node: x and dy are plain 2-D row-major matrices, the radial-MLP case): a loader
// warp issues three tensor-map loads per stage (two [8 x 256] boxes of x, one [8 x nout] box of dy; rows and columns past the
// ends are zero-filled by the hardware) into the raw ring.  TMA = false: the transform warps issue per-thread cp.async instead.
template <bool TMA>
__global__ void __launch_bounds__(WU_THREADS, 1)
k_wgrad_umma(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ ws, const WuParams P,
             const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_y) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* const raw = smem + WU_OS * WU_OP;
  uint64_t* op_full = reinterpret_cast<uint64_t*>(smem + WU_BAR_OFF);    // [OS] transform warps (16 arrivals) -> MMA
  uint64_t* op_empty = op_full + WU_OS;                                  // [OS] MMA commit -> transform warps
  uint64_t* acc_full = op_empty + WU_OS;                                 // MMA commit at the end of a pass -> drain
  uint64_t* acc_empty = acc_full + 1;                                    // drained (16 arrivals) -> MMA
  uint64_t* raw_full = acc_empty + 1;                                    // [RS] TMA bytes -> transform warps      (TMA only)
  uint64_t* raw_empty = raw_full + WU_RS;                                // [RS] transform warps (16) -> loader    (TMA only)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(raw_empty + WU_RS);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool prof = P.prof && blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == WU_MMA_WARP);
  long long pc[4] = {0, 0, 0, 0}, t_start = prof ? clock64() : 0, t0 = t_start;
  auto tick = [&](int k) {
    if (prof) {
      const long long t1 = clock64();
      pc[k] += t1 - t0;
      t0 = t1;
    }
  };

  int gi = 0;
#pragma unroll
  for (int q = 1; q < WU_MAXG; ++q)
    if (q < P.n_groups && (int)blockIdx.x >= P.g[q].cta_begin) gi = q;
  const WuGroup d = P.g[gi];
  const int slice = blockIdx.x - d.cta_begin;
  const int64_t cta_rows = (int64_t)d.pps * d.spp * WU_ROWS, r_begin = slice * cta_rows;
  const int64_t r_end = r_begin + cta_rows < d.total_rows ? r_begin + cta_rows : d.total_rows;
  const int n_stages = (int)((r_end - r_begin + WU_ROWS - 1) / WU_ROWS);
  const int width = d.tiles * 128;

  if (tid == 0) {
    for (int s = 0; s < WU_OS; ++s) { mbar_init(&op_full[s], WU_XFORM / 32); mbar_init(&op_empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, WU_XFORM / 32);
    for (int s = 0; s < WU_RS; ++s) { mbar_init(&raw_full[s], 1); mbar_init(&raw_empty[s], WU_XFORM / 32); }
    mbar_fence_init();
  }
  if (warp == WU_MMA_WARP) tmem_alloc(tmem_slot, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < WU_XFORM / 32) {
    // ============================================================ loader + transform warps
    // Copy chunks of a stage (16 bytes each): x rows -> id = tid + 512 i over [row][width / 4]; dy rows -> threads < rows * nout / 4.
    // A tasks: (chunk q of 4 rows, weight row c), c < width -> idx = q * width + c = tid + 512 i.  B tasks: (q, column n).
    const int w4 = width >> 2, n4 = d.nout >> 2;
    int crow[WU_CPT], ccol[WU_CPT];
#pragma unroll
    for (int i = 0; i < WU_CPT; ++i) {
      const int id = tid + WU_XFORM * i;
      crow[i] = id / w4;                    // >= WU_ROWS: no such chunk (narrow groups)
      ccol[i] = (id - crow[i] * w4) * 4;
    }
    const bool has_y = tid < WU_ROWS * n4;
    const int yrow = has_y ? tid / n4 : 0, ycol = has_y ? (tid - yrow * n4) * 4 : 0;
    auto issue = [&](int st) {   // cp.async of stage st into raw slot st % RS; rows / columns that do not exist are zero-filled
      if (st < n_stages) {
        uint8_t* slot = raw + (size_t)(st % WU_RS) * WU_RAW;
        const int64_t r0 = r_begin + (int64_t)st * WU_ROWS;
#pragma unroll
        for (int i = 0; i < WU_CPT; ++i) {
          if (crow[i] < WU_ROWS) {
            const int64_t r = r0 + crow[i];
            const bool ok = r < r_end && ccol[i] < d.ncols;
            const uint32_t r32 = ok ? (uint32_t)r : (uint32_t)r_begin;
            const uint32_t n = d.rows == 1 ? r32 : __umulhi(r32, d.inv_rows), m = r32 - n * (uint32_t)d.rows;
            const float* src = x + (int64_t)n * P.in_stride + d.in_off + (int64_t)m * d.in_ld + d.col0 + (ok ? ccol[i] : 0);
            wu_cp16(slot + (crow[i] * WU_XLD + ccol[i]) * 4, src, ok ? 16u : 0u);
          }
        }
        if (has_y) {
          const int64_t r = r0 + yrow;
          const bool ok = r < r_end;
          const uint32_t r32 = ok ? (uint32_t)r : (uint32_t)r_begin;
          const uint32_t n = d.rows == 1 ? r32 : __umulhi(r32, d.inv_rows), m = r32 - n * (uint32_t)d.rows;
          wu_cp16(slot + WU_RAWX + (yrow * WU_MAXN + ycol) * 4, dy + (int64_t)n * P.out_stride + d.out_off + (int64_t)m * d.nout + ycol,
                  ok ? 16u : 0u);
        }
      }
      wu_cp_commit();   // one group per stage, also when nothing was issued: the wait below counts groups
    };
    const bool has_b = tid < WU_CH * d.nout;
    const int bq = has_b ? tid / d.nout : 0, bn = has_b ? tid - bq * d.nout : 0;
    int aq[WU_TPT], ac[WU_TPT];
#pragma unroll
    for (int i = 0; i < WU_TPT; ++i) {
      const int idx = tid + WU_XFORM * i;
      aq[i] = idx / width;
      ac[i] = idx - aq[i] * width;
    }
    const int quarter = warp & 3, sub = warp >> 2, pieces = d.nout / 16;
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    if (!TMA) {
#pragma unroll 1
      for (int st = 0; st < WU_RS - 1; ++st) issue(st);
    }
    tick(2);
    for (int st = 0; st < n_stages; ++st) {
      const int os = st % WU_OS;
      const float* rx = reinterpret_cast<const float*>(raw + (size_t)(st % WU_RS) * WU_RAW);
      const float* ry = rx + WU_ROWS * WU_XLD;
      uint8_t* a_st = smem + (size_t)os * WU_OP;
      uint8_t* b_st = a_st + WU_A_BYTES;
      if (TMA) {
        wu_wait<20>(&raw_full[st % WU_RS], (st / WU_RS) & 1);
        tick(0);
      } else {
        wu_cp_wait<WU_RS - 2>();   // this thread's chunks of stage st have landed
        wu_bar_xform();            // ... everyone's have, and everyone has finished reading stage st - 1
        tick(0);
        issue(st + WU_RS - 1);     // refill the slot of stage st - 1
        tick(2);
      }
      wu_wait<20>(&op_empty[os], ((st / WU_OS) & 1) ^ 1);
      tick(1);
#pragma unroll
      for (int i = 0; i < WU_TPT; ++i) {
        if (aq[i] < WU_CH) {
          const int q = aq[i], c = ac[i];
          float v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)   // TMA: [box][row][256 columns]; cp.async: [row][512 columns]
            v[j] = TMA ? rx[(c >> 8) * (WU_ROWS * WU_BOXC) + (4 * q + j) * WU_BOXC + (c & (WU_BOXC - 1))] : rx[(4 * q + j) * WU_XLD + c];
          float4 hi, lo;
          hi.x = wu_hi(v[0]); hi.y = wu_hi(v[1]); hi.z = wu_hi(v[2]); hi.w = wu_hi(v[3]);
          lo.x = v[0] - hi.x; lo.y = v[1] - hi.y; lo.z = v[2] - hi.z; lo.w = v[3] - hi.w;
          uint8_t* dst = a_st + (c >> 7) * WU_TILE + q * WU_LBO + (c & 127) * 16;
          *reinterpret_cast<float4*>(dst) = hi;
          *reinterpret_cast<float4*>(dst + WU_TERM) = lo;
        }
      }
      if (has_b) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = ry[(4 * bq + j) * WU_MAXN + bn];
        float4 hi, lo;
        hi.x = wu_hi(v[0]); hi.y = wu_hi(v[1]); hi.z = wu_hi(v[2]); hi.w = wu_hi(v[3]);
        lo.x = v[0] - hi.x; lo.y = v[1] - hi.y; lo.z = v[2] - hi.z; lo.w = v[3] - hi.w;
        uint8_t* dst = b_st + bq * (2 * d.nout * 16) + bn * 16;
        *reinterpret_cast<float4*>(dst) = hi;
        *reinterpret_cast<float4*>(dst + d.nout * 16) = lo;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&op_full[os]);
        if (TMA) mbar_arrive(&raw_empty[st % WU_RS]);   // this warp's raw values are in registers / operand tiles
      }
      tick(2);
      if ((st + 1) % d.spp == 0 || st + 1 == n_stages) {
        // ---- end of a pass: drain.  TMEM lane = weight row inside the tile; warp w reads lanes 32 (w % 4) ..., the 4 warps of a
        // quarter share the (tile, 16-column piece) pairs.  Partial layout: [w][tile * 128 + row] (lanes write consecutive floats).
        const int pass = st / d.spp;
        wu_wait<32>(acc_full, pass & 1);
        tc_fence_after_sync();
        float* wsi = ws + d.ws_off + ((int64_t)slice * d.pps + pass) * width * d.nout;
        for (int p = sub; p < d.tiles * pieces; p += 4) {
          const int t = p / pieces, c0 = (p - t * pieces) * 16;
          uint32_t vh[16], vl[16];
          tmem_ld_32x32b_x16(tmem_base + lane_base + t * 2 * d.nout + c0, vh);
          tmem_ld_32x32b_x16(tmem_base + lane_base + t * 2 * d.nout + d.nout + c0, vl);
          tmem_ld_wait();
          const int row = t * 128 + quarter * 32 + lane;
#pragma unroll
          for (int j = 0; j < 16; ++j) wsi[(int64_t)(c0 + j) * width + row] = __uint_as_float(vh[j]) + __uint_as_float(vl[j]);
        }
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty);
        tick(3);
      }
    }
    wu_cp_wait<0>();
    if (prof)
      for (int k = 0; k < 4; ++k) atomicAdd(&wu_prof[k], (unsigned long long)pc[k]);
  } else if (warp == WU_MMA_WARP && lane == 0) {
    {
      // ============================================================ MMA issuer
      const uint32_t idesc_n = make_idesc_tf32(128, d.nout), idesc_2n = make_idesc_tf32(128, 2 * d.nout);
      const uint32_t b_lbo = (uint32_t)d.nout * 2 * 16;
      for (int st = 0; st < n_stages; ++st) {
        const int s = st % WU_OS, sp = st % d.spp;
        if (sp == 0 && st > 0) {   // the previous pass has been drained
          wu_wait<0>(acc_empty, ((st / d.spp) - 1) & 1);
          tc_fence_after_sync();
          tick(2);
        }
        wu_wait<0>(&op_full[s], (st / WU_OS) & 1);
        tc_fence_after_sync();
        tick(0);
        const uint32_t a0 = smem_u32(smem + (size_t)s * WU_OP), b0 = a0 + WU_A_BYTES;
#pragma unroll
        for (int ks = 0; ks < WU_ROWS / 8; ++ks) {
          const uint64_t bd = make_desc_kmajor(b0 + ks * 2 * b_lbo, b_lbo, 128);
          for (int t = 0; t < d.tiles; ++t) {
            const uint32_t at = a0 + t * WU_TILE + ks * 2 * WU_LBO;
            const uint32_t dcol = tmem_base + t * 2 * d.nout;
            mma_tf32(dcol, make_desc_kmajor(at, WU_LBO, 128), bd, idesc_2n, (sp | ks) ? 1u : 0u);   // A_hi x [B_hi | B_lo]
            mma_tf32(dcol, make_desc_kmajor(at + WU_TERM, WU_LBO, 128), bd, idesc_n, 1u);         // A_lo x B_hi -> first half
          }
        }
        mma_commit(&op_empty[s]);
        if (sp == d.spp - 1 || st + 1 == n_stages) mma_commit(acc_full);
        tick(1);
      }
      if (prof) {
        for (int k = 0; k < 3; ++k) atomicAdd(&wu_prof[4 + k], (unsigned long long)pc[k]);
        atomicAdd(&wu_prof[7], (unsigned long long)(clock64() - t_start));
      }
    }
  }
  else if (TMA && warp == WU_TMA_WARP && lane == 0) {
    // ============================================================ TMA loader
    const int boxes = (d.ncols + WU_BOXC - 1) / WU_BOXC;   // 1 or 2
    const uint32_t bytes = (uint32_t)(boxes * WU_ROWS * WU_BOXC + WU_ROWS * d.nout) * 4u;
    for (int st = 0; st < n_stages; ++st) {
      const int rs = st % WU_RS;
      const int r0 = (int)(r_begin + (int64_t)st * WU_ROWS);
      uint8_t* slot = raw + (size_t)rs * WU_RAW;
      wu_wait<20>(&raw_empty[rs], ((st / WU_RS) & 1) ^ 1);
      mbar_arrive_expect_tx(&raw_full[rs], bytes);
      for (int b = 0; b < boxes; ++b) eeg_tma::load_2d(slot + b * (WU_ROWS * WU_BOXC * 4), &map_x, d.col0 + b * WU_BOXC, r0, &raw_full[rs]);
      eeg_tma::load_2d(slot + WU_RAWX, &map_y, 0, r0, &raw_full[rs]);
    }
  }
  __syncthreads();
  tc_fence_after_sync();
  if (warp == WU_MMA_WARP) tmem_dealloc(tmem_base, 512);
}

// dW[element] += scale * (sum over the passes of the element's group).  Block = 32 elements x 32 partial lanes: lane y adds the
// passes y, y + 32, ... (four independent chains), then the 32 lane sums are added in lane order: a fixed order, so the
// result is bitwise repeatable; a single thread per element would chain n_pass dependent loads (120 us at n_pass = 120).
__global__ void __launch_bounds__(1024) k_wu_reduce(const float* __restrict__ ws, float* __restrict__ dW, const WuParams P, int total) {
  __shared__ float part[32][33];
  const int i = blockIdx.x * 32 + threadIdx.x, y = threadIdx.y;
  int gi = 0;
#pragma unroll
  for (int q = 1; q < WU_MAXG; ++q)
    if (q < P.n_groups && i >= P.g[q].elem_begin) gi = q;
  const WuGroup& d = P.g[gi];
  const int e = i - d.elem_begin, n = e / d.ncols, u = e - n * d.ncols, width = d.tiles * 128;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (i < total) {
    const float* p = ws + d.ws_off + (int64_t)n * width + u;
    const int64_t stride = (int64_t)width * d.nout;
    int sl = y;
    for (; sl + 96 < d.n_pass; sl += 128) {
      s0 += p[sl * stride]; s1 += p[(sl + 32) * stride]; s2 += p[(sl + 64) * stride]; s3 += p[(sl + 96) * stride];
    }
    for (; sl < d.n_pass; sl += 32) s0 += p[sl * stride];
  }
  part[y][threadIdx.x] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (y == 0 && i < total) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 32; ++q) s += part[q][threadIdx.x];
    const int64_t dst = d.w_off + (P.transB ? (int64_t)n * d.k + d.col0 + u : (int64_t)(d.col0 + u) * d.nout + n);
    dW[dst] += s * d.scale;
  }
}

// Plans the work items.  Returns the workspace floats needed, or -1 when the shape is outside this kernel: every non-empty
// block must have nout = 64 and >= 128 weight rows, and the problem must be large enough to fill the machine.
int64_t wu_plan(const EegLayout* lin, const EegLayout* lout, const int64_t* w_off, const float* scale, int64_t N, WuParams* P,
                int* total_ctas, int* total_elems) {
  struct G { int b, tiles, col0, ncols; double cost; int slices; };
  G g[WU_MAXG];
  int ng = 0;
  double total_cost = 0.0, volume = 0.0;
  for (int b = 0; b < lin->n_blocks; ++b) {
    const int k = lin->mul[b], nout = lout->mul[b];
    if (k == 0 || nout == 0) continue;
    const int64_t R = N * lin->rows[b];
    // nout = 32 (the path-stacked interaction.linear, 2-3 tiles per group) measured 0.39 ms here against 0.37 ms on CUDA cores
    if (nout != 64 || k < 128 || lin->rows[b] != lout->rows[b] || R >= (1LL << 28) || R < 1) return -1;
    // bulk copies: 16-byte aligned row segments
    if (k % 4 || lin->stride % 4 || lout->stride % 4 || lin->off[b] % 4 || lout->off[b] % 4) return -1;
    const int T = (k + 127) / 128, parts = (T + WU_MAXT - 1) / WU_MAXT;
    int col0 = 0;
    for (int p = 0; p < parts; ++p) {
      if (ng >= WU_MAXG) return -1;
      const int tiles = T / parts + (p < T % parts ? 1 : 0);
      g[ng].b = b;
      g[ng].tiles = tiles;
      g[ng].col0 = col0;
      g[ng].ncols = k - col0 < tiles * 128 ? k - col0 : tiles * 128;
      g[ng].cost = (double)R * (tiles + 0.5);
      total_cost += g[ng].cost;
      col0 += tiles * 128;
      ++ng;
    }
    volume += (double)R * k;
  }
  if (ng == 0 || volume < (double)(1 << 22)) return -1;
  // row slices in proportion to cost, EEG_NUM_SMS items in total (one wave of one CTA per SM)
  int used = 0;
  for (int i = 0; i < ng; ++i) {
    g[i].slices = (int)(EEG_NUM_SMS * g[i].cost / total_cost);
    if (g[i].slices < 1) g[i].slices = 1;
    used += g[i].slices;
  }
  while (used < EEG_NUM_SMS) {
    int best = 0;
    for (int i = 1; i < ng; ++i)
      if (g[i].cost / g[i].slices > g[best].cost / g[best].slices) best = i;
    ++g[best].slices;
    ++used;
  }
  int64_t ws_floats = 0;
  int ctas = 0, elems = 0;
  for (int i = 0; i < ng; ++i) {
    const int b = g[i].b;
    const int64_t R = N * lin->rows[b];
    // rows per CTA -> passes of equal length (<= WU_PASS_ROWS rows, whole stages)
    const int64_t want = (R + g[i].slices - 1) / g[i].slices;
    const int64_t pps = (want + WU_PASS_ROWS - 1) / WU_PASS_ROWS;
    const int64_t spp = ((want + pps - 1) / pps + WU_ROWS - 1) / WU_ROWS;
    const int64_t pass_rows = spp * WU_ROWS;
    const int64_t n_pass = (R + pass_rows - 1) / pass_rows;
    const int n_slices = (int)((n_pass + pps - 1) / pps);
    if (P) {
      WuGroup& d = P->g[i];
      d.rows = lin->rows[b];
      d.nout = lout->mul[b];
      d.in_off = lin->off[b];
      d.in_ld = lin->mul[b];
      d.out_off = lout->off[b];
      d.k = lin->mul[b];
      d.col0 = g[i].col0;
      d.ncols = g[i].ncols;
      d.tiles = g[i].tiles;
      d.cta_begin = ctas;
      d.n_pass = (int)n_pass;
      d.pps = (int)pps;
      d.spp = (int)spp;
      d.elem_begin = elems;
      d.inv_rows = d.rows > 1 ? (unsigned)((0x100000000ULL + d.rows - 1) / d.rows) : 0u;
      d.total_rows = R;
      d.ws_off = ws_floats;
      d.w_off = w_off ? w_off[b] : 0;
      d.scale = scale ? scale[b] : 1.0f;
    }
    ws_floats += n_pass * g[i].tiles * 128 * lout->mul[b];
    ctas += n_slices;
    elems += g[i].ncols * lout->mul[b];
  }
  if (P) {
    P->n_groups = ng;
    P->in_stride = lin->stride;
    P->out_stride = lout->stride;
  }
  if (total_ctas) *total_ctas = ctas;
  if (total_elems) *total_elems = elems;
  return ws_floats;
}

}  // namespace

// Workspace floats of the tensor-core weight gradient, -1 when the shape is outside it (eeg_bwd.cu sizes its workspace with this).
int64_t eeg_wgrad_umma_ws_floats(const EegLayout* lin, const EegLayout* lout, int64_t N) {
  return wu_plan(lin, lout, nullptr, nullptr, N, nullptr, nullptr, nullptr);
}

// 0 = launched (dW += the gradient; bias gradients are the caller's), -1 = shape outside this kernel or workspace too small
int eeg_wgrad_umma_launch(const float* x, const EegLayout* lin, const float* dy, const EegLayout* lout, const int64_t* host_w_off,
                          const float* host_scale, int transB, float* dW, int64_t N, float* ws, int64_t ws_floats, cudaStream_t stream) {
  WuParams P;
  int ctas = 0, elems = 0;
  const int64_t need = wu_plan(lin, lout, host_w_off, host_scale, N, &P, &ctas, &elems);
  if (need < 0 || ws == nullptr || ws_floats < need) return -1;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy)) & 15) return -1;
  P.transB = transB;
  static const bool prof_on = getenv("EEG_WGRAD_PROF") != nullptr;
  P.prof = prof_on;
  static EegDevFlags attr_dev; bool& attr_set = eeg_dev_flag(attr_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_wgrad_umma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WU_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_wgrad_umma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WU_SMEM);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_block_linear_wgrad (tcgen05): cannot opt in to %u B shared memory: %s", WU_SMEM, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  // one irrep block with one row per node: x and dy are 2-D matrices -> tensor-map (TMA) loads; EEG_WGRAD_TMA=0 forces cp.async
  static const bool tma_off = getenv("EEG_WGRAD_TMA") && atoi(getenv("EEG_WGRAD_TMA")) == 0;
  alignas(64) CUtensorMap map_x, map_y;
  memset(&map_x, 0, sizeof(map_x));
  memset(&map_y, 0, sizeof(map_y));
  int nb = 0, b0 = 0;
  for (int b = 0; b < lin->n_blocks; ++b)
    if (lin->mul[b] > 0 && lout->mul[b] > 0) { ++nb; b0 = b; }
  bool tma = !tma_off && nb == 1 && lin->rows[b0] == 1;
  if (tma)
    tma = eeg_tma::make_map_2d(&map_x, x + lin->off[b0], N, lin->mul[b0], lin->stride, WU_BOXC, WU_ROWS) &&
          eeg_tma::make_map_2d(&map_y, dy + lout->off[b0], N, lout->mul[b0], lout->stride, lout->mul[b0], WU_ROWS);
  if (tma) k_wgrad_umma<true><<<(unsigned)ctas, WU_THREADS, WU_SMEM, stream>>>(x, dy, ws, P, map_x, map_y);
  else k_wgrad_umma<false><<<(unsigned)ctas, WU_THREADS, WU_SMEM, stream>>>(x, dy, ws, P, map_x, map_y);
  EEG_CHECK_LAUNCH("eeg_block_linear_wgrad (tcgen05)");
  k_wu_reduce<<<(unsigned)((elems + 31) / 32), dim3(32, 32), 0, stream>>>(ws, dW, P, elems);
  EEG_CHECK_LAUNCH("eeg_block_linear_wgrad (tcgen05 reduce)");
  return 0;
}

// Reads and clears the stall counters of k_wgrad_umma (EEG_WGRAD_PROF=1): out[8], see wu_prof.
extern "C" int eeg_wgrad_umma_prof_read(unsigned long long* out) {
  cudaError_t e = cudaMemcpyFromSymbol(out, wu_prof, sizeof(unsigned long long) * 8);
  if (e != cudaSuccess) { eeg_set_error("eeg_wgrad_umma_prof_read: %s", cudaGetErrorString(e)); return 2; }
  unsigned long long z[8] = {};
  e = cudaMemcpyToSymbol(wu_prof, z, sizeof(z));
  if (e != cudaSuccess) { eeg_set_error("eeg_wgrad_umma_prof_read: %s", cudaGetErrorString(e)); return 2; }
  return 0;
}
