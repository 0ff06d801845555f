// Channelwise equivariant tensor-product message + scatter-sum on the 5th-gen tensor cores (C = 32 channels).
// Same contract as eeg_tp_scatter_fwd in eeg_tp.cu (which it replaces for the shapes below); replaces
// node_feats[sender] gather + e3nn o3.TensorProduct('uvu') + torch_scatter.scatter(sum)/agg_norm_const
// (gnn/blocks.py:591-597).
//
// Per edge e (sender s, receiver n) and path p = (l1, l2, l3) the message row is
//     m[e, (p,k), u] = w[e,p,u] * T[e, (p,k), u],     T[e, (p,k), u] = sum_i G_e[(p,k), i] * x[s, i, u],
//     G_e[(p,k), i] = sum_j C_p[i,j,k] Y_j(e)                     (sparse: 974 non-zeros of 230 x 25 for lmax = 4)
// T is a [230 x 25] x [25 x 32] product whose left operand depends only on the edge.  The CUDA-core kernel evaluates it
// with the lane = channel and G broadcast from shared memory (1 LDS.128 per 4 FMAs: issue / shared-memory bound, 24 % of
// the FMA pipe).  Here it is ONE small GEMM per edge on tcgen05: rows (p,k) are split into row tiles of <= 128 TMEM lanes
// by l1 (tools/gen_kernels.py: tpu_layout), each with K = 16 input components, N = 32 channels, 3xTF32 (eeg_umma.cuh).
//
// One persistent CTA per SM owns a contiguous, edge-balanced range of receiver nodes and streams its CSR edges through a
// ring of TU_SLOTS shared-memory / TMEM slots:
//   fill warps (8)  : A = G_e^T: each thread owns NIT fixed non-zeros (coefficients in registers), value = c0 Y[j0] + c1 Y[j1]
//                     (Y held one value per lane, fetched by shuffles), split hi/lo, scattered into the zero-initialised
//                     K-major tile; B = x[s]^T: coalesced rows of the sender, split hi/lo, K-major [32 x 16] tiles.
//   loader (1 warp) : per-path weight rows by cp.async.bulk (TMA engine), TU_WE consecutive edges per copy (w_t is path-major,
//                     so the rows of consecutive CSR slots are contiguous) into a ring of TU_WG group buffers: the stream
//                     from HBM runs 2-3 groups ahead of the compute (a 128-byte copy per (edge, path) was issue bound).
//   MMA (1 lane)    : 2 tcgen05.mma kind::tf32 per (tile, K-step of 8): A_hi [B_hi | B_lo] (N = 64) + A_lo B_hi (N = 32): the A tile,
//                     which dominates the operand traffic of these skinny products, is read twice instead of three times.
//   epilogue (8)    : thread = TMEM lane = message row (p,k): tcgen05.ld the 32 channels, multiply by w[e,p,:], accumulate
//                     over the receiver's edges in registers (CSR order: deterministic), and at the end of a node write
//                     the row scaled by 1/agg_norm_const through a per-warp transpose tile with full-line stores.
#include <stdlib.h>
#include "eeg_common.cuh"
#include "eeg_umma.cuh"
#include "generated/tpu_3_3.cuh"
#include "generated/tpu_4_4.cuh"

namespace {
using namespace eeg_umma;

constexpr int TU_C = 32;                       // channels (= N of the MMA = lanes of a feature row)
constexpr int TU_FILL_WARPS = EEG_TPU_FILL_THREADS / 32;
constexpr int TU_SLOTS = 3;
constexpr uint32_t TU_B_LBO = 2 * TU_C * 16;   // bytes between 4-float K-chunks of a [64 x KT] B tile (rows 0-31 x_hi, 32-63 x_lo)
constexpr int TU_WE = 4;                       // edges per weight copy (one bulk copy = TU_WE x 128 B of one path)
constexpr int TU_WG = 3;                       // weight group buffers in flight
constexpr int TU_W_PITCH = TU_WE * 128 + 16;   // bytes per path in a group buffer (+16: rows of different paths -> different banks)
constexpr int TU_STG_LD = 36;                  // floats per row of the flush transpose tile

template <int LIN, int LMAX>
struct TuInfo;
#define EEG_TU_INFO(LIN, LMAX)                                                                                       \
  template <>                                                                                                        \
  struct TuInfo<LIN, LMAX> {                                                                                         \
    static constexpr int NT = EEG_TPU_##LIN##_##LMAX##_NT, KT = EEG_TPU_##LIN##_##LMAX##_KT, NIT = EEG_TPU_##LIN##_##LMAX##_NIT, \
                         NPATHS = EEG_TPU_##LIN##_##LMAX##_NPATHS, ROWS = EEG_TPU_##LIN##_##LMAX##_ROWS;               \
    static __device__ __forceinline__ const int2* rows() { return eeg_tpu_##LIN##_##LMAX##_rows; }                    \
    static __device__ __forceinline__ const int4* gtab() { return eeg_tpu_##LIN##_##LMAX##_gtab; }                    \
    static __device__ __forceinline__ int xoff(int t) { return eeg_tpu_##LIN##_##LMAX##_xoff[t]; }                    \
    static __device__ __forceinline__ int xn(int t) { return eeg_tpu_##LIN##_##LMAX##_xn[t]; }                        \
  };
EEG_TU_INFO(3, 3) EEG_TU_INFO(4, 4)
#undef EEG_TU_INFO

template <int LIN, int LMAX>
struct TuCfg {
  using I = TuInfo<LIN, LMAX>;
  static constexpr int NT = I::NT, KT = I::KT, KCH = KT / 4;
  static constexpr int EPI_WARPS = 4 * NT;
  static constexpr int FILL_WARP0 = EPI_WARPS, MMA_WARP = EPI_WARPS + TU_FILL_WARPS, LOAD_WARP = MMA_WARP + 1;
  static constexpr int THREADS = (LOAD_WARP + 1) * 32;
  static constexpr uint32_t A_TILE = KCH * EEG_TPU_A_LBO, A_TERM = NT * A_TILE, A_BYTES = 2 * A_TERM;
  static constexpr uint32_t B_TILE = KCH * TU_B_LBO, B_BYTES = NT * B_TILE, B_LO = TU_C * 16;   // lo rows follow the hi rows
  static constexpr uint32_t W_BYTES = (I::NPATHS * TU_W_PITCH + 127) / 128 * 128;   // one weight group buffer
  static constexpr uint32_t SLOT_BYTES = A_BYTES + B_BYTES;
  static constexpr uint32_t STG_BYTES = EPI_WARPS * 32 * TU_STG_LD * 4;
  static constexpr uint32_t ROWTAB_BYTES = NT * 128 * 8;
  static constexpr uint32_t W_OFF = TU_SLOTS * SLOT_BYTES;
  static constexpr uint32_t STG_OFF = W_OFF + TU_WG * W_BYTES;
  static constexpr uint32_t BAR_OFF = STG_OFF + STG_BYTES + ROWTAB_BYTES;
  static constexpr uint32_t SMEM_BYTES = BAR_OFF + 256;
  static constexpr int D_COLS = NT * 2 * TU_C;             // TMEM columns per slot: per tile [hi.hi + lo.hi | hi.lo]
  static constexpr int TMEM_COLS = TU_SLOTS * D_COLS <= 128 ? 128 : (TU_SLOTS * D_COLS <= 256 ? 256 : 512);
  static constexpr int XCHUNKS = NT * KCH;                 // [tile][K-chunk of 4 components] pieces of the B operand
  static_assert(TU_SLOTS * D_COLS <= 512, "TMEM budget");
  static_assert(A_TERM % 16 == 0 && SLOT_BYTES % 128 == 0, "alignment");
  static_assert(XCHUNKS <= TU_FILL_WARPS, "one B chunk per fill warp");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

// Optional per-role stall accounting (EEG_TP_PROF=1; read back with eeg_tp_umma_prof_read): clock64 spent in each wait site
// of CTA b, by ONE lane per role.  Slots: 0 fill.wait_slot 1 fill.total 2 mma.wait_filled 3 mma.total 4 epi.wait_w
// 5 epi.wait_d 6 epi.flush 7 epi.total 8 load.wait_free 9 load.total 10 edges 11 nodes
__device__ unsigned long long eeg_tpu_prof[EEG_NUM_SMS * 16];
template <bool PROF>
struct TuTimer {
  long long t0 = 0;
  __device__ __forceinline__ void start() { if (PROF) t0 = clock64(); }
  __device__ __forceinline__ void stop(unsigned long long& acc) { if (PROF) acc += (unsigned long long)(clock64() - t0); }
};

template <int SLEEP_NS>
__device__ __forceinline__ void tu_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
    if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
    if (spins > (1u << 22)) __trap();   // protocol bug: launch error on the host instead of a hung GPU
  }
}
__device__ __forceinline__ float tu_rna(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

// first node whose CSR range starts at or after edge `target` (row_ptr is non-decreasing, row_ptr[N] = E)
__device__ __forceinline__ int64_t tu_lower_bound(const int32_t* __restrict__ row_ptr, int64_t N, int64_t target) {
  int64_t lo = 0, hi = N;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (__ldg(row_ptr + mid) < target) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

template <int LIN, int LMAX, bool PROF>
__global__ void __launch_bounds__(TuCfg<LIN, LMAX>::THREADS, 1)
k_tp_umma(const float* __restrict__ h, const float* __restrict__ Y, int ystride, const float* __restrict__ w_t, int64_t E,
          const int32_t* __restrict__ sender, const int32_t* __restrict__ row_ptr, float inv_norm, float* __restrict__ M, int64_t N) {
  using CF = TuCfg<LIN, LMAX>;
  using I = TuInfo<LIN, LMAX>;
  constexpr int NT = CF::NT, KCH = CF::KCH, S = TU_SLOTS;
  constexpr int DIN = (LIN + 1) * (LIN + 1);
  extern __shared__ __align__(1024) uint8_t tu_smem[];
  uint8_t* const slots = tu_smem;
  uint8_t* const wbufs = tu_smem + CF::W_OFF;
  float* const stg_all = reinterpret_cast<float*>(tu_smem + CF::STG_OFF);
  int2* const rowtab = reinterpret_cast<int2*>(tu_smem + CF::STG_OFF + CF::STG_BYTES);
  uint64_t* const bars = reinterpret_cast<uint64_t*>(tu_smem + CF::BAR_OFF);
  uint64_t* const filled = bars;              // [S] fill threads -> MMA
  uint64_t* const d_full = bars + S;          // [S] MMA commit -> epilogue
  uint64_t* const slot_free = bars + 2 * S;   // [S] epilogue -> fill
  uint64_t* const w_full = bars + 3 * S;      // [TU_WG] loader (expect_tx + bulk copies) -> epilogue
  uint64_t* const w_free = w_full + TU_WG;    // [TU_WG] epilogue -> loader
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(w_free + TU_WG);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned long long pw0 = 0, pw1 = 0, pw2 = 0, ptot = 0;   // PROF accumulators of this thread
  TuTimer<PROF> tw, tt;
  unsigned long long* const prof = eeg_tpu_prof + (blockIdx.x % EEG_NUM_SMS) * 16;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      // one arrival per WARP (lane 0 after __syncwarp): 256 per-thread arrivals on one barrier word serialise in the
      // shared-memory atomic unit (~32 clk per warp instruction) and cost ~250 clk per edge in the first version
      mbar_init(&filled[s], TU_FILL_WARPS);
      mbar_init(&d_full[s], 1);
      mbar_init(&slot_free[s], CF::EPI_WARPS);
    }
    for (int b = 0; b < TU_WG; ++b) {
      mbar_init(&w_full[b], 1);
      mbar_init(&w_free[b], CF::EPI_WARPS);
    }
    mbar_fence_init();
  }
  if (warp == CF::MMA_WARP) tmem_alloc(tmem_slot, CF::TMEM_COLS);
  // zero the operand tiles once: the sparsity pattern of A is fixed, so only its non-zeros are rewritten per edge
  for (int s = 0; s < S; ++s) {
    float4* z = reinterpret_cast<float4*>(slots + (size_t)s * CF::SLOT_BYTES);
    for (uint32_t i = tid; i < (CF::A_BYTES + CF::B_BYTES) / 16; i += CF::THREADS) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int i = tid; i < NT * 128; i += CF::THREADS) rowtab[i] = I::rows()[i];
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  // this CTA's receiver nodes: contiguous, balanced by edge count
  const int64_t n_lo = blockIdx.x == 0 ? 0 : tu_lower_bound(row_ptr, N, E * blockIdx.x / gridDim.x);
  const int64_t n_hi = blockIdx.x + 1 == gridDim.x ? N : tu_lower_bound(row_ptr, N, E * (blockIdx.x + 1) / gridDim.x);
  const int e_lo = n_lo < n_hi ? __ldg(row_ptr + n_lo) : 0, e_hi = n_lo < n_hi ? __ldg(row_ptr + n_hi) : 0;
  const uint32_t n_e = (uint32_t)(e_hi - e_lo);
  // weight groups: TU_WE consecutive CSR slots, aligned to absolute slot index (a group may start before e_lo)
  const int wg0 = e_lo / TU_WE;
  const uint32_t n_wg = n_e ? (uint32_t)((e_hi - 1) / TU_WE - wg0 + 1) : 0u;

  if (warp < CF::EPI_WARPS) {
    // ======================================================= epilogue: thread = TMEM lane = message row (p, k)
    const int tile = warp >> 2, quad = warp & 3;
    const int2 info = rowtab[tile * 128 + quad * 32 + lane];
    const uint32_t wrow = (uint32_t)(info.x < 0 ? 0 : info.x) * TU_W_PITCH;
    int wk_cur = -1;   // weight group (relative to wg0) this thread has acquired
    const uint32_t taddr0 = tmem_base + ((uint32_t)(quad * 32) << 16) + tile * 2 * TU_C;
    float* const stg = stg_all + warp * (32 * TU_STG_LD);
    const int r4 = lane >> 3, q = lane & 7;
    float acc[TU_C];
#pragma unroll
    for (int j = 0; j < TU_C; ++j) acc[j] = 0.f;
    uint32_t it = 0;
    int ee = e_lo;
    tt.start();
    for (int64_t n = n_lo; n < n_hi; ++n) {
      const int eb = ee;
      ee = __ldg(row_ptr + n + 1);
      for (int e = eb; e < ee; ++e, ++it) {
        const uint32_t s = it % S, ph = (it / S) & 1;
        const int wk = e / TU_WE - wg0, wb = wk % TU_WG;
        if (wk != wk_cur) {   // first edge of a weight group: the bulk copies land on w_full
          tw.start();
          tu_wait<0>(&w_full[wb], (uint32_t)(wk / TU_WG) & 1);
          tw.stop(pw0);
          wk_cur = wk;
        }
        const uint8_t* wsm = wbufs + (size_t)wb * CF::W_BYTES + wrow + (e % TU_WE) * 128;
        tw.start();
        tu_wait<0>(&d_full[s], ph);
        tw.stop(pw1);
        tc_fence_after_sync();
        // two halves of 16 channels: [hi.hi + lo.hi] and [hi.lo] accumulator columns are summed on the way (keeps the live
        // registers at 32 + the 32 accumulators; loading both 32-column pieces at once spilled)
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t v[16], v2[16];
          tmem_ld_32x32b_x16(taddr0 + s * CF::D_COLS + hf * 16, v);
          tmem_ld_32x32b_x16(taddr0 + s * CF::D_COLS + TU_C + hf * 16, v2);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 wv = *reinterpret_cast<const float4*>(wsm + 64 * hf + 16 * j);
            acc[16 * hf + 4 * j + 0] = fmaf(wv.x, __uint_as_float(v[4 * j + 0]) + __uint_as_float(v2[4 * j + 0]), acc[16 * hf + 4 * j + 0]);
            acc[16 * hf + 4 * j + 1] = fmaf(wv.y, __uint_as_float(v[4 * j + 1]) + __uint_as_float(v2[4 * j + 1]), acc[16 * hf + 4 * j + 1]);
            acc[16 * hf + 4 * j + 2] = fmaf(wv.z, __uint_as_float(v[4 * j + 2]) + __uint_as_float(v2[4 * j + 2]), acc[16 * hf + 4 * j + 2]);
            acc[16 * hf + 4 * j + 3] = fmaf(wv.w, __uint_as_float(v[4 * j + 3]) + __uint_as_float(v2[4 * j + 3]), acc[16 * hf + 4 * j + 3]);
          }
        }
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&slot_free[s]);
          if (e % TU_WE == TU_WE - 1 || e == e_hi - 1) mbar_arrive(&w_free[wb]);   // last use of the group in this CTA
        }
      }
      // end of node n: rows without edges store zeros.  thread = row -> (transpose tile) -> 8 lanes per 128-byte row
      tw.start();
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        *reinterpret_cast<float4*>(stg + lane * TU_STG_LD + 4 * j) =
            make_float4(acc[4 * j] * inv_norm, acc[4 * j + 1] * inv_norm, acc[4 * j + 2] * inv_norm, acc[4 * j + 3] * inv_norm);
        acc[4 * j] = acc[4 * j + 1] = acc[4 * j + 2] = acc[4 * j + 3] = 0.f;
      }
      __syncwarp();
      float* const mn = M + n * (int64_t)(I::ROWS * TU_C);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = 4 * i + r4;
        const int orow = rowtab[tile * 128 + quad * 32 + r].y;
        if (orow >= 0)
          __stcs(reinterpret_cast<float4*>(mn + orow * TU_C) + q, *reinterpret_cast<const float4*>(stg + r * TU_STG_LD + 4 * q));
      }
      __syncwarp();
      tw.stop(pw2);
    }
    tt.stop(ptot);
    if (PROF && warp == 0 && lane == 0) {
      prof[4] = pw0; prof[5] = pw1; prof[6] = pw2; prof[7] = ptot; prof[10] = n_e; prof[11] = (unsigned long long)(n_hi - n_lo);
    }
  } else if (warp < CF::MMA_WARP) {
    // ======================================================= fill: A = G_e^T (sparse scatter), B = x[sender]^T
    const int fw = warp - CF::FILL_WARP0, ft = fw * 32 + lane;
    uint32_t g_off[I::NIT];
    int g_j0[I::NIT], g_j1[I::NIT];
    float g_c0[I::NIT], g_c1[I::NIT];
#pragma unroll
    for (int qi = 0; qi < I::NIT; ++qi) {
      const int4 t = I::gtab()[qi * EEG_TPU_FILL_THREADS + ft];
      g_off[qi] = ((uint32_t)t.x & 0xffffu) * 4u;
      g_j0[qi] = (t.x >> 16) & 0xff;
      g_j1[qi] = (t.x >> 24) & 0xff;
      g_c0[qi] = __int_as_float(t.y);
      g_c1[qi] = __int_as_float(t.z);
    }
    // B chunk of this warp: tile xt, K-chunk xc = components xi0 .. xi0 + 3 of the sender row (zero beyond the tile's xn)
    const bool has_x = fw < CF::XCHUNKS;
    const int xt = has_x ? fw / KCH : 0, xc = has_x ? fw % KCH : 0;
    const int xi0 = I::xoff(xt) + 4 * xc;
    int xcnt = I::xn(xt) - 4 * xc;
    xcnt = !has_x ? 0 : (xcnt < 0 ? 0 : (xcnt > 4 ? 4 : xcnt));
    const uint32_t b_off = CF::A_BYTES + xt * CF::B_TILE + xc * TU_B_LBO + lane * 16;
    auto load_x = [&](int snd, float (&xv)[4]) {
#pragma unroll
      for (int c = 0; c < 4; ++c) xv[c] = c < xcnt ? __ldg(h + ((int64_t)snd * DIN + xi0 + c) * TU_C + lane) : 0.f;
    };
    auto load_y = [&](int e) { return lane < ystride ? __ldg(Y + (int64_t)e * ystride + lane) : 0.f; };
    // register prefetch ring, two edges deep (the sender rows come from L2 / HBM: ~1-2 edge periods of latency)
    auto snd_at = [&](uint32_t i) { return i < n_e ? __ldg(sender + e_lo + i) : 0; };
    float yv = 0.f, xv[4] = {0.f, 0.f, 0.f, 0.f}, y1v = 0.f, x1v[4] = {0.f, 0.f, 0.f, 0.f};
    int s2 = snd_at(2), s3 = snd_at(3);
    if (n_e > 0) {
      yv = load_y(e_lo);
      load_x(snd_at(0), xv);
    }
    if (n_e > 1) {
      y1v = load_y(e_lo + 1);
      load_x(snd_at(1), x1v);
    }
    tt.start();
    for (uint32_t it = 0; it < n_e; ++it) {
      const uint32_t s = it % S, ph = (it / S) & 1;
      // issue the loads of edge it + 2 before blocking on the slot
      float y2v = 0.f, x2v[4] = {0.f, 0.f, 0.f, 0.f};
      if (it + 2 < n_e) {
        y2v = load_y(e_lo + (int)it + 2);
        load_x(s2, x2v);
      }
      s2 = s3;
      s3 = snd_at(it + 4);
      tw.start();
      tu_wait<0>(&slot_free[s], ph ^ 1);
      tw.stop(pw0);
      uint8_t* const sl = slots + (size_t)s * CF::SLOT_BYTES;
#pragma unroll
      for (int qi = 0; qi < I::NIT; ++qi) {
        const float y0 = __shfl_sync(0xffffffffu, yv, g_j0[qi]);
        const float y1 = __shfl_sync(0xffffffffu, yv, g_j1[qi]);
        const float g = fmaf(g_c1[qi], y1, g_c0[qi] * y0);
        const float hi = tu_rna(g);
        *reinterpret_cast<float*>(sl + g_off[qi]) = hi;
        *reinterpret_cast<float*>(sl + CF::A_TERM + g_off[qi]) = g - hi;
      }
      if (has_x) {
        float4 hi, lo;
        hi.x = tu_rna(xv[0]); hi.y = tu_rna(xv[1]); hi.z = tu_rna(xv[2]); hi.w = tu_rna(xv[3]);
        lo.x = xv[0] - hi.x; lo.y = xv[1] - hi.y; lo.z = xv[2] - hi.z; lo.w = xv[3] - hi.w;
        *reinterpret_cast<float4*>(sl + b_off) = hi;
        *reinterpret_cast<float4*>(sl + b_off + CF::B_LO) = lo;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&filled[s]);
      yv = y1v;
      y1v = y2v;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        xv[c] = x1v[c];
        x1v[c] = x2v[c];
      }
    }
    tt.stop(ptot);
    if (PROF && fw == 0 && lane == 0) {
      prof[0] = pw0; prof[1] = ptot;
    }
  } else if (warp == CF::MMA_WARP) {
    // ======================================================= MMA issuer (one lane)
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(128, TU_C), idesc2 = make_idesc_tf32(128, 2 * TU_C);
      const uint32_t desc_hi = (128u >> 4) | (1u << 14);   // SBO = 128 B, descriptor version 1
      auto desc = [desc_hi](uint32_t lo) {
        uint64_t dd;
        asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(desc_hi));
        return dd;
      };
      const uint32_t a_lo0 = ((smem_u32(slots) >> 4) & 0x3FFFu) | (((uint32_t)EEG_TPU_A_LBO >> 4) << 16);
      const uint32_t b_lo0 = (((smem_u32(slots) + CF::A_BYTES) >> 4) & 0x3FFFu) | ((TU_B_LBO >> 4) << 16);
      tt.start();
      for (uint32_t it = 0; it < n_e; ++it) {
        const uint32_t s = it % S, ph = (it / S) & 1;
        tw.start();
        tu_wait<0>(&filled[s], ph);
        tw.stop(pw0);
        tc_fence_after_sync();
        const uint32_t so = s * (CF::SLOT_BYTES >> 4);
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          const uint32_t dcol = tmem_base + s * CF::D_COLS + t * 2 * TU_C;
#pragma unroll
          for (int ks = 0; ks < CF::KT / 8; ++ks) {
            const uint32_t ad = a_lo0 + so + ((t * CF::A_TILE + ks * 2 * EEG_TPU_A_LBO) >> 4);
            const uint32_t bd = b_lo0 + so + ((t * CF::B_TILE + ks * 2 * TU_B_LBO) >> 4);
            mma_tf32(dcol, desc(ad), desc(bd), idesc2, ks ? 1u : 0u);                      // A_hi x [B_hi | B_lo]  (N = 64)
            mma_tf32(dcol, desc(ad + (CF::A_TERM >> 4)), desc(bd), idesc, 1u);             // A_lo x B_hi -> first half
          }
        }
        mma_commit(&d_full[s]);
      }
      tt.stop(ptot);
      if (PROF) { prof[2] = pw0; prof[3] = ptot; }
    }
  } else {
    // ======================================================= weight loader: lane l copies paths l, l + 32, ...
    tt.start();
    for (uint32_t k = 0; k < n_wg; ++k) {
      const int wb = (int)(k % TU_WG);
      const int64_t e0 = (int64_t)(wg0 + (int)k) * TU_WE;
      const uint32_t bytes = (uint32_t)((E - e0 < TU_WE ? E - e0 : TU_WE) * 128);
      if (lane == 0) {
        tw.start();
        tu_wait<64>(&w_free[wb], ((k / TU_WG) & 1) ^ 1);
        tw.stop(pw0);
        mbar_arrive_expect_tx(&w_full[wb], I::NPATHS * bytes);
      }
      __syncwarp();
      uint8_t* const wsm = wbufs + (size_t)wb * CF::W_BYTES;
      for (int p = lane; p < I::NPATHS; p += 32)
        bulk_g2s(wsm + p * TU_W_PITCH, w_t + ((int64_t)p * E + e0) * TU_C, bytes, &w_full[wb]);
    }
    tt.stop(ptot);
    if (PROF && lane == 0) { prof[8] = pw0; prof[9] = ptot; }
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (warp == CF::MMA_WARP) tmem_dealloc(tmem_base, CF::TMEM_COLS);
}

template <int LIN, int LMAX>
int launch_tp_umma(const float* h, const float* Y, int ystride, const float* w_t, int64_t E, const int32_t* sender,
                   const int32_t* row_ptr, float inv_norm, float* M, int64_t N, cudaStream_t s) {
  using CF = TuCfg<LIN, LMAX>;
  static EegDevFlags attr_dev;
  bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_tp_umma<LIN, LMAX, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CF::SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_tp_umma<LIN, LMAX, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CF::SMEM_BYTES);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_tp_scatter_fwd (tcgen05): cannot opt in to %u B shared memory: %s", CF::SMEM_BYTES, cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = EEG_NUM_SMS;
  int64_t grid = sms;
  if (grid > N) grid = N;
  static const bool prof = getenv("EEG_TP_PROF") && atoi(getenv("EEG_TP_PROF")) != 0;
  if (prof) k_tp_umma<LIN, LMAX, true><<<(unsigned)grid, CF::THREADS, CF::SMEM_BYTES, s>>>(h, Y, ystride, w_t, E, sender, row_ptr, inv_norm, M, N);
  else k_tp_umma<LIN, LMAX, false><<<(unsigned)grid, CF::THREADS, CF::SMEM_BYTES, s>>>(h, Y, ystride, w_t, E, sender, row_ptr, inv_norm, M, N);
  EEG_CHECK_LAUNCH("eeg_tp_scatter_fwd (tcgen05)");
  return 0;
}

}  // namespace

// Debug: copies the per-CTA stall counters of the last EEG_TP_PROF=1 launch (EEG_NUM_SMS x 16 uint64) to host memory.
extern "C" int eeg_tp_umma_prof_read(unsigned long long* host_out) {
  cudaError_t e = cudaMemcpyFromSymbol(host_out, eeg_tpu_prof, sizeof(unsigned long long) * EEG_NUM_SMS * 16);
  if (e != cudaSuccess) {
    eeg_set_error("eeg_tp_umma_prof_read: %s", cudaGetErrorString(e));
    return 2;
  }
  return 0;
}

// Returns 0 when launched, -1 when the shape is outside this kernel (the caller then runs the CUDA-core kernel).
int eeg_tp_umma_launch(int lmax_in, int lmax, int C, const float* h, const float* Y, int ystride, const float* w_t, int64_t E,
                       const int32_t* sender_sorted, const int32_t* row_ptr, float inv_norm, float* M, int64_t N, cudaStream_t s) {
  if (C != TU_C || E >= (1LL << 31) - 8 || N * (int64_t)EEG_TPU_4_4_ROWS * TU_C >= (1LL << 40)) return -1;
  if ((reinterpret_cast<uintptr_t>(w_t) & 15) != 0) return -1;   // bulk copies need 16-byte aligned sources
  if (lmax_in == 4 && lmax == 4) return launch_tp_umma<4, 4>(h, Y, ystride, w_t, E, sender_sorted, row_ptr, inv_norm, M, N, s);
  if (lmax_in == 3 && lmax == 3) return launch_tp_umma<3, 3>(h, Y, ystride, w_t, E, sender_sorted, row_ptr, inv_norm, M, N, s);
  return -1;
}
