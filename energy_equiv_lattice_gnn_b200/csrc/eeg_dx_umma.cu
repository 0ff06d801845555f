// Input gradient of the last radial-MLP layer on the 5th-gen tensor cores:  out[r, j] = scale * sum_k in[r, k] W[k, j]
// with a LONG reduction and 64 outputs (gnn/blocks.py:538-549 backward: ds = dw W2, [E x 1344] x [1344 x 64]; the CUDA-core
// GEMM of eeg_linear.cu ran it at 30 TFLOP/s).  `in` and W are plain row-major 2-D matrices (one irrep block, one row per
// node), so both operands arrive by tensor-map TMA loads (eeg_tma.cuh).
//
// One CTA owns DX_TILES = 4 tiles of 128 rows (all 512 TMEM columns: [hi | lo] x 64 outputs per tile) and walks K in stages of
// 16: per stage the loader warp issues four [128 rows x 16 k] boxes of `in` and one [16 k x 64] box of W into a raw ring;
// the transform warps split into tf32 hi / lo (3xTF32, eeg_umma.cuh) and write the K-major UMMA operand tiles (`in` is
// already K-major: float4 in, float4 out; W is transposed on the way like the operands of eeg_wgrad_umma.cu), one lane issues
// A_hi x [W_hi | W_lo] and A_lo x W_hi per tile and K-step.  The tensor core's fp32 accumulation truncates (DESIGN.md, round-2
// addendum), so the accumulators are drained every 64 K-steps (512 k) and the passes are added in fp32 registers-to-memory by
// the CTA that owns the rows: fixed order, no atomics.
//   warp 17    : TMA loader        warps 0-15 : transform + drain        warp 16 : MMA issuer
#include <stdlib.h>
#include <string.h>
#include "eeg_common.cuh"
#include "eeg_umma.cuh"
#include "eeg_tma.cuh"

namespace {
using namespace eeg_umma;

constexpr int DX_TILES = 4, DX_KS = 16, DX_N = 64;
constexpr int DX_RS = 2, DX_OS = 2;                 // raw stages, operand stages
constexpr int DX_SPP = 32;                          // stages per pass (64 K-steps)
constexpr int DX_XFORM = 512;
constexpr int DX_MMA_WARP = DX_XFORM / 32, DX_TMA_WARP = DX_MMA_WARP + 1;
constexpr int DX_THREADS = DX_XFORM + 64;
constexpr uint32_t DX_LBO = 128 * 16 + 32;          // bytes between 4-k chunks of an A tile; +32 keeps the float4 stores conflict free
constexpr uint32_t DX_TERM = (DX_KS / 4) * DX_LBO;  // 8320 B: one term (hi or lo) of one tile
constexpr uint32_t DX_TILE = 2 * DX_TERM;           // 16640 B
constexpr uint32_t DX_A_BYTES = DX_TILES * DX_TILE; // 66560 B
constexpr uint32_t DX_BLBO = 2 * DX_N * 16;         // 2048 B between 4-k chunks of the W operand ([hi rows | lo rows])
constexpr uint32_t DX_B_BYTES = (DX_KS / 4) * DX_BLBO;   // 8192 B
constexpr uint32_t DX_OP = DX_A_BYTES + DX_B_BYTES; // 74752 B
constexpr uint32_t DX_RAWA = 128 * DX_KS * 4;       // 8192 B: one [128 x 16] box
constexpr uint32_t DX_RAW = DX_TILES * DX_RAWA + DX_KS * DX_N * 4;   // 36864 B
constexpr uint32_t DX_RAW_OFF = (DX_OS * DX_OP + 1023) / 1024 * 1024;   // raw ring 1024-byte aligned
constexpr uint32_t DX_BAR_OFF = DX_RAW_OFF + DX_RS * DX_RAW;
constexpr uint32_t DX_SMEM = DX_BAR_OFF + 256;

template <int SLEEP_NS>
__device__ __forceinline__ void dx_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
    if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
    if (spins > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ float dx_hi(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

__global__ void __launch_bounds__(DX_THREADS, 1)
k_dx_umma(float* __restrict__ out, int64_t R, int K, int out_stride, float scale, const __grid_constant__ CUtensorMap map_in,
          const __grid_constant__ CUtensorMap map_w) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* const raw = smem + DX_RAW_OFF;
  uint64_t* raw_full = reinterpret_cast<uint64_t*>(smem + DX_BAR_OFF);   // [RS] TMA bytes -> transform warps
  uint64_t* raw_empty = raw_full + DX_RS;                                // [RS] transform warps (16) -> loader
  uint64_t* op_full = raw_empty + DX_RS;                                 // [OS] transform warps (16) -> MMA
  uint64_t* op_empty = op_full + DX_OS;                                  // [OS] MMA commit -> transform warps
  uint64_t* acc_full = op_empty + DX_OS;                                 // MMA commit at the end of a pass -> drain
  uint64_t* acc_empty = acc_full + 1;                                    // drained (16 arrivals) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t row0 = (int64_t)blockIdx.x * (DX_TILES * 128);
  const int n_stages = K / DX_KS;

  if (tid == 0) {
    for (int s = 0; s < DX_RS; ++s) { mbar_init(&raw_full[s], 1); mbar_init(&raw_empty[s], DX_XFORM / 32); }
    for (int s = 0; s < DX_OS; ++s) { mbar_init(&op_full[s], DX_XFORM / 32); mbar_init(&op_empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, DX_XFORM / 32);
    mbar_fence_init();
  }
  if (warp == DX_MMA_WARP) tmem_alloc(tmem_slot, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < DX_XFORM / 32) {
    // ============================================================ transform warps
    // A task i (tile i): thread -> (row = tid >> 2, 4-k chunk kq = tid & 3): LDS.128 of the raw box, two STS.128.
    // W task (threads < 256): (chunk q = tid >> 6, output j = tid & 63): four LDS.32 down a column, two STS.128.
    const int kq = tid & 3, arow = tid >> 2;
    const bool has_b = tid < (DX_KS / 4) * DX_N;
    const int bq = tid >> 6, bj = tid & 63;
    const int quarter = warp & 3, sub = warp >> 2;
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    for (int st = 0; st < n_stages; ++st) {
      const int rs = st % DX_RS, os = st % DX_OS;
      const uint8_t* rw = raw + (size_t)rs * DX_RAW;
      uint8_t* a_st = smem + (size_t)os * DX_OP;
      uint8_t* b_st = a_st + DX_A_BYTES;
      dx_wait<20>(&raw_full[rs], (st / DX_RS) & 1);
      dx_wait<20>(&op_empty[os], ((st / DX_OS) & 1) ^ 1);
#pragma unroll
      for (int t = 0; t < DX_TILES; ++t) {
        const float4 v = *reinterpret_cast<const float4*>(rw + t * DX_RAWA + arow * (DX_KS * 4) + kq * 16);
        float4 hi, lo;
        hi.x = dx_hi(v.x); hi.y = dx_hi(v.y); hi.z = dx_hi(v.z); hi.w = dx_hi(v.w);
        lo.x = v.x - hi.x; lo.y = v.y - hi.y; lo.z = v.z - hi.z; lo.w = v.w - hi.w;
        uint8_t* dst = a_st + t * DX_TILE + kq * DX_LBO + arow * 16;
        *reinterpret_cast<float4*>(dst) = hi;
        *reinterpret_cast<float4*>(dst + DX_TERM) = lo;
      }
      if (has_b) {
        const float* wb = reinterpret_cast<const float*>(rw + DX_TILES * DX_RAWA);
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = wb[(4 * bq + j) * DX_N + bj];
        float4 hi, lo;
        hi.x = dx_hi(v[0]); hi.y = dx_hi(v[1]); hi.z = dx_hi(v[2]); hi.w = dx_hi(v[3]);
        lo.x = v[0] - hi.x; lo.y = v[1] - hi.y; lo.z = v[2] - hi.z; lo.w = v[3] - hi.w;
        uint8_t* dst = b_st + bq * DX_BLBO + bj * 16;
        *reinterpret_cast<float4*>(dst) = hi;
        *reinterpret_cast<float4*>(dst + DX_N * 16) = lo;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&op_full[os]);
        mbar_arrive(&raw_empty[rs]);
      }
      if ((st + 1) % DX_SPP == 0 || st + 1 == n_stages) {
        // ---- end of a pass: drain.  TMEM lane = row inside the tile; the 4 warps of a lane quarter share the (tile, 16-column
        // piece) pairs; pass 0 stores, later passes add to what this CTA stored before.
        const int pass = st / DX_SPP;
        dx_wait<32>(acc_full, pass & 1);
        tc_fence_after_sync();
        for (int p = sub; p < DX_TILES * (DX_N / 16); p += 4) {
          const int t = p >> 2, c0 = (p & 3) * 16;
          uint32_t vh[16], vl[16];
          tmem_ld_32x32b_x16(tmem_base + lane_base + t * 2 * DX_N + c0, vh);
          tmem_ld_32x32b_x16(tmem_base + lane_base + t * 2 * DX_N + DX_N + c0, vl);
          tmem_ld_wait();
          const int64_t r = row0 + t * 128 + quarter * 32 + lane;
          if (r < R) {
            float4* o = reinterpret_cast<float4*>(out + r * out_stride + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float4 y = make_float4((__uint_as_float(vh[4 * j]) + __uint_as_float(vl[4 * j])) * scale,
                                     (__uint_as_float(vh[4 * j + 1]) + __uint_as_float(vl[4 * j + 1])) * scale,
                                     (__uint_as_float(vh[4 * j + 2]) + __uint_as_float(vl[4 * j + 2])) * scale,
                                     (__uint_as_float(vh[4 * j + 3]) + __uint_as_float(vl[4 * j + 3])) * scale);
              if (pass > 0) {
                const float4 prev = o[j];
                y.x += prev.x; y.y += prev.y; y.z += prev.z; y.w += prev.w;
              }
              o[j] = y;
            }
          }
        }
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty);
      }
    }
  } else if (warp == DX_MMA_WARP && lane == 0) {
    // ============================================================ MMA issuer
    const uint32_t idesc_n = make_idesc_tf32(128, DX_N), idesc_2n = make_idesc_tf32(128, 2 * DX_N);
    for (int st = 0; st < n_stages; ++st) {
      const int s = st % DX_OS, sp = st % DX_SPP;
      if (sp == 0 && st > 0) {   // the previous pass has been drained
        dx_wait<0>(acc_empty, ((st / DX_SPP) - 1) & 1);
        tc_fence_after_sync();
      }
      dx_wait<0>(&op_full[s], (st / DX_OS) & 1);
      tc_fence_after_sync();
      const uint32_t a0 = smem_u32(smem + (size_t)s * DX_OP), b0 = a0 + DX_A_BYTES;
#pragma unroll
      for (int ks = 0; ks < DX_KS / 8; ++ks) {
        const uint64_t bd = make_desc_kmajor(b0 + ks * 2 * DX_BLBO, DX_BLBO, 128);
#pragma unroll
        for (int t = 0; t < DX_TILES; ++t) {
          const uint32_t at = a0 + t * DX_TILE + ks * 2 * DX_LBO;
          const uint32_t dcol = tmem_base + t * 2 * DX_N;
          mma_tf32(dcol, make_desc_kmajor(at, DX_LBO, 128), bd, idesc_2n, (sp | ks) ? 1u : 0u);   // A_hi x [W_hi | W_lo]
          mma_tf32(dcol, make_desc_kmajor(at + DX_TERM, DX_LBO, 128), bd, idesc_n, 1u);         // A_lo x W_hi -> first half
        }
      }
      mma_commit(&op_empty[s]);
      if (sp == DX_SPP - 1 || st + 1 == n_stages) mma_commit(acc_full);
    }
  } else if (warp == DX_TMA_WARP && lane == 0) {
    // ============================================================ TMA loader
    for (int st = 0; st < n_stages; ++st) {
      const int rs = st % DX_RS;
      uint8_t* slot = raw + (size_t)rs * DX_RAW;
      dx_wait<20>(&raw_empty[rs], ((st / DX_RS) & 1) ^ 1);
      mbar_arrive_expect_tx(&raw_full[rs], DX_RAW);
#pragma unroll
      for (int t = 0; t < DX_TILES; ++t) eeg_tma::load_2d(slot + t * DX_RAWA, &map_in, st * DX_KS, (int)(row0 + t * 128), &raw_full[rs]);
      eeg_tma::load_2d(slot + DX_TILES * DX_RAWA, &map_w, 0, st * DX_KS, &raw_full[rs]);
    }
  }
  __syncthreads();
  tc_fence_after_sync();
  if (warp == DX_MMA_WARP) tmem_dealloc(tmem_base, 512);
}

}  // namespace

// 0 = launched, -1 = shape outside this kernel (caller falls back to the CUDA-core GEMM), > 0 = error.
// Shape: ONE non-empty block with one row per node, K a multiple of 16 and >= 128, 64 outputs, W stored [K][64] (transB = 0),
// no bias / activation / residual, at least 4096 rows.
int eeg_dx_umma_launch(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off, const float* host_scale,
                       int transB, float* out, const EegLayout* lout, int64_t N, cudaStream_t stream) {
  int nb = 0, b0 = 0;
  for (int b = 0; b < lin->n_blocks; ++b)
    if (lin->mul[b] > 0 && lout->mul[b] > 0) { ++nb; b0 = b; }
  if (nb != 1 || transB != 0 || lin->rows[b0] != 1 || lout->rows[b0] != 1) return -1;
  const int K = lin->mul[b0];
  if (lout->mul[b0] != DX_N || K < 128 || K % DX_KS != 0 || N < 4096 || N >= (1LL << 31)) return -1;
  if (lin->stride % 4 || lout->stride % 4 || lin->off[b0] % 4 || lout->off[b0] % 4 || host_w_off[b0] % 4) return -1;
  const float* a = in + lin->off[b0];
  const float* w = W + host_w_off[b0];
  float* o = out + lout->off[b0];
  if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(o)) & 15) return -1;
  static const bool off = getenv("EEG_DX_UMMA") && atoi(getenv("EEG_DX_UMMA")) == 0;   // A/B timing
  if (off) return -1;
  alignas(64) CUtensorMap map_in, map_w;
  if (!eeg_tma::make_map_2d(&map_in, a, N, K, lin->stride, DX_KS, 128) || !eeg_tma::make_map_2d(&map_w, w, K, DX_N, DX_N, DX_N, DX_KS)) return -1;
  static EegDevFlags attr_dev; bool& attr_set = eeg_dev_flag(attr_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_dx_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DX_SMEM);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_block_linear_fwd (tcgen05, long K): cannot opt in to %u B shared memory: %s", DX_SMEM, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const unsigned grid = (unsigned)((N + DX_TILES * 128 - 1) / (DX_TILES * 128));
  k_dx_umma<<<grid, DX_THREADS, DX_SMEM, stream>>>(o, N, K, lout->stride, host_scale[b0], map_in, map_w);
  EEG_CHECK_LAUNCH("eeg_block_linear_fwd (tcgen05, long K)");
  return 0;
}
