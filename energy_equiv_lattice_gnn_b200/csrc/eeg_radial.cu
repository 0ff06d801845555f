// Radial MLP -> per-edge tensor-product weights, fused, with the wide last layer on the 5th-gen tensor cores.
// Replaces `conv_tp_weights` (gnn/blocks.py:538-549,590): Linear(2*nbasis,64) SiLU [Linear(64,64) SiLU]* Linear(64,W).
//
// Persistent CTAs (one per SM), each looping over tiles of 128 consecutive CSR edge slots.
//   warps 0-15 (producer / epilogue, thread = (edge, quarter of the columns) for the hidden layers, lane = channel for
//       the output; four warps per SM sub-partition, one alone is latency bound):
//       first layer on CUDA cores -> hi/lo split (3xTF32, eeg_umma.cuh) -> shared memory in the UMMA K-major layout;
//       hidden 64x64 layers: tcgen05.mma (M = 128 edges, N = 64), TMEM -> bias + SiLU -> split -> same smem tile;
//       last layer transposed: D^T[(path,channel), edge] = W[(path,channel), :] . H[edge, :]^T, M = 128 rows =
//       4 paths x 32 channels, N = 128 edges, so that a TMEM lane is a channel and every st.global of the epilogue
//       writes one full 128-byte line of the path-major weight tensor  w_t[tile][E][32].
//   warp 16 (one lane): streams packed weight stages (32 KB, hi+lo) global(L2) -> smem with cp.async.bulk (TMA engine)
//       through a 4-stage mbarrier ring.
//   warp 17 (one lane): issues tcgen05.mma kind::tf32 (3 MMAs per K-step), fp32 accumulators in TMEM
//       (3 x 128 columns for the output groups + 64 columns for the hidden layers), tcgen05.commit -> mbarriers.
#include "eeg_common.cuh"
#include "eeg_umma.cuh"

namespace {
using namespace eeg_umma;

constexpr int RM_EDGES = 128;   // edges per tile
constexpr int RM_H = 64;        // hidden width (inter_MLP_dim) == K
constexpr int RM_NSTAGE = 4;    // weight stages in flight
constexpr int RM_NACC = 3;      // output accumulator stages (128 TMEM columns each)
constexpr int RM_TMEM_COLS = 512;
constexpr int RM_HID_COL = RM_NACC * RM_EDGES;  // 64 columns for the hidden-layer accumulator
constexpr int RM_MAX_IN = 16;
constexpr int RM_MAX_HIDDEN = 4;
constexpr uint32_t RM_TERM_BYTES = RM_EDGES * RM_H * 4;   // 32 KB: one term (hi or lo) of the activation tile
constexpr uint32_t RM_STAGE_BYTES = 32768;                // hidden: [2][16][64][4] floats; output: [2][8][128][4] floats
constexpr uint32_t RM_STAGE_FLOATS = RM_STAGE_BYTES / 4;
constexpr int RM_WORKERS = 512;   // 16 producer / epilogue warps: thread = (edge row, quarter of the columns)
constexpr int RM_THREADS = RM_WORKERS + 64;

struct RadialSmem {  // byte offsets into dynamic shared memory
  static constexpr uint32_t HT = 0;                                    // activation tile [2 terms][16 chunks][128 edges][4]
  static constexpr uint32_t WST = HT + 2 * RM_TERM_BYTES;              // weight stages
  static constexpr uint32_t W0 = WST + RM_NSTAGE * RM_STAGE_BYTES;     // [64][16] first-layer weight (input dim padded)
  static constexpr uint32_t B0 = W0 + RM_H * RM_MAX_IN * 4;            // [64]
  static constexpr uint32_t BH = B0 + RM_H * 4;                        // [RM_MAX_HIDDEN][64]
  static constexpr uint32_t BAR = BH + RM_MAX_HIDDEN * RM_H * 4;       // mbarriers + TMEM base slot
  static constexpr uint32_t TOTAL = BAR + 256;
};

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }

// Weight image = sequence of 32 KB stages in the order the kernel consumes them for every edge tile:
//   n_hidden stages  [term][chunk 16][row 64][4]  : Wh[l][row][chunk*4+i]
//   2*G stages (g, half) [term][chunk 8][row 128][4] : W2[128 g + row][half*32 + chunk*4 + i]  (rows >= n_out: 0)
__global__ void k_radial_pack(const float* __restrict__ Wh, int n_hidden, const float* __restrict__ W2, int n_out, int n_stages,
                              float* __restrict__ img) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= (int64_t)n_stages * RM_STAGE_FLOATS) return;
  const int st = (int)(i / RM_STAGE_FLOATS), idx = (int)(i % RM_STAGE_FLOATS);
  const int term = idx / (RM_STAGE_FLOATS / 2), rem = idx % (RM_STAGE_FLOATS / 2);
  float x = 0.f;
  if (st < n_hidden) {
    const int chunk = rem / 256, row = (rem % 256) / 4, q = rem % 4;
    x = __ldg(Wh + ((int64_t)st * RM_H + row) * RM_H + chunk * 4 + q);
  } else {
    const int g = (st - n_hidden) / 2, half = (st - n_hidden) % 2;
    const int chunk = rem / 512, row = (rem % 512) / 4, q = rem % 4;
    const int n = g * 128 + row;
    if (n < n_out) x = __ldg(W2 + (int64_t)n * RM_H + half * 32 + chunk * 4 + q);
  }
  const float hi = tf32_hi(x);
  img[i] = term == 0 ? hi : x - hi;
}

__global__ void __launch_bounds__(RM_THREADS, 1)
k_radial_mlp(const float* __restrict__ f, int n_in, const float* __restrict__ W0g, const float* __restrict__ b0g,
             const float* __restrict__ bhg, int n_hidden, const float* __restrict__ img, int n_ptiles,
             float* __restrict__ w_t, int64_t E) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + RadialSmem::BAR);
  uint64_t* w_full = bars;                        // [RM_NSTAGE]  loader -> MMA (tx bytes)
  uint64_t* w_empty = w_full + RM_NSTAGE;         // [RM_NSTAGE]  MMA (tcgen05.commit) -> loader
  uint64_t* acc_full = w_empty + RM_NSTAGE;       // [RM_NACC]    MMA -> epilogue
  uint64_t* acc_empty = acc_full + RM_NACC;       // [RM_NACC]    epilogue (512 threads) -> MMA
  uint64_t* hid_full = acc_empty + RM_NACC;       // [1]          MMA -> epilogue (hidden-layer accumulator)
  uint64_t* h_full = hid_full + 1;                // [1]          producers (512 threads) -> MMA (activation tile ready)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(h_full + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = (n_ptiles + 3) / 4;               // output groups of 4 path tiles (128 rows)
  const int n_stages = n_hidden + 2 * G;
  const int64_t n_etiles = (E + RM_EDGES - 1) / RM_EDGES;

  if (tid == 0) {
    for (int s = 0; s < RM_NSTAGE; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < RM_NACC; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], RM_WORKERS); }
    mbar_init(hid_full, 1);
    mbar_init(h_full, RM_WORKERS);
    mbar_fence_init();
  }
  if (warp == 17) tmem_alloc(tmem_slot, RM_TMEM_COLS);
  float* W0s = reinterpret_cast<float*>(smem + RadialSmem::W0);
  float* b0s = reinterpret_cast<float*>(smem + RadialSmem::B0);
  float* bhs = reinterpret_cast<float*>(smem + RadialSmem::BH);
  for (int i = tid; i < RM_H * RM_MAX_IN; i += RM_THREADS) {
    const int n = i / RM_MAX_IN, k = i % RM_MAX_IN;
    W0s[i] = (k < n_in) ? __ldg(W0g + n * n_in + k) : 0.f;
  }
  for (int i = tid; i < RM_H; i += RM_THREADS) b0s[i] = __ldg(b0g + i);
  for (int i = tid; i < n_hidden * RM_H; i += RM_THREADS) bhs[i] = __ldg(bhg + i);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 16) {
    // =========================================================== producer / epilogue warps
    // four warps per SM sub-partition (one alone is latency bound: dependent FMA / MUFU chains, IPC ~0.25)
    const int r = tid & 127;  // edge row of the tile == TMEM lane for the hidden layers
    const int hf = tid >> 7;  // which quarter of the 64 hidden columns / of the 128 edge columns this thread handles
    float4* slot_hi = reinterpret_cast<float4*>(smem + RadialSmem::HT + r * 16);
    float4* slot_lo = reinterpret_cast<float4*>(smem + RadialSmem::HT + RM_TERM_BYTES + r * 16);
    constexpr int CH = RM_EDGES;  // float4 stride between K-chunks
    auto store_split = [&](int c, float x0, float x1, float x2, float x3) {
      float4 hi, lo;
      hi.x = tf32_hi(x0); hi.y = tf32_hi(x1); hi.z = tf32_hi(x2); hi.w = tf32_hi(x3);
      lo.x = x0 - hi.x; lo.y = x1 - hi.y; lo.z = x2 - hi.z; lo.w = x3 - hi.w;
      slot_hi[c * CH] = hi;
      slot_lo[c * CH] = lo;
    };
    const uint32_t lane_base = ((uint32_t)((warp & 3) * 32)) << 16;
    uint32_t hid_it = 0, acc_it = 0;
    for (int64_t tile = blockIdx.x; tile < n_etiles; tile += gridDim.x) {
      const int64_t e0 = tile * RM_EDGES;
      const int64_t e = e0 + r;
      {  // first layer on CUDA cores: thread = edge
        float fin[RM_MAX_IN];
#pragma unroll
        for (int k = 0; k < RM_MAX_IN; ++k) fin[k] = (e < E && k < n_in) ? __ldg(f + e * n_in + k) : 0.f;
#pragma unroll
        for (int cq = 0; cq < RM_H / 16; ++cq) {
          const int c = hf * (RM_H / 16) + cq;
          float o[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int n = 4 * c + q;
            float s = b0s[n];
            const float4* wr = reinterpret_cast<const float4*>(W0s + n * RM_MAX_IN);
#pragma unroll
            for (int k4 = 0; k4 < RM_MAX_IN / 4; ++k4) {
              const float4 w4 = wr[k4];
              s = fmaf(fin[4 * k4], w4.x, s); s = fmaf(fin[4 * k4 + 1], w4.y, s);
              s = fmaf(fin[4 * k4 + 2], w4.z, s); s = fmaf(fin[4 * k4 + 3], w4.w, s);
            }
            o[q] = silu_f(s);
          }
          store_split(c, o[0], o[1], o[2], o[3]);
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(h_full);
      // hidden 64x64 layers: accumulator row = edge; bias + SiLU, re-split, overwrite the activation tile
      for (int l = 0; l < n_hidden; ++l) {
        mbar_wait(hid_full, hid_it & 1);
        ++hid_it;
        tc_fence_after_sync();
        uint32_t v0[16];
        tmem_ld_32x32b_x16(tmem_base + lane_base + RM_HID_COL + hf * 16, v0);
        tmem_ld_wait();
        tc_fence_before_sync();
        const float* bl = bhs + l * RM_H + hf * 16;
#pragma unroll
        for (int c = 0; c < 4; ++c)
          store_split(hf * 4 + c, silu_f(__uint_as_float(v0[4 * c]) + bl[4 * c]), silu_f(__uint_as_float(v0[4 * c + 1]) + bl[4 * c + 1]),
                      silu_f(__uint_as_float(v0[4 * c + 2]) + bl[4 * c + 2]), silu_f(__uint_as_float(v0[4 * c + 3]) + bl[4 * c + 3]));
        fence_proxy_async_smem();
        mbar_arrive(h_full);
      }
      // output groups: TMEM lane = (path tile q = warp, channel = lane), columns = edges
      for (int g = 0; g < G; ++g) {
        const int a = acc_it % RM_NACC;
        mbar_wait(&acc_full[a], (acc_it / RM_NACC) & 1);
        ++acc_it;
        tc_fence_after_sync();
        const int p = 4 * g + (warp & 3);
        float* dst = w_t + ((int64_t)p * E + e0) * 32 + lane;
#pragma unroll 1
        for (int cc = hf; cc < hf + 1; ++cc) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(tmem_base + lane_base + (uint32_t)(a * RM_EDGES + cc * 32), v);
          tmem_ld_wait();
          if (p < n_ptiles) {
            if (e0 + cc * 32 + 32 <= E) {
#pragma unroll
              for (int j = 0; j < 32; ++j) dst[(cc * 32 + j) * 32] = __uint_as_float(v[j]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (e0 + cc * 32 + j < E) dst[(cc * 32 + j) * 32] = __uint_as_float(v[j]);
            }
          }
        }
        tc_fence_before_sync();
        mbar_arrive(&acc_empty[a]);
      }
    }
  } else if (warp == 16) {
    // =========================================================== weight-stage loader (TMA engine bulk copies)
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < n_etiles; tile += gridDim.x) {
        for (int st = 0; st < n_stages; ++st, ++it) {
          const int s = it % RM_NSTAGE;
          mbar_wait(&w_empty[s], ((it / RM_NSTAGE) & 1) ^ 1);
          mbar_arrive_expect_tx(&w_full[s], RM_STAGE_BYTES);
          bulk_g2s(smem + RadialSmem::WST + s * RM_STAGE_BYTES, img + (int64_t)st * RM_STAGE_FLOATS, RM_STAGE_BYTES, &w_full[s]);
        }
      }
    }
  } else {
    // =========================================================== MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_hid = make_idesc_tf32(128, RM_H);       // M = 128 edges, N = 64 hidden outputs
      constexpr uint32_t idesc_out = make_idesc_tf32(128, RM_EDGES);   // M = 128 (path,channel) rows, N = 128 edges
      const uint32_t ht = smem_u32(smem + RadialSmem::HT);
      uint32_t it = 0, acc_it = 0, h_it = 0;
      for (int64_t tile = blockIdx.x; tile < n_etiles; tile += gridDim.x) {
        for (int l = 0; l < n_hidden; ++l, ++it) {
          mbar_wait(h_full, h_it & 1);
          ++h_it;
          const int s = it % RM_NSTAGE;
          mbar_wait(&w_full[s], (it / RM_NSTAGE) & 1);
          tc_fence_after_sync();
          const uint32_t ws = smem_u32(smem + RadialSmem::WST + s * RM_STAGE_BYTES);
#pragma unroll
          for (int j = 0; j < RM_H / 8; ++j) {
            const uint64_t ahi = make_desc_kmajor(ht + 2 * j * 2048, 2048, 128);
            const uint64_t alo = make_desc_kmajor(ht + RM_TERM_BYTES + 2 * j * 2048, 2048, 128);
            const uint64_t bhi = make_desc_kmajor(ws + 2 * j * 1024, 1024, 128);
            const uint64_t blo = make_desc_kmajor(ws + RM_STAGE_BYTES / 2 + 2 * j * 1024, 1024, 128);
            mma_tf32(tmem_base + RM_HID_COL, alo, bhi, idesc_hid, j > 0 ? 1u : 0u);
            mma_tf32(tmem_base + RM_HID_COL, ahi, blo, idesc_hid, 1u);
            mma_tf32(tmem_base + RM_HID_COL, ahi, bhi, idesc_hid, 1u);
          }
          mma_commit(&w_empty[s]);
          mma_commit(hid_full);
        }
        mbar_wait(h_full, h_it & 1);
        ++h_it;
        for (int g = 0; g < G; ++g) {
          const int a = acc_it % RM_NACC;
          mbar_wait(&acc_empty[a], ((acc_it / RM_NACC) & 1) ^ 1);
          ++acc_it;
          const uint32_t d = tmem_base + (uint32_t)(a * RM_EDGES);
          for (int half = 0; half < 2; ++half, ++it) {
            const int s = it % RM_NSTAGE;
            mbar_wait(&w_full[s], (it / RM_NSTAGE) & 1);
            tc_fence_after_sync();
            const uint32_t ws = smem_u32(smem + RadialSmem::WST + s * RM_STAGE_BYTES);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t kc = (uint32_t)(half * 8 + 2 * j);  // K-chunk of the activation tile
              const uint64_t whi = make_desc_kmajor(ws + 2 * j * 2048, 2048, 128);
              const uint64_t wlo = make_desc_kmajor(ws + RM_STAGE_BYTES / 2 + 2 * j * 2048, 2048, 128);
              const uint64_t hhi = make_desc_kmajor(ht + kc * 2048, 2048, 128);
              const uint64_t hlo = make_desc_kmajor(ht + RM_TERM_BYTES + kc * 2048, 2048, 128);
              mma_tf32(d, wlo, hhi, idesc_out, (half | j) ? 1u : 0u);
              mma_tf32(d, whi, hlo, idesc_out, 1u);
              mma_tf32(d, whi, hhi, idesc_out, 1u);
            }
            mma_commit(&w_empty[s]);
          }
          mma_commit(&acc_full[a]);
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (warp == 17) tmem_dealloc(tmem_base, RM_TMEM_COLS);
}

}  // namespace

extern "C" int64_t eeg_radial_image_floats(int n_hidden, int n_out) {
  const int G = (n_out / 32 + 3) / 4;
  return (int64_t)(n_hidden + 2 * G) * RM_STAGE_FLOATS;
}

extern "C" int eeg_radial_pack(const float* Wh, int n_hidden, int hidden_dim, const float* W2, int n_out, float* img, void* stream) {
  EEG_REQUIRE(hidden_dim == RM_H, "eeg_radial_pack: inter_MLP_dim must be %d (got %d)", RM_H, hidden_dim);
  EEG_REQUIRE(n_hidden >= 0 && n_hidden <= RM_MAX_HIDDEN, "eeg_radial_pack: at most %d hidden layers (got %d)", RM_MAX_HIDDEN, n_hidden);
  EEG_REQUIRE(n_out > 0 && n_out % 32 == 0, "eeg_radial_pack: output width must be a multiple of 32 (got %d)", n_out);
  const int n_stages = (int)(eeg_radial_image_floats(n_hidden, n_out) / RM_STAGE_FLOATS);
  const int64_t tot = (int64_t)n_stages * RM_STAGE_FLOATS;
  k_radial_pack<<<eeg_div_up(tot, 256), 256, 0, (cudaStream_t)stream>>>(Wh, n_hidden, W2, n_out, n_stages, img);
  EEG_CHECK_LAUNCH("eeg_radial_pack");
  return 0;
}

extern "C" int eeg_radial_mlp_fwd(const float* f, int n_in, const float* W0, const float* b0, const float* bh, int n_hidden,
                                  int hidden_dim, const float* img, int n_out, float* w_t, int64_t E, void* stream) {
  EEG_REQUIRE(hidden_dim == RM_H, "eeg_radial_mlp_fwd: inter_MLP_dim must be %d (got %d)", RM_H, hidden_dim);
  EEG_REQUIRE(n_in > 0 && n_in <= RM_MAX_IN, "eeg_radial_mlp_fwd: 2*num_edge_bases must be <= %d (got %d)", RM_MAX_IN, n_in);
  EEG_REQUIRE(n_hidden >= 0 && n_hidden <= RM_MAX_HIDDEN, "eeg_radial_mlp_fwd: at most %d hidden layers (got %d)", RM_MAX_HIDDEN, n_hidden);
  EEG_REQUIRE(n_out > 0 && n_out % 32 == 0, "eeg_radial_mlp_fwd: output width must be a multiple of 32 (got %d)", n_out);
  if (E == 0) return 0;
  static EegDevFlags attr_set_dev; bool& attr_set = eeg_dev_flag(attr_set_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_radial_mlp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RadialSmem::TOTAL);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_radial_mlp_fwd: cannot opt in to %u B shared memory: %s", RadialSmem::TOTAL, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int64_t n_etiles = (E + RM_EDGES - 1) / RM_EDGES;
  const unsigned grid = (unsigned)(n_etiles < EEG_NUM_SMS ? n_etiles : EEG_NUM_SMS);
  k_radial_mlp<<<grid, RM_THREADS, RadialSmem::TOTAL, (cudaStream_t)stream>>>(f, n_in, W0, b0, bh, n_hidden, img, n_out / 32, w_t, E);
  EEG_CHECK_LAUNCH("eeg_radial_mlp_fwd");
  return 0;
}
