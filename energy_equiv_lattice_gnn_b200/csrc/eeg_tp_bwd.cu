// Backward of the channelwise tensor-product message + scatter (gnn/blocks.py:591-597).
//   dw[e, p*C + c]      = inv_norm * sum_k dM[recv(e), row(p,k), c] * t_{p,k}(h'[send(e)], Y_e)          (edge-major, CSR order)
//   dx[e, i, c]         = inv_norm * sum_p w[e,p,c] sum_k G_p[i,k](Y_e) dM[recv(e), row(p,k), c]          (edge-major, CSR order)
//   dh'[s, i, c]        = sum_{e : send(e) = s} dx[e, i, c]   by eeg_segment_sum_rows over the SENDER-sorted CSR (fixed order: the
//                         gradient is bitwise repeatable; the first version used atomicAdd into the sender rows)
// One warp per run of 16 receiver rows (contiguous CSR edges), lane = channel; per edge the lane-uniform matrices
// G_p[i,k] = sum_j C_p[i,j,k] Y_j are staged once per warp in shared memory (the forward's table), then a generated
// straight-line function (tools/gen_kernels.py: gen_tp2_bwd) evaluates every path.
#include <stdlib.h>
#include "eeg_common.cuh"
#include "generated/all_tp2.cuh"
#include "generated/all_tp2b.cuh"

namespace {

constexpr int WARPS = 4, NPW = 4;   // small row ranges: the kernel is bound by global-load latency, it needs many warps in flight

template <int LIN, int LMAX>
struct TpB;
#define EEG_TPB(LIN, LMAX)                                                                                          \
  template <>                                                                                                       \
  struct TpB<LIN, LMAX> {                                                                                           \
    static constexpr int NTAB = EEG_TP2_##LIN##_##LMAX##_NTAB, ROWS = EEG_TP2_##LIN##_##LMAX##_ROWS,                 \
                         NPATHS = EEG_TP2_##LIN##_##LMAX##_NPATHS;                                                  \
    static __device__ __forceinline__ const int4* tab() { return eeg_tp2_##LIN##_##LMAX##_tab; }                    \
    static __device__ __forceinline__ void run(const float* x, const float4* G4, const float* wp, long ws, const float* dMp, \
                                               long ms, long cs, float* dwp, float* dx) {                           \
      eeg_tp2b_##LIN##_##LMAX(x, G4, wp, ws, dMp, ms, cs, dwp, dx);                                                 \
    }                                                                                                               \
  };
EEG_TP_COMBOS(EEG_TPB)
#undef EEG_TPB

// Software prefetch into L1 (no registers, no shared memory): the generated per-edge program issues ~300 dependent global
// loads (230 dM rows, 42 weight rows, 25 sender rows) between ~2000 FMAs and ran at 72 % long-scoreboard stalls
// (profiles/r01_ncu_tp_bwd_summary.txt); the lines of the NEXT edge / node are requested while the current one is evaluated.
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

template <int LIN, int LMAX>
__global__ void __launch_bounds__(WARPS * 32) k_tp_bwd(const float* __restrict__ h, const float* __restrict__ Y, int ystride,
                                                       const float* __restrict__ w_t, int64_t E, const int32_t* __restrict__ sender,
                                                       const int32_t* __restrict__ row_ptr, float inv_norm, const float* __restrict__ dM,
                                                       float* __restrict__ dw, float* __restrict__ dh, int64_t N, int C, int chunks) {
  using I = TpB<LIN, LMAX>;
  constexpr int DIN = (LIN + 1) * (LIN + 1);
  extern __shared__ __align__(16) uint8_t tpb_smem[];
  int4* tab_s = reinterpret_cast<int4*>(tpb_smem);
  for (int i = threadIdx.x; i < I::NTAB; i += WARPS * 32) tab_s[i] = I::tab()[i];
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* Gs = reinterpret_cast<float*>(tpb_smem + (size_t)I::NTAB * 16) + (size_t)wid * I::NTAB;
  const int chunk = blockIdx.x % chunks;
  const int64_t n0 = ((int64_t)(blockIdx.x / chunks) * WARPS + wid) * NPW;
  if (n0 >= N) return;
  const int64_t n1 = n0 + NPW < N ? n0 + NPW : N;
  const int c = chunk * 32 + lane;
  const int64_t wstride = (int64_t)chunks * E * 32;   // floats between consecutive paths of w_t
  int ee = __ldg(row_ptr + n0);
  const int e_last = __ldg(row_ptr + n1);
  const float* const wbase = w_t + (int64_t)chunk * E * 32;
  auto prefetch_edge = [&](int e) {   // weight rows (one 128-byte line per path) and sender rows of CSR slot e
    for (int p = lane; p < I::NPATHS; p += 32) prefetch_l1(wbase + (int64_t)p * wstride + (int64_t)e * 32);
    if (lane < DIN) prefetch_l1(h + ((int64_t)__ldg(sender + e) * DIN + lane) * C + chunk * 32);
  };
  auto prefetch_node = [&](int64_t n) {   // the node's dM rows for this 32-channel chunk
    const float* base = dM + n * (int64_t)I::ROWS * C + chunk * 32;
    for (int r = lane; r < I::ROWS; r += 32) prefetch_l1(base + (int64_t)r * C);
  };
  if (ee < e_last) prefetch_edge(ee);
  for (int64_t n = n0; n < n1; ++n) {
    const int eb = ee;
    ee = __ldg(row_ptr + n + 1);
    if (n + 1 < n1) prefetch_node(n + 1);
    const float* dMp = dM + n * (int64_t)I::ROWS * C + c;
    for (int e = eb; e < ee; ++e) {
      if (e + 1 < e_last) prefetch_edge(e + 1);
      const int s = __ldg(sender + e);
      const float yv = lane < ystride ? __ldg(Y + (int64_t)e * ystride + lane) : 0.f;
      __syncwarp();
      for (int q = lane; q < I::NTAB; q += 32) {
        const int4 t = tab_s[q];
        const float y0 = __shfl_sync(0xffffffffu, yv, t.x);
        const float y1 = __shfl_sync(0xffffffffu, yv, t.z);
        Gs[q] = fmaf(__int_as_float(t.y), y0, __int_as_float(t.w) * y1);
      }
      float x[DIN], dx[DIN];
#pragma unroll
      for (int d = 0; d < DIN; ++d) {
        x[d] = __ldg(h + ((int64_t)s * DIN + d) * C + c);
        dx[d] = 0.f;
      }
      __syncwarp();
      float* dwp = dw + (int64_t)e * I::NPATHS * C + c;
      I::run(x, reinterpret_cast<const float4*>(Gs), w_t + (int64_t)chunk * E * 32 + (int64_t)e * 32 + lane, wstride, dMp, C, C, dwp, dx);
      // fold 1 / agg_norm_const (callers that pre-scale dM pass inv_norm = 1 and skip the read-modify-write of dw)
      if (inv_norm != 1.0f) {
#pragma unroll
        for (int p = 0; p < I::NPATHS; ++p) dwp[(int64_t)p * C] *= inv_norm;
      }
#pragma unroll
      for (int d = 0; d < DIN; ++d) __stcs(dh + ((int64_t)e * DIN + d) * C + c, dx[d] * inv_norm);   // dh = per-edge rows [E, DIN, C]
    }
  }
}


// ---- variant 2 (default): the 4 warps of a CTA share ONE receiver node at a time.  The node's dM rows (ROWS x 32 floats of
// this channel chunk, 29 KB for lmax 4) are staged in shared memory by cp.async, double buffered across nodes, and every
// warp evaluates a different edge of the node from that tile: 230 of the ~300 loads per edge become LDS instead of global
// loads that 16 resident warps evicted from L1 before use (variant 1: 72 % long-scoreboard stalls, an L1 prefetch did not help).
constexpr int NPC2 = 16;   // receiver nodes per CTA
__device__ __forceinline__ void cp_async16_b(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}

template <int LIN, int LMAX>
__global__ void __launch_bounds__(WARPS * 32, 3) k_tp_bwd2(const float* __restrict__ h, const float* __restrict__ Y, int ystride,
                                                        const float* __restrict__ w_t, int64_t E, const int32_t* __restrict__ sender,
                                                        const int32_t* __restrict__ row_ptr, float inv_norm, const float* __restrict__ dM,
                                                        float* __restrict__ dw, float* __restrict__ dxe, int64_t N, int C, int chunks) {
  using I = TpB<LIN, LMAX>;
  constexpr int DIN = (LIN + 1) * (LIN + 1);
  constexpr int TILE = I::ROWS * 32;   // floats
  extern __shared__ __align__(16) uint8_t tpb_smem[];
  const int4* __restrict__ tab_s = I::tab();   // read through L1 (coalesced int4): keeping it out of shared memory lets 3 CTAs share an SM
  float* Gall = reinterpret_cast<float*>(tpb_smem);
  float* dMs = Gall + (size_t)WARPS * I::NTAB;   // [2][ROWS][32]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* Gs = Gall + (size_t)wid * I::NTAB;
  const int chunk = blockIdx.x % chunks;
  const int64_t n0 = (int64_t)(blockIdx.x / chunks) * NPC2;
  if (n0 >= N) return;
  const int64_t n1 = n0 + NPC2 < N ? n0 + NPC2 : N;
  const int c = chunk * 32 + lane;
  const int64_t wstride = (int64_t)chunks * E * 32;
  auto stage = [&](int64_t n, int buf) {
    const float* src = dM + n * (int64_t)I::ROWS * C + chunk * 32;
    float* dst = dMs + buf * TILE;
    for (int i = threadIdx.x; i < I::ROWS * 8; i += WARPS * 32) cp_async16_b(dst + i * 4, src + (int64_t)(i >> 3) * C + (i & 7) * 4);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  stage(n0, 0);
  for (int64_t n = n0; n < n1; ++n) {
    const int buf = (int)(n - n0) & 1;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ONE barrier per node: tile n (staged during node n - 1) is visible to every warp, and every warp has finished reading
    // tile n - 1, whose buffer the prefetch of tile n + 1 below overwrites.  (The first version used two barriers per node
    // and sat 27 % of its stall samples there: profiles/r02_ncu_tp_bwd2_summary.txt.)
    __syncthreads();
    if (n + 1 < n1) stage(n + 1, buf ^ 1);
    const int eb = __ldg(row_ptr + n), ee = __ldg(row_ptr + n + 1);
    const float* dMp = dMs + buf * TILE + lane;
    for (int e = eb + wid; e < ee; e += WARPS) {
      const int s = __ldg(sender + e);
      const float yv = lane < ystride ? __ldg(Y + (int64_t)e * ystride + lane) : 0.f;
      // all per-path weights of the edge first: 42 independent loads in flight while G is staged, instead of one exposed
      // global-load latency per path inside the generated program (52 % long-scoreboard stalls)
      float wreg[I::NPATHS];
      const float* wsrc = w_t + (int64_t)chunk * E * 32 + (int64_t)e * 32 + lane;
#pragma unroll
      for (int p = 0; p < I::NPATHS; ++p) wreg[p] = __ldg(wsrc + (int64_t)p * wstride);
      float x[DIN], dx[DIN];
#pragma unroll
      for (int d = 0; d < DIN; ++d) {
        x[d] = __ldg(h + ((int64_t)s * DIN + d) * C + c);
        dx[d] = 0.f;
      }
      __syncwarp();
      for (int q = lane; q < I::NTAB; q += 32) {
        const int4 t = __ldg(tab_s + q);
        const float y0 = __shfl_sync(0xffffffffu, yv, t.x);
        const float y1 = __shfl_sync(0xffffffffu, yv, t.z);
        Gs[q] = fmaf(__int_as_float(t.y), y0, __int_as_float(t.w) * y1);
      }
      __syncwarp();
      float* dwp = dw + (int64_t)e * I::NPATHS * C + c;
      I::run(x, reinterpret_cast<const float4*>(Gs), wreg, 1, dMp, 32, C, dwp, dx);   // weights from registers (stride 1)
      if (inv_norm != 1.0f) {
#pragma unroll
        for (int p = 0; p < I::NPATHS; ++p) dwp[(int64_t)p * C] *= inv_norm;
      }
#pragma unroll
      for (int d = 0; d < DIN; ++d) __stcs(dxe + ((int64_t)e * DIN + d) * C + c, dx[d] * inv_norm);
    }
  }
}

template <int LIN, int LMAX>
int launch_tpb(const float* h, const float* Y, int ystride, const float* w_t, int64_t E, const int32_t* sender, const int32_t* row_ptr,
               float inv_norm, const float* dM, float* dw, float* dh, int64_t N, int C, cudaStream_t s) {
  using I = TpB<LIN, LMAX>;
  // EEG_TPB_V=1: variant 1 (warp = 4 nodes, all loads global); default: variant 2 (CTA-shared node tile in shared memory)
  static const int variant = getenv("EEG_TPB_V") ? atoi(getenv("EEG_TPB_V")) : 2;
  const size_t smem1 = (size_t)I::NTAB * 16 + (size_t)WARPS * I::NTAB * 4;
  const size_t smem = variant == 1 ? smem1 : (size_t)WARPS * I::NTAB * 4 + (size_t)2 * I::ROWS * 32 * 4;
  static EegDevFlags attr_set_dev; bool& attr_set = eeg_dev_flag(attr_set_dev);
  if (!attr_set && smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(k_tp_bwd<LIN, LMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_tp_bwd2<LIN, LMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_tp_scatter_bwd: cannot opt in to %zu B shared memory: %s", smem, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int chunks = C / 32;
  if (variant == 1)
    k_tp_bwd<LIN, LMAX><<<(unsigned)(eeg_div_up(N, WARPS * NPW) * chunks), WARPS * 32, smem, s>>>(h, Y, ystride, w_t, E, sender, row_ptr,
                                                                                                  inv_norm, dM, dw, dh, N, C, chunks);
  else
    k_tp_bwd2<LIN, LMAX><<<(unsigned)(eeg_div_up(N, NPC2) * chunks), WARPS * 32, smem, s>>>(h, Y, ystride, w_t, E, sender, row_ptr, inv_norm,
                                                                                           dM, dw, dh, N, C, chunks);
  EEG_CHECK_LAUNCH("eeg_tp_scatter_bwd");
  return 0;
}

// out[n, :] = sum over j in [ptr[n], ptr[n+1]) of in[perm[j], :], one thread per float4 column, rows added in CSR order
__global__ void __launch_bounds__(256) k_segment_sum_rows(const float4* __restrict__ in, const int32_t* __restrict__ perm,
                                                          const int32_t* __restrict__ ptr, int64_t N, int w4, float4* __restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N * w4) return;
  const int64_t n = i / w4;
  const int c = (int)(i - n * w4);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  const int e1 = __ldg(ptr + n + 1);
  for (int j = __ldg(ptr + n); j < e1; ++j) {
    const float4 v = __ldcs(in + (int64_t)__ldg(perm + j) * w4 + c);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  out[i] = s;
}

}  // namespace

extern "C" int eeg_segment_sum_rows(const float* in, const int32_t* perm, const int32_t* ptr, int64_t N, int64_t width, float* out,
                                    void* stream) {
  EEG_REQUIRE(width > 0 && width % 4 == 0, "eeg_segment_sum_rows: width must be a positive multiple of 4 (got %lld)", (long long)width);
  if (N == 0) return 0;
  const int64_t total = N * (width / 4);
  k_segment_sum_rows<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(in), perm, ptr, N, (int)(width / 4), reinterpret_cast<float4*>(out));
  EEG_CHECK_LAUNCH("eeg_segment_sum_rows");
  return 0;
}

// dw [E, n_paths*C] and dx_edge [E, (lmax_in+1)^2 * C] (both in CSR order) are fully written; the sender-side sum of dx_edge
// is eeg_segment_sum_rows (below) over a CSR of the sender_sorted array.
extern "C" int eeg_tp_scatter_bwd(int lmax_in, int lmax, int C, const float* h, const float* Y, int ystride, const float* w_t,
                                  int64_t E, const int32_t* sender_sorted, const int32_t* row_ptr, float inv_norm, const float* dM,
                                  float* dw, float* dx_edge, int64_t N, void* stream) {
  float* dh = dx_edge;
  EEG_REQUIRE(C > 0 && C % 32 == 0, "eeg_tp_scatter_bwd: C must be a multiple of 32 (got %d)", C);
  if (N == 0 || E == 0) return 0;
  EEG_REQUIRE(ystride % 4 == 0 && ystride <= 32, "eeg_tp_scatter_bwd: ystride must be a multiple of 4 and <= 32 (got %d)", ystride);
  cudaStream_t s = (cudaStream_t)stream;
#define EEG_X(LIN, LMAX) \
  if (lmax_in == LIN && lmax == LMAX) return launch_tpb<LIN, LMAX>(h, Y, ystride, w_t, E, sender_sorted, row_ptr, inv_norm, dM, dw, dh, N, C, s);
  EEG_TP_COMBOS(EEG_X)
#undef EEG_X
  eeg_set_error("eeg_tp_scatter_bwd: unsupported (lmax_in=%d, lmax=%d)", lmax_in, lmax);
  return 1;
}
