// Symmetric contraction (gnn/mace.py:112-277), gradient with respect to the weights (the reference gets it from autograd
// through its einsum chain):
//   eeg_sc_coef_grad       : dT[z][slot][c] = sum_{n in node slice z} dout[n,M,c] x[n,i,c] [x[n,j,c]] [x[n,t,c]]  with
//                            (M, i, j, t) the monomial of the slot (tables.sc_slot_table).  Generated per-segment
//                            straight-line code (tools/gen_kernels.py: gen_sc_wgrad): a warp owns one segment of ~118
//                            consecutive slots, lane = channel, accumulators in registers, dout_M x_i x_j shared between
//                            neighbouring slots at compile time.  Node tiles are staged once per CTA (4 segments) in
//                            shared memory by cp.async, double buffered.  Fixed node order, eeg_sc_grad_slices() partial
//                            sums, no atomics: deterministic.
//   eeg_sc_coefficients_bwd: dW[k,c] = sum_slot A[slot,k] sum_z dT[z][slot][c]   (transpose of eeg_sc_coefficients); the
//                            slices are first summed into slice 0 in place (dT is scratch)
#include "eeg_common.cuh"
#include "eeg_umma.cuh"   // mbarrier + cp.async.bulk helpers
#include "generated/all_scw.cuh"

namespace {

// A CTA owns WG_SEGS segments; warp w runs segment (w & 3), so the warps of one SM sub-partition (w mod 4) all execute the
// same ~4 KB loop body (it stays in the L0 instruction cache; 8 different bodies per SM measured 76% no-instruction
// stalls).  The WG_WPS warps of a segment take the nodes of a tile round robin and are summed through shared memory at the end.
constexpr int WG_SEGS = 4, WG_WPS = 2, WG_WARPS = WG_SEGS * WG_WPS, WG_NB = 4 * WG_WPS, WG_SLICES = 74;

template <int LMAX, int NU> struct Scw;
#define EEG_X(L, N) \
  template <> struct Scw<L, N> { static constexpr int SEG = EEG_SCW_l##L##_nu##N##_SEG, NSEG = EEG_SCW_l##L##_nu##N##_NSEG, NCOEF = EEG_SCW_l##L##_nu##N##_NCOEF; };
EEG_SC_COMBOS(EEG_X)
#undef EEG_X

__device__ __forceinline__ void cp_async16_zfill(void* smem, const void* gmem, bool valid) {
  const unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 16 : 0;   // src-size 0: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(gmem), "r"(n) : "memory");
}

#define XS(i) xk[(i) * 32]
#define GS(M) gk[(M) * 32]
// one segment: its straight-line body once per staged node
#define WG_LOOP(BODY)                                       \
  _Pragma("unroll 1") for (int k = sub; k < WG_NB; k += WG_WPS) { \
    const float* xk = xt + k * (D * 32);                    \
    const float* gk = gt + k * (D * 32);                    \
    BODY                                                    \
  }

template <int LMAX, int NU>
__global__ void __launch_bounds__(WG_WARPS * 32, 1) k_sc_wgrad(const float* __restrict__ x, const float* __restrict__ dout, int C,
                                                                int64_t N, float* __restrict__ dT) {
  constexpr int D = (LMAX + 1) * (LMAX + 1), SEG = Scw<LMAX, NU>::SEG, NCOEF = Scw<LMAX, NU>::NCOEF;
  constexpr int TILE = WG_NB * D * 32;                 // floats per tensor per stage
  extern __shared__ __align__(16) float wg_smem[];     // [2 stages][x | dout][WG_NB][D][32]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int seg = blockIdx.x * WG_SEGS + (warp & (WG_SEGS - 1)), sub = warp / WG_SEGS, chunk = blockIdx.y;
  const int64_t per = (N + gridDim.z - 1) / gridDim.z;
  const int64_t n_begin = blockIdx.z * per, n_end = n_begin + per < N ? n_begin + per : N;
  const int n_tiles = n_begin < n_end ? (int)((n_end - n_begin + WG_NB - 1) / WG_NB) : 0;

  // C == 32: a full tile of x (and of dout) is ONE contiguous 25.6 KB block -> one bulk copy (TMA engine) each, issued by one
  // thread, completion on an mbarrier.  (The per-thread cp.async loop below - 12.5 chunks per thread and tile, each with its
  // own row / piece arithmetic - was 28 % of the kernel's instructions: profiles/r02_ncu_tp_bwd2_scw_summary.txt.)  Partial
  // tiles (only the last of a slice) and C > 32 keep the cp.async path with its zero fill.
  __shared__ uint64_t bulk_bar[2];
  if (threadIdx.x == 0) {
    eeg_umma::mbar_init(&bulk_bar[0], 1);
    eeg_umma::mbar_init(&bulk_bar[1], 1);
    eeg_umma::mbar_fence_init();
  }
  __syncthreads();
  uint32_t bulk_phase[2] = {0u, 0u};
  auto is_bulk = [&](int t) { return C == 32 && n_begin + (int64_t)(t + 1) * WG_NB <= n_end; };
  auto stage = [&](int t) {   // tile t: rows (node, d) are consecutive rows of the [N*D, C] matrices; 8 x 16 B per row chunk
    float* dst = wg_smem + (t & 1) * 2 * TILE;
    const int64_t row0 = (n_begin + (int64_t)t * WG_NB) * D, row_end = n_end * D;
    if (is_bulk(t)) {
      if (threadIdx.x == 0) {
        eeg_umma::fence_proxy_async_smem();   // the buffer's previous readers (generic proxy) are behind the barrier of the caller
        eeg_umma::mbar_arrive_expect_tx(&bulk_bar[t & 1], 2 * TILE * 4);
        eeg_umma::bulk_g2s(dst, x + row0 * 32, TILE * 4, &bulk_bar[t & 1]);
        eeg_umma::bulk_g2s(dst + TILE, dout + row0 * 32, TILE * 4, &bulk_bar[t & 1]);
      }
      return;
    }
    for (int e = threadIdx.x; e < 2 * WG_NB * D * 8; e += WG_WARPS * 32) {
      const int which = e >= WG_NB * D * 8, r = (which ? e - WG_NB * D * 8 : e) >> 3, piece = e & 7;
      const bool ok = row0 + r < row_end;
      const float* src = (which ? dout : x) + (ok ? (row0 + r) * C + chunk * 32 + piece * 4 : 0);
      cp_async16_zfill(dst + which * TILE + r * 32 + piece * 4, src, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  float A[SEG];
#pragma unroll
  for (int k = 0; k < SEG; ++k) A[k] = 0.f;
  if (n_tiles > 0) stage(0);
  for (int t = 0; t < n_tiles; ++t) {
    if (t + 1 < n_tiles) stage(t + 1);
    if (is_bulk(t)) {
      eeg_umma::mbar_wait(&bulk_bar[t & 1], bulk_phase[t & 1]);
      bulk_phase[t & 1] ^= 1u;
    } else if (t + 1 < n_tiles && !is_bulk(t + 1)) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");   // C > 32: the next tile's copies stay in flight
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const float* xt = wg_smem + (t & 1) * 2 * TILE + lane;
    const float* gt = xt + TILE;
#define EEG_X(L, N) \
    if constexpr (LMAX == L && NU == N) { EEG_SCW_l##L##_nu##N##_SWITCH(seg, WG_LOOP) }
    EEG_SC_COMBOS(EEG_X)
#undef EEG_X
    __syncthreads();   // the stage is refilled by the next iteration's prefetch
  }
  // fold the WG_WPS partial sums of a segment (fixed order), staging memory reused: [WG_SEGS][SEG][32]
  static_assert(WG_SEGS * SEG * 32 <= 4 * TILE, "reduction scratch must fit the staging buffers");
  float* red = wg_smem + ((warp & (WG_SEGS - 1)) * SEG) * 32 + lane;
  for (int r = WG_WPS - 1; r >= 1; --r) {
    if (sub == r) {
#pragma unroll
      for (int k = 0; k < SEG; ++k) red[k * 32] = A[k];
    }
    __syncthreads();
    if (sub == 0) {
#pragma unroll
      for (int k = 0; k < SEG; ++k) A[k] += red[k * 32];
    }
    __syncthreads();
  }
  if (sub != 0) return;
  float* o = dT + ((int64_t)blockIdx.z * NCOEF + (int64_t)seg * SEG) * C + chunk * 32 + lane;
#pragma unroll
  for (int k = 0; k < SEG; ++k)
    if (seg * SEG + k < NCOEF) o[(int64_t)k * C] = A[k];
}

// slice 0 <- sum over slices in a fixed order: block = 32 float4 elements x 8 partial lanes (lane y adds slices y, y + 8, ...;
// the lane sums are added in lane order).  One thread per element chained n_slices dependent loads.
__global__ void __launch_bounds__(256) k_sc_slice_sum(float4* __restrict__ dT, int64_t n4, int n_slices) {
  __shared__ float4 part[8][33];
  const int64_t e = (int64_t)blockIdx.x * 32 + threadIdx.x;
  const int y = threadIdx.y;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e < n4)
    for (int z = y; z < n_slices; z += 8) {
      const float4 v = dT[z * n4 + e];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  part[y][threadIdx.x] = s;
  __syncthreads();
  if (y == 0 && e < n4) {
    float4 r = part[0][threadIdx.x];
#pragma unroll
    for (int q = 1; q < 8; ++q) {
      const float4 v = part[q][threadIdx.x];
      r.x += v.x; r.y += v.y; r.z += v.z; r.w += v.w;
    }
    dT[e] = r;
  }
}

__global__ void k_sc_coefficients_bwd(const int32_t* __restrict__ tp, const int32_t* __restrict__ trow, const float* __restrict__ tval,
                                      int n_wrows, const float* __restrict__ dT, int C, float* __restrict__ dW) {
  const int k = blockIdx.x * blockDim.y + threadIdx.y;
  if (k >= n_wrows) return;
  for (int c = threadIdx.x; c < C; c += 32) {
    float s = 0.f;
    for (int q = tp[k]; q < tp[k + 1]; ++q) {
      s = fmaf(__ldg(tval + q), dT[(int64_t)__ldg(trow + q) * C + c], s);
    }
    dW[(int64_t)k * C + c] = s;
  }
}

template <int LMAX, int NU>
int launch_wgrad(int C, const float* x, const float* dout, int64_t N, float* dT, cudaStream_t s) {
  constexpr int D = (LMAX + 1) * (LMAX + 1);
  constexpr size_t smem = (size_t)2 * 2 * WG_NB * D * 32 * sizeof(float);
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_sc_wgrad<LMAX, NU>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_sc_coef_grad: cannot opt in to %zu B shared memory: %s", smem, cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  static_assert(Scw<LMAX, NU>::NSEG % WG_SEGS == 0, "segment count");
  dim3 grid(Scw<LMAX, NU>::NSEG / WG_SEGS, C / 32, WG_SLICES);   // N == 0: every slice is written as zeros
  k_sc_wgrad<LMAX, NU><<<grid, WG_WARPS * 32, smem, s>>>(x, dout, C, N, dT);
  EEG_CHECK_LAUNCH("eeg_sc_coef_grad");
  return 0;
}

}  // namespace

extern "C" int eeg_sc_grad_slices(void) { return WG_SLICES; }

extern "C" int eeg_sc_coef_grad(int lmax, int nu, int C, const float* x, const float* dout, int n_coef, int64_t N, float* dT,
                                void* stream) {
  EEG_REQUIRE(nu >= 1 && nu <= 3 && lmax >= 1 && lmax <= 4, "eeg_sc_coef_grad: unsupported (lmax=%d, correlation=%d)", lmax, nu);
  EEG_REQUIRE(C > 0 && C % 32 == 0, "eeg_sc_coef_grad: C must be a multiple of 32");
#define EEG_X(L, NN)                                                                                                               \
  if (lmax == L && nu == NN) {                                                                                                     \
    constexpr int want = Scw<L, NN>::NCOEF;                                                                                        \
    EEG_REQUIRE(n_coef == want, "eeg_sc_coef_grad: n_coef %d does not match the generated program", n_coef);                       \
    return launch_wgrad<L, NN>(C, x, dout, N, dT, (cudaStream_t)stream);                                                           \
  }
  EEG_SC_COMBOS(EEG_X)
#undef EEG_X
  return 1;
}

extern "C" int eeg_sc_coefficients_bwd(const int32_t* t_ptr, const int32_t* t_row, const float* t_val, int n_wrows, int n_coef,
                                       float* dT, int C, float* dW, void* stream) {
  if (n_wrows == 0 || n_coef == 0) return 0;
  EEG_REQUIRE(C % 4 == 0, "eeg_sc_coefficients_bwd: C must be a multiple of 4");
  const int64_t n4 = (int64_t)n_coef * C / 4;
  k_sc_slice_sum<<<eeg_div_up(n4, 32), dim3(32, 8), 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(dT), n4, eeg_sc_grad_slices());
  EEG_CHECK_LAUNCH("eeg_sc_coefficients_bwd(slice sum)");
  dim3 block(32, 4), grid(eeg_div_up(n_wrows, 4));
  k_sc_coefficients_bwd<<<grid, block, 0, (cudaStream_t)stream>>>(t_ptr, t_row, t_val, n_wrows, dT, C, dW);
  EEG_CHECK_LAUNCH("eeg_sc_coefficients_bwd");
  return 0;
}
