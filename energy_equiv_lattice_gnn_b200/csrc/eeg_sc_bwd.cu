// Symmetric contraction (gnn/mace.py:112-277), backward.  The reference differentiates its einsum chain with autograd;
// here the gradients are explicit:
//   eeg_sym_contract_bwd_x : dx[n,d,c] = sum_M dout[n,M,c] dP_M/dx_d, reverse-mode sweep of the same Horner tree and
//                            coefficient stream as the forward (generated macro program, packed FFMA2 on node pairs)
// (weight gradient: eeg_sc_wgrad.cu)
#include "eeg_common.cuh"
#include "generated/all_scb.cuh"

namespace {

#ifndef EEG_SCB_TPC
#define EEG_SCB_TPC 192
#endif
// 2 channels x (2 * TPC) nodes per tile, node pairs per thread.  TPC = 192 (12 warps per SM; 164 registers x 384 threads and
// 214 KB of shared memory still fit one SM): the kernel is latency bound on its instruction stream (45 % no_inst stalls with 8
// warps, profiles/r02_ncu_full_bwd_summary.txt), so more warps sharing the same stream pay off.
constexpr int CH = 2, TPC = EEG_SCB_TPC, THREADS = CH * TPC, NT = 2 * TPC;

typedef unsigned long long p2;  // two fp32 lanes (node A, node B)
__device__ __forceinline__ p2 pk(float a, float b) { p2 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ p2 ffma2(p2 a, p2 b, p2 c) { p2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ p2 fmul2(p2 a, p2 b) { p2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ void unpk(p2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }

#define Q(n) const float4 q##n = T4[n];
#define GM(M) g = pk(dsin[(M) * NT], dsin[(M) * NT + TPC]);
#define GI(i) gi = fmul2(g, x[i]);
#define S1(c) m1 = pk((c), (c));
#define Z1() m1 = 0ull;
#define GJ(j) gij = fmul2(gi, x[j]);
#define S2(c) m2 = pk((c), (c));
#define Z2() m2 = 0ull;
#define B3(c, t) m2 = ffma2(pk((c), (c)), x[t], m2); dx[t] = ffma2(pk((c), (c)), gij, dx[t]);
#define E2(j) m1 = ffma2(m2, x[j], m1); dx[j] = ffma2(gi, m2, dx[j]);
#define E1(i) dx[i] = ffma2(g, m1, dx[i]);

template <int LMAX, int NU>
struct ProgB;
#define EEG_X(L, N)                                                                                                        \
  template <>                                                                                                              \
  struct ProgB<L, N> {                                                                                                     \
    static constexpr int NCOEF = EEG_SCB_l##L##_nu##N##_NCOEF, D = (L + 1) * (L + 1);                                        \
    static __device__ __forceinline__ void run(const p2* __restrict__ x, p2* __restrict__ dx, const float4* __restrict__ T4, \
                                               const float* __restrict__ dsin) {                                           \
      p2 g, gi, gij, m1, m2;                                                                                               \
      EEG_SCB_l##L##_nu##N##_PROGRAM                                                                                       \
    }                                                                                                                      \
  };
EEG_SC_COMBOS(EEG_X)
#undef EEG_X

template <int LMAX, int NU>
__global__ void __launch_bounds__(THREADS, 1) k_sc_bwd_x(const float* __restrict__ xg, const float* __restrict__ dog,
                                                         const float* __restrict__ T, float* __restrict__ dxg, int64_t N, int C,
                                                         int tiles_per_cta) {
  using F = ProgB<LMAX, NU>;
  constexpr int D = F::D, NCOEF = F::NCOEF;
  extern __shared__ __align__(16) float smem[];
  float* Ts = smem;                     // [CH][NCOEF]
  float* xs = Ts + CH * NCOEF;          // [CH][D][NT]   x tile, then the dx tile
  float* ds = xs + CH * D * NT;         // [CH][D][NT]   dout tile
  const int t = threadIdx.x;
  const int c0 = blockIdx.x * CH;
  {
    const float4* src = reinterpret_cast<const float4*>(T + (int64_t)c0 * NCOEF);
    float4* dst = reinterpret_cast<float4*>(Ts);
    for (int i = t; i < CH * NCOEF / 4; i += THREADS) dst[i] = __ldg(src + i);
  }
  const int nl = t % TPC, cc = t / TPC;
  const int64_t n_tiles = (N + NT - 1) / NT;
  const int64_t tile_begin = (int64_t)blockIdx.y * tiles_per_cta;
  for (int64_t tile = tile_begin; tile < tile_begin + tiles_per_cta && tile < n_tiles; ++tile) {
    const int64_t n0 = tile * NT;
    __syncthreads();
    for (int idx = t; idx < NT * D; idx += THREADS) {
      const int n_l = idx % NT, d = idx / NT;
      float2 v = make_float2(0.f, 0.f), w = make_float2(0.f, 0.f);
      if (n0 + n_l < N) {
        v = __ldg(reinterpret_cast<const float2*>(xg + ((n0 + n_l) * D + d) * (int64_t)C + c0));
        w = __ldg(reinterpret_cast<const float2*>(dog + ((n0 + n_l) * D + d) * (int64_t)C + c0));
      }
      xs[(0 * D + d) * NT + n_l] = v.x;
      xs[(1 * D + d) * NT + n_l] = v.y;
      ds[(0 * D + d) * NT + n_l] = w.x;
      ds[(1 * D + d) * NT + n_l] = w.y;
    }
    __syncthreads();
    p2 xv[D], dxv[D];
    float* xin = xs + cc * D * NT + nl;
#pragma unroll
    for (int d = 0; d < D; ++d) {
      xv[d] = pk(xin[d * NT], xin[d * NT + TPC]);
      dxv[d] = 0ull;
    }
    F::run(xv, dxv, reinterpret_cast<const float4*>(Ts + cc * NCOEF), ds + cc * D * NT + nl);
#pragma unroll
    for (int d = 0; d < D; ++d) {  // this thread's own x slots become its dx
      float a, b;
      unpk(dxv[d], a, b);
      xin[d * NT] = a;
      xin[d * NT + TPC] = b;
    }
    __syncthreads();
    for (int idx = t; idx < NT * D; idx += THREADS) {
      const int n_l = idx % NT, d = idx / NT;
      if (n0 + n_l < N)
        *reinterpret_cast<float2*>(dxg + ((n0 + n_l) * D + d) * (int64_t)C + c0) =
            make_float2(xs[(0 * D + d) * NT + n_l], xs[(1 * D + d) * NT + n_l]);
    }
  }
}

template <int LMAX, int NU>
int launch_bwd_x(int C, const float* x, const float* dout, const float* T, float* dx, int64_t N, cudaStream_t s) {
  using F = ProgB<LMAX, NU>;
  const size_t smem = (size_t)(CH * F::NCOEF + 2 * CH * F::D * NT) * sizeof(float);
  static EegDevFlags attr_set_dev; bool& attr_set = eeg_dev_flag(attr_set_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_sc_bwd_x<LMAX, NU>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_sym_contract_bwd_x: cannot opt in to %zu B shared memory: %s", smem, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int groups = C / CH;
  const int64_t n_tiles = (N + NT - 1) / NT;
  int slices = EEG_NUM_SMS / groups;
  if (slices > n_tiles) slices = (int)n_tiles;
  if (slices < 1) slices = 1;
  const int tiles_per_cta = (int)((n_tiles + slices - 1) / slices);
  slices = (int)((n_tiles + tiles_per_cta - 1) / tiles_per_cta);
  k_sc_bwd_x<LMAX, NU><<<dim3(groups, slices), THREADS, smem, s>>>(x, dout, T, dx, N, C, tiles_per_cta);
  EEG_CHECK_LAUNCH("eeg_sym_contract_bwd_x");
  return 0;
}

}  // namespace

extern "C" int eeg_sym_contract_bwd_x(int lmax, int nu, int C, const float* x, const float* dout, const float* T, float* dx,
                                      int64_t N, void* stream) {
  EEG_REQUIRE(nu >= 1 && nu <= 3 && lmax >= 1 && lmax <= 4, "eeg_sym_contract_bwd_x: unsupported (lmax=%d, correlation=%d)", lmax, nu);
  EEG_REQUIRE(C > 0 && C % CH == 0, "eeg_sym_contract_bwd_x: C must be a multiple of %d", CH);
  if (N == 0) return 0;
#define EEG_X(L, NN) \
  if (lmax == L && nu == NN) return launch_bwd_x<L, NN>(C, x, dout, T, dx, N, (cudaStream_t)stream);
  EEG_SC_COMBOS(EEG_X)
#undef EEG_X
  return 1;
}
