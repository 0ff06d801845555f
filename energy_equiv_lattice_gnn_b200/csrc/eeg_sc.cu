// Symmetric contraction (MACE product basis), gnn/mace.py:112-277, as a per-(node, channel) cubic polynomial map
// R^D -> R^D in Horner form.  Never materialises the reference's [N, C, D, D, D] intermediate.
//   eeg_sc_coefficients : T[c, slot] = sum_k A[slot, k] * Wflat[k, c]    (A = symmetrised generalised CG, CSR)
//   eeg_sym_contract_fwd: CTA = 4 channels x 128-node tiles; lane = node, so the coefficient stream is warp-uniform
//                         and is read from shared memory as broadcast float4 (4 coefficients per LDS).
#include <stdlib.h>
#include "eeg_common.cuh"
#include "generated/sc_l3_nu3.cuh"
#include "generated/sc_l4_nu3.cuh"

namespace {

constexpr int SC_CH = 4;      // channels per CTA
constexpr int SC_TPC = 128;   // threads per channel
constexpr int SC_THREADS = SC_CH * SC_TPC;

__global__ void k_sc_coefficients(const int32_t* __restrict__ rp, const int32_t* __restrict__ col, const float* __restrict__ val,
                                  int n_coef, const float* __restrict__ Wflat, int C, float* __restrict__ T) {
  int slot = blockIdx.x * blockDim.y + threadIdx.y;
  if (slot >= n_coef) return;
  for (int c = threadIdx.x; c < C; c += 32) {
    float s = 0.f;
    for (int q = rp[slot]; q < rp[slot + 1]; ++q) s = fmaf(__ldg(val + q), __ldg(Wflat + (int64_t)__ldg(col + q) * C + c), s);
    T[(int64_t)c * n_coef + slot] = s;
  }
}

// The generated macro program is expanded once per NPT (nodes per thread): every coefficient (one broadcast LDS.128
// per 4) feeds NPT independent Horner chains.  Statements are spelled out per chain (no loops: 10^4 tiny unrolled
// loops make the front end crawl).
template <int LMAX, int NU>
struct ScFn;
#define Q(n) const float4 q##n = T4[n];

// ---- NPT = 1 ----
#define S2(c) m20 = (c);
#define F3I(c, t) m20 = (c) * x0[t];
#define F3(c, t) m20 = fmaf((c), x0[t], m20);
#define S1(c) m10 = (c);
#define F2I(j) m10 = m20 * x0[j];
#define F2(j) m10 = fmaf(m20, x0[j], m10);
#define Z() a0 = 0.0f;
#define F1(i) a0 = fmaf(m10, x0[i], a0);
#define O(o) out[(o) * ostride] = a0;
template <int LMAX>
__device__ __forceinline__ void sc_run1(const float* x0, const float4* __restrict__ T4, float* __restrict__ out, int ostride);
template <>
__device__ __forceinline__ void sc_run1<4>(const float* x0, const float4* __restrict__ T4, float* __restrict__ out, int ostride) {
  float a0, m10, m20;
  EEG_SC_l4_nu3_PROGRAM
}
template <>
__device__ __forceinline__ void sc_run1<3>(const float* x0, const float4* __restrict__ T4, float* __restrict__ out, int ostride) {
  float a0, m10, m20;
  EEG_SC_l3_nu3_PROGRAM
}
#undef S2
#undef F3I
#undef F3
#undef S1
#undef F2I
#undef F2
#undef Z
#undef F1
#undef O
// ---- NPT = 2 ----
#define S2(c) m20 = (c); m21 = (c);
#define F3I(c, t) m20 = (c) * x0[t]; m21 = (c) * x1[t];
#define F3(c, t) m20 = fmaf((c), x0[t], m20); m21 = fmaf((c), x1[t], m21);
#define S1(c) m10 = (c); m11 = (c);
#define F2I(j) m10 = m20 * x0[j]; m11 = m21 * x1[j];
#define F2(j) m10 = fmaf(m20, x0[j], m10); m11 = fmaf(m21, x1[j], m11);
#define Z() a0 = 0.0f; a1 = 0.0f;
#define F1(i) a0 = fmaf(m10, x0[i], a0); a1 = fmaf(m11, x1[i], a1);
#define O(o) out[(o) * ostride] = a0; out[(o) * ostride + vstride] = a1;
template <int LMAX>
__device__ __forceinline__ void sc_run2(const float* x0, const float* x1, const float4* __restrict__ T4, float* __restrict__ out,
                                        int ostride, int vstride);
template <>
__device__ __forceinline__ void sc_run2<4>(const float* x0, const float* x1, const float4* __restrict__ T4, float* __restrict__ out,
                                           int ostride, int vstride) {
  float a0, a1, m10, m11, m20, m21;
  EEG_SC_l4_nu3_PROGRAM
}
template <>
__device__ __forceinline__ void sc_run2<3>(const float* x0, const float* x1, const float4* __restrict__ T4, float* __restrict__ out,
                                           int ostride, int vstride) {
  float a0, a1, m10, m11, m20, m21;
  EEG_SC_l3_nu3_PROGRAM
}
#undef S2
#undef F3I
#undef F3
#undef S1
#undef F2I
#undef F2
#undef Z
#undef F1
#undef O
#undef Q

template <>
struct ScFn<4, 3> {
  static constexpr int NCOEF = EEG_SC_l4_nu3_NCOEF, D = 25, L = 4;
};
template <>
struct ScFn<3, 3> {
  static constexpr int NCOEF = EEG_SC_l3_nu3_NCOEF, D = 16, L = 3;
};

// CTA = SC_CH channels x (SC_TPC * NPT) nodes per tile; thread = (channel, NPT nodes); lane = node so the coefficient
// stream is warp-uniform.  x tile is transposed through shared memory (coalesced float4 global access both ways).
template <int LMAX, int NU, int NPT>
__global__ void __launch_bounds__(SC_THREADS, 1) k_sym_contract(const float* __restrict__ x, const float* __restrict__ T,
                                                                float* __restrict__ out, int64_t N, int C, int tiles_per_cta) {
  using F = ScFn<LMAX, NU>;
  constexpr int D = F::D, NCOEF = F::NCOEF, NT = SC_TPC * NPT;
  extern __shared__ __align__(16) float smem[];
  float* Ts = smem;                     // [SC_CH][NCOEF]
  float* xs = smem + SC_CH * NCOEF;     // [SC_CH][D][NT]
  const int t = threadIdx.x;
  const int c0 = blockIdx.x * SC_CH;
  {
    const float4* src = reinterpret_cast<const float4*>(T + (int64_t)c0 * NCOEF);
    float4* dst = reinterpret_cast<float4*>(Ts);
    for (int i = t; i < SC_CH * NCOEF / 4; i += SC_THREADS) dst[i] = __ldg(src + i);
  }
  const int nl = t % SC_TPC, cc = t / SC_TPC;
  const int64_t n_tiles = (N + NT - 1) / NT;
  const int64_t tile_begin = (int64_t)blockIdx.y * tiles_per_cta;
  for (int64_t tile = tile_begin; tile < tile_begin + tiles_per_cta && tile < n_tiles; ++tile) {
    const int64_t n0 = tile * NT;
    __syncthreads();  // previous tile's stores done / Ts visible
    for (int idx = t; idx < NT * D; idx += SC_THREADS) {
      int n_l = idx % NT, d = idx / NT;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n0 + n_l < N) v = __ldg(reinterpret_cast<const float4*>(x + ((n0 + n_l) * D + d) * (int64_t)C + c0));
      xs[(0 * D + d) * NT + n_l] = v.x;
      xs[(1 * D + d) * NT + n_l] = v.y;
      xs[(2 * D + d) * NT + n_l] = v.z;
      xs[(3 * D + d) * NT + n_l] = v.w;
    }
    __syncthreads();
    float xv[NPT][D];
#pragma unroll
    for (int v = 0; v < NPT; ++v)
#pragma unroll
      for (int d = 0; d < D; ++d) xv[v][d] = xs[(cc * D + d) * NT + v * SC_TPC + nl];
    // outputs overwrite this thread's own x slots (its inputs are in registers)
    if constexpr (NPT == 1) sc_run1<F::L>(xv[0], reinterpret_cast<const float4*>(Ts + cc * NCOEF), xs + cc * D * NT + nl, NT);
    else sc_run2<F::L>(xv[0], xv[1], reinterpret_cast<const float4*>(Ts + cc * NCOEF), xs + cc * D * NT + nl, NT, SC_TPC);
    __syncthreads();
    for (int idx = t; idx < NT * D; idx += SC_THREADS) {
      int n_l = idx % NT, d = idx / NT;
      if (n0 + n_l < N) {
        float4 v = make_float4(xs[(0 * D + d) * NT + n_l], xs[(1 * D + d) * NT + n_l], xs[(2 * D + d) * NT + n_l],
                               xs[(3 * D + d) * NT + n_l]);
        *reinterpret_cast<float4*>(out + ((n0 + n_l) * D + d) * (int64_t)C + c0) = v;
      }
    }
  }
}

template <int LMAX, int NU, int NPT>
int launch_sc(int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s) {
  using F = ScFn<LMAX, NU>;
  constexpr int NT = SC_TPC * NPT;
  const size_t smem = (size_t)(SC_CH * F::NCOEF + SC_CH * F::D * NT) * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_sym_contract<LMAX, NU, NPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_sym_contract_fwd: cannot opt in to %zu B shared memory: %s", smem, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int groups = C / SC_CH;
  const int64_t n_tiles = (N + NT - 1) / NT;
  int slices = (EEG_NUM_SMS + groups - 1) / groups;
  if (slices > n_tiles) slices = (int)n_tiles;
  if (slices < 1) slices = 1;
  const int tiles_per_cta = (int)((n_tiles + slices - 1) / slices);
  slices = (int)((n_tiles + tiles_per_cta - 1) / tiles_per_cta);
  dim3 grid(groups, slices);
  k_sym_contract<LMAX, NU, NPT><<<grid, SC_THREADS, smem, s>>>(x, T, out, N, C, tiles_per_cta);
  EEG_CHECK_LAUNCH("eeg_sym_contract_fwd");
  return 0;
}

}  // namespace

extern "C" int eeg_sc_coefficients(const int32_t* a_row_ptr, const int32_t* a_col, const float* a_val, int n_coef,
                                   const float* Wflat, int C, float* T, void* stream) {
  if (n_coef == 0) return 0;
  dim3 block(32, 8), grid(eeg_div_up(n_coef, 8));
  k_sc_coefficients<<<grid, block, 0, (cudaStream_t)stream>>>(a_row_ptr, a_col, a_val, n_coef, Wflat, C, T);
  EEG_CHECK_LAUNCH("eeg_sc_coefficients");
  return 0;
}

extern "C" int eeg_sym_contract_fwd(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, void* stream) {
  EEG_REQUIRE(C > 0 && C % SC_CH == 0, "eeg_sym_contract_fwd: C must be a multiple of %d", SC_CH);
  if (N == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  static const int npt = getenv("EEG_SC_NPT") ? atoi(getenv("EEG_SC_NPT")) : 2;  // tuning knob (nodes per thread)
  if (lmax == 4 && nu == 3) return npt == 1 ? launch_sc<4, 3, 1>(C, x, T, out, N, s) : launch_sc<4, 3, 2>(C, x, T, out, N, s);
  if (lmax == 3 && nu == 3) return npt == 1 ? launch_sc<3, 3, 1>(C, x, T, out, N, s) : launch_sc<3, 3, 2>(C, x, T, out, N, s);
  eeg_set_error("eeg_sym_contract_fwd: unsupported (lmax=%d, correlation=%d)", lmax, nu);
  return 1;
}
