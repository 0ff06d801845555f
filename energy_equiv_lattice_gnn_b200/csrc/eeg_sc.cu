// Symmetric contraction (MACE product basis), gnn/mace.py:112-277, as a per-(node, channel) cubic polynomial map
// R^D -> R^D in Horner form.  Never materialises the reference's [N, C, D, D, D] intermediate.
//   eeg_sc_coefficients : T[c, slot] = sum_k A[slot, k] * Wflat[k, c]    (A = symmetrised generalised CG, CSR)
//   eeg_sym_contract_fwd: CTA = 4 channels x 128-node tiles; lane = node, so the coefficient stream is warp-uniform
//                         and is read from shared memory as broadcast float4 (4 coefficients per LDS).
#include "eeg_common.cuh"
#include "generated/sc_l3_nu3.cuh"
#include "generated/sc_l4_nu3.cuh"

namespace {

constexpr int SC_CH = 4;      // channels per CTA
constexpr int SC_NT = 128;    // nodes per tile
constexpr int SC_THREADS = SC_CH * SC_NT;

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

template <int LMAX, int NU>
struct ScFn;
template <>
struct ScFn<4, 3> {
  static constexpr int NCOEF = EEG_SC_l4_nu3_NCOEF, D = 25;
  static __device__ __forceinline__ void run(const float* x, const float4* T4, float* o) { eeg_sc_l4_nu3(x, T4, o); }
};
template <>
struct ScFn<3, 3> {
  static constexpr int NCOEF = EEG_SC_l3_nu3_NCOEF, D = 16;
  static __device__ __forceinline__ void run(const float* x, const float4* T4, float* o) { eeg_sc_l3_nu3(x, T4, o); }
};

template <int LMAX, int NU>
__global__ void __launch_bounds__(SC_THREADS, 1) k_sym_contract(const float* __restrict__ x, const float* __restrict__ T,
                                                                float* __restrict__ out, int64_t N, int C, int tiles_per_cta) {
  using F = ScFn<LMAX, NU>;
  constexpr int D = F::D, NCOEF = F::NCOEF;
  extern __shared__ __align__(16) float smem[];
  float* Ts = smem;                     // [SC_CH][NCOEF]
  float* xs = smem + SC_CH * NCOEF;     // [SC_CH][D][SC_NT]
  const int t = threadIdx.x;
  const int c0 = blockIdx.x * SC_CH;
  // coefficient streams of this CTA's channels (contiguous in T[c][slot])
  {
    const float4* src = reinterpret_cast<const float4*>(T + (int64_t)c0 * NCOEF);
    float4* dst = reinterpret_cast<float4*>(Ts);
    for (int i = t; i < SC_CH * NCOEF / 4; i += SC_THREADS) dst[i] = __ldg(src + i);
  }
  const int nl = t % SC_NT, cc = t / SC_NT;
  const int64_t n_tiles = (N + SC_NT - 1) / SC_NT;
  const int64_t tile_begin = (int64_t)blockIdx.y * tiles_per_cta;
  for (int64_t tile = tile_begin; tile < tile_begin + tiles_per_cta && tile < n_tiles; ++tile) {
    const int64_t n0 = tile * SC_NT;
    __syncthreads();  // previous tile's stores done / Ts visible
    for (int idx = t; idx < SC_NT * D; idx += SC_THREADS) {
      int n_l = idx % SC_NT, d = idx / SC_NT;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n0 + n_l < N) v = __ldg(reinterpret_cast<const float4*>(x + ((n0 + n_l) * D + d) * (int64_t)C + c0));
      xs[(0 * D + d) * SC_NT + n_l] = v.x;
      xs[(1 * D + d) * SC_NT + n_l] = v.y;
      xs[(2 * D + d) * SC_NT + n_l] = v.z;
      xs[(3 * D + d) * SC_NT + n_l] = v.w;
    }
    __syncthreads();
    float xv[D], ov[D];
#pragma unroll
    for (int d = 0; d < D; ++d) xv[d] = xs[(cc * D + d) * SC_NT + nl];
    F::run(xv, reinterpret_cast<const float4*>(Ts + cc * NCOEF), ov);
#pragma unroll
    for (int d = 0; d < D; ++d) xs[(cc * D + d) * SC_NT + nl] = ov[d];
    __syncthreads();
    for (int idx = t; idx < SC_NT * D; idx += SC_THREADS) {
      int n_l = idx % SC_NT, d = idx / SC_NT;
      if (n0 + n_l < N) {
        float4 v = make_float4(xs[(0 * D + d) * SC_NT + n_l], xs[(1 * D + d) * SC_NT + n_l], xs[(2 * D + d) * SC_NT + n_l],
                               xs[(3 * D + d) * SC_NT + n_l]);
        *reinterpret_cast<float4*>(out + ((n0 + n_l) * D + d) * (int64_t)C + c0) = v;
      }
    }
  }
}

template <int LMAX, int NU>
int launch_sc(int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s) {
  using F = ScFn<LMAX, NU>;
  const size_t smem = (size_t)(SC_CH * F::NCOEF + SC_CH * F::D * SC_NT) * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_sym_contract<LMAX, NU>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_sym_contract_fwd: cannot opt in to %zu B shared memory: %s", smem, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int groups = C / SC_CH;
  const int64_t n_tiles = (N + SC_NT - 1) / SC_NT;
  int slices = (EEG_NUM_SMS + groups - 1) / groups;
  if (slices > n_tiles) slices = (int)n_tiles;
  if (slices < 1) slices = 1;
  const int tiles_per_cta = (int)((n_tiles + slices - 1) / slices);
  slices = (int)((n_tiles + tiles_per_cta - 1) / tiles_per_cta);
  dim3 grid(groups, slices);
  k_sym_contract<LMAX, NU><<<grid, SC_THREADS, smem, s>>>(x, T, out, N, C, tiles_per_cta);
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
  if (lmax == 4 && nu == 3) return launch_sc<4, 3>(C, x, T, out, N, s);
  if (lmax == 3 && nu == 3) return launch_sc<3, 3>(C, x, T, out, N, s);
  eeg_set_error("eeg_sym_contract_fwd: unsupported (lmax=%d, correlation=%d)", lmax, nu);
  return 1;
}
