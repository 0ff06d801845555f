// Symmetric contraction (MACE product basis), gnn/mace.py:112-277, as a per-(node, channel) cubic polynomial map
// R^D -> R^D in Horner form.  Never materialises the reference's [N, C, D, D, D] intermediate.
//   eeg_sc_coefficients : T[c, slot] = sum_k A[slot, k] * Wflat[k, c]    (A = symmetrised generalised CG, CSR)
//   eeg_sym_contract_fwd: kernel body in eeg_sc_impl.cuh, compiled once per variant (eeg_sc_v_*.cu): lane = node, the
//                         coefficient stream is warp-uniform (broadcast float4 from shared memory); the default
//                         variant keeps node pairs in 64-bit registers and issues packed FFMA2 (fma.rn.f32x2).
#include <stdlib.h>
#include <string.h>
#include "eeg_common.cuh"

namespace {

// block = 32 slots x 32 channels (C <= 32 per pass): thread (x = channel, y) evaluates slots y, y + 8, ..; the tile is then
// written transposed so that a warp stores 32 consecutive slots of one channel (T is channel-major: the first version's
// lane = channel store touched 32 lines per instruction and took 0.33 ms per layer and training step)
__global__ void __launch_bounds__(256) k_sc_coefficients(const int32_t* __restrict__ rp, const int32_t* __restrict__ col,
                                                         const float* __restrict__ val, int n_coef, const float* __restrict__ Wflat, int C,
                                                         float* __restrict__ T) {
  __shared__ float tile[32][33];
  const int slot0 = blockIdx.x * 32;
  for (int c0 = 0; c0 < C; c0 += 32) {
    const int c = c0 + threadIdx.x;
    for (int sl = threadIdx.y; sl < 32; sl += 8) {
      const int slot = slot0 + sl;
      float s = 0.f;
      if (slot < n_coef && c < C)
        for (int q = __ldg(rp + slot); q < __ldg(rp + slot + 1); ++q) s = fmaf(__ldg(val + q), __ldg(Wflat + (int64_t)__ldg(col + q) * C + c), s);
      tile[sl][threadIdx.x] = s;
    }
    __syncthreads();
    for (int cc = threadIdx.y; cc < 32; cc += 8)
      if (slot0 + threadIdx.x < n_coef && c0 + cc < C) T[(int64_t)(c0 + cc) * n_coef + slot0 + threadIdx.x] = tile[threadIdx.x][cc];
    __syncthreads();
  }
}

}  // namespace

int eeg_sc_launch_s1(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s);
int eeg_sc_launch_p2(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s);
int eeg_sc_launch_p2a(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s);

extern "C" int eeg_sc_coefficients(const int32_t* a_row_ptr, const int32_t* a_col, const float* a_val, int n_coef,
                                   const float* Wflat, int C, float* T, void* stream) {
  if (n_coef == 0) return 0;
  dim3 block(32, 8), grid(eeg_div_up(n_coef, 32));
  k_sc_coefficients<<<grid, block, 0, (cudaStream_t)stream>>>(a_row_ptr, a_col, a_val, n_coef, Wflat, C, T);
  EEG_CHECK_LAUNCH("eeg_sc_coefficients");
  return 0;
}

extern "C" int eeg_sym_contract_fwd(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, void* stream) {
  EEG_REQUIRE(nu >= 1 && nu <= 3, "eeg_sym_contract_fwd: unsupported correlation=%d (generated: 1..3)", nu);
  EEG_REQUIRE(lmax >= 1 && lmax <= 4, "eeg_sym_contract_fwd: unsupported lmax=%d (generated: 1..4)", lmax);
  if (N == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  // Variants (eeg_sc_v_*.cu) run the same fp32 FMA sequence per node => bit-identical results.  Measured on B200
  // (N = 30720, C = 32): lmax 4: s1 647 us, p2 525 us, p2a 492 us; lmax 3: s1 203 us, p2 148 us, p2a 196 us.
  static const char* env = getenv("EEG_SC_VARIANT");  // tuning / A-B knob
  const int v = !env ? (lmax == 4 && nu == 3 ? 2 : 1) : !strcmp(env, "s1") ? 0 : !strcmp(env, "p2") ? 1 : 2;
  if (v == 0) return eeg_sc_launch_s1(lmax, nu, C, x, T, out, N, s);
  if (v == 1) return eeg_sc_launch_p2(lmax, nu, C, x, T, out, N, s);
  return eeg_sc_launch_p2a(lmax, nu, C, x, T, out, N, s);
}
