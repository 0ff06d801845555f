// Validation / test metrics on device (scripts/train_utils.py:66-89 validation_step, :149-172 obtain_errors): per lattice
//   mse, l1        mean over the 36 Mandel entries of (pred - target)^2 and |pred - target|          (:157-158, :76)
//   dir_loss       mean over P directions of |c(d) - c_pred(d)|, c(d) = C_ijkl d_i d_j d_k d_l          (:161-163, :80-84)
//   mean_stiffness mean over P directions of the target's c(d)                                         (:164)
//   eigenvalues    torch.linalg.eigvalsh of both 6x6 matrices (ascending, lower triangle)               (:166-170)
// The reference expands Mandel -> Cartesian rank 4 (lattices/lattices/utils/elasticity_func.py:430-463: upper triangle of
// the 6x6, fully symmetrised, off-diagonal Mandel rows divided by sqrt2) and contracts 81 terms per direction; the same
// number is v^T U v with v = (d0^2, d1^2, d2^2, sqrt2 d1 d2, sqrt2 d0 d2, sqrt2 d0 d1) and U the upper triangle counted
// twice off the diagonal, 21 terms.
#include "eeg_common.cuh"
#include "eeg_eig6.cuh"

namespace {

constexpr int MW = 4;  // lattices (warps) per CTA

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// warp = lattice; lanes stride over directions and over the 36 entries
__global__ void __launch_bounds__(MW * 32) k_validation_metrics(const float* __restrict__ pred, const float* __restrict__ target,
                                                                const float* __restrict__ dirs, int P, int64_t G,
                                                                float* __restrict__ per_lattice, float* __restrict__ eig_pred,
                                                                float* __restrict__ eig_target) {
  __shared__ float sm[MW][2][36];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t g = (int64_t)blockIdx.x * MW + w;
  if (g >= G) return;
  float se = 0.f, ae = 0.f;
  for (int t = lane; t < 36; t += 32) {
    const float p = pred[g * 36 + t], q = target[g * 36 + t];
    sm[w][0][t] = p;
    sm[w][1][t] = q;
    se = fmaf(p - q, p - q, se);
    ae += fabsf(p - q);
  }
  __syncwarp();
  const float* Cp = sm[w][0];
  const float* Ct = sm[w][1];
  float dl = 0.f, ms = 0.f;
  for (int p = lane; p < P; p += 32) {
    const float d0 = dirs[3 * p], d1 = dirs[3 * p + 1], d2 = dirs[3 * p + 2];
    const float r2 = 1.41421356237309515f;
    const float v[6] = {d0 * d0, d1 * d1, d2 * d2, r2 * d1 * d2, r2 * d0 * d2, r2 * d0 * d1};
    float cp = 0.f, ct = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = i; j < 6; ++j) {
        const float vv = (i == j ? 1.f : 2.f) * v[i] * v[j];
        cp = fmaf(Cp[i * 6 + j], vv, cp);
        ct = fmaf(Ct[i * 6 + j], vv, ct);
      }
    dl += fabsf(ct - cp);
    ms += ct;
  }
  se = warp_sum(se);
  ae = warp_sum(ae);
  dl = warp_sum(dl);
  ms = warp_sum(ms);
  if (lane == 0) {
    float* o = per_lattice + g * 4;
    o[0] = se * (1.f / 36.f);
    o[1] = ae * (1.f / 36.f);
    o[2] = P > 0 ? dl / (float)P : 0.f;
    o[3] = P > 0 ? ms / (float)P : 0.f;
  }
  if (lane == 0 && eig_pred) eeg_eigvalsh6(Cp, eig_pred + g * 6);
  if (lane == 1 && eig_target) eeg_eigvalsh6(Ct, eig_target + g * 6);
}

}  // namespace

extern "C" int eeg_validation_metrics(const float* pred, const float* target, const float* directions, int P, int64_t G,
                                      float* per_lattice, float* eig_pred, float* eig_target, void* stream) {
  if (G == 0) return 0;
  EEG_REQUIRE(pred && target && per_lattice && (P == 0 || directions), "eeg_validation_metrics: null pointer");
  EEG_REQUIRE(P >= 0, "eeg_validation_metrics: negative direction count %d", P);
  k_validation_metrics<<<eeg_div_up(G, MW), MW * 32, 0, (cudaStream_t)stream>>>(pred, target, directions, P, G, per_lattice, eig_pred,
                                                                             eig_target);
  EEG_CHECK_LAUNCH("eeg_validation_metrics");
  return 0;
}
