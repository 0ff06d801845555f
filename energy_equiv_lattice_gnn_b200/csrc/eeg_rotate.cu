// RotateLat on device (scripts/train_utils.py:114-146): one rotation Q_g per lattice applied to a collated batch.
//   positions'[n] = Q[batch[n]] . positions[n]               ('ij,...j->...i', :124)
//   shifts'[e]    = Q[batch[sender(e)]] . shifts[e]           (:125)
//   C'_abcd       = Q_ai Q_bj Q_ck Q_dl C_ijkl -> Mandel 6x6  (:122-123, :133-134;
//                   lattices/lattices/utils/elasticity_func.py:406-428 picks (a,b,c,d) for the upper triangle, mirrors it
//                   and scales by {1, sqrt2, 2})
// The reference does this per lattice on the host inside the DataLoader; here the whole batch is rotated by two tiny
// launches after the upload.
#include "eeg_common.cuh"

namespace {

// thread = row; graph id = graph_of[via ? via[i] : i]
__global__ void k_rotate_vec3(const float* __restrict__ Q, const int64_t* __restrict__ graph_of, const int64_t* __restrict__ via,
                              const float* __restrict__ in, float* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t g = graph_of[via ? via[i] : i];
  const float* q = Q + g * 9;
  const float x = in[3 * i], y = in[3 * i + 1], z = in[3 * i + 2];
  out[3 * i] = fmaf(__ldg(q + 0), x, fmaf(__ldg(q + 1), y, __ldg(q + 2) * z));
  out[3 * i + 1] = fmaf(__ldg(q + 3), x, fmaf(__ldg(q + 4), y, __ldg(q + 5) * z));
  out[3 * i + 2] = fmaf(__ldg(q + 6), x, fmaf(__ldg(q + 7), y, __ldg(q + 8) * z));
}

constexpr int RW = 4;  // warps (lattices) per CTA

// warp = lattice: four single-index rotations of the 81-entry tensor through shared memory (4 x 243 FMAs instead of
// 81 x 81), then the 36 Mandel entries
__global__ void __launch_bounds__(RW * 32) k_rotate_cart4_to_mandel(const float* __restrict__ Q, const float* __restrict__ C4,
                                                                    float* __restrict__ out, int64_t G) {
  __shared__ float buf[RW][2][81];
  __shared__ float qs[RW][9];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t g = (int64_t)blockIdx.x * RW + w;
  if (g >= G) return;
  if (lane < 9) qs[w][lane] = Q ? Q[g * 9 + lane] : (lane % 4 == 0 ? 1.f : 0.f);
  for (int t = lane; t < 81; t += 32) buf[w][0][t] = C4[g * 81 + t];
  __syncwarp();
  // pass p rotates index p (stride 27, 9, 3, 1): dst[.., a, ..] = sum_i Q[a,i] src[.., i, ..]
  int stride = 27;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const float* src = buf[w][p & 1];
    float* dst = buf[w][(p & 1) ^ 1];
    for (int t = lane; t < 81; t += 32) {
      const int a = (t / stride) % 3, base = t - a * stride;
      dst[t] = fmaf(qs[w][a * 3], src[base], fmaf(qs[w][a * 3 + 1], src[base + stride], qs[w][a * 3 + 2] * src[base + 2 * stride]));
    }
    __syncwarp();
    stride /= 3;
  }
  const float* T = buf[w][0];
  for (int t = lane; t < 36; t += 32) {
    int i = t / 6, j = t % 6;
    const float m = (i < 3 && j < 3) ? 1.f : ((i >= 3 && j >= 3) ? 2.f : 1.41421356237309515f);
    if (i > j) { const int s = i; i = j; j = s; }  // lower triangle mirrors the upper one
    // Mandel pair of index 0..5: (0,0) (1,1) (2,2) (1,2) (0,2) (0,1)
    const int a = i < 3 ? i : (i == 3 ? 1 : 0), b = i < 3 ? i : (i == 5 ? 1 : 2);
    const int c = j < 3 ? j : (j == 3 ? 1 : 0), d = j < 3 ? j : (j == 5 ? 1 : 2);
    out[g * 36 + t] = T[a * 27 + b * 9 + c * 3 + d] * m;
  }
}

}  // namespace

extern "C" int eeg_rotate_vec3(const float* Q, const int64_t* graph_of, const int64_t* via, const float* in, float* out, int64_t n,
                               void* stream) {
  if (n == 0) return 0;
  EEG_REQUIRE(Q && graph_of && in && out, "eeg_rotate_vec3: null pointer");
  k_rotate_vec3<<<eeg_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(Q, graph_of, via, in, out, n);
  EEG_CHECK_LAUNCH("eeg_rotate_vec3");
  return 0;
}

extern "C" int eeg_rotate_cart4_to_mandel(const float* Q, const float* C4, float* out, int64_t G, void* stream) {
  if (G == 0) return 0;
  EEG_REQUIRE(C4 && out, "eeg_rotate_cart4_to_mandel: null pointer");
  k_rotate_cart4_to_mandel<<<eeg_div_up(G, RW), RW * 32, 0, (cudaStream_t)stream>>>(Q, C4, out, G);
  EEG_CHECK_LAUNCH("eeg_rotate_cart4_to_mandel");
  return 0;
}
