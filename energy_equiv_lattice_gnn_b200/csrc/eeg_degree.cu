// interaction_reduction = 'mean' (gnn/blocks.py:595-597, torch_scatter.scatter(reduce='mean')): the summed messages of a
// receiver are divided by its in-degree (clamped to >= 1, as torch_scatter does).  Non-default variant of the reference
// (train_main.py:31 uses 'sum'), so it is a separate in-place pass over M[N, row_floats] after eeg_tp_scatter_fwd instead
// of another specialisation of the generated TP kernel; the same diagonal scaling applied to dM is its backward.
#include "eeg_common.cuh"

namespace {

__global__ void __launch_bounds__(256) k_scale_rows_inv_degree(float4* __restrict__ x, const int32_t* __restrict__ row_ptr, int64_t N,
                                                               int row_vec4) {
  for (int64_t n = blockIdx.x; n < N; n += gridDim.x) {
    const int deg = __ldg(row_ptr + n + 1) - __ldg(row_ptr + n);
    if (deg <= 1) continue;
    const float sc = 1.0f / (float)deg;
    float4* row = x + n * row_vec4;
    for (int i = threadIdx.x; i < row_vec4; i += blockDim.x) {
      float4 v = row[i];
      v.x *= sc;
      v.y *= sc;
      v.z *= sc;
      v.w *= sc;
      row[i] = v;
    }
  }
}

}  // namespace

extern "C" int eeg_scale_rows_inv_degree(float* x, const int32_t* row_ptr, int64_t N, int64_t row_floats, void* stream) {
  if (N == 0 || row_floats == 0) return 0;
  EEG_REQUIRE(x && row_ptr, "eeg_scale_rows_inv_degree: null pointer");
  EEG_REQUIRE(row_floats % 4 == 0 && ((uintptr_t)x & 15) == 0, "eeg_scale_rows_inv_degree: rows must be 16-byte aligned multiples of 4 floats");
  const int64_t grid = N < 148 * 16 ? N : 148 * 16;
  k_scale_rows_inv_degree<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(x), row_ptr, N, (int)(row_floats / 4));
  EEG_CHECK_LAUNCH("eeg_scale_rows_inv_degree");
  return 0;
}

// [E, B, 32] (edge-major per-edge tensor-product weights, as a generic radial MLP writes them) -> [B, E, 32] (path-major w_t,
// the layout eeg_tp_scatter_fwd / _bwd read).  Only used when the radial MLP is outside the fused tcgen05 kernel's shape
// (inter_MLP_dim != 64, more than 16 edge bases, a single layer).
namespace {
__global__ void k_blocks32_transpose(const float4* __restrict__ in, float4* __restrict__ out, int64_t E, int B, int to_path_major) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;   // float4 index of the destination
  if (i >= E * B * 8) return;
  const int q = (int)(i & 7);
  const int64_t r = i >> 3;
  if (to_path_major) {   // dst [B, E]: r = b * E + e
    const int64_t b = r / E, e = r - b * E;
    out[i] = __ldg(in + (e * B + b) * 8 + q);
  } else {               // dst [E, B]: r = e * B + b
    const int64_t e = r / B, b = r - e * B;
    out[i] = __ldg(in + (b * E + e) * 8 + q);
  }
}
}  // namespace

extern "C" int eeg_blocks32_transpose(const float* in, float* out, int64_t E, int B, int to_path_major, void* stream) {
  if (E == 0 || B == 0) return 0;
  const int64_t tot = E * B * 8;
  k_blocks32_transpose<<<(unsigned)eeg_div_up(tot, 256), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(in),
                                                                                          reinterpret_cast<float4*>(out), E, B, to_path_major);
  EEG_CHECK_LAUNCH("eeg_blocks32_transpose");
  return 0;
}
