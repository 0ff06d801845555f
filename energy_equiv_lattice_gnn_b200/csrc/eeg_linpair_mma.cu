// Two chained per-irrep channel mixes of multiplicity-32 node features in one pass (inference):
//   y1 = W_a x (+ residual)      EquivariantProductBlock.linear of layer i (gnn/blocks.py:471-476,488-490; model.py:95-96)
//   y2 = W_b y1                  TensorProductInteractionBlock.linear_up of layer i+1 (gnn/blocks.py:516-521,589)
// Both results are written ([N, D, 32] each); y1 is not read back: its accumulator fragment becomes the A fragment of the
// second mix by quad shuffles.  Node-major warp-level MMAs, 3xTF32 (eeg_mma.cuh, eeg_readout_mma.cu for the scheme).
#include "eeg_common.cuh"
#include "eeg_mma.cuh"

namespace {

constexpr int LP_WARPS = 16, LP_NB = 32, LP_MAXB = 5;

struct LpBlock {
  int rows, off;
  unsigned inv_rows;
  int wa_sm, wb_sm;
  int64_t wa_off, wb_off;
  float sa, sb;
};
struct LpParams {
  LpBlock b[LP_MAXB];
  int n_groups, stride;
  unsigned char warp_group[LP_WARPS], warp_rank[LP_WARPS], group_warps[LP_MAXB];
};

__global__ void __launch_bounds__(LP_WARPS * 32, 1) k_linpair_mma(const float* __restrict__ X, const float* __restrict__ res,
                                                                 const float* __restrict__ Wa, const float* __restrict__ Wb,
                                                                 float* __restrict__ Y1, float* __restrict__ Y2, int64_t N,
                                                                 LpParams P, int n_chunks) {
  extern __shared__ __align__(16) float lp_smem[];
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3, warp = threadIdx.x >> 5;
  for (int gi = 0; gi < P.n_groups; ++gi) {   // weight fragments: (tile j, k-step s, hh, lane): W[k = 8 s + t + 4 hh][n = 8 j + g] * scale
    const LpBlock& d = P.b[gi];
    for (int e = threadIdx.x; e < 8 * 4 * 2 * 32; e += LP_WARPS * 32) {
      const int ln = e & 31, hh = (e >> 5) & 1, s = (e >> 6) & 3, jt = e >> 8;
      const int k = 8 * s + (ln & 3) + 4 * hh, nn = ln >> 2, j = jt & 3;
      const float w = jt < 4 ? d.sa * __ldg(Wa + d.wa_off + k * 32 + 8 * j + nn) : d.sb * __ldg(Wb + d.wb_off + k * 32 + 8 * j + nn);
      float* dst = lp_smem + (jt < 4 ? d.wa_sm : d.wb_sm) + (j * 4 + s) * CM_FRAG;
      const float hi = __uint_as_float(cm_tf32(w));
      dst[(hh * 2 + 0) * 32 + ln] = hi;
      dst[(hh * 2 + 1) * 32 + ln] = w - hi;
    }
  }
  __syncthreads();

  const int gi = P.warp_group[warp], rank = P.warp_rank[warp], nw = P.group_warps[gi];
  const LpBlock d = P.b[gi];
  const float* fa = lp_smem + d.wa_sm + lane;
  const float* fb = lp_smem + d.wb_sm + lane;

  int chunk = blockIdx.x, shift = 0;
  unsigned r0 = rank * 16;
  auto rows_in = [&](int c) {
    const int64_t left = N - (int64_t)c * LP_NB;
    return (unsigned)((left < LP_NB ? left : LP_NB) * d.rows);
  };
  auto settle = [&]() {
    while (chunk < n_chunks && r0 >= rows_in(chunk)) {
      shift = (shift + (int)((rows_in(chunk) + 15) / 16)) % nw;
      chunk += gridDim.x;
      r0 = (unsigned)((rank + nw - shift) % nw) * 16;
    }
  };
  auto load = [&](float (&x)[16], float2 (&rr)[4][2], int64_t (&oo)[2], bool (&ok)[2]) {
    const unsigned rows_here = rows_in(chunk), r_base = (unsigned)((int64_t)chunk * LP_NB * d.rows);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const unsigned rl = r0 + g + 8 * h, r = r_base + rl;
      ok[h] = rl < rows_here;
      const unsigned n = d.rows == 1 ? r : __umulhi(r, d.inv_rows), m = r - n * d.rows;
      oo[h] = ok[h] ? (int64_t)n * P.stride + d.off + m * 32 : 0;
      const float* p = X + oo[h] + t;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        x[(s * 2 + 0) * 2 + h] = ok[h] ? __ldg(p + 8 * s) : 0.f;
        x[(s * 2 + 1) * 2 + h] = ok[h] ? __ldg(p + 8 * s + 4) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        rr[j][h] = (res != nullptr && ok[h]) ? __ldg(reinterpret_cast<const float2*>(res + oo[h] + 8 * j + 2 * t)) : make_float2(0.f, 0.f);
    }
  };

  settle();
  float xn[16];
  float2 rn[4][2];
  int64_t on[2];
  bool okn[2];
  if (chunk < n_chunks) load(xn, rn, on, okn);
  while (chunk < n_chunks) {
    float x[16];
    int64_t oo[2];
    bool ok[2];
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = xn[i];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      oo[h] = on[h];
      ok[h] = okn[h];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[j][2 * h + 0] = rn[j][h].x;
        acc[j][2 * h + 1] = rn[j][h].y;
      }
    }
    r0 += nw * 16;
    settle();
    if (chunk < n_chunks) load(xn, rn, on, okn);   // next tile's loads fly during the two mixes below
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float av[4] = {x[(s * 2) * 2], x[(s * 2) * 2 + 1], x[(s * 2 + 1) * 2], x[(s * 2 + 1) * 2 + 1]};
      cm_kstep<4>(acc, av, fa, s);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (ok[h]) *reinterpret_cast<float2*>(Y1 + oo[h] + 8 * j + 2 * t) = make_float2(acc[j][2 * h], acc[j][2 * h + 1]);
    float y2[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) y2[j][q] = 0.f;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      float a[4];
      cm_c2a(acc[s], a, lane);
      cm_kstep<4>(y2, a, fb, s);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (ok[h]) *reinterpret_cast<float2*>(Y2 + oo[h] + 8 * j + 2 * t) = make_float2(y2[j][2 * h], y2[j][2 * h + 1]);
  }
}

}  // namespace

extern "C" int eeg_linear_pair_fwd(const float* x, const float* residual, const EegLayout* lh, const float* Wa,
                                   const int64_t* host_wa_off, const float* host_sa, const float* Wb, const int64_t* host_wb_off,
                                   const float* host_sb, float* y1, float* y2, int64_t N, void* stream) {
  EEG_REQUIRE(lh && lh->n_blocks >= 1 && lh->n_blocks <= LP_MAXB && !(lh->stride & 1), "eeg_linear_pair_fwd: unsupported layout");
  if (N == 0) return 0;
  LpParams P;
  P.stride = lh->stride;
  P.n_groups = lh->n_blocks;
  int sm = 0;
  for (int b = 0; b < lh->n_blocks; ++b) {
    EEG_REQUIRE(lh->mul[b] == 32 && lh->rows[b] >= 1 && !(lh->off[b] & 1), "eeg_linear_pair_fwd: block %d: multiplicity must be 32", b);
    EEG_REQUIRE(N * lh->rows[b] < (1LL << 28), "eeg_linear_pair_fwd: too many rows");
    LpBlock& d = P.b[b];
    d.rows = lh->rows[b];
    d.off = lh->off[b];
    d.inv_rows = d.rows > 1 ? (unsigned)(((1ULL << 32) + d.rows - 1) / d.rows) : 0u;
    d.wa_off = host_wa_off[b];
    d.wb_off = host_wb_off[b];
    d.sa = host_sa[b];
    d.sb = host_sb[b];
    d.wa_sm = sm;
    sm += 16 * CM_FRAG;
    d.wb_sm = sm;
    sm += 16 * CM_FRAG;
  }
  int wcount[LP_MAXB];
  for (int q = 0; q < P.n_groups; ++q) wcount[q] = 1;
  for (int w = P.n_groups; w < LP_WARPS; ++w) {
    int best = 0;
    for (int q = 1; q < P.n_groups; ++q)
      if (P.b[q].rows * wcount[best] > P.b[best].rows * wcount[q]) best = q;
    ++wcount[best];
  }
  for (int q = 0, w = 0; q < P.n_groups; ++q) {
    P.group_warps[q] = (unsigned char)wcount[q];
    for (int r = 0; r < wcount[q]; ++r, ++w) {
      P.warp_group[w] = (unsigned char)q;
      P.warp_rank[w] = (unsigned char)r;
    }
  }
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_linpair_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, LP_MAXB * 32 * CM_FRAG * 4);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_linear_pair_fwd: cannot opt in to shared memory: %s", cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  const int64_t chunks = eeg_div_up(N, LP_NB);
  EEG_REQUIRE(chunks < (1LL << 31), "eeg_linear_pair_fwd: too many nodes");
  const unsigned ctas = (unsigned)(chunks < EEG_NUM_SMS ? chunks : EEG_NUM_SMS);
  k_linpair_mma<<<ctas, LP_WARPS * 32, (size_t)sm * 4, (cudaStream_t)stream>>>(x, residual, Wa, Wb, y1, y2, N, P, (int)chunks);
  EEG_CHECK_LAUNCH("eeg_linear_pair_fwd");
  return 0;
}
