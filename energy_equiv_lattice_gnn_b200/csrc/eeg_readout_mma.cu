// Inference tail (see eeg_readout.cu for the math and the reference citations) on warp-level tensor-core MMAs:
//   h = product.linear(B) + h_in -> z = linear_1(h) -> Gate -> o = linear_2(g), every mix a 3xTF32 mma.sync.m16n8k8 chain.
// A warp handles 16-row tiles of one irrep block (warps are bound to blocks in proportion to the rows per node, a CTA walks
// chunks of whole nodes: eeg_lin32mma.cu).  The accumulator fragment of one mix becomes the A fragment of the next one with
// quad shuffles (no shared memory round trip); the three weight matrices of every block sit in shared memory as hi/lo B
// fragments.  Pass 0 = the 0e block (scalars + the gate activations, written to `gates`), pass 1 = the gated blocks.
#include "eeg_common.cuh"
#include "eeg_mma.cuh"

namespace {

constexpr int CM_WARPS = 16, CM_NB = 32, CM_MAXB = 5;

struct CmBlock {
  int rows, nout, in_off, out_off, blk;
  unsigned inv_rows;
  int wp_sm, w1_sm, w2_sm;   // float offsets of the fragment images in shared memory
  int n1_tiles;              // 8-column tiles of linear_1 (4, or 4 (1 + gate blocks) for the 0e block)
  int64_t wp_off, w1_off, w2_off;
  float sp, s1, s2;
};
struct CmParams {
  CmBlock b[CM_MAXB];
  int n_groups, in_stride, out_stride, gstride;
  float act_const;
  unsigned char warp_group[CM_WARPS], warp_rank[CM_WARPS], group_warps[CM_MAXB];
};

template <int PASS>
__global__ void __launch_bounds__(CM_WARPS * 32, 1) k_chain_mma(const float* __restrict__ Bx, const float* __restrict__ res,
                                                               const float* __restrict__ Wp, const float* __restrict__ W1,
                                                               const float* __restrict__ W2, float* __restrict__ gates,
                                                               float* __restrict__ out, int64_t N, CmParams P, int n_chunks) {
  extern __shared__ __align__(16) float cm_smem[];
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3, warp = threadIdx.x >> 5;
  // ---- weight fragments of every group: element (tile j, k-step s, hh, lane): W[k = 8 s + t + 4 hh][n = 8 j + g] * scale
  for (int gi = 0; gi < P.n_groups; ++gi) {
    const CmBlock& d = P.b[gi];
    const int n1 = d.n1_tiles * 8;
    const int tot = (4 + d.n1_tiles + d.nout / 8) * 4 * 2 * 32;   // (tile, k-step, hh, lane) over the three matrices
    for (int e = threadIdx.x; e < tot; e += CM_WARPS * 32) {
      const int ln = e & 31, hh = (e >> 5) & 1, s = (e >> 6) & 3, jt = e >> 8;
      const int k = 8 * s + (ln & 3) + 4 * hh, nn = ln >> 2;
      float w;
      float* dst;
      if (jt < 4) {
        w = d.sp * __ldg(Wp + d.wp_off + k * 32 + 8 * jt + nn);
        dst = cm_smem + d.wp_sm + (jt * 4 + s) * CM_FRAG;
      } else if (jt < 4 + d.n1_tiles) {
        const int j = jt - 4;
        w = d.s1 * __ldg(W1 + d.w1_off + (int64_t)k * n1 + 8 * j + nn);
        dst = cm_smem + d.w1_sm + (j * 4 + s) * CM_FRAG;
      } else {
        const int j = jt - 4 - d.n1_tiles;
        w = d.s2 * __ldg(W2 + d.w2_off + k * d.nout + 8 * j + nn);
        dst = cm_smem + d.w2_sm + (j * 4 + s) * CM_FRAG;
      }
      const float hi = __uint_as_float(cm_tf32(w));
      dst[(hh * 2 + 0) * 32 + ln] = hi;
      dst[(hh * 2 + 1) * 32 + ln] = w - hi;
    }
  }
  __syncthreads();

  const int gi = P.warp_group[warp], rank = P.warp_rank[warp], nw = P.group_warps[gi];
  const CmBlock d = P.b[gi];
  const float* fp = cm_smem + d.wp_sm + lane;
  const float* f1 = cm_smem + d.w1_sm + lane;
  const float* f2 = cm_smem + d.w2_sm + lane;

  int chunk = blockIdx.x, shift = 0;
  unsigned r0 = rank * 16;
  auto rows_in = [&](int c) {
    const int64_t left = N - (int64_t)c * CM_NB;
    return (unsigned)((left < CM_NB ? left : CM_NB) * d.rows);
  };
  auto settle = [&]() {
    while (chunk < n_chunks && r0 >= rows_in(chunk)) {
      shift = (shift + (int)((rows_in(chunk) + 15) / 16)) % nw;
      chunk += gridDim.x;
      r0 = (unsigned)((rank + nw - shift) % nw) * 16;
    }
  };
  // one tile's operands: A fragments of B (x), residual in accumulator layout (rr), node and output offset of rows g / g + 8
  auto load = [&](float (&x)[16], float2 (&rr)[4][2], unsigned (&nd)[2], int64_t (&oo)[2], bool (&ok)[2]) {
    const unsigned rows_here = rows_in(chunk), r_base = (unsigned)((int64_t)chunk * CM_NB * d.rows);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const unsigned rl = r0 + g + 8 * h, r = r_base + rl;
      ok[h] = rl < rows_here;
      const unsigned n = d.rows == 1 ? r : __umulhi(r, d.inv_rows), m = r - n * d.rows;
      nd[h] = ok[h] ? n : 0;
      const int64_t oi = ok[h] ? (int64_t)n * P.in_stride + d.in_off + m * 32 : 0;
      oo[h] = ok[h] ? (int64_t)n * P.out_stride + d.out_off + m * d.nout : 0;
      const float* p = Bx + oi + t;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        x[(s * 2 + 0) * 2 + h] = ok[h] ? __ldg(p + 8 * s) : 0.f;
        x[(s * 2 + 1) * 2 + h] = ok[h] ? __ldg(p + 8 * s + 4) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        rr[j][h] = (res != nullptr && ok[h]) ? __ldg(reinterpret_cast<const float2*>(res + oi + 8 * j + 2 * t)) : make_float2(0.f, 0.f);
    }
  };

  settle();
  float xn[16];
  float2 rn[4][2];
  unsigned ndn[2];
  int64_t on[2];
  bool okn[2];
  if (chunk < n_chunks) load(xn, rn, ndn, on, okn);
  while (chunk < n_chunks) {
    float x[16];
    unsigned nd[2];
    int64_t oo[2];
    bool ok[2];
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = xn[i];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      nd[h] = ndn[h];
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
    if (chunk < n_chunks) load(xn, rn, ndn, on, okn);   // next tile's loads fly during the three mixes below
    // ---- h = product.linear(B) + h_in
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float av[4] = {x[(s * 2) * 2], x[(s * 2) * 2 + 1], x[(s * 2 + 1) * 2], x[(s * 2 + 1) * 2 + 1]};
      cm_kstep<4>(acc, av, fp, s);
    }
    float ha[4][4];   // h as A fragments, one per k-step
#pragma unroll
    for (int s = 0; s < 4; ++s) cm_c2a(acc[s], ha[s], lane);
    // ---- z = linear_1(h); Gate
    float z[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) z[j][q] = 0.f;
#pragma unroll
    for (int s = 0; s < 4; ++s) cm_kstep<4>(z, ha[s], f1, s);
    if (PASS == 0) {
      // 0e block: columns 0..31 are the scalars; every further group of 32 columns holds the gates of one block l >= 1
      for (int cg = 1; cg * 4 < d.n1_tiles; ++cg) {
        float zg[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int q = 0; q < 4; ++q) zg[j][q] = 0.f;
#pragma unroll
        for (int s = 0; s < 4; ++s) cm_kstep<4>(zg, ha[s], f1 + cg * 16 * CM_FRAG, s);
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (ok[h]) {
            float* gp = gates + (int64_t)nd[h] * P.gstride + (cg - 1) * 32 + 2 * t;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<float2*>(gp + 8 * j) = make_float2(P.act_const * eeg_silu(zg[j][2 * h]), P.act_const * eeg_silu(zg[j][2 * h + 1]));
          }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int q = 0; q < 4; ++q) z[j][q] = P.act_const * eeg_silu(z[j][q]);
    } else {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float* gp = gates + (int64_t)nd[h] * P.gstride + (d.blk - 1) * 32 + 2 * t;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 gv = ok[h] ? __ldg(reinterpret_cast<const float2*>(gp + 8 * j)) : make_float2(0.f, 0.f);
          z[j][2 * h] *= gv.x;
          z[j][2 * h + 1] *= gv.y;
        }
      }
    }
    // ---- o = linear_2(g)
    float o[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) o[j][q] = 0.f;
    const int nt2 = d.nout / 8;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      float ga[4];
      cm_c2a(z[s], ga, lane);
      if (nt2 == 2) cm_kstep<2>(o, ga, f2, s);
      else cm_kstep<4>(o, ga, f2, s);   // weight image is zero-padded to four tiles
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < nt2) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (ok[h]) *reinterpret_cast<float2*>(out + oo[h] + 8 * j + 2 * t) = make_float2(o[j][2 * h], o[j][2 * h + 1]);
      }
  }
}

}  // namespace

// 0 = launched, -1 = shape outside this kernel
int eeg_readout_chain_mma_launch(const float* B, const float* residual, const EegLayout* lh, const float* Wp,
                                 const int64_t* host_wp_off, const float* host_sp, const float* W1, const int64_t* host_w1_off,
                                 const float* host_s1, const float* W2, const int64_t* host_w2_off, const float* host_s2,
                                 float act_const, float* gates, float* out, const EegLayout* lout, int64_t N, cudaStream_t stream) {
  const int nb = lh->n_blocks;
  if (nb < 1 || nb > CM_MAXB || ((lh->stride | lout->stride) & 1)) return -1;
  for (int b = 0; b < nb; ++b) {
    if (lh->mul[b] != 32 || lh->rows[b] != 2 * b + 1 || lout->rows[b] != lh->rows[b]) return -1;
    if (lout->mul[b] % 8 || lout->mul[b] < 8 || lout->mul[b] > 32 || (lout->off[b] & 1) || (lh->off[b] & 1)) return -1;
    if (N * lh->rows[b] >= (1LL << 28)) return -1;
  }
  static EegDevFlags attr_dev; bool& attr = eeg_dev_flag(attr_dev);
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_chain_mma<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_chain_mma<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_readout_chain_fwd: cannot opt in to shared memory: %s", cudaGetErrorString(e));
      return 2;
    }
    attr = true;
  }
  const int64_t chunks = eeg_div_up(N, CM_NB);
  if (chunks >= (1LL << 31)) return -1;
  const unsigned ctas = (unsigned)(chunks < EEG_NUM_SMS ? chunks : EEG_NUM_SMS);
  for (int pass = 0; pass < 2; ++pass) {
    const int b0 = pass == 0 ? 0 : 1, b1 = pass == 0 ? 1 : nb;
    if (b0 >= b1) continue;
    CmParams P;
    P.in_stride = lh->stride;
    P.out_stride = lout->stride;
    P.gstride = 32 * (nb - 1);
    P.act_const = act_const;
    P.n_groups = b1 - b0;
    int sm = 0;
    for (int b = b0; b < b1; ++b) {
      CmBlock& d = P.b[b - b0];
      d.rows = lh->rows[b];
      d.nout = lout->mul[b];
      d.in_off = lh->off[b];
      d.out_off = lout->off[b];
      d.blk = b;
      d.inv_rows = d.rows > 1 ? (unsigned)(((1ULL << 32) + d.rows - 1) / d.rows) : 0u;
      d.n1_tiles = b == 0 ? 4 * nb : 4;
      d.wp_off = host_wp_off[b];
      d.w1_off = host_w1_off[b];
      d.w2_off = host_w2_off[b];
      d.sp = host_sp[b];
      d.s1 = host_s1[b];
      d.s2 = host_s2[b];
      d.wp_sm = sm;
      sm += 16 * CM_FRAG;
      d.w1_sm = sm;
      sm += d.n1_tiles * 4 * CM_FRAG;
      d.w2_sm = sm;
      sm += 16 * CM_FRAG;   // padded to four tiles (zero beyond nout: the staging loop leaves them untouched, so clear below)
    }
    if ((size_t)sm * 4 > 200 * 1024) return -1;
    // warps per group in proportion to the rows per node
    int wcount[CM_MAXB];
    for (int q = 0; q < P.n_groups; ++q) wcount[q] = 1;
    for (int w = P.n_groups; w < CM_WARPS; ++w) {
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
    if (pass == 0)
      k_chain_mma<0><<<ctas, CM_WARPS * 32, (size_t)sm * 4, stream>>>(B, residual, Wp, W1, W2, gates, out, N, P, (int)chunks);
    else
      k_chain_mma<1><<<ctas, CM_WARPS * 32, (size_t)sm * 4, stream>>>(B, residual, Wp, W1, W2, gates, out, N, P, (int)chunks);
    EEG_CHECK_LAUNCH("eeg_readout_chain_fwd(mma)");
  }
  return 0;
}
