// Gated non-linearity (e3nn.nn.Gate, gnn/blocks.py:268-282) and the graph-level tail:
// scatter(mean) pooling (gnn/model.py:100-106) -> final o3.Linear (:108) -> Spherical_to_Cartesian + Cart_4_to_Mandel
// (gnn/blocks.py:392-442, folded into one [n_q, 36] matrix) -> PositiveLayer (gnn/blocks.py:185-229).
#include "eeg_common.cuh"
#include "eeg_eig6.cuh"

namespace {

struct GateParams {
  EegLayout lz, lo;
  int n_scalars;
  int gate_off[EEG_MAX_L];  // offset of block b's gates inside z block 0
  float c;
};

__global__ void k_gate(const float* __restrict__ z, float* __restrict__ out, int64_t N, GateParams P) {
  const int64_t n = blockIdx.x;
  const float* zr = z + n * P.lz.stride;
  float* orow = out + n * P.lo.stride;
  for (int i = threadIdx.x; i < P.lo.stride; i += blockDim.x) {
    int b = 0;
#pragma unroll
    for (int q = 1; q < EEG_MAX_L; ++q)
      if (q < P.lo.n_blocks && i >= P.lo.off[q]) b = q;
    float v;
    if (b == 0) {
      v = P.c * eeg_silu(zr[i]);
    } else {
      int r = i - P.lo.off[b];
      int u = r % P.lo.mul[b];
      float g = P.c * eeg_silu(zr[P.gate_off[b] + u]);
      v = zr[P.lz.off[b] + r] * g;
    }
    orow[i] = v;
  }
}

struct TailParams {
  EegLayout lo;
  int qmul[EEG_MAX_L];
  int qoff[EEG_MAX_L];   // offset into q vector
  int woff[EEG_MAX_L];   // offset into Wq
  int poff[EEG_MAX_L];   // offset into pooled smem
  float qscale[EEG_MAX_L];
  int nq, npool, positive, reduce;
};

__device__ void mat6_mul(const float* A, const float* B, float* Cm, int t) {
  if (t < 36) {
    int r = t / 6, c = t % 6;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) s = fmaf(A[r * 6 + k], B[k * 6 + c], s);
    Cm[t] = s;
  }
}

__global__ void __launch_bounds__(128) k_pool_tail(const float* __restrict__ o, const int32_t* __restrict__ gp,
                                                   const float* __restrict__ Wq, const float* __restrict__ bias,
                                                   const float* __restrict__ Pm, float* __restrict__ q_out,
                                                   float* __restrict__ A_out, float* __restrict__ C_out, TailParams P) {
  extern __shared__ float sm[];
  float* pooled = sm;              // [npool]
  float* q = pooled + P.npool;     // [nq]
  float* A = q + P.nq;             // [36]
  float* B = A + 36;               // [36]
  float* Cm = B + 36;              // [36]
  const int g = blockIdx.x, t = threadIdx.x;
  const int nb = gp[g], ne = gp[g + 1];
  const float inv = (P.reduce == 0) ? 1.0f / (float)max(ne - nb, 1) : 1.0f;
  // pooled features of the blocks that feed the final linear (fixed node order -> deterministic)
  for (int i = t; i < P.npool; i += blockDim.x) {
    int b = 0;
#pragma unroll
    for (int k = 1; k < EEG_MAX_L; ++k)
      if (k < P.lo.n_blocks && P.qmul[k] > 0 && i >= P.poff[k]) b = k;
    if (P.qmul[0] == 0 && b == 0) {  // first pooled block is not block 0
      for (int k = 0; k < P.lo.n_blocks; ++k)
        if (P.qmul[k] > 0) { b = k; break; }
      for (int k = b + 1; k < P.lo.n_blocks; ++k)
        if (P.qmul[k] > 0 && i >= P.poff[k]) b = k;
    }
    const int r = i - P.poff[b];
    const float* src = o + (int64_t)nb * P.lo.stride + P.lo.off[b] + r;
    float s = 0.f;
    for (int n = nb; n < ne; ++n, src += P.lo.stride) s += __ldg(src);
    pooled[i] = s * inv;
  }
  __syncthreads();
  // q[b][w][m] = scale_b * sum_u pooled_b[m][u] W_b[u][w] (+ bias for block 0)
  for (int i = t; i < P.nq; i += blockDim.x) {
    int b = -1;
    for (int k = 0; k < P.lo.n_blocks; ++k)
      if (P.qmul[k] > 0 && i >= P.qoff[k]) b = k;
    const int rows = P.lo.rows[b], mul = P.lo.mul[b];
    const int r = i - P.qoff[b];
    const int w = r / rows, m = r % rows;
    float s = 0.f;
    for (int u = 0; u < mul; ++u) s = fmaf(pooled[P.poff[b] + m * mul + u], __ldg(Wq + P.woff[b] + u * P.qmul[b] + w), s);
    s *= P.qscale[b];
    if (b == 0 && bias != nullptr) s += __ldg(bias + w);
    q[i] = s;
    if (q_out) q_out[(int64_t)g * P.nq + i] = s;
  }
  __syncthreads();
  if (t < 36) {
    float s = 0.f;
    for (int k = 0; k < P.nq; ++k) s = fmaf(q[k], __ldg(Pm + k * 36 + t), s);
    A[t] = s;
    if (A_out) A_out[(int64_t)g * 36 + t] = s;
  }
  __syncthreads();
  const float* res = A;
  if (P.positive == 1) {  // A^2
    mat6_mul(A, A, Cm, t);
    res = Cm;
  } else if (P.positive == 2) {  // A^4
    mat6_mul(A, A, B, t);
    __syncthreads();
    mat6_mul(B, B, Cm, t);
    res = Cm;
  } else if (P.positive == 3 || P.positive == 4) {  // (I + A/k)^k
    const float kk = (P.positive == 3) ? 2.f : 4.f;
    if (t < 36) A[t] = A[t] / kk + ((t / 6 == t % 6) ? 1.f : 0.f);
    __syncthreads();
    mat6_mul(A, A, B, t);
    res = B;
    if (P.positive == 4) {
      __syncthreads();
      mat6_mul(B, B, Cm, t);
      res = Cm;
    }
  }
  else if (P.positive == 5) {  // exp(A), A symmetric: V diag(exp(ev)) V^T
    if (t == 0) eeg_eigh6(A, Cm, B);   // Cm[0..5] = eigenvalues, B = V
    __syncthreads();
    float s = 0.f;
    if (t < 36) {
#pragma unroll
      for (int k = 0; k < 6; ++k) s = fmaf(B[(t / 6) * 6 + k] * expf(Cm[k]), B[(t % 6) * 6 + k], s);
    }
    __syncthreads();
    if (t < 36) Cm[t] = s;
    res = Cm;
  }
  __syncthreads();
  if (t < 36) C_out[(int64_t)g * 36 + t] = res[t];
}

}  // namespace

extern "C" int eeg_gate_fwd(const float* z, const EegLayout* lz, float* out, const EegLayout* lout, int n_scalars,
                            float act_const, int64_t N, void* stream) {
  EEG_REQUIRE(lz && lout && lz->n_blocks == lout->n_blocks, "eeg_gate_fwd: layout mismatch");
  if (N == 0) return 0;
  GateParams P;
  P.lz = *lz;
  P.lo = *lout;
  P.n_scalars = n_scalars;
  P.c = act_const;
  int goff = n_scalars;
  EEG_REQUIRE(lout->mul[0] == n_scalars, "eeg_gate_fwd: out block 0 must hold the scalars");
  for (int b = 0; b < lz->n_blocks; ++b) {
    P.gate_off[b] = 0;
    if (b == 0) continue;
    EEG_REQUIRE(lz->mul[b] == lout->mul[b] && lz->rows[b] == lout->rows[b], "eeg_gate_fwd: block %d mismatch", b);
    P.gate_off[b] = lz->off[0] + goff;
    goff += lz->mul[b];
  }
  EEG_REQUIRE(goff == lz->mul[0] && lz->rows[0] == 1, "eeg_gate_fwd: z block 0 must be scalars + gates");
  k_gate<<<(unsigned)N, 256, 0, (cudaStream_t)stream>>>(z, out, N, P);
  EEG_CHECK_LAUNCH("eeg_gate_fwd");
  return 0;
}

extern "C" int eeg_pool_tail_fwd(const float* o, const EegLayout* lo, const int32_t* graph_ptr, const float* Wq,
                                 const int32_t* host_qmul, const float* host_qscale, const float* bias, const float* P,
                                 int positive, int reduce, float* q_out, float* A_out, float* C_out, int64_t G, void* stream) {
  EEG_REQUIRE(lo && lo->n_blocks <= EEG_MAX_L, "eeg_pool_tail_fwd: bad layout");
  EEG_REQUIRE(positive >= 0 && positive <= 5, "eeg_pool_tail_fwd: unsupported positive function %d", positive);
  if (G == 0) return 0;
  TailParams T;
  T.lo = *lo;
  int nq = 0, npool = 0, woff = 0;
  for (int b = 0; b < EEG_MAX_L; ++b) {
    T.qmul[b] = 0; T.qoff[b] = 0; T.woff[b] = 0; T.poff[b] = 0; T.qscale[b] = 0.f;
  }
  for (int b = 0; b < lo->n_blocks; ++b) {
    T.qmul[b] = host_qmul[b];
    T.qscale[b] = host_qscale[b];
    T.qoff[b] = nq;
    T.woff[b] = woff;
    T.poff[b] = npool;
    if (host_qmul[b] > 0) {
      nq += host_qmul[b] * lo->rows[b];
      woff += host_qmul[b] * lo->mul[b];
      npool += lo->rows[b] * lo->mul[b];
    }
  }
  EEG_REQUIRE(host_qmul[0] > 0, "eeg_pool_tail_fwd: block 0 (scalars) must feed the final linear");
  T.nq = nq; T.npool = npool; T.positive = positive; T.reduce = reduce;
  size_t smem = (size_t)(npool + nq + 36 * 3) * sizeof(float);
  EEG_REQUIRE(smem <= 48 * 1024, "eeg_pool_tail_fwd: pooled row too large");
  k_pool_tail<<<(unsigned)G, 128, smem, (cudaStream_t)stream>>>(o, graph_ptr, Wq, bias, P, q_out, A_out, C_out, T);
  EEG_CHECK_LAUNCH("eeg_pool_tail_fwd");
  return 0;
}

// node embedding: torch.nn.Linear(1, C) on node_attrs [N,1] (gnn/model.py:122,142): out[n,c] = a[n] * W[c] + b[c]
namespace {
__global__ void k_node_embed(const float* __restrict__ a, const float* __restrict__ W, const float* __restrict__ b,
                             float* __restrict__ out, int64_t N, int C) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  int64_t n = i / C;
  int c = (int)(i - n * C);
  out[i] = fmaf(__ldg(a + n), __ldg(W + c), __ldg(b + c));
}
}  // namespace

extern "C" int eeg_node_embed_fwd(const float* node_attrs, const float* W, const float* b, float* out, int64_t N, int C,
                                  void* stream) {
  if (N == 0) return 0;
  k_node_embed<<<eeg_div_up(N * C, 256), 256, 0, (cudaStream_t)stream>>>(node_attrs, W, b, out, N, C);
  EEG_CHECK_LAUNCH("eeg_node_embed_fwd");
  return 0;
}
