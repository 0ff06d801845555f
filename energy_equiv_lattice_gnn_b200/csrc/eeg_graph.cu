// Graph indexing (CSR by receiver, graph ranges) and fused periodic edge featurisation.
// Replaces: torch_scatter's unsorted atomics (gnn/blocks.py:595-597), get_edge_vectors_and_lengths
// (gnn/mace.py:338-352), soft_one_hot_linspace x2 and o3.SphericalHarmonics (gnn/model.py:146-157).
#include <stdarg.h>
#include "eeg_common.cuh"
#include "generated/all_sh.cuh"

static thread_local char g_err[512] = "";
void eeg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* eeg_last_error(void) { return g_err; }
extern "C" int eeg_abi_version(void) { return 1; }

// ---------------------------------------------------------------------------------------------------------------
// CSR: histogram -> single-CTA-per-chunk exclusive scan (decoupled: 3 small kernels) -> fill -> per-row sort.
// Integer atomics only, followed by an in-row sort by edge id => result is deterministic and equals a stable sort.
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_count(const int64_t* __restrict__ recv, int64_t E, int32_t* __restrict__ cnt) {
  int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e < E) atomicAdd(&cnt[recv[e]], 1);
}

// block-level inclusive scan of 1024-element chunks; chunk sums to `sums`
__global__ void k_scan_chunks(const int32_t* __restrict__ in, int64_t n, int32_t* __restrict__ out, int32_t* __restrict__ sums) {
  __shared__ int32_t warp_tot[32];
  int64_t i = blockIdx.x * 1024LL + threadIdx.x;
  int v = (i < n) ? in[i] : 0;
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int s = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, s, o);
    if (lane >= o) s += t;
  }
  if (lane == 31) warp_tot[wid] = s;
  __syncthreads();
  if (wid == 0) {
    int t = warp_tot[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int u = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t += u;
    }
    warp_tot[lane] = t;
  }
  __syncthreads();
  int incl = s + (wid ? warp_tot[wid - 1] : 0);
  if (i < n) out[i] = incl - v;  // exclusive within chunk
  if (threadIdx.x == 1023) sums[blockIdx.x] = incl;
}

// single block: exclusive scan of chunk sums (n_chunks arbitrary, looped)
__global__ void k_scan_sums(int32_t* __restrict__ sums, int n_chunks) {
  __shared__ int32_t carry_s;
  __shared__ int32_t warp_tot[32];
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int base = 0; base < n_chunks; base += 1024) {
    int i = base + threadIdx.x;
    int v = (i < n_chunks) ? sums[i] : 0;
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += t;
    }
    if (lane == 31) warp_tot[wid] = s;
    __syncthreads();
    if (wid == 0) {
      int t = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int u = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += u;
      }
      warp_tot[lane] = t;
    }
    __syncthreads();
    int incl = s + (wid ? warp_tot[wid - 1] : 0) + carry_s;
    if (i < n_chunks) sums[i] = incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = incl;
    __syncthreads();
  }
}

__global__ void k_add_offsets(int32_t* __restrict__ row_ptr, const int32_t* __restrict__ sums, int64_t n, int32_t total,
                              int32_t* __restrict__ cursor) {
  int64_t i = blockIdx.x * 1024LL + threadIdx.x;
  if (i < n) {
    int v = row_ptr[i] + sums[blockIdx.x];
    row_ptr[i] = v;
    cursor[i] = v;
  }
  if (i == n) row_ptr[n] = total;
}

__global__ void k_fill(const int64_t* __restrict__ recv, int64_t E, int32_t* __restrict__ cursor, int32_t* __restrict__ perm) {
  int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e < E) perm[atomicAdd(&cursor[recv[e]], 1)] = (int32_t)e;
}

// one thread per row: insertion sort of the row's edge ids (in-degree is small: ~4, <= a few dozen)
__global__ void k_sort_rows(const int32_t* __restrict__ row_ptr, int64_t N, int32_t* __restrict__ perm) {
  int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (n >= N) return;
  int b = row_ptr[n], e = row_ptr[n + 1];
  for (int i = b + 1; i < e; ++i) {
    int v = perm[i], j = i - 1;
    while (j >= b && perm[j] > v) {
      perm[j + 1] = perm[j];
      --j;
    }
    perm[j + 1] = v;
  }
}

extern "C" size_t eeg_csr_workspace_bytes(int64_t E, int64_t N) {
  (void)E;
  int64_t chunks = (N + 1 + 1023) / 1024;
  return (size_t)((N + 1) + chunks + 8) * sizeof(int32_t);
}

extern "C" int eeg_csr_build(const int64_t* edge_index, int64_t E, int64_t N, int32_t* row_ptr, int32_t* perm, void* ws,
                             size_t ws_bytes, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  EEG_REQUIRE(E >= 0 && N >= 0 && E < (1LL << 31) && N < (1LL << 31), "eeg_csr_build: sizes out of range");
  EEG_REQUIRE(ws_bytes >= eeg_csr_workspace_bytes(E, N), "eeg_csr_build: workspace too small");
  int32_t* cursor = (int32_t*)ws;          // [N+1], first used as the histogram
  int64_t chunks = (N + 1 + 1023) / 1024;
  int32_t* sums = cursor + (N + 1);        // [chunks]
  const int64_t* recv = edge_index + E;
  cudaMemsetAsync(cursor, 0, (size_t)(N + 1) * sizeof(int32_t), s);
  if (E > 0) k_count<<<eeg_div_up(E, 256), 256, 0, s>>>(recv, E, cursor);
  k_scan_chunks<<<(int)chunks, 1024, 0, s>>>(cursor, N + 1, row_ptr, sums);
  k_scan_sums<<<1, 1024, 0, s>>>(sums, (int)chunks);
  k_add_offsets<<<(int)chunks, 1024, 0, s>>>(row_ptr, sums, N, (int32_t)E, cursor);
  if (E > 0) {
    k_fill<<<eeg_div_up(E, 256), 256, 0, s>>>(recv, E, cursor, perm);
    k_sort_rows<<<eeg_div_up(N, 128), 128, 0, s>>>(row_ptr, N, perm);
  }
  EEG_CHECK_LAUNCH("eeg_csr_build");
  return 0;
}

// graph_ptr[g] = first node n with batch[n] >= g  (batch non-decreasing); graph_ptr[G] = N
__global__ void k_graph_ptr(const int64_t* __restrict__ batch, int64_t N, int64_t G, int32_t* __restrict__ gp) {
  int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (n > N) return;
  int64_t prev = (n == 0) ? -1 : batch[n - 1];
  int64_t cur = (n == N) ? G : batch[n];
  if (cur > G) cur = G;
  for (int64_t g = prev + 1; g <= cur; ++g) gp[g] = (int32_t)n;
}

extern "C" int eeg_graph_ptr(const int64_t* batch, int64_t N, int64_t G, int32_t* graph_ptr, void* stream) {
  EEG_REQUIRE(N >= 0 && G >= 0 && N < (1LL << 31), "eeg_graph_ptr: sizes out of range");
  k_graph_ptr<<<eeg_div_up(N + 1, 256), 256, 0, (cudaStream_t)stream>>>(batch, N, G, graph_ptr);
  EEG_CHECK_LAUNCH("eeg_graph_ptr");
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Edge featurisation: one thread per CSR slot; float4 stores (Y row padded to a multiple of 4 floats).
// ---------------------------------------------------------------------------------------------------------------
template <int LMAX>
__device__ __forceinline__ void sh_eval(float x, float y, float z, float* Y);
#define EEG_X(L) \
  template <>    \
  __device__ __forceinline__ void sh_eval<L>(float x, float y, float z, float* Y) { eeg_sh_l##L(x, y, z, Y); }
EEG_LMAX_LIST(EEG_X)
#undef EEG_X

template <int LMAX, int NB>
__global__ void __launch_bounds__(128) k_edge_featurize(const float* __restrict__ pos, const int64_t* __restrict__ edge_index,
                                                        const float* __restrict__ shifts, const float* __restrict__ radius,
                                                        const int32_t* __restrict__ perm, int64_t E, float len_end, float rmax,
                                                        float* __restrict__ Yout, int ystride, float* __restrict__ fout,
                                                        int32_t* __restrict__ sender_sorted, int nb_rt) {
  constexpr int D = (LMAX + 1) * (LMAX + 1);
  constexpr int DP = (D + 3) / 4 * 4;
  int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= E) return;
  const int e = perm[j];
  const int64_t snd = edge_index[e], rcv = edge_index[E + e];
  // gnn/mace.py:345-347
  float vx = __ldg(pos + 3 * rcv + 0) - __ldg(pos + 3 * snd + 0) + __ldg(shifts + 3 * (int64_t)e + 0);
  float vy = __ldg(pos + 3 * rcv + 1) - __ldg(pos + 3 * snd + 1) + __ldg(shifts + 3 * (int64_t)e + 1);
  float vz = __ldg(pos + 3 * rcv + 2) - __ldg(pos + 3 * snd + 2) + __ldg(shifts + 3 * (int64_t)e + 2);
  float len = sqrtf(vx * vx + vy * vy + vz * vz);
  float inv = 1.0f / fmaxf(len, 1e-12f);  // F.normalize(eps=1e-12)
  float Y[DP];
  sh_eval<LMAX>(vx * inv, vy * inv, vz * inv, Y);
#pragma unroll
  for (int k = D; k < DP; ++k) Y[k] = 0.0f;
  float4* yo = reinterpret_cast<float4*>(Yout + j * (int64_t)ystride);
#pragma unroll
  for (int k = 0; k < DP / 4; ++k) yo[k] = make_float4(Y[4 * k], Y[4 * k + 1], Y[4 * k + 2], Y[4 * k + 3]);
  // soft_one_hot_linspace(gaussian): exp(-((x - k*step)/step)^2) / 1.12, step = end/(NB-1)
  if constexpr (NB == 0) {   // any number of bases (num_edge_bases is a free hyper-parameter, gnn/model.py:146-152)
    const float r = __ldg(radius + e);
    const float sl = len_end / (float)(nb_rt - 1), sr = rmax / (float)(nb_rt - 1);
    float* fo = fout + j * (int64_t)(2 * nb_rt);
    for (int k = 0; k < nb_rt; ++k) {
      const float dl = (len - (float)k * sl) / sl, dr = (r - (float)k * sr) / sr;
      fo[k] = expf(-dl * dl) / 1.12f;
      fo[nb_rt + k] = expf(-dr * dr) / 1.12f;
    }
    sender_sorted[j] = (int32_t)snd;
    return;
  }
  float f[2 * (NB ? NB : 1)];
  const float r = __ldg(radius + e);
  const float sl = len_end / (float)(NB - 1), sr = rmax / (float)(NB - 1);
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    float dl = (len - (float)k * sl) / sl;
    float dr = (r - (float)k * sr) / sr;
    f[k] = expf(-dl * dl) / 1.12f;
    f[NB + k] = expf(-dr * dr) / 1.12f;
  }
  float* fo = fout + j * (int64_t)(2 * NB);
  if constexpr ((2 * NB) % 4 == 0) {
#pragma unroll
    for (int k = 0; k < 2 * NB / 4; ++k)
      reinterpret_cast<float4*>(fo)[k] = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
  } else {
#pragma unroll
    for (int k = 0; k < 2 * NB; ++k) fo[k] = f[k];
  }
  sender_sorted[j] = (int32_t)snd;
}

extern "C" int eeg_edge_featurize(const float* pos, const int64_t* edge_index, const float* shifts, const float* radius,
                                  const int32_t* perm, int64_t E, int lmax, int nbasis, float len_end, float rmax, float* Y,
                                  int ystride, float* f, int32_t* sender_sorted, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int D = (lmax + 1) * (lmax + 1);
  EEG_REQUIRE(ystride == (D + 3) / 4 * 4, "eeg_edge_featurize: ystride must be D rounded up to 4");
  if (E == 0) return 0;
  dim3 g(eeg_div_up(E, 128)), b(128);
#define EEG_EF(L, NBV)                                                                                              \
  if (lmax == L && nbasis == NBV) {                                                                                 \
    k_edge_featurize<L, NBV><<<g, b, 0, s>>>(pos, edge_index, shifts, radius, perm, E, len_end, rmax, Y, ystride, f, \
                                              sender_sorted, nbasis);                                               \
    EEG_CHECK_LAUNCH("eeg_edge_featurize");                                                                         \
    return 0;                                                                                                       \
  }
  EEG_REQUIRE(nbasis >= 2, "eeg_edge_featurize: num_edge_bases must be >= 2 (got %d)", nbasis);
#define EEG_X(L) EEG_EF(L, 6) EEG_EF(L, 8) EEG_EF(L, 4)
  EEG_LMAX_LIST(EEG_X)
#undef EEG_X
#undef EEG_EF
#define EEG_X(L)                                                                                                       \
  if (lmax == L) {                                                                                                     \
    k_edge_featurize<L, 0><<<g, b, 0, s>>>(pos, edge_index, shifts, radius, perm, E, len_end, rmax, Y, ystride, f, sender_sorted, nbasis); \
    EEG_CHECK_LAUNCH("eeg_edge_featurize");                                                                            \
    return 0;                                                                                                          \
  }
  EEG_LMAX_LIST(EEG_X)
#undef EEG_X
  eeg_set_error("eeg_edge_featurize: unsupported (lmax=%d, nbasis=%d)", lmax, nbasis);
  return 1;
}

// ---------------------------------------------------------------------------------------------------------------
// Input validation (opt-in): the CSR build and the edge featurisation index device memory with edge_index / batch as given
// (torch_scatter and PyG would raise a device-side assert on a bad index; here a malformed batch - an index outside [0, N), a
// graph id outside [0, G), a decreasing `batch` - would silently corrupt memory).  flags (one int32, zeroed here) collects
//   bit 0: edge_index entry < 0 or >= N      bit 1: batch entry < 0 or >= G      bit 2: batch not non-decreasing
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_validate(const int64_t* __restrict__ edge_index, int64_t E, int64_t N, const int64_t* __restrict__ batch, int64_t G,
                           int32_t* __restrict__ flags) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int f = 0;
  if (i < 2 * E) {
    const int64_t v = edge_index[i];
    if (v < 0 || v >= N) f |= 1;
  }
  if (batch != nullptr && i < N) {
    const int64_t b = batch[i];
    if (b < 0 || b >= G) f |= 2;
    if (i > 0 && batch[i - 1] > b) f |= 4;
  }
  if (f) atomicOr(flags, f);
}

extern "C" int eeg_graph_validate(const int64_t* edge_index, int64_t E, int64_t N, const int64_t* batch, int64_t G, int32_t* flags,
                                  void* stream) {
  EEG_REQUIRE(flags != nullptr && E >= 0 && N >= 0, "eeg_graph_validate: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(flags, 0, sizeof(int32_t), s);
  if (e != cudaSuccess) {
    eeg_set_error("eeg_graph_validate: memset failed: %s", cudaGetErrorString(e));
    return 2;
  }
  const int64_t n = 2 * E > N ? 2 * E : N;
  if (n == 0) return 0;
  k_validate<<<eeg_div_up(n, 256), 256, 0, s>>>(edge_index, E, N, batch, G, flags);
  EEG_CHECK_LAUNCH("eeg_graph_validate");
  return 0;
}
