// Batch construction on the device: the PyG-equivalent collation of
// scripts/train_main.py:67-78 (DataLoader -> Batch.from_data_list) as ONE gather kernel over a dataset
// that is resident in HBM in flat form (flatdata.FlatDataset), and the rotation sampling of RotateLat
// (scripts/train_utils.py:121, e3nn o3.rand_matrix) without a host round trip.
#include "eeg_common.cuh"

namespace {

// largest g with ptr[g] <= i   (ptr non-decreasing, ptr[0] = 0 <= i < ptr[G])
__device__ __forceinline__ int64_t seg_of(const int64_t* __restrict__ ptr, int64_t G, int64_t i) {
  int64_t lo = 0, hi = G;
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (__ldg(ptr + mid) <= i) lo = mid;
    else hi = mid;
  }
  return lo;
}

struct CollateArgs {
  const float *pos, *nattr, *shifts, *eattr;
  const int32_t *esnd, *ercv;
  const int64_t *src_node, *src_edge, *out_node, *out_edge;   // [G], [G], [G+1], [G+1]
  float *o_pos, *o_nattr, *o_shifts, *o_eattr;
  int64_t *o_ei, *o_batch;
  const float* g_src;    // per-lattice rows (targets), may be NULL
  const int64_t* g_idx;  // [G] dataset index of each lattice of the batch
  float* g_dst;
  int64_t g_width;
};

__global__ void k_collate(CollateArgs a, int64_t G, int64_t N, int64_t E) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < N) {
    const int64_t g = seg_of(a.out_node, G, i);
    const int64_t src = __ldg(a.src_node + g) + (i - __ldg(a.out_node + g));
    a.o_pos[3 * i + 0] = __ldg(a.pos + 3 * src + 0);
    a.o_pos[3 * i + 1] = __ldg(a.pos + 3 * src + 1);
    a.o_pos[3 * i + 2] = __ldg(a.pos + 3 * src + 2);
    a.o_nattr[i] = __ldg(a.nattr + src);
    a.o_batch[i] = g;
  } else if (i < N + E) {
    const int64_t e = i - N;
    const int64_t g = seg_of(a.out_edge, G, e);
    const int64_t src = __ldg(a.src_edge + g) + (e - __ldg(a.out_edge + g));
    const int64_t noff = __ldg(a.out_node + g);
    a.o_ei[e] = noff + __ldg(a.esnd + src);
    a.o_ei[E + e] = noff + __ldg(a.ercv + src);
    a.o_shifts[3 * e + 0] = __ldg(a.shifts + 3 * src + 0);
    a.o_shifts[3 * e + 1] = __ldg(a.shifts + 3 * src + 1);
    a.o_shifts[3 * e + 2] = __ldg(a.shifts + 3 * src + 2);
    a.o_eattr[e] = __ldg(a.eattr + src);
  } else if (a.g_src != nullptr) {
    const int64_t k = i - N - E;
    if (k < G * a.g_width) {
      const int64_t g = k / a.g_width, c = k - g * a.g_width;
      a.g_dst[k] = __ldg(a.g_src + __ldg(a.g_idx + g) * a.g_width + c);
    }
  }
}

// counter-based generator: splitmix64 of (seed, counter) -> uniform in [0, 1) with 24 random bits
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ float u01(uint64_t seed, uint64_t ctr) {
  return (float)(splitmix64(seed ^ splitmix64(ctr)) >> 40) * (1.0f / 16777216.0f);
}

// o3.rand_matrix(): alpha, gamma ~ U[0, 2 pi), cos(beta) ~ U[-1, 1], R = Ry(alpha) Rx(beta) Ry(gamma)   (e3nn angles_to_matrix)
__global__ void k_rand_rot(uint64_t seed, uint64_t offset, int64_t G, float* __restrict__ Q) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= G) return;
  const uint64_t c = 3 * (offset + (uint64_t)g);
  const float alpha = 6.283185307179586f * u01(seed, c), gamma = 6.283185307179586f * u01(seed, c + 2);
  const float cb = 2.0f * u01(seed, c + 1) - 1.0f, sb = sqrtf(fmaxf(1.0f - cb * cb, 0.0f));
  float sa, ca, sg, cg;
  sincosf(alpha, &sa, &ca);
  sincosf(gamma, &sg, &cg);
  // Ry(a) = [[c,0,s],[0,1,0],[-s,0,c]], Rx(b) = [[1,0,0],[0,c,-s],[0,s,c]]
  // M = Ry(alpha) Rx(beta): rows (ca, sa sb, sa cb), (0, cb, -sb), (-sa, ca sb, ca cb)
  const float m00 = ca, m01 = sa * sb, m02 = sa * cb, m10 = 0.f, m11 = cb, m12 = -sb, m20 = -sa, m21 = ca * sb, m22 = ca * cb;
  float* q = Q + 9 * g;
  q[0] = m00 * cg - m02 * sg; q[1] = m01; q[2] = m00 * sg + m02 * cg;
  q[3] = m10 * cg - m12 * sg; q[4] = m11; q[5] = m10 * sg + m12 * cg;
  q[6] = m20 * cg - m22 * sg; q[7] = m21; q[8] = m20 * sg + m22 * cg;
}

}  // namespace

extern "C" int eeg_collate_gather(const float* pos, const float* node_attrs, const int32_t* edge_snd, const int32_t* edge_rcv,
                                  const float* shifts, const float* edge_attr, const int64_t* src_node_off,
                                  const int64_t* src_edge_off, const int64_t* out_node_ptr, const int64_t* out_edge_ptr,
                                  int64_t G, int64_t N, int64_t E, float* o_pos, float* o_node_attrs, int64_t* o_edge_index,
                                  float* o_shifts, float* o_edge_attr, int64_t* o_batch, const float* graph_rows,
                                  const int64_t* graph_idx, int64_t graph_width, float* o_graph_rows, void* stream) {
  EEG_REQUIRE(G >= 0 && N >= 0 && E >= 0, "eeg_collate_gather: negative size");
  EEG_REQUIRE(graph_rows == nullptr || (graph_idx != nullptr && o_graph_rows != nullptr && graph_width > 0),
              "eeg_collate_gather: graph_rows needs graph_idx, o_graph_rows and a positive width");
  if (G == 0) return 0;
  CollateArgs a;
  a.pos = pos; a.nattr = node_attrs; a.shifts = shifts; a.eattr = edge_attr; a.esnd = edge_snd; a.ercv = edge_rcv;
  a.src_node = src_node_off; a.src_edge = src_edge_off; a.out_node = out_node_ptr; a.out_edge = out_edge_ptr;
  a.o_pos = o_pos; a.o_nattr = o_node_attrs; a.o_shifts = o_shifts; a.o_eattr = o_edge_attr; a.o_ei = o_edge_index; a.o_batch = o_batch;
  a.g_src = graph_rows; a.g_idx = graph_idx; a.g_dst = o_graph_rows; a.g_width = graph_rows ? graph_width : 0;
  const int64_t total = N + E + G * a.g_width;
  if (total == 0) return 0;
  k_collate<<<(unsigned)eeg_div_up(total, 256), 256, 0, (cudaStream_t)stream>>>(a, G, N, E);
  EEG_CHECK_LAUNCH("eeg_collate_gather");
  return 0;
}

extern "C" int eeg_rand_rotations(uint64_t seed, uint64_t offset, int64_t G, float* Q, void* stream) {
  if (G <= 0) return 0;
  k_rand_rot<<<(unsigned)eeg_div_up(G, 128), 128, 0, (cudaStream_t)stream>>>(seed, offset, G, Q);
  EEG_CHECK_LAUNCH("eeg_rand_rotations");
  return 0;
}
