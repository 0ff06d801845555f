// Weight gradients of the per-irrep linears with 32 output channels on warp-level tensor-core MMAs:
//   dW_b[u, w] = scale_b * sum over rows r = (node, m) of block b of  x[r, u] * dy[r, w],   w < 32, u < K_b (multiple of 32)
// i.e. the path-stacked interaction.linear (gnn/blocks.py:552-559, K = 160..320; blocks with K = 32 stay on the CUDA-core
// kernel).  The CUDA-core kernel (eeg_bwd.cu k_linear_wgrad) runs these at 17-22 TFLOP/s, bound by shared-memory loads per FMA
// and two barriers per 32 rows.
//
// Work item = (irrep block, 32-row slab of the weight matrix, slice of <= 512 rows), one WARP per item, no shared memory:
// D[u, w] (32 x 32: 2 x 4 mma tiles) += A[u, r] B[r, w] with the reduction index r as the K of mma.sync.m16n8k8.  Both
// fragments come straight from global memory (A[u][r] = x[r][u0 + u]: a 32-byte run of one row per 8 lanes; the 128-byte line
// of a row is consumed by four consecutive loads and stays in L1), are split into tf32 hi / lo (3xTF32) and the next 8 rows
// are loaded while the current 24 MMAs run.  Slices are at most 512 rows = 64 K-steps, which bounds the effect of the tensor
// core's truncating fp32 accumulation (DESIGN.md round-2 addendum) at 2.5e-6; the slice partials go to the workspace and are
// added in slice order (bitwise repeatable, no atomics).  The 8 warps of a CTA take consecutive slabs of the same row slice, so
// they share the dy rows in L1.
#include "eeg_common.cuh"

namespace {

constexpr int WM_WARPS = 8, WM_MAXB = 8, WM_SLICE = 512;

struct WmBlock {
  int rows, k, slabs, in_off, out_off;
  unsigned inv_rows;      // ceil(2^32 / rows) for rows > 1
  int slices;             // row slices
  int item_begin;         // first work item
  int64_t total_rows, w_off, ws_off;
  float scale;
};
struct WmParams {
  WmBlock b[WM_MAXB];
  int n_blocks, in_stride, out_stride, transB, n_items;
};

__device__ __forceinline__ uint32_t wm_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void wm_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(WM_WARPS * 32, 2) k_wgrad_mma(const float* __restrict__ x, const float* __restrict__ dy,
                                                               float* __restrict__ ws, const WmParams P) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int item = blockIdx.x * WM_WARPS + (threadIdx.x >> 5);
  if (item >= P.n_items) return;
  int bi = 0;
#pragma unroll
  for (int q = 1; q < WM_MAXB; ++q)
    if (q < P.n_blocks && item >= P.b[q].item_begin) bi = q;
  const WmBlock d = P.b[bi];
  const int local = item - d.item_begin, slice = local / d.slabs, slab = local - slice * d.slabs;
  const int64_t r_begin = (int64_t)slice * WM_SLICE;
  const int64_t r_end = r_begin + WM_SLICE < d.total_rows ? r_begin + WM_SLICE : d.total_rows;
  const int u0 = slab * 32;

  // fragment values of one K-step (8 rows): lane holds rows r0 + t and r0 + t + 4; xa[q] = x[row][u0 + g + 8 q], ya[j] = dy[row][g + 8 j]
  auto load = [&](int64_t r0, float (&xa)[4], float (&xb)[4], float (&ya)[4], float (&yb)[4]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int64_t r = r0 + t + 4 * h;
      const bool ok = r < r_end;
      const unsigned r32 = ok ? (unsigned)r : (unsigned)r_begin;
      const unsigned n = d.rows == 1 ? r32 : __umulhi(r32, d.inv_rows), m = r32 - n * d.rows;
      const float* px = x + (int64_t)n * P.in_stride + d.in_off + (int64_t)m * d.k + u0 + g;
      const float* py = dy + (int64_t)n * P.out_stride + d.out_off + m * 32 + g;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float vx = ok ? __ldg(px + 8 * q) : 0.f, vy = ok ? __ldg(py + 8 * q) : 0.f;
        if (h == 0) { xa[q] = vx; ya[q] = vy; } else { xb[q] = vx; yb[q] = vy; }
      }
    }
  };
  float acc[2][4][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.f;
  float nxa[4], nxb[4], nya[4], nyb[4];
  load(r_begin, nxa, nxb, nya, nyb);
  for (int64_t r0 = r_begin; r0 < r_end; r0 += 8) {
    // A tile i (rows of D = u0 + 16 i + ..): a0 = (u g, r t), a1 = (u g + 8, r t), a2 = (u g, r t + 4), a3 = (u g + 8, r t + 4)
    uint32_t ah[2][4], al[2][4], bh[4][2], bl[4][2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float v[4] = {nxa[2 * i], nxa[2 * i + 1], nxb[2 * i], nxb[2 * i + 1]};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        ah[i][q] = wm_tf32(v[q]);
        al[i][q] = __float_as_uint(v[q] - __uint_as_float(ah[i][q]));
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {   // B tile j (columns 8 j + ..): b0 = (r t, w g), b1 = (r t + 4, w g)
      bh[j][0] = wm_tf32(nya[j]); bl[j][0] = __float_as_uint(nya[j] - __uint_as_float(bh[j][0]));
      bh[j][1] = wm_tf32(nyb[j]); bl[j][1] = __float_as_uint(nyb[j] - __uint_as_float(bh[j][1]));
    }
    if (r0 + 8 < r_end) load(r0 + 8, nxa, nxb, nya, nyb);   // the next K-step's loads fly during the MMAs
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        wm_mma(acc[i][j], al[i], bh[j][0], bh[j][1]);
        wm_mma(acc[i][j], ah[i], bl[j][0], bl[j][1]);
        wm_mma(acc[i][j], ah[i], bh[j][0], bh[j][1]);
      }
  }
  // partial of this item: [32 u][32 w]; c0 = (u g, w 2t), c1 = (u g, w 2t + 1), c2 = (u g + 8, w 2t), c3 = (u g + 8, w 2t + 1)
  float* o = ws + d.ws_off + ((int64_t)slab * d.slices + slice) * 1024;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      *reinterpret_cast<float2*>(o + (16 * i + g) * 32 + 8 * j + 2 * t) = make_float2(acc[i][j][0], acc[i][j][1]);
      *reinterpret_cast<float2*>(o + (16 * i + g + 8) * 32 + 8 * j + 2 * t) = make_float2(acc[i][j][2], acc[i][j][3]);
    }
}

// dW[element] += scale * (sum over the slices of the element's slab, in slice order): 32 elements x 32 partial lanes per block
__global__ void __launch_bounds__(1024) k_wm_reduce(const float* __restrict__ ws, float* __restrict__ dW, const WmParams P, int64_t total) {
  __shared__ float part[32][33];
  const int64_t i = blockIdx.x * 32LL + threadIdx.x;   // element index over all blocks: block-major, then [u][w]
  const int y = threadIdx.y;
  int bi = -1;
  int64_t e = i;
  if (i < total)
    for (int q = 0; q < P.n_blocks; ++q) {
      const int64_t n = (int64_t)P.b[q].k * 32;
      if (e < n) { bi = q; break; }
      e -= n;
    }
  float s = 0.f;
  int u = 0, w = 0;
  if (bi >= 0) {
    const WmBlock& d = P.b[bi];
    u = (int)(e >> 5);
    w = (int)(e & 31);
    const float* p = ws + d.ws_off + (int64_t)(u >> 5) * d.slices * 1024 + (u & 31) * 32 + w;
    for (int sl = y; sl < d.slices; sl += 32) s += p[(int64_t)sl * 1024];
  }
  part[y][threadIdx.x] = s;
  __syncthreads();
  if (y == 0 && bi >= 0) {
    const WmBlock& d = P.b[bi];
    float r = 0.f;
#pragma unroll
    for (int q = 0; q < 32; ++q) r += part[q][threadIdx.x];
    dW[d.w_off + (P.transB ? (int64_t)w * d.k + u : (int64_t)u * 32 + w)] += r * d.scale;
  }
}

// fills P (when not NULL); returns the workspace floats, or -1 when the shape is outside this kernel: every non-empty block
// has 32 output channels and a multiple of 32 (>= 64) input channels, and there are enough rows to fill the machine
int64_t wm_plan(const EegLayout* lin, const EegLayout* lout, const int64_t* w_off, const float* scale, int64_t N, WmParams* P,
                int64_t* total_elems) {
  int nb = 0, items = 0;
  int64_t ws = 0, elems = 0, slab_rows = 0;
  for (int b = 0; b < lin->n_blocks; ++b) {
    const int k = lin->mul[b], nout = lout->mul[b], rows = lin->rows[b];
    if (k == 0 || nout == 0) continue;
    // K = 32 (one slab per block: the five 32 -> 32 mixes) measured 0.071 ms here against 0.056 ms on the CUDA-core kernel
    if (nout != 32 || k % 32 || k < 64 || rows != lout->rows[b] || rows <= 0 || nb >= WM_MAXB) return -1;
    const int64_t R = N * rows;
    if (R >= (1LL << 28)) return -1;
    const int slabs = k / 32, slices = (int)((R + WM_SLICE - 1) / WM_SLICE);
    if ((int64_t)items + (int64_t)slabs * slices >= (1LL << 30)) return -1;
    if (P) {
      WmBlock& d = P->b[nb];
      d.rows = rows;
      d.k = k;
      d.slabs = slabs;
      d.in_off = lin->off[b];
      d.out_off = lout->off[b];
      d.inv_rows = rows > 1 ? (unsigned)(((1ULL << 32) + rows - 1) / rows) : 0u;
      d.slices = slices;
      d.item_begin = items;
      d.total_rows = R;
      d.w_off = w_off ? w_off[b] : 0;
      d.ws_off = ws;
      d.scale = scale ? scale[b] : 1.0f;
    }
    items += slabs * slices;
    ws += (int64_t)slabs * slices * 1024;
    elems += (int64_t)k * 32;
    slab_rows += R * slabs;
    ++nb;
  }
  if (nb == 0 || slab_rows < (1 << 18)) return -1;   // small problems stay on the CUDA-core kernel
  if (P) {
    P->n_blocks = nb;
    P->in_stride = lin->stride;
    P->out_stride = lout->stride;
    P->n_items = items;
  }
  if (total_elems) *total_elems = elems;
  return ws;
}

}  // namespace

int64_t eeg_wgrad_mma_ws_floats(const EegLayout* lin, const EegLayout* lout, int64_t N) { return wm_plan(lin, lout, nullptr, nullptr, N, nullptr, nullptr); }

// 0 = launched (dW += the gradient; bias gradients are the caller's), -1 = shape outside this kernel or workspace too small
int eeg_wgrad_mma_launch(const float* x, const EegLayout* lin, const float* dy, const EegLayout* lout, const int64_t* host_w_off,
                         const float* host_scale, int transB, float* dW, int64_t N, float* ws, int64_t ws_floats, cudaStream_t stream) {
  WmParams P;
  int64_t elems = 0;
  const int64_t need = wm_plan(lin, lout, host_w_off, host_scale, N, &P, &elems);
  if (need < 0 || ws == nullptr || ws_floats < need) return -1;
  P.transB = transB;
  k_wgrad_mma<<<(unsigned)((P.n_items + WM_WARPS - 1) / WM_WARPS), WM_WARPS * 32, 0, stream>>>(x, dy, ws, P);
  EEG_CHECK_LAUNCH("eeg_block_linear_wgrad (mma)");
  k_wm_reduce<<<(unsigned)((elems + 31) / 32), dim3(32, 32), 0, stream>>>(ws, dW, P, elems);
  EEG_CHECK_LAUNCH("eeg_block_linear_wgrad (mma reduce)");
  return 0;
}
