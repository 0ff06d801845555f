// Symmetric-contraction kernel body, compiled once per variant (eeg_sc_v*.cu define EEG_SC_V_* before including).
//   EEG_SC_V_NAME   launch function name suffix
//   EEG_SC_V_NPT    nodes per thread (1, 2 or 4)
//   EEG_SC_V_PACKED 1: node pairs live in 64-bit registers and every Horner step is one packed FFMA2 / FMUL2
//                   (fma.rn.f32x2, sm_100a) whose coefficient operand is the scalar-broadcast form
//   EEG_SC_V_TPC    threads per channel, EEG_SC_V_CH channels per CTA
//   EEG_SC_V_ASYNC  1: the next tile's x is prefetched into the (padded) shared tile with 4-byte cp.async while the
//                   current tile is evaluated from registers, and outputs go straight to global memory
// CTA = CH channels x (TPC * NPT) nodes per tile; lane = node, so the coefficient stream is warp-uniform and is read
// from shared memory as broadcast float4.  The x tile is transposed through shared memory (coalesced float4 global
// access both ways).  Never materialises the reference's [N, C, D, D, D] intermediate (gnn/mace.py:261-277).
#include "eeg_common.cuh"
#ifndef EEG_SC_V_ASYNC
#define EEG_SC_V_ASYNC 0
#endif
#ifndef EEG_SC_V_MINB
#define EEG_SC_V_MINB 1  // resident CTAs per SM (2: one CTA does tile I/O while the other computes)
#endif
#include "generated/all_sc.cuh"

namespace {

constexpr int NPT = EEG_SC_V_NPT, TPC = EEG_SC_V_TPC, CH = EEG_SC_V_CH, THREADS = TPC * CH, NT = TPC * NPT;
static_assert(CH == 4 || CH == 2, "x tile transposition assumes 2 or 4 channels per CTA");

typedef unsigned long long p2;  // two fp32 lanes (node A, node B)
__device__ __forceinline__ p2 pk(float a, float b) { p2 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ p2 ffma2(p2 a, p2 b, p2 c) { p2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ p2 fmul2(p2 a, p2 b) { p2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ void unpk(p2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }

#define Q(n) const float4 q##n = T4[n];
#if !EEG_SC_V_PACKED && EEG_SC_V_NPT == 1
#define S2(c) m20 = (c);
#define F3I(c, t) m20 = (c) * x[t];
#define F3(c, t) m20 = fmaf((c), x[t], m20);
#define S1(c) m10 = (c);
#define F2I(j) m10 = m20 * x[j];
#define F2(j) m10 = fmaf(m20, x[j], m10);
#define Z() a0 = 0.0f;
#define F1(i) a0 = fmaf(m10, x[i], a0);
#define O(o) out[(o) * NT] = a0;
#define SC_DECL float a0, m10, m20;
typedef float xreg;
#define SC_NX 1
#elif !EEG_SC_V_PACKED && EEG_SC_V_NPT == 2
#define S2(c) m20 = (c); m21 = (c);
#define F3I(c, t) m20 = (c) * x[t]; m21 = (c) * x[D + t];
#define F3(c, t) m20 = fmaf((c), x[t], m20); m21 = fmaf((c), x[D + t], m21);
#define S1(c) m10 = (c); m11 = (c);
#define F2I(j) m10 = m20 * x[j]; m11 = m21 * x[D + j];
#define F2(j) m10 = fmaf(m20, x[j], m10); m11 = fmaf(m21, x[D + j], m11);
#define Z() a0 = 0.0f; a1 = 0.0f;
#define F1(i) a0 = fmaf(m10, x[i], a0); a1 = fmaf(m11, x[D + i], a1);
#define O(o) out[(o) * NT] = a0; out[(o) * NT + TPC] = a1;
#define SC_DECL float a0, a1, m10, m11, m20, m21;
typedef float xreg;
#define SC_NX 2
#elif EEG_SC_V_PACKED && EEG_SC_V_NPT == 2
#define S2(c) m20 = pk((c), (c));
#define F3I(c, t) m20 = fmul2(pk((c), (c)), x[t]);
#define F3(c, t) m20 = ffma2(pk((c), (c)), x[t], m20);
#define S1(c) m10 = pk((c), (c));
#define F2I(j) m10 = fmul2(m20, x[j]);
#define F2(j) m10 = ffma2(m20, x[j], m10);
#define Z() a0 = 0ull;
#define F1(i) a0 = ffma2(m10, x[i], a0);
#if EEG_SC_V_ASYNC
#define O(o) { float u0, u1; unpk(a0, u0, u1); if (ok0) out[(o) * ostride] = u0; if (ok1) out[(o) * ostride + vstride] = u1; }
#else
#define O(o) { float u0, u1; unpk(a0, u0, u1); out[(o) * NT] = u0; out[(o) * NT + TPC] = u1; }
#endif
#define SC_DECL p2 a0, m10, m20;
typedef p2 xreg;
#define SC_NX 1
#elif EEG_SC_V_PACKED && EEG_SC_V_NPT == 4
#define S2(c) m20 = pk((c), (c)); m21 = m20;
#define F3I(c, t) m20 = fmul2(pk((c), (c)), x[t]); m21 = fmul2(pk((c), (c)), x[D + t]);
#define F3(c, t) m20 = ffma2(pk((c), (c)), x[t], m20); m21 = ffma2(pk((c), (c)), x[D + t], m21);
#define S1(c) m10 = pk((c), (c)); m11 = m10;
#define F2I(j) m10 = fmul2(m20, x[j]); m11 = fmul2(m21, x[D + j]);
#define F2(j) m10 = ffma2(m20, x[j], m10); m11 = ffma2(m21, x[D + j], m11);
#define Z() a0 = 0ull; a1 = 0ull;
#define F1(i) a0 = ffma2(m10, x[i], a0); a1 = ffma2(m11, x[D + i], a1);
#define O(o) { float u0, u1, u2, u3; unpk(a0, u0, u1); unpk(a1, u2, u3); out[(o) * NT] = u0; out[(o) * NT + TPC] = u1; \
               out[(o) * NT + 2 * TPC] = u2; out[(o) * NT + 3 * TPC] = u3; }
#define SC_DECL p2 a0, a1, m10, m11, m20, m21;
typedef p2 xreg;
#define SC_NX 2
#else
#error "unsupported symmetric-contraction variant"
#endif

template <int LMAX, int NU>
struct Prog;
// one specialisation per generated (lmax, correlation) program (generated/registry.cuh)
#define EEG_X(L, N)                                                                                                          \
  template <>                                                                                                                \
  struct Prog<L, N> {                                                                                                        \
    static constexpr int NCOEF = EEG_SC_l##L##_nu##N##_NCOEF, D = (L + 1) * (L + 1);                                           \
    static __device__ __forceinline__ void run(const xreg* __restrict__ x, const float4* __restrict__ T4, float* __restrict__ out, \
                                               int ostride = 0, int64_t vstride = 0, bool ok0 = true, bool ok1 = true) {      \
      SC_DECL                                                                                                                \
      EEG_SC_l##L##_nu##N##_PROGRAM                                                                                          \
    }                                                                                                                        \
  };
EEG_SC_COMBOS(EEG_X)
#undef EEG_X

#if EEG_SC_V_ASYNC
static_assert(EEG_SC_V_PACKED && NPT == 2 && CH == 4, "async variant: packed node pairs, 4 channels per CTA");
constexpr int XPAD = 8;  // floats between channel planes of the x tile: the (node, channel) cp.async pattern hits 32 banks

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gmem_src, bool valid) {
  const int sz = valid ? 4 : 0;  // src-size 0: zero-fill
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src), "r"(sz)
               : "memory");
}

// x tile [CH][D][NT] (+XPAD per channel plane) <- x[n0 .. n0+NT) : one warp instruction covers 8 nodes x 4 channels
// (8 half-sectors of 16 B) and writes 32 distinct banks.
template <int D>
__device__ __forceinline__ void prefetch_tile(float* xs, const float* __restrict__ xg, int64_t n0, int64_t N, int C, int c0, int t) {
  const int c = t & 3;
  float* dst = xs + c * (D * NT + XPAD);
  const float* src = xg + c0 + c;
#pragma unroll 5
  for (int q = t >> 2; q < D * NT; q += THREADS / 4) {
    const int d = q / NT, n_l = q % NT;
    const bool ok = n0 + n_l < N;
    cp_async4(dst + q, src + ((ok ? n0 + n_l : 0) * D + d) * (int64_t)C, ok);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}

template <int LMAX, int NU>
__global__ void __launch_bounds__(THREADS, EEG_SC_V_MINB) k_sym_contract(const float* __restrict__ xg, const float* __restrict__ T,
                                                             float* __restrict__ outg, int64_t N, int C, int tiles_per_cta) {
  using F = Prog<LMAX, NU>;
  constexpr int D = F::D, NCOEF = F::NCOEF;
  extern __shared__ __align__(16) float smem[];
  float* Ts = smem;                 // [CH][NCOEF]
  float* xs = smem + CH * NCOEF;    // [CH][D*NT + XPAD]
  const int t = threadIdx.x;
  const int c0 = blockIdx.x * CH;
  const int64_t n_tiles = (N + NT - 1) / NT;
  const int64_t tile_begin = (int64_t)blockIdx.y * tiles_per_cta;
  const int64_t tile_end = tile_begin + tiles_per_cta < n_tiles ? tile_begin + tiles_per_cta : n_tiles;
  if (tile_begin < tile_end) prefetch_tile<D>(xs, xg, tile_begin * NT, N, C, c0, t);
  {
    const float4* src = reinterpret_cast<const float4*>(T + (int64_t)c0 * NCOEF);
    float4* dst = reinterpret_cast<float4*>(Ts);
    for (int i = t; i < CH * NCOEF / 4; i += THREADS) dst[i] = __ldg(src + i);
  }
  const int nl = t % TPC, cc = t / TPC;
  const float* xin = xs + cc * (D * NT + XPAD) + nl;
  for (int64_t tile = tile_begin; tile < tile_end; ++tile) {
    const int64_t n0 = tile * NT;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // tile (and, first time, Ts) visible to every thread
    xreg xv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) xv[d] = pk(xin[d * NT], xin[d * NT + TPC]);
    __syncthreads();  // everyone holds its inputs in registers: the tile buffer is free again
    if (tile + 1 < tile_end) prefetch_tile<D>(xs, xg, n0 + NT, N, C, c0, t);
    F::run(xv, reinterpret_cast<const float4*>(Ts + cc * NCOEF), outg + (n0 + nl) * D * (int64_t)C + c0 + cc, C,
           (int64_t)TPC * D * C, n0 + nl < N, n0 + nl + TPC < N);
  }
}
#define SC_SMEM_FLOATS(F) (CH * F::NCOEF + CH * (F::D * NT + XPAD))
#else
#define SC_SMEM_FLOATS(F) (CH * F::NCOEF + CH * F::D * NT)
template <int LMAX, int NU>
__global__ void __launch_bounds__(THREADS, EEG_SC_V_MINB) k_sym_contract(const float* __restrict__ xg, const float* __restrict__ T,
                                                             float* __restrict__ outg, int64_t N, int C, int tiles_per_cta) {
  using F = Prog<LMAX, NU>;
  constexpr int D = F::D, NCOEF = F::NCOEF;
  extern __shared__ __align__(16) float smem[];
  float* Ts = smem;                 // [CH][NCOEF]
  float* xs = smem + CH * NCOEF;    // [CH][D][NT]
  const int t = threadIdx.x;
  const int c0 = blockIdx.x * CH;
  {
    const float4* src = reinterpret_cast<const float4*>(T + (int64_t)c0 * NCOEF);
    float4* dst = reinterpret_cast<float4*>(Ts);
    for (int i = t; i < CH * NCOEF / 4; i += THREADS) dst[i] = __ldg(src + i);
  }
  const int nl = t % TPC, cc = t / TPC;
  const int64_t n_tiles = (N + NT - 1) / NT;
  const int64_t tile_begin = (int64_t)blockIdx.y * tiles_per_cta;
  for (int64_t tile = tile_begin; tile < tile_begin + tiles_per_cta && tile < n_tiles; ++tile) {
    const int64_t n0 = tile * NT;
    __syncthreads();  // previous tile's stores done / Ts visible
    for (int idx = t; idx < NT * D; idx += THREADS) {
      const int n_l = idx % NT, d = idx / NT;
      if constexpr (CH == 4) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (n0 + n_l < N) v = __ldg(reinterpret_cast<const float4*>(xg + ((n0 + n_l) * D + d) * (int64_t)C + c0));
        xs[(0 * D + d) * NT + n_l] = v.x;
        xs[(1 * D + d) * NT + n_l] = v.y;
        xs[(2 * D + d) * NT + n_l] = v.z;
        xs[(3 * D + d) * NT + n_l] = v.w;
      } else {
        float2 v = make_float2(0.f, 0.f);
        if (n0 + n_l < N) v = __ldg(reinterpret_cast<const float2*>(xg + ((n0 + n_l) * D + d) * (int64_t)C + c0));
        xs[(0 * D + d) * NT + n_l] = v.x;
        xs[(1 * D + d) * NT + n_l] = v.y;
      }
    }
    __syncthreads();
    xreg xv[SC_NX * D];
    const float* xin = xs + cc * D * NT + nl;
#if EEG_SC_V_PACKED
#pragma unroll
    for (int v = 0; v < SC_NX; ++v)
#pragma unroll
      for (int d = 0; d < D; ++d) xv[v * D + d] = pk(xin[d * NT + (2 * v) * TPC], xin[d * NT + (2 * v + 1) * TPC]);
#else
#pragma unroll
    for (int v = 0; v < SC_NX; ++v)
#pragma unroll
      for (int d = 0; d < D; ++d) xv[v * D + d] = xin[d * NT + v * TPC];
#endif
    // outputs overwrite this thread's own x slots (its inputs are in registers)
    F::run(xv, reinterpret_cast<const float4*>(Ts + cc * NCOEF), xs + cc * D * NT + nl);
    __syncthreads();
    for (int idx = t; idx < NT * D; idx += THREADS) {
      const int n_l = idx % NT, d = idx / NT;
      if (n0 + n_l < N) {
        if constexpr (CH == 4) {
          float4 v = make_float4(xs[(0 * D + d) * NT + n_l], xs[(1 * D + d) * NT + n_l], xs[(2 * D + d) * NT + n_l],
                                 xs[(3 * D + d) * NT + n_l]);
          *reinterpret_cast<float4*>(outg + ((n0 + n_l) * D + d) * (int64_t)C + c0) = v;
        } else {
          float2 v = make_float2(xs[(0 * D + d) * NT + n_l], xs[(1 * D + d) * NT + n_l]);
          *reinterpret_cast<float2*>(outg + ((n0 + n_l) * D + d) * (int64_t)C + c0) = v;
        }
      }
    }
  }
}

#endif  // EEG_SC_V_ASYNC

template <int LMAX, int NU>
int launch_sc(int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s) {
  using F = Prog<LMAX, NU>;
  const size_t smem = (size_t)SC_SMEM_FLOATS(F) * sizeof(float);
  static EegDevFlags attr_set_dev; bool& attr_set = eeg_dev_flag(attr_set_dev);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_sym_contract<LMAX, NU>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      eeg_set_error("eeg_sym_contract_fwd: cannot opt in to %zu B shared memory: %s", smem, cudaGetErrorString(e));
      return 2;
    }
    attr_set = true;
  }
  const int groups = C / CH;
  const int64_t n_tiles = (N + NT - 1) / NT;
  int slices = EEG_NUM_SMS * EEG_SC_V_MINB / groups;  // all CTAs resident: a single wave
  if (slices > n_tiles) slices = (int)n_tiles;
  if (slices < 1) slices = 1;
  const int tiles_per_cta = (int)((n_tiles + slices - 1) / slices);
  slices = (int)((n_tiles + tiles_per_cta - 1) / tiles_per_cta);
  dim3 grid(groups, slices);
  k_sym_contract<LMAX, NU><<<grid, THREADS, smem, s>>>(x, T, out, N, C, tiles_per_cta);
  EEG_CHECK_LAUNCH("eeg_sym_contract_fwd");
  return 0;
}

}  // namespace

#define EEG_SC_CAT2(a, b) a##b
#define EEG_SC_CAT(a, b) EEG_SC_CAT2(a, b)
int EEG_SC_CAT(eeg_sc_launch_, EEG_SC_V_NAME)(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, cudaStream_t s) {
  if (C % CH != 0) {
    eeg_set_error("eeg_sym_contract_fwd: C must be a multiple of %d", CH);
    return 1;
  }
#define EEG_X(L, NN) \
  if (lmax == L && nu == NN) return launch_sc<L, NN>(C, x, T, out, N, s);
  EEG_SC_COMBOS(EEG_X)
#undef EEG_X
  eeg_set_error("eeg_sym_contract_fwd: unsupported (lmax=%d, correlation=%d)", lmax, nu);
  return 1;
}
