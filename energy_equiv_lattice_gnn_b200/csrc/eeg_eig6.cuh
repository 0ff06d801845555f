// Eigenvalues of a 6x6 symmetric matrix by cyclic Jacobi (one thread), ascending - the device side of
// torch.linalg.eigvalsh in scripts/train_utils.py:166-170.  Only the lower triangle of `src` is read (LAPACK 'L').
// Symmetric update in the Numerical-Recipes form: only the entries a rotation touches are rewritten, a[p][q] is set to
// zero exactly, all indices are compile-time after unrolling (the matrix lives in registers).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define EEG_HD __host__ __device__
#else
#define EEG_HD
#endif

EEG_HD inline void eeg_eigvalsh6(const float* src, float* ev) {
  float a[6][6];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) a[i][j] = a[j][i] = src[i * 6 + j];
  for (int sweep = 0; sweep < 12; ++sweep) {
    float off = 0.f, diag = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      diag = fmaf(a[i][i], a[i][i], diag);
#pragma unroll
      for (int j = 0; j < i; ++j) off = fmaf(a[i][j], a[i][j], off);
    }
    if (off <= 1e-14f * diag) break;
#pragma unroll
    for (int p = 0; p < 5; ++p)
#pragma unroll
      for (int q = p + 1; q < 6; ++q) {
        const float apq = a[p][q];
        if (fabsf(apq) > 1e-30f) {
          const float theta = (a[q][q] - a[p][p]) / (2.f * apq);
          const float at = fabsf(theta);
          float t = 1.f / (at + sqrtf(fmaf(at, at, 1.f)));
          if (at > 1e18f) t = 0.5f / at;  // theta^2 would overflow
          if (theta < 0.f) t = -t;
          const float c = 1.f / sqrtf(fmaf(t, t, 1.f)), s = t * c;
          a[p][p] -= t * apq;
          a[q][q] += t * apq;
          a[p][q] = a[q][p] = 0.f;
#pragma unroll
          for (int k = 0; k < 6; ++k)
            if (k != p && k != q) {
              const float akp = a[k][p], akq = a[k][q];
              a[k][p] = a[p][k] = c * akp - s * akq;
              a[k][q] = a[q][k] = s * akp + c * akq;
            }
        }
      }
  }
  float e[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) e[i] = a[i][i];
  // sorting network-free insertion by compare-exchange passes (static indices)
#pragma unroll
  for (int pass = 0; pass < 6; ++pass)
#pragma unroll
    for (int i = pass & 1; i + 1 < 6; i += 2) {
      const float lo = fminf(e[i], e[i + 1]), hi = fmaxf(e[i], e[i + 1]);
      e[i] = lo;
      e[i + 1] = hi;
    }
#pragma unroll
  for (int i = 0; i < 6; ++i) ev[i] = e[i];
}

// Eigen-decomposition A = V diag(ev) V^T of a 6x6 symmetric matrix (cyclic Jacobi with accumulated rotations, one thread;
// eigenvalues unsorted, V row-major with eigenvectors in columns).  Device side of torch.linalg.matrix_exp for the
// symmetric Mandel matrices of PositiveLayer('matrix_exp') (gnn/blocks.py:195-196,217-218) and of its adjoint.
EEG_HD inline void eeg_eigh6(const float* src, float* ev, float* V) {
  float a[6][6], v[6][6];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      a[i][j] = 0.5f * (src[i * 6 + j] + src[j * 6 + i]);
      v[i][j] = i == j ? 1.f : 0.f;
    }
  for (int sweep = 0; sweep < 16; ++sweep) {
    float off = 0.f, diag = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      diag = fmaf(a[i][i], a[i][i], diag);
#pragma unroll
      for (int j = 0; j < i; ++j) off = fmaf(a[i][j], a[i][j], off);
    }
    if (off <= 1e-15f * diag || off == 0.f) break;
#pragma unroll
    for (int p = 0; p < 5; ++p)
#pragma unroll
      for (int q = p + 1; q < 6; ++q) {
        const float apq = a[p][q];
        if (fabsf(apq) > 1e-30f) {
          const float theta = (a[q][q] - a[p][p]) / (2.f * apq);
          const float at = fabsf(theta);
          float t = 1.f / (at + sqrtf(fmaf(at, at, 1.f)));
          if (at > 1e18f) t = 0.5f / at;
          if (theta < 0.f) t = -t;
          const float c = 1.f / sqrtf(fmaf(t, t, 1.f)), s = t * c;
          a[p][p] -= t * apq;
          a[q][q] += t * apq;
          a[p][q] = a[q][p] = 0.f;
#pragma unroll
          for (int k = 0; k < 6; ++k) {
            if (k != p && k != q) {
              const float akp = a[k][p], akq = a[k][q];
              a[k][p] = a[p][k] = c * akp - s * akq;
              a[k][q] = a[q][k] = s * akp + c * akq;
            }
            const float vkp = v[k][p], vkq = v[k][q];
            v[k][p] = c * vkp - s * vkq;
            v[k][q] = s * vkp + c * vkq;
          }
        }
      }
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    ev[i] = a[i][i];
#pragma unroll
    for (int j = 0; j < 6; ++j) V[i * 6 + j] = v[i][j];
  }
}

// Divided difference (exp(a) - exp(b)) / (a - b), the kernel of the Frechet derivative of exp at a symmetric matrix.
EEG_HD inline float eeg_exp_divdiff(float a, float b) {
  const float d = a - b;
  if (fabsf(d) < 1e-2f) return expf(b) * (1.f + d * (0.5f + d * (1.f / 6.f + d * (1.f / 24.f))));
  return (expf(a) - expf(b)) / d;
}
