// linalg.cuh -- the dense 6x6 / 5x3 / 3x3 linear algebra of the LM loops, host+device.
// Defines, for the product, the arithmetic the reference obtains from OpenCV:
//   gemm        CV_32F Mat*Mat: double accumulation, one rounding   (laserOdometry.cpp:765-767)
//   qr_solve    cv::solve(DECOMP_QR): Householder QR, hal::QR32f     (:768; laserMapping.cpp:832,925)
//   jacobi_eigen cv::eigen on symmetric input, JacobiImpl_           (:775; laserMapping.cpp:767,932)
//   lu_inverse  Mat::inv(): LU with partial pivoting, hal::LU32f     (:790; laserMapping.cpp:947)
// Restated from the published OpenCV >= 3.2 algorithms; tests/test_linalg_golden.py pins them
// bit-for-bit against cv2 4.13 outputs.
#pragma once
#include <float.h>
#include <math.h>
#include "loam/detmath.h"

namespace loamgpu {
namespace la {

// C(m x n) = A(m x k) * B(k x n), double accumulation in ascending k, one rounding
LOAM_HD void gemm(const float* A, const float* B, float* C, int m, int k, int n) {
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      double s = 0.0;
      for (int p = 0; p < k; p++) s += (double)A[i * k + p] * (double)B[p * n + j];
      C[i * n + j] = (float)s;
    }
}

// Householder QR solve of A(m x n) x = b(m x nb), m >= n.  A and b are overwritten; the
// solution is the first n rows of b.  Returns 0 if a diagonal entry of R is < eps.
LOAM_HD int qr_solve(float* A, int m, int n, float* b, int nb) {
  const float eps = FLT_EPSILON * 10;  // hal::QR32f
  float vl[16];
  float hf[16];
  for (int l = 0; l < n; l++) {
    int vs = m - l;
    float vn = 0.f;
    for (int i = 0; i < vs; i++) {
      vl[i] = A[(l + i) * n + l];
      vn += vl[i] * vl[i];
    }
    float tmpV = vl[0];
    vl[0] = vl[0] + (float)((vl[0] >= 0.f) - (vl[0] < 0.f)) * loam::dm::sqrt(vn);
    vn = loam::dm::sqrt(vn + vl[0] * vl[0] - tmpV * tmpV);
    for (int i = 0; i < vs; i++) vl[i] /= vn;
    for (int j = l; j < n; j++) {
      float vA = 0.f;
      for (int i = l; i < m; i++) vA += vl[i - l] * A[i * n + j];
      for (int i = l; i < m; i++) A[i * n + j] -= 2 * vl[i - l] * vA;
    }
    hf[l] = vl[0] * vl[0];
    for (int i = 1; i < vs; i++) A[(l + i) * n + l] = vl[i] / vl[0];
  }
  for (int l = 0; l < n; l++) {
    vl[0] = 1.f;
    for (int j = 1; j < m - l; j++) vl[j] = A[(j + l) * n + l];
    for (int j = 0; j < nb; j++) {
      float vB = 0.f;
      for (int i = l; i < m; i++) vB += vl[i - l] * b[i * nb + j];
      for (int i = l; i < m; i++) b[i * nb + j] -= 2 * vl[i - l] * vB * hf[l];
    }
  }
  for (int i = n - 1; i >= 0; i--) {
    for (int j = n - 1; j > i; j--)
      for (int p = 0; p < nb; p++) b[i * nb + p] -= b[j * nb + p] * A[i * n + j];
    if (fabsf(A[i * n + i]) < eps) return 0;
    for (int p = 0; p < nb; p++) b[i * nb + p] /= A[i * n + i];
  }
  return 1;
}

LOAM_HD float hypot_cv(float a, float b) {
  a = fabsf(a);
  b = fabsf(b);
  if (a > b) {
    b /= a;
    return a * loam::dm::sqrt(1 + b * b);
  }
  if (b > 0) {
    a /= b;
    return b * loam::dm::sqrt(1 + a * a);
  }
  return 0;
}

// Jacobi eigen-decomposition of symmetric A (n x n, overwritten).  W: n eigenvalues descending,
// V: eigenvectors as rows (n x n).  n <= 8.
LOAM_HD void jacobi_eigen(float* A, float* W, float* V, int n) {
  const float eps = FLT_EPSILON;
  int indR[8], indC[8];
  int i, j, k, m;
  for (i = 0; i < n; i++) {
    for (j = 0; j < n; j++) V[i * n + j] = 0.f;
    V[i * n + i] = 1.f;
  }
  int iters, maxIters = n * n * 30;
  float mv = 0.f;
  for (k = 0; k < n; k++) {
    W[k] = A[(n + 1) * k];
    if (k < n - 1) {
      for (m = k + 1, mv = fabsf(A[n * k + m]), i = k + 2; i < n; i++) {
        float val = fabsf(A[n * k + i]);
        if (mv < val) mv = val, m = i;
      }
      indR[k] = m;
    }
    if (k > 0) {
      for (m = 0, mv = fabsf(A[k]), i = 1; i < k; i++) {
        float val = fabsf(A[n * i + k]);
        if (mv < val) mv = val, m = i;
      }
      indC[k] = m;
    }
  }
  if (n > 1)
    for (iters = 0; iters < maxIters; iters++) {
      for (k = 0, mv = fabsf(A[indR[0]]), i = 1; i < n - 1; i++) {
        float val = fabsf(A[n * i + indR[i]]);
        if (mv < val) mv = val, k = i;
      }
      int l = indR[k];
      for (i = 1; i < n; i++) {
        float val = fabsf(A[n * indC[i] + i]);
        if (mv < val) mv = val, k = indC[i], l = i;
      }
      float p = A[n * k + l];
      if (fabsf(p) <= eps) break;
      float y = (float)((W[l] - W[k]) * 0.5);
      float t = fabsf(y) + hypot_cv(p, y);
      float s = hypot_cv(p, t);
      float c = t / s;
      s = p / s;
      t = (p / t) * p;
      if (y < 0) s = -s, t = -t;
      A[n * k + l] = 0;
      W[k] -= t;
      W[l] += t;
      float a0, b0;
#define CVL_ROT(v0, v1) a0 = v0, b0 = v1, v0 = a0 * c - b0 * s, v1 = a0 * s + b0 * c
      for (i = 0; i < k; i++) CVL_ROT(A[n * i + k], A[n * i + l]);
      for (i = k + 1; i < l; i++) CVL_ROT(A[n * k + i], A[n * i + l]);
      for (i = l + 1; i < n; i++) CVL_ROT(A[n * k + i], A[n * l + i]);
      for (i = 0; i < n; i++) CVL_ROT(V[n * k + i], V[n * l + i]);
#undef CVL_ROT
      for (j = 0; j < 2; j++) {
        int idx = j == 0 ? k : l;
        if (idx < n - 1) {
          for (m = idx + 1, mv = fabsf(A[n * idx + m]), i = idx + 2; i < n; i++) {
            float val = fabsf(A[n * idx + i]);
            if (mv < val) mv = val, m = i;
          }
          indR[idx] = m;
        }
        if (idx > 0) {
          for (m = 0, mv = fabsf(A[idx]), i = 1; i < idx; i++) {
            float val = fabsf(A[n * i + idx]);
            if (mv < val) mv = val, m = i;
          }
          indC[idx] = m;
        }
      }
    }
  for (k = 0; k < n - 1; k++) {
    m = k;
    for (i = k + 1; i < n; i++)
      if (W[m] < W[i]) m = i;
    if (k != m) {
      float tw = W[m]; W[m] = W[k]; W[k] = tw;
      for (i = 0; i < n; i++) {
        float tv = V[n * m + i]; V[n * m + i] = V[n * k + i]; V[n * k + i] = tv;
      }
    }
  }
}

// ---- register-resident specialisations (device hot paths) -------------------------------------
// Same operations in the same order as the generic routines above (bit-identical results,
// checked by tests/test_gpu_parity.py::test_device_linalg_fixed_equals_generic); every loop bound
// is a compile-time constant so the arrays live in registers instead of local memory.
template <int M, int N>
LOAM_HD int qr_solve_fixed(float* A, float* b) {   // one right-hand side
  const float eps = FLT_EPSILON * 10;
  float hf[N];
#pragma unroll
  for (int l = 0; l < N; l++) {
    float vl[M];
    float vn = 0.f;
#pragma unroll
    for (int i = 0; i < M - l; i++) {
      vl[i] = A[(l + i) * N + l];
      vn += vl[i] * vl[i];
    }
    float tmpV = vl[0];
    vl[0] = vl[0] + (float)((vl[0] >= 0.f) - (vl[0] < 0.f)) * loam::dm::sqrt(vn);
    vn = loam::dm::sqrt(vn + vl[0] * vl[0] - tmpV * tmpV);
#pragma unroll
    for (int i = 0; i < M - l; i++) vl[i] /= vn;
#pragma unroll
    for (int j = l; j < N; j++) {
      float vA = 0.f;
#pragma unroll
      for (int i = l; i < M; i++) vA += vl[i - l] * A[i * N + j];
#pragma unroll
      for (int i = l; i < M; i++) A[i * N + j] -= 2 * vl[i - l] * vA;
    }
    hf[l] = vl[0] * vl[0];
#pragma unroll
    for (int i = 1; i < M - l; i++) A[(l + i) * N + l] = vl[i] / vl[0];
  }
#pragma unroll
  for (int l = 0; l < N; l++) {
    float vl[M];
    vl[0] = 1.f;
#pragma unroll
    for (int j = 1; j < M - l; j++) vl[j] = A[(j + l) * N + l];
    float vB = 0.f;
#pragma unroll
    for (int i = l; i < M; i++) vB += vl[i - l] * b[i];
#pragma unroll
    for (int i = l; i < M; i++) b[i] -= 2 * vl[i - l] * vB * hf[l];
  }
  int ok = 1;
#pragma unroll
  for (int i = N - 1; i >= 0; i--) {
    if (ok) {
#pragma unroll
      for (int j = N - 1; j > i; j--) b[i] -= b[j] * A[i * N + j];
      if (fabsf(A[i * N + i]) < eps) ok = 0;
      else b[i] /= A[i * N + i];
    }
  }
  return ok;
}

// 3x3 symmetric Jacobi with the matrix in registers.  Only the strict upper triangle is ever read
// by the generic routine; element access by (runtime row, runtime col) is a 3-way select.
struct Sym3 {
  float a01, a02, a12;
  LOAM_HD float get(int r, int c) const { return r == 0 ? (c == 1 ? a01 : a02) : a12; }
  LOAM_HD void set(int r, int c, float v) { if (r == 0) { if (c == 1) a01 = v; else a02 = v; } else a12 = v; }
};
LOAM_HD void jacobi_eigen3(const float* Ain, float* W, float* V) {
  const float eps = FLT_EPSILON;
  Sym3 A;
  A.a01 = Ain[1]; A.a02 = Ain[2]; A.a12 = Ain[5];
  float w0 = Ain[0], w1 = Ain[4], w2 = Ain[8];
  float v[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
  // indR[0] in {1,2}, indR[1] = 2, indC[1] = 0, indC[2] in {0,1}
  int indR0 = 1;
  if (fabsf(A.a01) < fabsf(A.a02)) indR0 = 2;
  int indC2 = 0;
  if (fabsf(A.a02) < fabsf(A.a12)) indC2 = 1;
  for (int iters = 0; iters < 270; iters++) {
    int k = 0;
    float mv = fabsf(A.get(0, indR0));
    { float val = fabsf(A.a12); if (mv < val) mv = val, k = 1; }
    int l = k == 0 ? indR0 : 2;
    { float val = fabsf(A.a01); if (mv < val) mv = val, k = 0, l = 1; }
    { float val = fabsf(A.get(indC2, 2)); if (mv < val) mv = val, k = indC2, l = 2; }
    float p = A.get(k, l);
    if (fabsf(p) <= eps) break;
    float wk = k == 0 ? w0 : w1;           // k in {0,1}
    float wl = l == 1 ? w1 : w2;           // l in {1,2}
    float y = (float)((wl - wk) * 0.5);
    float t = fabsf(y) + hypot_cv(p, y);
    float s = hypot_cv(p, t);
    float c = t / s;
    s = p / s;
    t = (p / t) * p;
    if (y < 0) s = -s, t = -t;
    A.set(k, l, 0.f);
    wk -= t; wl += t;
    if (k == 0) w0 = wk; else w1 = wk;
    if (l == 1) w1 = wl; else w2 = wl;
    float a0, b0;
    // the one remaining index m (not k, not l): exactly one of the three generic loops runs once
    if (k == 0 && l == 1) {        // i = 2 > l : rot(A[k][2], A[l][2])
      a0 = A.a02; b0 = A.a12; A.a02 = a0 * c - b0 * s; A.a12 = a0 * s + b0 * c;
    } else if (k == 0 && l == 2) { // i = 1 in (k,l) : rot(A[k][1], A[1][l])
      a0 = A.a01; b0 = A.a12; A.a01 = a0 * c - b0 * s; A.a12 = a0 * s + b0 * c;
    } else {                       // k == 1, l == 2; i = 0 < k : rot(A[0][k], A[0][l])
      a0 = A.a01; b0 = A.a02; A.a01 = a0 * c - b0 * s; A.a02 = a0 * s + b0 * c;
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
      float vk = k == 0 ? v[i] : v[3 + i];
      float vl_ = l == 1 ? v[3 + i] : v[6 + i];
      float nk = vk * c - vl_ * s, nl = vk * s + vl_ * c;
      if (k == 0) v[i] = nk; else v[3 + i] = nk;
      if (l == 1) v[3 + i] = nl; else v[6 + i] = nl;
    }
    // recompute indR / indC of rows/cols k and l (the only ones with a choice are indR[0], indC[2])
    if (k == 0) { indR0 = 1; if (fabsf(A.a01) < fabsf(A.a02)) indR0 = 2; }
    if (l == 2) { indC2 = 0; if (fabsf(A.a02) < fabsf(A.a12)) indC2 = 1; }
  }
  W[0] = w0; W[1] = w1; W[2] = w2;
#pragma unroll
  for (int i = 0; i < 9; i++) V[i] = v[i];
  // selection sort, descending (same swaps as the generic routine)
#pragma unroll
  for (int k = 0; k < 2; k++) {
    int m = k;
#pragma unroll
    for (int i = k + 1; i < 3; i++)
      if (W[m] < W[i]) m = i;
    if (k != m) {
      float tw = W[m]; W[m] = W[k]; W[k] = tw;
#pragma unroll
      for (int i = 0; i < 3; i++) {
        float tv = V[3 * m + i]; V[3 * m + i] = V[3 * k + i]; V[3 * k + i] = tv;
      }
    }
  }
}

// true when every eigenvalue of the symmetric 6x6 matrix A is provably > tau: LDL^T of A - tau*I
// in fp64 has only positive pivots.  Used to skip the Jacobi eigen-decomposition of the
// degeneracy test when its outcome ("not degenerate") is certain.
LOAM_HD bool spd6_above(const float* A, double tau) {
  double L[36];
#pragma unroll
  for (int i = 0; i < 36; i++) L[i] = (double)A[i];
#pragma unroll
  for (int i = 0; i < 6; i++) L[i * 6 + i] -= tau;
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 6; j++) {
    double dj = L[j * 6 + j];
    if (!(dj > 0.0)) ok = false;
    double inv = ok ? 1.0 / dj : 0.0;
#pragma unroll
    for (int i = j + 1; i < 6; i++) {
      double f = L[i * 6 + j] * inv;
#pragma unroll
      for (int k = j + 1; k <= i; k++) L[i * 6 + k] -= f * L[k * 6 + j];
    }
  }
  return ok;
}

// LU inverse with partial pivoting (hal::LU32f on [A | I]).  A overwritten; returns 0 if singular
// (then inv is all zeros, as cv::invert does).
LOAM_HD int lu_inverse(float* A, float* inv, int m) {
  const float eps = FLT_EPSILON * 10;
  int i, j, k;
  for (i = 0; i < m; i++)
    for (j = 0; j < m; j++) inv[i * m + j] = i == j ? 1.f : 0.f;
  for (i = 0; i < m; i++) {
    k = i;
    for (j = i + 1; j < m; j++)
      if (fabsf(A[j * m + i]) > fabsf(A[k * m + i])) k = j;
    if (fabsf(A[k * m + i]) < eps) {
      for (j = 0; j < m * m; j++) inv[j] = 0.f;
      return 0;
    }
    if (k != i) {
      for (j = i; j < m; j++) { float t = A[i * m + j]; A[i * m + j] = A[k * m + j]; A[k * m + j] = t; }
      for (j = 0; j < m; j++) { float t = inv[i * m + j]; inv[i * m + j] = inv[k * m + j]; inv[k * m + j] = t; }
    }
    float d = -1 / A[i * m + i];
    for (j = i + 1; j < m; j++) {
      float alpha = A[j * m + i] * d;
      for (k = i + 1; k < m; k++) A[j * m + k] += alpha * A[i * m + k];
      for (k = 0; k < m; k++) inv[j * m + k] += alpha * inv[i * m + k];
    }
  }
  for (i = m - 1; i >= 0; i--)
    for (j = 0; j < m; j++) {
      float s = inv[i * m + j];
      for (k = i + 1; k < m; k++) s -= A[i * m + k] * inv[k * m + j];
      inv[i * m + j] = s / A[i * m + i];
    }
  return 1;
}


}  // namespace la
}  // namespace loamgpu
