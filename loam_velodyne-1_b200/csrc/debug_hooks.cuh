// debug_hooks.cuh -- small test hooks exported through the C ABI so that the device versions of
// detmath, linalg, the radix sort and the voxel filter can be checked in isolation (tests/).
#pragma once
#include "sort_voxel.cuh"
#include "linalg.cuh"
#include "lo_kernels2.cuh"

namespace loamgpu {

__global__ void dbg_math_kernel(int fn, const float* x, const float* y, float* out, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float r = 0.f;
  switch (fn) {
    case 0: r = dm::sin(x[i]); break;
    case 1: r = dm::cos(x[i]); break;
    case 2: r = dm::atan(x[i]); break;
    case 3: r = dm::atan2(y[i], x[i]); break;
    case 4: r = dm::asin(x[i]); break;
    case 5: r = dm::sqrt(x[i]); break;
    case 6: r = x[i] / y[i]; break;
    case 7: r = x[i] * y[i] + x[i]; break;  // must NOT be contracted into an FMA
  }
  out[i] = r;
}

// op 0: qr_solve (m x n, 1 rhs) -> out[n]; 1: jacobi_eigen (n x n) -> out[n + n*n]; 2: lu_inverse -> out[n*n];
// 3: qr_solve_fixed<6,6> / <5,3>; 4: jacobi_eigen3
__global__ void dbg_linalg_kernel(int op, const float* A, const float* b, float* out, int m, int n, int batch) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= batch) return;
  float a[36], bb[8], w[8], v[36];
  const float* Ai = A + (size_t)t * m * n;
  for (int i = 0; i < m * n; i++) a[i] = Ai[i];
  if (op == 0) {
    for (int i = 0; i < m; i++) bb[i] = b[(size_t)t * m + i];
    int ok = la::qr_solve(a, m, n, bb, 1);
    for (int i = 0; i < n; i++) out[(size_t)t * n + i] = ok ? bb[i] : 0.f;
  } else if (op == 1) {
    la::jacobi_eigen(a, w, v, n);
    float* o = out + (size_t)t * (n + n * n);
    for (int i = 0; i < n; i++) o[i] = w[i];
    for (int i = 0; i < n * n; i++) o[n + i] = v[i];
  } else if (op == 2) {
    la::lu_inverse(a, v, n);
    for (int i = 0; i < n * n; i++) out[(size_t)t * n * n + i] = v[i];
  } else if (op == 3) {  // register-resident specialisations: same layout as op 0 / op 1
    for (int i = 0; i < m; i++) bb[i] = b[(size_t)t * m + i];
    int ok = 0;
    if (m == 6 && n == 6) ok = la::qr_solve_fixed<6, 6>(a, bb);
    else if (m == 5 && n == 3) ok = la::qr_solve_fixed<5, 3>(a, bb);
    for (int i = 0; i < n; i++) out[(size_t)t * n + i] = ok ? bb[i] : 0.f;
  } else if (op == 4 && n == 3) {
    la::jacobi_eigen3(a, w, v);
    float* o = out + (size_t)t * (n + n * n);
    for (int i = 0; i < n; i++) o[i] = w[i];
    for (int i = 0; i < n * n; i++) o[n + i] = v[i];
  }
}

// lm_solve_step (lo_kernels2.cuh) on `batch` given normal equations, one thread each.  mode 0 = as the
// LM kernels run it (fp64 LDL^T pre-test that may skip the Jacobi eigen-decomposition), mode 1 = the
// eigen-decomposition always (the reference's unconditional cv::eigen, laserOdometry.cpp:770-791)
__global__ void dbg_solve_step_kernel(const float* AtA, const float* AtB, int batch, float thr, int mode, int* deg, float* matP, float* X) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= batch) return;
  double tot[28];
  int k = 0;
  for (int i = 0; i < 6; i++)
    for (int j = i; j < 6; j++) tot[k++] = (double)AtA[(size_t)t * 36 + i * 6 + j];
  for (int i = 0; i < 6; i++) tot[21 + i] = (double)AtB[(size_t)t * 6 + i];
  tot[27] = 0.0;
  int d = 0;
  float P[36], x[6];
  if (mode == 0) lm_solve_step<false>(tot, true, thr, &d, P, x);
  else lm_solve_step<true>(tot, true, thr, &d, P, x);
  deg[t] = d;
  for (int i = 0; i < 36; i++) matP[(size_t)t * 36 + i] = P[i];
  for (int i = 0; i < 6; i++) X[(size_t)t * 6 + i] = x[i];
}

}  // namespace loamgpu
