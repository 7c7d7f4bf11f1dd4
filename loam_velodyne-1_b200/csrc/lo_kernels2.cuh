// lo_kernels2.cuh -- the persistent cooperative LM kernel of laserOdometry and its satellites.
#pragma once
#include "lo_kernels.cuh"

namespace loamgpu {

// Solve step shared by odometry and mapping: normal equations from the reduced sums, QR solve,
// (iteration 0) eigen-degeneracy analysis, projection.  laserOdometry.cpp:765-797 /
// laserMapping.cpp:922-954.  tot[0..20] upper triangle of AtA, tot[21..26] AtB.
__device__ __forceinline__ void lm_solve_step(const double* tot, bool first_iter, float eig_thresh, int* isDegenerate, float* matP, float* X) {
  float AtA[36], AtB[6], A2[36];
  int k = 0;
  for (int i = 0; i < 6; i++)
    for (int j = i; j < 6; j++) {
      float v = (float)tot[k++];
      AtA[i * 6 + j] = v;
      AtA[j * 6 + i] = v;
    }
  for (int i = 0; i < 6; i++) AtB[i] = (float)tot[21 + i];
  for (int i = 0; i < 36; i++) A2[i] = AtA[i];
  float b2[6];
  for (int i = 0; i < 6; i++) b2[i] = AtB[i];
  int ok = la::qr_solve(A2, 6, 6, b2, 1);
  for (int i = 0; i < 6; i++) X[i] = ok ? b2[i] : 0.f;
  if (first_iter) {
    float E[6], V[36], V2[36], Vc[36], Vinv[36];
    for (int i = 0; i < 36; i++) A2[i] = AtA[i];
    la::jacobi_eigen(A2, E, V, 6);
    for (int i = 0; i < 36; i++) V2[i] = V[i];
    *isDegenerate = 0;
    for (int i = 5; i >= 0; i--) {
      if (E[i] < eig_thresh) {
        for (int j = 0; j < 6; j++) V2[i * 6 + j] = 0;
        *isDegenerate = 1;
      } else {
        break;
      }
    }
    for (int i = 0; i < 36; i++) Vc[i] = V[i];
    la::lu_inverse(Vc, Vinv, 6);
    la::gemm(Vinv, V2, matP, 6, 6, 6);
  }
  if (*isDegenerate) {
    float X2[6];
    for (int i = 0; i < 6; i++) X2[i] = X[i];
    la::gemm(matP, X2, X, 6, 6, 1);
  }
}

__device__ __forceinline__ bool lm_converged(const float* X, double thr) {
  // :815-826 (pow(x,2) in double, rad2deg in double)
  double r0 = (double)X[0] * 180.0 / M_PI, r1 = (double)X[1] * 180.0 / M_PI, r2 = (double)X[2] * 180.0 / M_PI;
  float deltaR = (float)dm::dsqrt(r0 * r0 + r1 * r1 + r2 * r2);
  double t0 = (double)(X[3] * 100), t1 = (double)(X[4] * 100), t2 = (double)(X[5] * 100);
  float deltaT = (float)dm::dsqrt(t0 * t0 + t1 * t1 + t2 * t2);
  return deltaR < thr && deltaT < thr;
}

// pose integration :830-856
__device__ __forceinline__ void lo_integrate(LOState* st, const float* transform, const float* imu) {
  float* transformSum = st->transformSum;
  const float imuPitchStart = imu[0], imuYawStart = imu[1], imuRollStart = imu[2];
  const float imuPitchLast = imu[3], imuYawLast = imu[4], imuRollLast = imu[5];
  const float imuShiftFromStartX = imu[6], imuShiftFromStartY = imu[7], imuShiftFromStartZ = imu[8];
  float rx, ry, rz, tx, ty, tz;
  pm::AccumulateRotation(transformSum[0], transformSum[1], transformSum[2],
                         -transform[0], -transform[1] * 1.05, -transform[2], rx, ry, rz);
  float x1 = dm::cos(rz) * (transform[3] - imuShiftFromStartX) - dm::sin(rz) * (transform[4] - imuShiftFromStartY);
  float y1 = dm::sin(rz) * (transform[3] - imuShiftFromStartX) + dm::cos(rz) * (transform[4] - imuShiftFromStartY);
  float z1 = transform[5] * 1.05 - imuShiftFromStartZ;
  float x2 = x1;
  float y2 = dm::cos(rx) * y1 - dm::sin(rx) * z1;
  float z2 = dm::sin(rx) * y1 + dm::cos(rx) * z1;
  tx = transformSum[3] - (dm::cos(ry) * x2 + dm::sin(ry) * z2);
  ty = transformSum[4] - y2;
  tz = transformSum[5] - (-dm::sin(ry) * x2 + dm::cos(ry) * z2);
  pm::PluginIMURotation(rx, ry, rz, imuPitchStart, imuYawStart, imuRollStart,
                        imuPitchLast, imuYawLast, imuRollLast, rx, ry, rz);
  transformSum[0] = rx; transformSum[1] = ry; transformSum[2] = rz;
  transformSum[3] = tx; transformSum[4] = ty; transformSum[5] = tz;
}

// The whole per-sweep LM loop :458-856 as one cooperative kernel.
__global__ void __launch_bounds__(LO_THREADS) lo_optimize(LOArgs A) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float s_tr[6];
  __shared__ float s_trig[6];
  __shared__ float s_imu[12];
  __shared__ float s_matP[36];
  __shared__ int s_deg, s_stop;
  __shared__ double s_red[LO_WARPS][LO_NRED];
  __shared__ double s_tot[LO_NRED];
  __shared__ int s_rows;
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const int gwarp = blockIdx.x * LO_WARPS + warp, nwarps = gridDim.x * LO_WARPS;
  LOState* st = A.st;
  if (threadIdx.x == 0) {
    lo_load_imu(A.imu, s_imu);
    const float scanPeriod = 0.1;
    for (int i = 0; i < 6; i++) s_tr[i] = st->transform[i];
    s_tr[3] -= s_imu[9] * scanPeriod;   // :461-463
    s_tr[4] -= s_imu[10] * scanPeriod;
    s_tr[5] -= s_imu[11] * scanPeriod;
    for (int i = 0; i < 36; i++) s_matP[i] = st->matP[i];
    s_deg = st->isDegenerate;
    s_rows = 0;
  }
  __syncthreads();
  const int n_sharp = A.cnt->n_sharp, n_flat = A.cnt->n_flat;
  const int F = n_sharp + n_flat;
  const int n_corner_last = st->n_corner_last, n_surf_last = st->n_surf_last;
  int iters_done = 0;
  if (st->cornerLastNum > 10 && st->surfLastNum > 100) {  // :465
    for (int iterCount = 0; iterCount < A.max_iter; iterCount++) {
      iters_done = iterCount + 1;
      const bool refresh = (iterCount % 5 == 0);
      // ---- features: transform, (search), residual -> row slab of this iteration ----
      for (int f = gwarp; f < F; f += nwarps) {
        const bool is_corner = f < n_sharp;
        const float4 pt = is_corner ? A.sharp[f] : A.flat[f - n_sharp];
        const float4 sel = lo_transform_to_start(pt, s_tr);
        bool valid = false;
        float4 coeff = make_float4(0, 0, 0, 0);
        if (is_corner) {
          int i1, i2;
          if (refresh) {
            lo_corner_search(A, sel, n_sharp, n_corner_last, i1, i2);
            if (lane == 0) { A.corner_ind1[f] = i1; A.corner_ind2[f] = i2; }
          } else {
            i1 = A.corner_ind1[f]; i2 = A.corner_ind2[f];
          }
          if (lane == 0 && i2 >= 0) valid = lo_corner_residual(sel, A.corner_last[i1], A.corner_last[i2], iterCount, coeff);
        } else {
          int g = f - n_sharp;
          int i1, i2, i3;
          if (refresh) {
            lo_surf_search(A, sel, n_flat, n_surf_last, i1, i2, i3);
            if (lane == 0) { A.surf_ind1[g] = i1; A.surf_ind2[g] = i2; A.surf_ind3[g] = i3; }
          } else {
            i1 = A.surf_ind1[g]; i2 = A.surf_ind2[g]; i3 = A.surf_ind3[g];
          }
          if (lane == 0 && i2 >= 0 && i3 >= 0)
            valid = lo_surf_residual(sel, A.surf_last[i1], A.surf_last[i2], A.surf_last[i3], iterCount, coeff);
        }
        if (lane == 0) {
          A.coeff[(size_t)iterCount * A.fcap + f] = coeff;
          A.row_valid[(size_t)iterCount * A.fcap + f] = valid ? 1 : 0;
        }
      }
      __syncwarp();
      // ---- Jacobian of every row accumulated so far (rows are NOT cleared per iteration) ----
      if (threadIdx.x == 0) {
        s_trig[0] = dm::sin(s_tr[0]); s_trig[1] = dm::cos(s_tr[0]);
        s_trig[2] = dm::sin(s_tr[1]); s_trig[3] = dm::cos(s_tr[1]);
        s_trig[4] = dm::sin(s_tr[2]); s_trig[5] = dm::cos(s_tr[2]);
      }
      __syncthreads();
      double acc[LO_NRED];
#pragma unroll
      for (int k = 0; k < LO_NRED; k++) acc[k] = 0.0;
      for (int f = gwarp; f < F; f += nwarps) {
        if (lane <= iterCount) {
          size_t slot = (size_t)lane * A.fcap + f;
          if (A.row_valid[slot]) {
            const float4 pt = f < n_sharp ? A.sharp[f] : A.flat[f - n_sharp];
            lo_row_accumulate(pt, A.coeff[slot], s_trig, s_tr, acc);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < LO_NRED; k++) {
        double v = acc[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][k] = v;
      }
      __syncthreads();
      double* part = A.partials + (size_t)(iterCount & 1) * gridDim.x * LO_NRED;
      if (threadIdx.x < LO_NRED) {
        double v = 0.0;
        for (int w = 0; w < LO_WARPS; w++) v += s_red[w][threadIdx.x];
        part[(size_t)blockIdx.x * LO_NRED + threadIdx.x] = v;
      }
      __threadfence();
      grid.sync();
      if (threadIdx.x < LO_NRED) {
        double v = 0.0;
        for (int b = 0; b < (int)gridDim.x; b++) v += __ldcg(&part[(size_t)b * LO_NRED + threadIdx.x]);
        s_tot[threadIdx.x] = v;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        s_stop = 0;
        int pointSelNum = (int)s_tot[27];
        s_rows = pointSelNum;
        if (pointSelNum >= 10) {  // :697-700
          float X[6];
          lm_solve_step(s_tot, iterCount == 0, 10.f, &s_deg, s_matP, X);
          bool nan = isnan(X[0]) || isnan(X[1]) || isnan(X[2]) || isnan(X[3]) || isnan(X[4]) || isnan(X[5]);  // :800
          if (!nan) {
            for (int i = 0; i < 6; i++) s_tr[i] += X[i];
          }
          if (lm_converged(X, 0.1)) s_stop = 1;
        }
      }
      __syncthreads();
      if (s_stop) break;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    lo_integrate(st, s_tr, s_imu);
    for (int i = 0; i < 6; i++) st->transform[i] = s_tr[i];
    for (int i = 0; i < 36; i++) st->matP[i] = s_matP[i];
    st->isDegenerate = s_deg;
    st->iters = iters_done;
    st->rows = s_rows;
  }
}

// TransformToEnd over less-sharp -> corner_last, less-flat -> surf_last, (optionally) full ->
// full_end (:875-891); the buffer "swap" (:893-899) is the write into the *_last arrays.
struct ToEndArgs {
  LOState* st;
  const SRCounts* cnt;
  const ImuState* imu;
  const float4* less_sharp; const float4* less_flat; const float4* full;
  float4* corner_last; float4* surf_last; float4* full_end;
  int do_full;
};
__global__ void __launch_bounds__(256) lo_transform_to_end_kernel(ToEndArgs A) {
  __shared__ float s_tr[6];
  __shared__ float s_imu[12];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 6; i++) s_tr[i] = A.st->transform[i];
    lo_load_imu(A.imu, s_imu);
  }
  __syncthreads();
  const int n0 = A.cnt->n_less_sharp, n1 = A.cnt->n_less_flat, n2 = A.do_full ? A.cnt->n_full : 0;
  const int total = n0 + n1 + n2;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    if (i < n0) A.corner_last[i] = lo_transform_to_end(A.less_sharp[i], s_tr, s_imu);
    else if (i < n0 + n1) A.surf_last[i - n0] = lo_transform_to_end(A.less_flat[i - n0], s_tr, s_imu);
    else A.full_end[i - n0 - n1] = lo_transform_to_end(A.full[i - n0 - n1], s_tr, s_imu);
  }
}

// bookkeeping after the swap (:901-903); mode 1 = first sweep (:427-456: plain adoption of the
// clouds, transformSum seeded from the IMU, the *LastNum counters are NOT touched)
__global__ void lo_finish(LOState* st, const SRCounts* cnt, const ImuState* imu, int first_sweep, int do_full) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  st->n_corner_last = cnt->n_less_sharp;
  st->n_surf_last = cnt->n_less_flat;
  if (first_sweep) {
    float im[12];
    lo_load_imu(imu, im);
    st->transformSum[0] += im[0];  // imuPitchStart
    st->transformSum[2] += im[2];  // imuRollStart
    st->index_enable = 1;          // kd-trees are built unconditionally at :436-437
    st->iters = 0; st->rows = 0;
  } else {
    st->cornerLastNum = cnt->n_less_sharp;
    st->surfLastNum = cnt->n_less_flat;
    st->index_enable = (st->cornerLastNum > 10 && st->surfLastNum > 100) ? 1 : 0;
    if (do_full) st->n_full_end = cnt->n_full;
  }
}

__global__ void __launch_bounds__(256) lo_adopt_first(const SRCounts* cnt, const float4* less_sharp, const float4* less_flat,
                                                      float4* corner_last, float4* surf_last) {
  const int n0 = cnt->n_less_sharp, n1 = cnt->n_less_flat;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n0 + n1; i += gridDim.x * 256) {
    if (i < n0) corner_last[i] = less_sharp[i];
    else surf_last[i - n0] = less_flat[i - n0];
  }
}

}  // namespace loamgpu
