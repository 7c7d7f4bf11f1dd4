// lm_kernels2.cuh -- laserMapping on sm_100a, part 2: the cooperative sweep-to-map LM kernel
// (LM-8 .. LM-11) and the map update (LM-12 .. LM-14).
// The scalar expressions restated in this file (cited below) come from loam_velodyne:
//   Copyright 2013, Ji Zhang, Carnegie Mellon University
//   Further contributions copyright (c) 2016, Southwest Research Institute.  All rights reserved.
// BSD 3-clause; the full notice, conditions and disclaimer are retained in NOTICE.md at the repository root.
#pragma once
#include "lm_kernels.cuh"
#include "sort_voxel.cuh"
#include "lm_refilter.cuh"

namespace loamgpu {

struct LMOptArgs {
  LMState* st;
  const float4* stack[2];     // laserCloudCornerStack / SurfStack (sensor frame)
  const float4* map[2];       // laserCloudCornerFromMap / SurfFromMap
  GridView grid[2];
  CIView ci[2];               // per-cube index (lm_optimize_ci)
  const CIState* cis;
  double* partials;           // [2][gridDim][LO_NRED]
  int* nn_dbg[2];             // 5 neighbour indices per query of the last iteration (-1 if rejected)
  const LMImu* imu;
  int has_imu;
  double timeLaserOdometry;
  int max_iter;
};

// corner: 5-NN mean, 3x3 covariance, eigen line test, point-to-line residual (:719-818)
__device__ __forceinline__ bool lm_corner_fit(const float4* nb, const float4 pointSel, float4& coeff) {
  float cx = 0; float cy = 0; float cz = 0;
#pragma unroll
  for (int j = 0; j < 5; j++) {
    cx += nb[j].x;
    cy += nb[j].y;
    cz += nb[j].z;
  }
  cx /= 5; cy /= 5; cz /= 5;
  float a11 = 0; float a12 = 0; float a13 = 0; float a22 = 0; float a23 = 0; float a33 = 0;
#pragma unroll
  for (int j = 0; j < 5; j++) {
    float ax = nb[j].x - cx;
    float ay = nb[j].y - cy;
    float az = nb[j].z - cz;
    a11 += ax * ax; a12 += ax * ay; a13 += ax * az;
    a22 += ay * ay; a23 += ay * az; a33 += az * az;
  }
  a11 /= 5; a12 /= 5; a13 /= 5; a22 /= 5; a23 /= 5; a33 /= 5;
  float A1[9] = {a11, a12, a13, a12, a22, a23, a13, a23, a33};
  float D1[3], V1[9];
  la::jacobi_eigen3(A1, D1, V1);
  if (!(D1[0] > 3 * D1[1])) return false;
  float x0 = pointSel.x; float y0 = pointSel.y; float z0 = pointSel.z;
  float x1 = cx + 0.1 * V1[0];
  float y1 = cy + 0.1 * V1[1];
  float z1 = cz + 0.1 * V1[2];
  float x2 = cx - 0.1 * V1[0];
  float y2 = cy - 0.1 * V1[1];
  float z2 = cz - 0.1 * V1[2];
  float a012 = dm::sqrt(((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
             * ((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
             + ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
             * ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
             + ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))
             * ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1)));
  float l12 = dm::sqrt((x1 - x2)*(x1 - x2) + (y1 - y2)*(y1 - y2) + (z1 - z2)*(z1 - z2));
  float la_ = ((y1 - y2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
            + (z1 - z2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))) / a012 / l12;
  float lb = -((x1 - x2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
           - (z1 - z2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
  float lc = -((x1 - x2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
           + (y1 - y2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
  float ld2 = a012 / l12;
  float s = 1 - 0.9 * dm::fabs(ld2);
  coeff.x = s * la_;
  coeff.y = s * lb;
  coeff.z = s * lc;
  coeff.w = s * ld2;
  return s > 0.1;
}

// surf: 5x3 QR plane fit, 0.2 m validity, point-to-plane residual (:826-875)
__device__ __forceinline__ bool lm_surf_fit(const float4* nb, const float4 pointSel, float4& coeff) {
  float A0[15], B0[5];
#pragma unroll
  for (int j = 0; j < 5; j++) {
    A0[j * 3 + 0] = nb[j].x;
    A0[j * 3 + 1] = nb[j].y;
    A0[j * 3 + 2] = nb[j].z;
    B0[j] = -1.f;
  }
  int ok = la::qr_solve_fixed<5, 3>(A0, B0);
  float pa = ok ? B0[0] : 0.f;
  float pb = ok ? B0[1] : 0.f;
  float pc = ok ? B0[2] : 0.f;
  float pd = 1;
  float ps = dm::sqrt(pa * pa + pb * pb + pc * pc);
  pa /= ps; pb /= ps; pc /= ps; pd /= ps;
  bool planeValid = true;
#pragma unroll
  for (int j = 0; j < 5; j++) {
    if (dm::fabs(pa * nb[j].x + pb * nb[j].y + pc * nb[j].z + pd) > 0.2) {
      planeValid = false;
      break;
    }
  }
  if (!planeValid) return false;
  float pd2 = pa * pointSel.x + pb * pointSel.y + pc * pointSel.z + pd;
  float s = 1 - 0.9 * dm::fabs(pd2) / dm::sqrt(dm::sqrt(pointSel.x * pointSel.x + pointSel.y * pointSel.y + pointSel.z * pointSel.z));
  coeff.x = s * pa;
  coeff.y = s * pb;
  coeff.z = s * pc;
  coeff.w = s * pd2;
  return s > 0.1;
}

// Jacobian row :898-920
__device__ __forceinline__ void lm_row(const float4 pointOri, const float4 coeffv, const LMTrig& g, float* a, float& b) {
  float srx = g.srx, crx = g.crx, sry = g.sry, cry = g.cry, srz = g.srz, crz = g.crz;
  struct { float x, y, z; } coeff = {coeffv.x, coeffv.y, coeffv.z};
  float arx = (crx*sry*srz*pointOri.x + crx*crz*sry*pointOri.y - srx*sry*pointOri.z) * coeff.x
            + (-srx*srz*pointOri.x - crz*srx*pointOri.y - crx*pointOri.z) * coeff.y
            + (crx*cry*srz*pointOri.x + crx*cry*crz*pointOri.y - cry*srx*pointOri.z) * coeff.z;
  float ary = ((cry*srx*srz - crz*sry)*pointOri.x
            + (sry*srz + cry*crz*srx)*pointOri.y + crx*cry*pointOri.z) * coeff.x
            + ((-cry*crz - srx*sry*srz)*pointOri.x
            + (cry*srz - crz*srx*sry)*pointOri.y - crx*sry*pointOri.z) * coeff.z;
  float arz = ((crz*srx*sry - cry*srz)*pointOri.x + (-cry*crz-srx*sry*srz)*pointOri.y)*coeff.x
            + (crx*crz*pointOri.x - crx*srz*pointOri.y) * coeff.y
            + ((sry*srz + cry*crz*srx)*pointOri.x + (crz*sry-cry*srx*srz)*pointOri.y)*coeff.z;
  a[0] = arx; a[1] = ary; a[2] = arz; a[3] = coeff.x; a[4] = coeff.y; a[5] = coeff.z;
  b = -coeffv.w;
}

// transformUpdate :199-232
__device__ __forceinline__ void lm_transform_update(LMState* st, float* T, const LMImu* imu, int has_imu, double timeLaserOdometry) {
  const float scanPeriod = 0.1;
  if (has_imu && imu->pointer_last >= 0) {
    float imuRollLast = 0, imuPitchLast = 0;
    int front = st->imuPointerFront;
    const int last = imu->pointer_last;
    while (front != last) {
      if (timeLaserOdometry + scanPeriod < imu->time[front]) break;
      front = (front + 1) % IMU_Q;
    }
    st->imuPointerFront = front;
    if (timeLaserOdometry + scanPeriod > imu->time[front]) {
      imuRollLast = imu->roll[front];
      imuPitchLast = imu->pitch[front];
    } else {
      int back = (front + IMU_Q - 1) % IMU_Q;
      float ratioFront = (timeLaserOdometry + scanPeriod - imu->time[back]) / (imu->time[front] - imu->time[back]);
      float ratioBack = (imu->time[front] - timeLaserOdometry - scanPeriod) / (imu->time[front] - imu->time[back]);
      imuRollLast = imu->roll[front] * ratioFront + imu->roll[back] * ratioBack;
      imuPitchLast = imu->pitch[front] * ratioFront + imu->pitch[back] * ratioBack;
    }
    T[0] = 0.998 * T[0] + 0.002 * imuPitchLast;
    T[2] = 0.998 * T[2] + 0.002 * imuRollLast;
  }
  for (int i = 0; i < 6; i++) {
    st->transformBefMapped[i] = st->transformSum[i];
    st->transformAftMapped[i] = T[i];
  }
}

// K6: the whole LM loop :706-978 as one cooperative kernel.  Every CTA owns a contiguous slice of
// the queries.  Per iteration: (A) one warp per query: 5-NN on the grid index, neighbours parked
// in shared memory; (B) one THREAD per query: line / plane fit and Jacobian row (32 fits per warp
// in flight instead of one on lane 0); (C) 28 threads sum the CTA's rows in query order in fp64;
// then the per-CTA partials are exchanged through one grid barrier and every CTA redundantly
// finishes the sum (ascending CTA order) and the 6x6 solve, so no broadcast is needed.
// The block size is the launch's (LM_THREADS beside the other nodes' kernels, LM_THREADS_ALONE when the node has
// the GPU to itself); nothing in the arithmetic depends on it.
constexpr int LM_CH = 256;   // queries per CTA per pass
#ifndef LOAM_LM_KW
#define LOAM_LM_KW 16
#endif
constexpr int LM_KW = LOAM_LM_KW;    // lanes cooperating on one 5-NN query
#ifndef LOAM_LM_MAXREG
#define LOAM_LM_MAXREG 128
#endif
// CI: neighbours come from the persistent per-cube indexes (lm_cube_index.cuh) instead of the index built over
// the gathered local map; same candidates, same order, same ids after the conversion in the debug output
template <bool CI>
__device__ __forceinline__ void lm_optimize_body(const LMOptArgs& A) {
  cg::grid_group grid = cg::this_grid();
  __shared__ int s_slot[CI ? 2 : 1][CI ? CI_WIN : 1], s_rank[CI ? CI_WIN : 1], s_goff[CI ? 2 : 1][CI ? 126 : 1];
  __shared__ float s_org[3];
  __shared__ float s_T[6];
  __shared__ LMTrig s_trig;
  __shared__ float s_matP[36];
  __shared__ int s_deg, s_stop, s_rows;
  __shared__ double s_g[PS_GROUPS][LO_NRED];
  __shared__ double s_tot[LO_NRED];
  __shared__ float4 s_nb[LM_CH + 4][5];   // +4: scratch slots for the idle groups of a warp
  __shared__ float s_row[LM_CH][8];          // a[0..5], b, valid
  __shared__ unsigned char s_found[LM_CH];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  LMState* st = A.st;
  int clk = (blockIdx.x == 0 && threadIdx.x == 0 && PHASE_ON) ? 1 : -1;
#define PS() do { if (clk >= 0) phase_stamp(1, clk); } while (0)
  PS();
  if (threadIdx.x == 0) {
    for (int i = 0; i < 6; i++) s_T[i] = st->transformTobeMapped[i];
    for (int i = 0; i < 36; i++) s_matP[i] = st->matP[i];
    s_deg = st->isDegenerate;
    s_rows = 0;
  }
  if (CI) {
    for (int i = threadIdx.x; i < 2 * CI_WIN; i += blockDim.x) s_slot[i / CI_WIN][i % CI_WIN] = A.cis->win_slot[i / CI_WIN][i % CI_WIN];
    for (int i = threadIdx.x; i < CI_WIN; i += blockDim.x) s_rank[i] = A.cis->win_rank[i];
    for (int i = threadIdx.x; i < 2 * 126; i += blockDim.x) s_goff[i / 126][i % 126] = st->gatherOff[i / 126][i % 126];
    if (threadIdx.x < 3) s_org[threadIdx.x] = A.cis->org[threadIdx.x];
  }
  __syncthreads();
  const int Qc = st->stackN[0], Qs = st->stackN[1];
  const int Q = Qc + Qs;
  // CTA b owns queries b, b + G, b + 2G, ... : corner and surf queries (different fit cost, different
  // map density) are spread evenly over the CTAs
  const int G = (int)gridDim.x;
  const int nmine = (Q > (int)blockIdx.x) ? (Q - (int)blockIdx.x + G - 1) / G : 0;
  int iters_done = 0;
  const bool enabled = st->optimize_enable != 0;
  if (enabled) {
    GridDesc gd0, gd1;
    if (!CI) { gd0 = *A.grid[0].desc; gd1 = *A.grid[1].desc; }
    for (int iterCount = 0; iterCount < A.max_iter; iterCount++) {
      iters_done = iterCount + 1;
      if (threadIdx.x == 0) s_trig = lm_trig(s_T);
      __syncthreads();
      PS();  // 0: trig done
      const LMTrig g = s_trig;
      double acc = 0.0;
      for (int c0 = 0; c0 < nmine; c0 += LM_CH) {
        const int nq = min(LM_CH, nmine - c0);
        // ---- (A) 5-NN, LM_KW lanes per query (all groups of a warp run the same trip count) ----
        constexpr int GPW = 32 / LM_KW;   // groups per warp
        const int nq_pad = (nq + GPW - 1) / GPW * GPW;
        for (int i = warp * GPW + (lane / LM_KW); i < nq_pad; i += (blockDim.x >> 5) * GPW) {
          const bool act = i < nq;
          const int q = (int)blockIdx.x + (c0 + (act ? i : 0)) * G;
          const int T = q < Qc ? 0 : 1;
          const int k = T ? q - Qc : q;
          const float4 pointSel = lm_to_map(A.stack[T][k], g, s_T);
          // one chunk per CTA (always, at the configured sizes): s_nb[i] / s_found[i] still describe this
          // query's previous iteration, which bounds the new search
          const bool have_prev = act && iterCount > 0 && nmine <= LM_CH && s_found[i] != 0;
          float4* nb = s_nb[act ? i : LM_CH + (lane / LM_KW)];
          bool found;
          if (CI) found = ci_knn5_group<LM_KW>(A.ci[T], s_slot[T], s_org, pointSel.x, pointSel.y, pointSel.z, nb, have_prev);
          else found = lm_knn5_group<LM_KW>(A.grid[T], T ? gd1 : gd0, pointSel.x, pointSel.y, pointSel.z, nb, have_prev);
          if (act && (lane & (LM_KW - 1)) == 0) s_found[i] = found ? 1 : 0;
        }
        __syncthreads();
        if (c0 == 0) PS();  // 1: kNN done
        // ---- (B) fit + Jacobian row, one thread per query ----
        for (int i = threadIdx.x; i < nq; i += blockDim.x) {
          const int q = (int)blockIdx.x + (c0 + i) * G;
          const int T = q < Qc ? 0 : 1;
          const int k = T ? q - Qc : q;
          const bool found = s_found[i] != 0;
          float valid = 0.f;
          if (found) {
            const float4 pointOri = A.stack[T][k];
            const float4 pointSel = lm_to_map(pointOri, g, s_T);
            float4 coeff;
            bool ok = T == 0 ? lm_corner_fit(s_nb[i], pointSel, coeff) : lm_surf_fit(s_nb[i], pointSel, coeff);
            if (ok) {
              float a[6], b;
              lm_row(pointOri, coeff, g, a, b);
#pragma unroll
              for (int j = 0; j < 6; j++) s_row[i][j] = a[j];
              s_row[i][6] = b;
              valid = 1.f;
            }
          }
          s_row[i][7] = valid;
          int* dbg = A.nn_dbg[T] + (size_t)k * 5;
#pragma unroll
          for (int j = 0; j < 5; j++) {
            int id = found ? __float_as_int(s_nb[i][j].w) : -1;
            if (CI && found) id = s_goff[T][s_rank[id >> 24]] + (id & 0x00ffffff);   // position in the gathered local map
            dbg[j] = id;
          }
        }
        __syncthreads();
        if (c0 == 0) PS();  // 2: fits done
        // ---- (C) the CTA's share of AtA / AtB ----
        red_rows(s_row, nq, acc);
        __syncthreads();
      }
      PS();  // 3: CTA rows summed
      double* part = A.partials + (size_t)(iterCount & 1) * gridDim.x * LO_NRED;
      red_finish(grid, acc, part, s_g, s_tot, 1, clk);  // stamps 4 (before barrier), 5 (after)
      PS();  // 6: partials summed
      if (threadIdx.x == 0) {
        s_stop = 0;
        int laserCloudSelNum = (int)s_tot[27];
        s_rows = laserCloudSelNum;
        if (laserCloudSelNum >= 50) {  // :887
          float X[6];
          lm_solve_step(s_tot, iterCount == 0, 100.f, &s_deg, s_matP, X);
          for (int i = 0; i < 6; i++) s_T[i] += X[i];  // :956-961 (no NaN guard in this node)
          if (lm_converged(X, 0.05)) s_stop = 1;
        }
      }
      __syncthreads();
      PS();  // 7: solve done
      if (s_stop) break;
    }
  }
  if (!enabled) {  // no search happened: report "rejected" for every query
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < (Qc + Qs) * 5; q += gridDim.x * blockDim.x) {
      if (q < Qc * 5) A.nn_dbg[0][q] = -1; else A.nn_dbg[1][q - Qc * 5] = -1;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (enabled) lm_transform_update(st, s_T, A.imu, A.has_imu, A.timeLaserOdometry);  // :977
    for (int i = 0; i < 6; i++) st->transformTobeMapped[i] = s_T[i];
    for (int i = 0; i < 36; i++) st->matP[i] = s_matP[i];
    st->isDegenerate = s_deg;
    st->iters = iters_done;
    st->rows = s_rows;
  }
  if (clk >= 0) g_phase_clk[1][0] = clk - 1;
#undef PS
}
__global__ void __maxnreg__(LOAM_LM_MAXREG) lm_optimize(LMOptArgs A) { lm_optimize_body<false>(A); }
__global__ void __maxnreg__(LOAM_LM_MAXREG) lm_optimize_ci(LMOptArgs A) { lm_optimize_body<true>(A); }

// ---- map update: lm_refilter.cuh ----

// K8: registered full-resolution cloud (:1060-1063)
__global__ void __launch_bounds__(256) lm_registered(LMState* st, const float4* in, const int* n_ptr, float4* out) {
  __shared__ LMTrig s_g;
  __shared__ float s_T[6];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 6; i++) s_T[i] = st->insT[i];   // the refined pose of this pass (snapshot of lm_insert_ids)
    s_g = lm_trig(s_T);
  }
  __syncthreads();
  const LMTrig g = s_g;
  const int n = *n_ptr;
  if (blockIdx.x == 0 && threadIdx.x == 0) st->registeredN = n;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) out[i] = lm_to_map(in[i], g, s_T);
}

// ---- LM-14: surround map every mapFrameNum-th mapped sweep (:1042-1057) ----------------------
// laserCloudSurround2 = for every surround cube (i, j, k order): corner points then surf points;
// laserCloudSurround = downSizeFilterCorner (0.2 m) of that, a whole-cloud pcl::VoxelGrid.
struct LMSurroundArgs {
  LMState* st;
  const CubeDesc* cubes[2];
  const float4* arena[2];
  float4* S;           // gathered cloud
  int scap;
};
// one block: segment offsets (cube sizes fetched in parallel, one scan)
__global__ void __launch_bounds__(256) lm_surround_plan(LMSurroundArgs A) {
  LMState* st = A.st;
  __shared__ int s_cnt[256];
  const int ns = st->surroundNum;
  // element 2*i = corner cube i, 2*i+1 = surf cube i
  const int e = threadIdx.x;
  int c = 0;
  if (e < 2 * ns && e < 250) c = A.cubes[e & 1][st->surroundInd[e >> 1]].cnt;
  s_cnt[e] = c;
  __syncthreads();
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    int run = 0;
    for (int b = 0; b < 256; b += 32) {
      int v = s_cnt[b + lane], x = v;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      s_cnt[b + lane] = run + x - v;
      run += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) {
      if (run > A.scap) { atomicOr(&st->error, 8); run = 0; }
      st->surroundIn = run;
    }
  }
  __syncthreads();
  if (e < 250) st->surroundOff[e] = s_cnt[e];
}
// flat copy (see lm_gather_map): point i of the gathered cloud, its segment by binary search
__global__ void __launch_bounds__(256) lm_surround_gather(LMSurroundArgs A) {
  const LMState* st = A.st;
  __shared__ int s_off[256], s_src[256];
  const int ne = min(2 * st->surroundNum, 250);
  const int total = st->surroundIn;
  if (ne == 0 || total == 0) return;
  for (int e = threadIdx.x; e <= ne; e += 256) {
    s_off[e] = e < ne ? st->surroundOff[e] : total;
    s_src[e] = e < ne ? A.cubes[e & 1][st->surroundInd[e >> 1]].off : 0;
  }
  __syncthreads();
  for (int base = blockIdx.x * 1024; base < total; base += gridDim.x * 1024) {
    float4 p[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = min(base + u * 256 + (int)threadIdx.x, total - 1);
      int lo = 0, hi = ne;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_off[mid] <= i) lo = mid; else hi = mid;
      }
      p[u] = __ldg(&A.arena[lo & 1][s_src[lo] + (i - s_off[lo])]);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = base + u * 256 + (int)threadIdx.x;
      if (i < total) A.S[i] = p[u];
    }
  }
}

// arena compaction: copy every live cube segment into the other arena (host-triggered)
struct LMCompactArgs {
  LMState* st;
  CubeDesc* cubes[2];
  const float4* src[2];
  float4* dst[2];
  int* new_off[2];   // [CUBE_NUM]
};
__global__ void __launch_bounds__(1024) lm_compact_plan(LMCompactArgs A) {
  const int T = blockIdx.x;
  __shared__ int s_scan[32];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < CUBE_NUM; base += 1024) {
    int c = base + threadIdx.x;
    int v = c < CUBE_NUM ? A.cubes[T][c].cnt : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane_id() >= o) x += y;
    }
    if (lane_id() == 31) s_scan[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = s_scan[threadIdx.x], s = w;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane_id() >= o) s += y;
      }
      s_scan[threadIdx.x] = s - w;
    }
    __syncthreads();
    int excl = s_carry + s_scan[threadIdx.x >> 5] + x - v;
    if (c < CUBE_NUM) A.new_off[T][c] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) A.st->bump[T] = s_carry;
}
// flat copy: live point i of the new arena, its cube by binary search over the 4851 new offsets
// (19 KB per type, cache resident); the descriptors are switched to the new offsets by
// lm_compact_commit once every point has been copied
__global__ void __launch_bounds__(256) lm_compact_copy(LMCompactArgs A) {
  const int T = blockIdx.y;
  const int total = A.st->bump[T];
  const int* __restrict__ noff = A.new_off[T];
  const CubeDesc* __restrict__ cubes = A.cubes[T];
  for (int base = blockIdx.x * 1024; base < total; base += gridDim.x * 1024) {
    float4 p[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = min(base + u * 256 + (int)threadIdx.x, total - 1);
      int lo = 0, hi = CUBE_NUM;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(&noff[mid]) <= i) lo = mid; else hi = mid;
      }
      p[u] = __ldg(&A.src[T][cubes[lo].off + (i - __ldg(&noff[lo]))]);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = base + u * 256 + (int)threadIdx.x;
      if (i < total) A.dst[T][i] = p[u];
    }
  }
}
__global__ void __launch_bounds__(256) lm_compact_commit(LMCompactArgs A) {
  const int T = blockIdx.y;
  for (int c = blockIdx.x * 256 + threadIdx.x; c < CUBE_NUM; c += gridDim.x * 256) A.cubes[T][c].off = A.new_off[T][c];
}

}  // namespace loamgpu
