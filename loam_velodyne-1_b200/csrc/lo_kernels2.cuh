// lo_kernels2.cuh -- the persistent cooperative LM kernel of laserOdometry and its satellites.
#pragma once
#include "lo_kernels.cuh"

namespace loamgpu {

// Solve step shared by odometry and mapping: normal equations from the reduced sums, QR solve,
// (iteration 0) eigen-degeneracy analysis, projection.  laserOdometry.cpp:765-797 /
// laserMapping.cpp:922-954.  tot[0..20] upper triangle of AtA, tot[21..26] AtB.
template <bool ALWAYS_EIGEN = false>
__device__ __forceinline__ void lm_solve_step(const double* tot, bool first_iter, float eig_thresh, int* isDegenerate, float* matP, float* X) {
  float AtA[36], A2[36], b2[6];
  {
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
      for (int j = i; j < 6; j++) {
        float v = (float)tot[k++];
        AtA[i * 6 + j] = v;
        AtA[j * 6 + i] = v;
      }
  }
#pragma unroll
  for (int i = 0; i < 6; i++) b2[i] = (float)tot[21 + i];
#pragma unroll
  for (int i = 0; i < 36; i++) A2[i] = AtA[i];
  int ok = la::qr_solve_fixed<6, 6>(A2, b2);
#pragma unroll
  for (int i = 0; i < 6; i++) X[i] = ok ? b2[i] : 0.f;
  if (first_iter) {
    // cv::eigen + threshold scan only decides isDegenerate (and matP, which is read only when
    // isDegenerate is set).  If A - tau*I is positive definite with tau comfortably above the
    // threshold (margin >> the fp32 Jacobi's eigenvalue error ~ 1e-6 * trace), every computed
    // eigenvalue is >= threshold and the outcome is known without the decomposition.
    double trace = 0.0;
#pragma unroll
    for (int i = 0; i < 6; i++) trace += (double)AtA[i * 6 + i];
    if (!ALWAYS_EIGEN && la::spd6_above(AtA, (double)eig_thresh + 1.0e-4 * trace)) {
      *isDegenerate = 0;
      for (int i = 0; i < 36; i++) matP[i] = (i % 7 == 0) ? 1.f : 0.f;
    } else {
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

// sines and cosines of up to 16 angles, one evaluation per lane: lane l < 16 writes sn[l] = sin(angles[l]),
// lane 16 + l writes cs[l] = cos(angles[l])
__device__ __forceinline__ void trig_table_warp(const float* angles, int count, float* sn, float* cs) {
  const int lane = lane_id(), j = lane & 15;
  if (j < count) {
    const float a = angles[j];
    if (lane < 16) sn[j] = dm::sin(a); else cs[j] = dm::cos(a);
  }
  __syncwarp();
}

// lo_integrate (:830-856) on one warp: the sines / cosines of the input angles of AccumulateRotation and
// PluginIMURotation are evaluated by the lanes at once (a single thread needs ~0.5 us per fp64 fdlibm call and
// there are ~45 of them), lane 0 then runs the closed forms against the tables -- same expression trees, and
// every table entry is what the same function returns for the same argument
struct LOIntegrateSmem { float ang[16], sn[16], cs[16]; };
__device__ __forceinline__ void lo_integrate_warp(LOState* st, const float* transform, const float* imu, LOIntegrateSmem* W) {
  const int lane = lane_id();
  float* transformSum = st->transformSum;
  // table 1: cx cy cz lx ly lz of AccumulateRotation, then the six IMU angles (bl*, al* of PluginIMURotation)
  if (lane == 0) {
    W->ang[0] = transformSum[0]; W->ang[1] = transformSum[1]; W->ang[2] = transformSum[2];
    W->ang[3] = -transform[0]; W->ang[4] = -transform[1] * 1.05; W->ang[5] = -transform[2];
    for (int i = 0; i < 6; i++) W->ang[6 + i] = imu[i];
  }
  __syncwarp();
  trig_table_warp(W->ang, 12, W->sn, W->cs);
  if (lane == 0) {
    float rx, ry, rz;
    pm::AccumulateRotationV(W->sn, W->cs, transformSum[0], transformSum[1], transformSum[2],
                            -transform[0], -transform[1] * 1.05, -transform[2], rx, ry, rz);
    W->ang[12] = rx; W->ang[13] = ry; W->ang[14] = rz;
  }
  __syncwarp();
  trig_table_warp(W->ang + 12, 3, W->sn + 12, W->cs + 12);
  if (lane == 0) {
    const float imuPitchStart = imu[0], imuYawStart = imu[1], imuRollStart = imu[2];
    const float imuPitchLast = imu[3], imuYawLast = imu[4], imuRollLast = imu[5];
    const float imuShiftFromStartX = imu[6], imuShiftFromStartY = imu[7], imuShiftFromStartZ = imu[8];
    float rx = W->ang[12], ry = W->ang[13], rz = W->ang[14], tx, ty, tz;
    const float srx = W->sn[12], sry = W->sn[13], srz = W->sn[14], crx = W->cs[12], cry = W->cs[13], crz = W->cs[14];
    float x1 = crz * (transform[3] - imuShiftFromStartX) - srz * (transform[4] - imuShiftFromStartY);
    float y1 = srz * (transform[3] - imuShiftFromStartX) + crz * (transform[4] - imuShiftFromStartY);
    float z1 = transform[5] * 1.05 - imuShiftFromStartZ;
    float x2 = x1;
    float y2 = crx * y1 - srx * z1;
    float z2 = srx * y1 + crx * z1;
    tx = transformSum[3] - (cry * x2 + sry * z2);
    ty = transformSum[4] - y2;
    tz = transformSum[5] - (-sry * x2 + cry * z2);
    // PluginIMURotation(bc = rx ry rz, bl = imu start angles, al = imu last angles): its table is
    // [bcx bcy bcz blx bly blz alx aly alz] = table entries 12..14 followed by 6..11
    float sn9[9], cs9[9];
    for (int i = 0; i < 3; i++) { sn9[i] = W->sn[12 + i]; cs9[i] = W->cs[12 + i]; }
    for (int i = 0; i < 6; i++) { sn9[3 + i] = W->sn[6 + i]; cs9[3 + i] = W->cs[6 + i]; }
    pm::PluginIMURotationV(sn9, cs9, rx, ry, rz, imuPitchStart, imuYawStart, imuRollStart,
                           imuPitchLast, imuYawLast, imuRollLast, rx, ry, rz);
    transformSum[0] = rx; transformSum[1] = ry; transformSum[2] = rz;
    transformSum[3] = tx; transformSum[4] = ty; transformSum[5] = tz;
  }
  __syncwarp();
}

// ---- deterministic reductions shared by lo_optimize and lm_optimize ---------------------------
// rows: [n][8] floats = a[0..5], b, valid.  28 sums (21 upper AtA, 6 AtB, row count) in fp64.
constexpr int RED_GROUPS = 8;
__device__ __forceinline__ void red_entry(int e, int& ri, int& rj) {
  ri = 0; rj = 0;
  if (e < 21) {
    int k = e, i = 0;
    while (k >= 6 - i) { k -= 6 - i; i++; }
    ri = i; rj = i + k;
  } else if (e < 27) { ri = e - 21; rj = 6; }
}
// thread (e = t % 28, grp = t / 28), t < 28 * RED_GROUPS, adds rows grp*len .. (grp+1)*len-1 of
// this pass to its running sum `acc` (ascending row order inside a group)
__device__ __forceinline__ void red_rows(const float (*rows)[8], int n, double& acc) {
  const int t = threadIdx.x;
  if (t >= LO_NRED * RED_GROUPS) return;
  const int e = t % LO_NRED, grp = t / LO_NRED;
  int ri, rj;
  red_entry(e, ri, rj);
  const int len = (n + RED_GROUPS - 1) / RED_GROUPS;
  const int r0 = grp * len, r1 = min(n, r0 + len);
  for (int r = r0; r < r1; r++) {
    if (rows[r][7] != 0.f) {
      if (e < 27) acc += (double)rows[r][ri] * (double)rows[r][rj];
      else acc += 1.0;
    }
  }
}
// same for rows laid out [warp][32] of which only slots < used per warp can be valid; group grp
// covers warps grp*wpg .. grp*wpg+wpg-1
__device__ __forceinline__ void red_rows_warped(const float (*rows)[8], int nwarps, int used, double& acc) {
  const int t = threadIdx.x;
  if (t >= LO_NRED * RED_GROUPS) return;
  const int e = t % LO_NRED, grp = t / LO_NRED;
  int ri, rj;
  red_entry(e, ri, rj);
  const int wpg = (nwarps + RED_GROUPS - 1) / RED_GROUPS;
  for (int w = grp * wpg; w < min(nwarps, (grp + 1) * wpg); w++)
    for (int l = 0; l < used; l++) {
      const float* row = rows[w * 32 + l];
      if (row[7] != 0.f) {
        if (e < 27) acc += (double)row[ri] * (double)row[rj];
        else acc += 1.0;
      }
    }
}
// rows laid out [feature block][32 slabs], 2 * RED_GROUPS feature blocks per pass (two features per warp): group grp
// sums block grp, then block RED_GROUPS + grp -- the order in which two successive one-feature-per-warp passes
// (red_rows_warped twice) would have added them
__device__ __forceinline__ void red_rows_paired(const float (*rows)[8], int used, double& acc) {
  const int t = threadIdx.x;
  if (t >= LO_NRED * RED_GROUPS) return;
  const int e = t % LO_NRED, grp = t / LO_NRED;
  int ri, rj;
  red_entry(e, ri, rj);
#pragma unroll
  for (int p = 0; p < 2; p++)
    for (int l = 0; l < used; l++) {
      const float* row = rows[(p * RED_GROUPS + grp) * 32 + l];
      if (row[7] != 0.f) {
        if (e < 27) acc += (double)row[ri] * (double)row[rj];
        else acc += 1.0;
      }
    }
}
// CTA total = groups in ascending order -> partials[blockIdx]; one grid barrier; every CTA then
// sums all partials (16 interleaved lanes per entry, then ascending) into s_tot[28].
constexpr int PS_GROUPS = 16;
__device__ __forceinline__ void red_finish(cg::grid_group& grid, double acc, double* part, double (*s_g)[LO_NRED], double* s_tot, int which, int& clk) {
  const int t = threadIdx.x;
  if (t < LO_NRED * RED_GROUPS) s_g[t / LO_NRED][t % LO_NRED] = acc;
  __syncthreads();
  if (t < LO_NRED) {
    double v = 0.0;
#pragma unroll
    for (int g = 0; g < RED_GROUPS; g++) v += s_g[g][t];
    part[(size_t)blockIdx.x * LO_NRED + t] = v;
  }
  __threadfence();
  if (clk >= 0) phase_stamp(which, clk);
  grid.sync();
  if (clk >= 0) phase_stamp(which, clk);
  for (int u = t; u < LO_NRED * PS_GROUPS; u += blockDim.x) {
    const int e = u % LO_NRED, g = u / LO_NRED;
    // same ascending sum as `for (b = g; b < gridDim.x; b += PS_GROUPS) v += part[b][e]`, ten loads in flight at a time
    // (one dependent L2 round trip per partial was 5 us of every LM iteration)
    double v = 0.0;
    for (int b0 = g; b0 < (int)gridDim.x; b0 += 10 * PS_GROUPS) {
      double x[10];
#pragma unroll
      for (int k = 0; k < 10; k++) {
        const int b = b0 + k * PS_GROUPS;
        x[k] = b < (int)gridDim.x ? __ldcg(&part[(size_t)b * LO_NRED + e]) : 0.0;
      }
#pragma unroll
      for (int k = 0; k < 10; k++)
        if (b0 + k * PS_GROUPS < (int)gridDim.x) v += x[k];
    }
    s_g[g][e] = v;
  }
  __syncthreads();
  if (t < LO_NRED) {
    double v = 0.0;
#pragma unroll
    for (int g = 0; g < PS_GROUPS; g++) v += s_g[g][t];
    s_tot[t] = v;
  }
  __syncthreads();
}

// The whole per-sweep LM loop :458-856 as one cooperative kernel.  Every CTA owns a contiguous
// slice of the features, one warp per feature.
#ifndef LOAM_LO_MAXREG
#define LOAM_LO_MAXREG 128
#endif
__global__ void __maxnreg__(LOAM_LO_MAXREG) lo_optimize(LOArgs A) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float s_tr[6];
  __shared__ float s_trig[6];
  __shared__ float s_imu[12];
  __shared__ float s_matP[36];
  __shared__ int s_deg, s_stop;
  __shared__ double s_g[PS_GROUPS][LO_NRED];
  __shared__ double s_tot[LO_NRED];
  __shared__ int s_rows;
  __shared__ int s_roff[2][SR_MAX_RINGS + 1];
  __shared__ float s_row[2 * LO_WARPS * 32][8]; // one row per (feature of this pass, iteration slab); the paired form holds two features per warp
  __shared__ float4 s_tri[2 * LO_WARPS][3];     // tripod of the warp's first-pass feature(s)
  __shared__ int s_triok[2 * LO_WARPS];
  __shared__ float4 s_cur[2 * LO_WARPS];        // paired form: this iteration's coefficients, leader lane -> row lane
  __shared__ int s_curok[2 * LO_WARPS];
  static_assert(LO_WARPS == RED_GROUPS, "one reduction group per warp");
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  LOState* st = A.st;
  int clk = (blockIdx.x == 0 && threadIdx.x == 0 && PHASE_ON) ? 1 : -1;
#define PS() do { if (clk >= 0) phase_stamp(0, clk); } while (0)
  PS();
  const int R = A.n_scans;
  for (int i = threadIdx.x; i < 2 * (R + 1); i += LO_THREADS) {
    int T = i / (R + 1), r = i % (R + 1);
    s_roff[T][r] = T ? A.surf_roff[r] : A.corner_roff[r];
  }
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
  if (F > A.fcap) {   // cannot happen through the C ABI (set_features / sr_select bound the clouds); never index past the row slabs
    if (blockIdx.x == 0 && threadIdx.x == 0) { st->iters = 0; st->rows = 0; }
    return;
  }
  const int n_corner_last = st->n_corner_last, n_surf_last = st->n_surf_last;
  // CTA b owns features b, b + G, b + 2G, ...: the corner features (the first n_sharp ones; their two ring
  // scans make them the expensive ones) are spread evenly over the CTAs instead of filling the first few,
  // which is what the slowest CTA at the iteration's grid barrier was waiting for.  A.contiguous = contiguous
  // slices.
  // A.contiguous = 2: block-cyclic -- chunks of LO_WARPS consecutive features (one pass of the CTA), CTA b takes chunks
  // b, b + G, ...: neighbouring features still share a pass (and their cells in L1/L2), while every CTA gets a piece of
  // the cheap head of the list and a piece of the expensive tail.
  const int G = (int)gridDim.x;
  const int per = (F + G - 1) / G;
  const bool cyclic = A.contiguous == 2;
  const int nchunks = (F + LO_WARPS - 1) / LO_WARPS;
  const int f_lo = A.contiguous ? min((int)blockIdx.x * per, F) : (int)blockIdx.x;
  const int f_step = A.contiguous ? 1 : G;
  const int nmine = cyclic ? ((nchunks > (int)blockIdx.x) ? (nchunks - (int)blockIdx.x + G - 1) / G : 0) * LO_WARPS
                           : (A.contiguous ? min(f_lo + per, F) - f_lo : ((F > (int)blockIdx.x) ? (F - (int)blockIdx.x + G - 1) / G : 0));
  int iters_done = 0;
  const bool gate = st->cornerLastNum > 10 && st->surfLastNum > 100;  // :465
  if (gate) {
    for (int iterCount = 0; iterCount < A.max_iter; iterCount++) {
      iters_done = iterCount + 1;
      const bool refresh = (iterCount % 5 == 0);
      if (threadIdx.x == 0) {
        s_trig[0] = dm::sin(s_tr[0]); s_trig[1] = dm::cos(s_tr[0]);
        s_trig[2] = dm::sin(s_tr[1]); s_trig[3] = dm::cos(s_tr[1]);
        s_trig[4] = dm::sin(s_tr[2]); s_trig[5] = dm::cos(s_tr[2]);
      }
      __syncthreads();
      double acc = 0.0;
      long long dbg_t0 = 0;
      if (iterCount == 0 && PHASE_ON && threadIdx.x == 0) dbg_t0 = clock64();
      // ---- correspondence search, balanced: a search costs 2 .. 10 us depending on where the feature lies, and with
      // fixed slices the grid barrier of every fifth iteration waited 13 us for the slowest CTA.  Here the warps of
      // the whole grid claim features one by one; the correspondences and this iteration's coefficients go to global
      // memory (where they are kept anyway) and, behind one more grid barrier, the fixed slices below only rebuild
      // the Jacobian rows -- so every sum is still taken over the same rows in the same order. ----
      const bool searched = refresh && A.dynsearch != 0;
      if (searched) {
        int* ctr = &st->claim[iterCount / 5];
        int nxt = 0;
        if (lane == 0) nxt = atomicAdd(ctr, 1);
        nxt = __shfl_sync(0xffffffffu, nxt, 0);
        while (nxt < F) {
          const int f = nxt;
          if (lane == 0) nxt = atomicAdd(ctr, 1);   // the next claim travels while this feature is searched
          const bool is_corner = f < n_sharp;
          const float4 pt = is_corner ? A.sharp[f] : A.flat[f - n_sharp];
          const float4 sel = lo_transform_to_start(pt, s_tr);
          bool valid = false;
          float4 coeff = make_float4(0, 0, 0, 0);
          if (is_corner) {
            int i1 = -1, i2 = -1;
            lo_corner_search(A, s_roff[0], R, sel, n_sharp, n_corner_last, i1, i2);
            if (lane == 0) {
              A.corner_ind1[f] = i1; A.corner_ind2[f] = i2;
              if (i2 >= 0) valid = lo_corner_residual(sel, A.corner_last[i1], A.corner_last[i2], iterCount, coeff);
            }
          } else {
            const int g = f - n_sharp;
            int i1 = -1, i2 = -1, i3 = -1;
            lo_surf_search(A, s_roff[1], R, sel, n_flat, n_surf_last, i1, i2, i3);
            if (lane == 0) {
              A.surf_ind1[g] = i1; A.surf_ind2[g] = i2; A.surf_ind3[g] = i3;
              if (i2 >= 0 && i3 >= 0) valid = lo_surf_residual(sel, A.surf_last[i1], A.surf_last[i2], A.surf_last[i3], iterCount, coeff);
            }
          }
          if (lane == 0) {
            A.coeff[(size_t)iterCount * A.fcap + f] = coeff;
            A.row_valid[(size_t)iterCount * A.fcap + f] = valid ? 1 : 0;
          }
          nxt = __shfl_sync(0xffffffffu, nxt, 0);
        }
        __threadfence();
        grid.sync();
      }
      // ---- paired form (the shipped configuration: contiguous slices + claimed searches): TWO features per warp,
      // lanes 0-15 / 16-31.  The serial part of an iteration -- the residual on one lane -- runs for both at once and a
      // CTA's 16 features need one pass instead of two; the rows land in the same (feature, slab) slots and are added
      // in the same order as by two one-feature passes. ----
      const bool paired = A.dynsearch != 0 && A.contiguous == 1;
      if (paired) {
        const int half = lane >> 4, hl = lane & 15;
        const int blk = half * LO_WARPS + warp;          // feature block inside the pass
        for (int j0 = 0; j0 < nmine; j0 += 2 * LO_WARPS) {
          const int fi = j0 + blk;
          const bool act = fi < nmine;
          const int f = f_lo + fi;
          const bool first_pass = (j0 == 0);
          s_row[blk * 32 + hl][7] = 0.f;
          s_row[blk * 32 + 16 + hl][7] = 0.f;
          const bool is_corner = f < n_sharp;
          float4 pt = make_float4(0, 0, 0, 0);
          if (act) pt = is_corner ? A.sharp[f] : A.flat[f - n_sharp];
          if (act && hl == 0) {
            const int g = f - n_sharp;
            int i1 = -1, i2 = -1, i3 = -1;
            float4 t1 = pt, t2 = pt, t3 = pt;
            bool ok = false;
            const bool cached = first_pass && !refresh;
            if (cached) {
              ok = s_triok[blk] != 0;
              t1 = s_tri[blk][0]; t2 = s_tri[blk][1]; t3 = s_tri[blk][2];
            } else if (!refresh || first_pass) {   // after a search only the first pass needs the tripod (to park it)
              if (is_corner) { i1 = __ldcg(&A.corner_ind1[f]); i2 = __ldcg(&A.corner_ind2[f]); ok = i2 >= 0; }
              else { i1 = __ldcg(&A.surf_ind1[g]); i2 = __ldcg(&A.surf_ind2[g]); i3 = __ldcg(&A.surf_ind3[g]); ok = i2 >= 0 && i3 >= 0; }
              if (ok) {
                if (is_corner) { t1 = A.corner_last[i1]; t2 = A.corner_last[i2]; }
                else { t1 = A.surf_last[i1]; t2 = A.surf_last[i2]; t3 = A.surf_last[i3]; }
              }
              if (refresh) { s_triok[blk] = ok ? 1 : 0; if (ok) { s_tri[blk][0] = t1; s_tri[blk][1] = t2; s_tri[blk][2] = t3; } }
            }
            if (!refresh) {
              bool valid = false;
              float4 coeff = make_float4(0, 0, 0, 0);
              if (ok) {
                const float4 sel = lo_transform_to_start(pt, s_tr);
                valid = is_corner ? lo_corner_residual(sel, t1, t2, iterCount, coeff) : lo_surf_residual(sel, t1, t2, t3, iterCount, coeff);
              }
              A.coeff[(size_t)iterCount * A.fcap + f] = coeff;
              A.row_valid[(size_t)iterCount * A.fcap + f] = valid ? 1 : 0;
              s_cur[blk] = coeff;
              s_curok[blk] = valid ? 1 : 0;
            }
          }
          __syncwarp();
          // Jacobian of every row accumulated so far: lane hl takes the slabs hl and hl + 16 of its half's feature
          if (act) {
            for (int sl = hl; sl <= iterCount; sl += 16) {
              const size_t slot = (size_t)sl * A.fcap + f;
              bool v;
              float4 cf;
              if (sl == iterCount) {
                if (refresh) { v = __ldcg(&A.row_valid[slot]) != 0; cf = __ldcg(&A.coeff[slot]); }   // written by the claiming warp, maybe on another SM
                else { v = s_curok[blk] != 0; cf = s_cur[blk]; }
              } else {
                v = A.row_valid[slot] != 0; cf = A.coeff[slot];
              }
              if (v) {
                float a[6], b;
                lo_row(pt, cf, s_trig, s_tr, a, b);
                float* row = s_row[blk * 32 + sl];
#pragma unroll
                for (int j = 0; j < 6; j++) row[j] = a[j];
                row[6] = b;
                row[7] = 1.f;
              }
            }
          }
          __syncthreads();
          if (first_pass) PS();  // 0: features + rows of the first pass done
          if (first_pass && iterCount == 0 && PHASE_ON && threadIdx.x == 0) g_dbg_clk[64 + blockIdx.x] = clock64() - dbg_t0;
          red_rows_paired(s_row, iterCount + 1, acc);
          __syncthreads();
        }
      } else
      for (int j0 = 0; j0 < nmine; j0 += LO_WARPS) {
        const int f = cyclic ? ((j0 / LO_WARPS) * G + (int)blockIdx.x) * LO_WARPS + warp : f_lo + (j0 + warp) * f_step;
        const bool first_pass = (j0 == 0);
        s_row[warp * 32 + lane][7] = 0.f;
        if (cyclic ? f < F : j0 + warp < nmine) {
          // ---- transform, (search), residual -> row slab of this iteration ----
          const bool is_corner = f < n_sharp;
          const float4 pt = is_corner ? A.sharp[f] : A.flat[f - n_sharp];
          if (refresh && PHASE_ON && threadIdx.x == 0) g_dbg_clk[256 + blockIdx.x] = clock64();
          const float4 sel = lo_transform_to_start(pt, s_tr);
          if (refresh && PHASE_ON && threadIdx.x == 0) g_dbg_clk[832 + blockIdx.x] = clock64();
          bool valid = false;
          float4 coeff = make_float4(0, 0, 0, 0);
          if (is_corner) {
            int i1 = -1, i2 = -1;
            float4 t1, t2;
            bool have = false;
            if (searched) {
              if (first_pass) { i1 = __ldcg(&A.corner_ind1[f]); i2 = __ldcg(&A.corner_ind2[f]); }   // only to park the tripod in s_tri
            } else if (refresh) {
              lo_corner_search(A, s_roff[0], R, sel, n_sharp, n_corner_last, i1, i2);
              if (lane == 0) { A.corner_ind1[f] = i1; A.corner_ind2[f] = i2; }
            } else if (first_pass) {
              have = true;
              i2 = s_triok[warp] ? 0 : -1;
              t1 = s_tri[warp][0]; t2 = s_tri[warp][1];
            } else {
              i1 = __ldcg(&A.corner_ind1[f]); i2 = __ldcg(&A.corner_ind2[f]);   // L2: another SM may have written them (dynamic search)
            }
            if (lane == 0) {
              if (!have && i2 >= 0) { t1 = A.corner_last[i1]; t2 = A.corner_last[i2]; }
              if (refresh && first_pass) { s_triok[warp] = i2 >= 0 ? 1 : 0; if (i2 >= 0) { s_tri[warp][0] = t1; s_tri[warp][1] = t2; } }
              if (i2 >= 0 && !searched) valid = lo_corner_residual(sel, t1, t2, iterCount, coeff);
            }
          } else {
            int g = f - n_sharp;
            int i1 = -1, i2 = -1, i3 = -1;
            float4 t1, t2, t3;
            bool have = false;
            if (searched) {
              if (first_pass) { i1 = __ldcg(&A.surf_ind1[g]); i2 = __ldcg(&A.surf_ind2[g]); i3 = __ldcg(&A.surf_ind3[g]); }
            } else if (refresh) {
              lo_surf_search(A, s_roff[1], R, sel, n_flat, n_surf_last, i1, i2, i3);
              if (lane == 0) { A.surf_ind1[g] = i1; A.surf_ind2[g] = i2; A.surf_ind3[g] = i3; }
            } else if (first_pass) {
              have = true;
              i2 = i3 = s_triok[warp] ? 0 : -1;
              t1 = s_tri[warp][0]; t2 = s_tri[warp][1]; t3 = s_tri[warp][2];
            } else {
              i1 = __ldcg(&A.surf_ind1[g]); i2 = __ldcg(&A.surf_ind2[g]); i3 = __ldcg(&A.surf_ind3[g]);
            }
            if (lane == 0) {
              const bool ok = i2 >= 0 && i3 >= 0;
              if (!have && ok) { t1 = A.surf_last[i1]; t2 = A.surf_last[i2]; t3 = A.surf_last[i3]; }
              if (refresh && first_pass) { s_triok[warp] = ok ? 1 : 0; if (ok) { s_tri[warp][0] = t1; s_tri[warp][1] = t2; s_tri[warp][2] = t3; } }
              if (ok && !searched) valid = lo_surf_residual(sel, t1, t2, t3, iterCount, coeff);
            }
          }
          if (refresh && PHASE_ON && threadIdx.x == 0) g_dbg_clk[640 + blockIdx.x] = clock64();
          if (lane == 0 && !searched) {
            A.coeff[(size_t)iterCount * A.fcap + f] = coeff;
            A.row_valid[(size_t)iterCount * A.fcap + f] = valid ? 1 : 0;
          }
          __syncwarp();
          // ---- Jacobian of every row accumulated so far (rows are NOT cleared per iteration):
          // lane k re-evaluates the row this feature produced in iteration k ----
          if (lane <= iterCount) {
            size_t slot = (size_t)lane * A.fcap + f;
            // a slab is written once, in its own iteration, and fcap is a multiple of the L1 line: no SM can hold a
            // stale sector of it (the newest slab may come from another SM: fetched from L2 to make that explicit)
            const bool fresh = searched && lane == iterCount;
            if (fresh ? __ldcg(&A.row_valid[slot]) : A.row_valid[slot]) {
              float a[6], b;
              lo_row(pt, fresh ? __ldcg(&A.coeff[slot]) : A.coeff[slot], s_trig, s_tr, a, b);
              float* row = s_row[warp * 32 + lane];
#pragma unroll
              for (int j = 0; j < 6; j++) row[j] = a[j];
              row[6] = b;
              row[7] = 1.f;
            }
          }
        }
        __syncthreads();
        if (first_pass) PS();  // 0: features + rows of the first pass done
        if (first_pass && iterCount == 0 && PHASE_ON && threadIdx.x == 0) g_dbg_clk[64 + blockIdx.x] = clock64() - dbg_t0;
        red_rows_warped(s_row, LO_WARPS, iterCount + 1, acc);
        __syncthreads();
      }
      double* part = A.partials + (size_t)(iterCount & 1) * gridDim.x * LO_NRED;
      red_finish(grid, acc, part, s_g, s_tot, 0, clk);   // stamps 1 (before barrier), 2 (after)
      PS();  // 3: partials summed
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
      PS();  // 4: solve done
      if (s_stop) break;
    }
  }
  // every CTA is behind the last iteration's grid barrier: nobody claims any more
  if (gate && A.dynsearch && blockIdx.x == 0 && threadIdx.x < 8) st->claim[threadIdx.x] = 0;
  // without an iteration no grid barrier has been passed yet: the state written below (and by the epilogue)
  // is still being read by CTAs that are only now starting
  if (!gate) grid.sync();
  if (blockIdx.x == 0 && warp == 0) {
    __shared__ LOIntegrateSmem s_int;
    lo_integrate_warp(st, s_tr, s_imu, &s_int);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (int i = 0; i < 6; i++) st->transform[i] = s_tr[i];
    for (int i = 0; i < 36; i++) st->matP[i] = s_matP[i];
    st->isDegenerate = s_deg;
    st->iters = iters_done;
    st->rows = s_rows;
  }
  PS();
  if (!A.epilogue) {
    if (clk >= 0) g_phase_clk[0][0] = clk - 1;
    return;
  }
  // ---- epilogue: the node's tail on the grid that is resident anyway.  Every CTA left the loop behind the
  // same grid barrier, so nobody reads corner_last / surf_last / the two indices any more. ----
  {
    __shared__ float s_mn[LO_WARPS][3], s_mx[LO_WARPS][3];
    __shared__ ToEndConsts s_k;
    lo_to_end_consts(&s_k, threadIdx.x, s_tr, s_imu);
    __syncthreads();
    const ToEndConsts kc = s_k;
    const int tid = blockIdx.x * LO_THREADS + threadIdx.x, nthr = gridDim.x * LO_THREADS;
    const int n0 = A.cnt->n_less_sharp, n1 = A.cnt->n_less_flat, n2 = A.do_full ? A.cnt->n_full : 0;
    // TransformToEnd + swap (:875-899); the bounding boxes of the two new "last" clouds come for free
    {
      float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
      for (int i = tid; i < n0; i += nthr) {
        const float4 p = lo_transform_to_end_pre(A.less_sharp[i], s_tr, s_imu, kc);
        A.corner_last_out[i] = p;
        mn[0] = fminf(mn[0], p.x); mx[0] = fmaxf(mx[0], p.x);
        mn[1] = fminf(mn[1], p.y); mx[1] = fmaxf(mx[1], p.y);
        mn[2] = fminf(mn[2], p.z); mx[2] = fmaxf(mx[2], p.z);
      }
      bbox_block_commit(A.gb[0].bbox, mn, mx, s_mn, s_mx, LO_WARPS);
    }
    {
      float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
      for (int i = tid; i < n1; i += nthr) {
        const float4 p = lo_transform_to_end_pre(A.less_flat[i], s_tr, s_imu, kc);
        A.surf_last_out[i] = p;
        mn[0] = fminf(mn[0], p.x); mx[0] = fmaxf(mx[0], p.x);
        mn[1] = fminf(mn[1], p.y); mx[1] = fmaxf(mx[1], p.y);
        mn[2] = fminf(mn[2], p.z); mx[2] = fmaxf(mx[2], p.z);
      }
      bbox_block_commit(A.gb[1].bbox, mn, mx, s_mn, s_mx, LO_WARPS);
    }
    for (int i = tid; i < n2; i += nthr) A.full_end[i] = lo_transform_to_end_pre(A.full[i], s_tr, s_imu, kc);
    // bookkeeping (:901-903) and the ring offset tables of the clouds that just became the "last" ones
    if (blockIdx.x == 0 && warp == 0) {
      for (int T = 0; T < 2; T++) {
        int* roff = T ? A.surf_roff_out : A.corner_roff_out;
        const int col = T ? 3 : 1;
        int run = 0;
        for (int r0 = 0; r0 < R; r0 += 32) {
          const int r = r0 + lane;
          const int c = r < R ? A.ring_counts[r * 4 + col] : 0;
          int x = c;
          for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
          }
          if (r < R) roff[r] = run + x - c;
          run += __shfl_sync(0xffffffffu, x, 31);
        }
        if (lane == 0) roff[R] = run;
      }
      if (lane == 0) {
        st->n_corner_last = n0; st->n_surf_last = n1;
        st->cornerLastNum = n0; st->surfLastNum = n1;
        st->index_enable = (n0 > 10 && n1 > 100) ? 1 : 0;
        if (A.do_full) st->n_full_end = n2;
      }
    }
    PS();
    // index rebuild (:905-906: kdtree*Last->setInputCloud), skipped as in the reference's gate at :903
    if (n0 > 10 && n1 > 100) {
      __threadfence();
      grid.sync();
      const GridDesc d0 = gridp_setup(A.gb[0], n0, tid, nthr);
      const GridDesc d1 = gridp_setup(A.gb[1], n1, tid, nthr);
      __threadfence();
      grid.sync();
      gridp_count(A.gb[0], d0, tid, nthr);
      gridp_count(A.gb[1], d1, tid, nthr);
      __threadfence();
      grid.sync();
      const int wid = blockIdx.x * LO_WARPS + warp, nw = gridDim.x * LO_WARPS;
      gridp_alloc(A.gb[0], d0, wid, nw);
      gridp_alloc(A.gb[1], d1, wid, nw);
      __threadfence();
      grid.sync();
      gridp_scatter(A.gb[0], d0, tid, nthr);
      gridp_scatter(A.gb[1], d1, tid, nthr);
    } else if (blockIdx.x == 0 && threadIdx.x < 12) {
      // nothing was built: re-arm the two boxes for the next sweep
      A.gb[threadIdx.x / 6].bbox[threadIdx.x % 6] = (threadIdx.x % 6) < 3 ? 0xffffffffu : 0u;
    }
    PS();
  }
  if (clk >= 0) g_phase_clk[0][0] = clk - 1;
#undef PS
}

// what the pipeline reports per sweep (loam_pipeline.cuh), written straight into the pinned host ring
struct PipeRecord {
  SRCounts cnt;
  float imu_trans[12];
  float transform[6], transformSum[6];
  int lo_iters, lo_rows, lo_deg;
  float aft[6], bef[6];
  int lm_iters, lm_rows, lm_deg, lm_error;
  int mapN[2], stackN[2], bump[2];
  int surroundN, pad;
};
// odometry part of a sweep's record (one thread)
__device__ __forceinline__ void pipe_record_body(PipeRecord* out, const SRCounts* cnt, const ImuState* imu, const LOState* lo) {
  PipeRecord r;
  r.cnt = *cnt;
  lo_load_imu(imu, r.imu_trans);
  for (int i = 0; i < 6; i++) { r.transform[i] = lo->transform[i]; r.transformSum[i] = lo->transformSum[i]; }
  r.lo_iters = lo->iters; r.lo_rows = lo->rows; r.lo_deg = lo->isDegenerate;
  for (int i = 0; i < 6; i++) { r.aft[i] = 0.f; r.bef[i] = 0.f; }
  r.lm_iters = 0; r.lm_rows = 0; r.lm_deg = 0; r.lm_error = 0;
  for (int T = 0; T < 2; T++) { r.mapN[T] = 0; r.stackN[T] = 0; r.bump[T] = 0; }
  r.surroundN = -1; r.pad = 0;   // -1: no surround map on this sweep (pipe_record_surround fills it in)
  *out = r;
}

// bookkeeping after the swap (:901-903), one warp; mode 1 = first sweep (:427-456: plain adoption of the
// clouds, transformSum seeded from the IMU, the *LastNum counters are NOT touched)
__device__ __forceinline__ void lo_finish_body(LOState* st, const SRCounts* cnt, const ImuState* imu, int first_sweep, int do_full,
                                               const int* ring_counts, int n_scans, int* corner_roff, int* surf_roff) {
  // ring offset tables of the clouds that just became corner_last / surf_last (ring-major)
  const int lane = threadIdx.x & 31;
  for (int T = 0; T < 2; T++) {
    int* roff = T ? surf_roff : corner_roff;
    const int col = T ? 3 : 1;
    int run = 0;
    for (int r0 = 0; r0 < n_scans; r0 += 32) {
      const int r = r0 + lane;
      const int c = r < n_scans ? ring_counts[r * 4 + col] : 0;
      int x = c;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      if (r < n_scans) roff[r] = run + x - c;
      run += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) roff[n_scans] = run;
  }
  if (lane != 0) return;
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
__global__ void lo_finish(LOState* st, const SRCounts* cnt, const ImuState* imu, int first_sweep, int do_full,
                          const int* ring_counts, int n_scans, int* corner_roff, int* surf_roff) {
  if (blockIdx.x == 0 && threadIdx.x < 32) lo_finish_body(st, cnt, imu, first_sweep, do_full, ring_counts, n_scans, corner_roff, surf_roff);
}

// TransformToEnd over less-sharp -> corner_last, less-flat -> surf_last, (optionally) full ->
// full_end (:875-891); the buffer "swap" (:893-899) is the write into the *_last arrays.  The same launch carries
// the node's one-warp bookkeeping (CTA 0; nothing in this kernel reads what it writes) and, for a pipeline, the
// sweep's pose record (last CTA, one thread; it reads what lo_optimize left, not what the bookkeeping writes):
// two one-warp launches fewer on the odometry chain.
struct ToEndArgs {
  LOState* st;
  const SRCounts* cnt;
  const ImuState* imu;
  const float4* less_sharp; const float4* less_flat; const float4* full;
  float4* corner_last; float4* surf_last; float4* full_end;
  int do_full;
  const int* ring_counts; int n_scans; int* corner_roff; int* surf_roff;
  PipeRecord* record;        // nullptr: no record (stand-alone node)
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
  if (A.record && blockIdx.x == gridDim.x - 1 && threadIdx.x == 255) pipe_record_body(A.record, A.cnt, A.imu, A.st);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    if (i < n0) A.corner_last[i] = lo_transform_to_end(A.less_sharp[i], s_tr, s_imu);
    else if (i < n0 + n1) A.surf_last[i - n0] = lo_transform_to_end(A.less_flat[i - n0], s_tr, s_imu);
    else A.full_end[i - n0 - n1] = lo_transform_to_end(A.full[i - n0 - n1], s_tr, s_imu);
  }
  // the bookkeeping overwrites n_full_end / *_last counters that only LATER kernels read; st->transform, read above by
  // every CTA, is not among the fields it writes
  if (blockIdx.x == 0 && threadIdx.x < 32) lo_finish_body(A.st, A.cnt, A.imu, 0, A.do_full, A.ring_counts, A.n_scans, A.corner_roff, A.surf_roff);
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
