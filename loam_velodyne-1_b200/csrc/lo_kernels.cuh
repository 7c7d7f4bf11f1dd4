// lo_kernels.cuh -- laserOdometry on sm_100a (SURVEY.md §8a rows LO-1 .. LO-12).
//
// The reference's LM loop (laserOdometry.cpp:470-827) is <= 25 strictly sequential iterations of
// {transform every feature, (every 5th) kd-tree 1-NN + ring-window scans, residuals, Jacobian of
// ALL rows accumulated so far, 6x6 normal equations, QR solve}.  Launch latency would dominate,
// so the whole loop is ONE cooperative kernel: one warp per feature, one grid-wide barrier per
// iteration, and every block redundantly finishes the (deterministic) 27-term reduction and the
// 6x6 solve so no second barrier or broadcast is needed.
// The scalar expressions restated in this file (cited below) come from loam_velodyne:
//   Copyright 2013, Ji Zhang, Carnegie Mellon University
//   Further contributions copyright (c) 2016, Southwest Research Institute.  All rights reserved.
// BSD 3-clause; the full notice, conditions and disclaimer are retained in NOTICE.md at the repository root.
#pragma once
#include <cooperative_groups.h>
#include "grid_index.cuh"
#include "linalg.cuh"
#include "pose_math.cuh"
#include "sr_kernels.cuh"

namespace loamgpu {
namespace cg = cooperative_groups;

constexpr int LO_THREADS = 256;   // half an SM's registers: lm_optimize (or other streams' kernels) runs beside it
constexpr int LO_WARPS = LO_THREADS / 32;
constexpr int LO_NRED = 28;  // 21 (AtA upper) + 6 (AtB) + 1 (row count)

struct LOState {
  float transform[6];      // laserOdometry.cpp:93
  float transformSum[6];   // :94
  int isDegenerate;        // :404
  float matP[36];          // :405
  int cornerLastNum;       // :81 (zero until first set at :901)
  int surfLastNum;         // :82
  int iters, rows;         // stats of the last sweep
  int index_enable;        // cornerLastNum > 10 && surfLastNum > 100 (:903)
  int n_corner_last, n_surf_last, n_full_end;  // sizes of the device clouds
  int sum_decoded;         // laserMapping input only: transformSum was already decoded from the /laser_odom_to_init
                           // message by the caller (a ROS node's laserOdometryHandler, laserMapping.cpp:304-321)
  int claim[8];            // lo_optimize: next unclaimed feature of the k-th correspondence search (zero between launches)
};

struct LOArgs {
  LOState* st;
  const SRCounts* cnt;          // n_sharp, n_flat of this sweep
  const float4* sharp;
  const float4* flat;
  const float4* corner_last;    // original order (indices of the reference)
  const float4* surf_last;
  GridView corner_grid, surf_grid;
  int* corner_ind1; int* corner_ind2;           // pointSearchCornerInd1/2 (:85-86)
  int* surf_ind1; int* surf_ind2; int* surf_ind3;  // :89-91
  float4* coeff;                // [max_iter][fcap] rows of every iteration (coeffSel accumulates, :458-459)
  unsigned char* row_valid;     // [max_iter][fcap]
  double* partials;             // [2][gridDim][LO_NRED]
  const ImuState* imu;          // imuTrans of this sweep
  const int* corner_roff;       // [n_scans + 1] first index of every ring inside corner_last
  const int* surf_roff;         // same for surf_last
  int n_scans;
  int fcap;
  int max_iter;
  int contiguous;               // 1 = contiguous feature slices per CTA (the round-1 assignment), 0 = strided
  int dynsearch;                // 1 = the correspondence searches (iterCount % 5 == 0) are claimed feature by feature
  // epilogue of the same kernel: TransformToEnd + swap (:875-899), bookkeeping (:901-903) and the rebuild of
  // the two indices (:905-906) -- the grid is resident anyway, so the node's tail costs four more grid
  // barriers instead of seven launches
  const float4* less_sharp; const float4* less_flat; const float4* full;
  float4* corner_last_out; float4* surf_last_out; float4* full_end;
  int do_full;
  int epilogue;                 // 0 = stop after the pose integration (the tail then runs as separate launches; A/B switch)
  GridBuf gb[2];
  const int* ring_counts;       // [n_scans][4] per-ring sizes of the four feature clouds of this sweep
  int* corner_roff_out; int* surf_roff_out;
};

// TransformToStart :101-124
__device__ __forceinline__ float4 lo_transform_to_start(const float4 pi, const float* transform) {
  float s = 10 * (pi.w - int(pi.w));
  float rx = s * transform[0];
  float ry = s * transform[1];
  float rz = s * transform[2];
  float tx = s * transform[3];
  float ty = s * transform[4];
  float tz = s * transform[5];
  float x1 = dm::cos(rz) * (pi.x - tx) + dm::sin(rz) * (pi.y - ty);
  float y1 = -dm::sin(rz) * (pi.x - tx) + dm::cos(rz) * (pi.y - ty);
  float z1 = (pi.z - tz);
  float x2 = x1;
  float y2 = dm::cos(rx) * y1 + dm::sin(rx) * z1;
  float z2 = -dm::sin(rx) * y1 + dm::cos(rx) * z1;
  float4 po;
  po.x = dm::cos(ry) * x2 - dm::sin(ry) * z2;
  po.y = y2;
  po.z = dm::sin(ry) * x2 + dm::cos(ry) * z2;
  po.w = pi.w;
  return po;
}

// TransformToEnd :126-194
__device__ __forceinline__ float4 lo_transform_to_end(const float4 pi, const float* transform, const float* imu /*12*/) {
  float s = 10 * (pi.w - int(pi.w));
  float rx = s * transform[0];
  float ry = s * transform[1];
  float rz = s * transform[2];
  float tx = s * transform[3];
  float ty = s * transform[4];
  float tz = s * transform[5];
  float x1 = dm::cos(rz) * (pi.x - tx) + dm::sin(rz) * (pi.y - ty);
  float y1 = -dm::sin(rz) * (pi.x - tx) + dm::cos(rz) * (pi.y - ty);
  float z1 = (pi.z - tz);
  float x2 = x1;
  float y2 = dm::cos(rx) * y1 + dm::sin(rx) * z1;
  float z2 = -dm::sin(rx) * y1 + dm::cos(rx) * z1;
  float x3 = dm::cos(ry) * x2 - dm::sin(ry) * z2;
  float y3 = y2;
  float z3 = dm::sin(ry) * x2 + dm::cos(ry) * z2;
  rx = transform[0];
  ry = transform[1];
  rz = transform[2];
  tx = transform[3];
  ty = transform[4];
  tz = transform[5];
  float x4 = dm::cos(ry) * x3 + dm::sin(ry) * z3;
  float y4 = y3;
  float z4 = -dm::sin(ry) * x3 + dm::cos(ry) * z3;
  float x5 = x4;
  float y5 = dm::cos(rx) * y4 - dm::sin(rx) * z4;
  float z5 = dm::sin(rx) * y4 + dm::cos(rx) * z4;
  float x6 = dm::cos(rz) * x5 - dm::sin(rz) * y5 + tx;
  float y6 = dm::sin(rz) * x5 + dm::cos(rz) * y5 + ty;
  float z6 = z5 + tz;
  const float imuPitchStart = imu[0], imuYawStart = imu[1], imuRollStart = imu[2];
  const float imuPitchLast = imu[3], imuYawLast = imu[4], imuRollLast = imu[5];
  const float imuShiftFromStartX = imu[6], imuShiftFromStartY = imu[7], imuShiftFromStartZ = imu[8];
  float x7 = dm::cos(imuRollStart) * (x6 - imuShiftFromStartX) - dm::sin(imuRollStart) * (y6 - imuShiftFromStartY);
  float y7 = dm::sin(imuRollStart) * (x6 - imuShiftFromStartX) + dm::cos(imuRollStart) * (y6 - imuShiftFromStartY);
  float z7 = z6 - imuShiftFromStartZ;
  float x8 = x7;
  float y8 = dm::cos(imuPitchStart) * y7 - dm::sin(imuPitchStart) * z7;
  float z8 = dm::sin(imuPitchStart) * y7 + dm::cos(imuPitchStart) * z7;
  float x9 = dm::cos(imuYawStart) * x8 + dm::sin(imuYawStart) * z8;
  float y9 = y8;
  float z9 = -dm::sin(imuYawStart) * x8 + dm::cos(imuYawStart) * z8;
  float x10 = dm::cos(imuYawLast) * x9 - dm::sin(imuYawLast) * z9;
  float y10 = y9;
  float z10 = dm::sin(imuYawLast) * x9 + dm::cos(imuYawLast) * z9;
  float x11 = x10;
  float y11 = dm::cos(imuPitchLast) * y10 + dm::sin(imuPitchLast) * z10;
  float z11 = -dm::sin(imuPitchLast) * y10 + dm::cos(imuPitchLast) * z10;
  float4 po;
  po.x = dm::cos(imuRollLast) * x11 + dm::sin(imuRollLast) * y11;
  po.y = -dm::sin(imuRollLast) * x11 + dm::cos(imuRollLast) * y11;
  po.z = z11;
  po.w = int(pi.w);
  return po;
}

struct ToEndConsts { float srx, crx, sry, cry, srz, crz; float s[6], c[6]; };
// angle j of lo_to_end_consts: 0..2 = transform rx, ry, rz; 3..8 = imuRollStart, imuPitchStart, imuYawStart, imuYawLast, imuPitchLast, imuRollLast
__device__ __forceinline__ float lo_to_end_angle(int j, const float* transform, const float* imu) {
  switch (j) {
    case 0: return transform[0];
    case 1: return transform[1];
    case 2: return transform[2];
    case 3: return imu[2];
    case 4: return imu[0];
    case 5: return imu[1];
    case 6: return imu[4];
    case 7: return imu[3];
    default: return imu[5];
  }
}
// 18 threads t = 0..17 of a CTA fill the shared ToEndConsts: t < 9 the sine of angle t, else the cosine of angle t - 9
__device__ __forceinline__ void lo_to_end_consts(ToEndConsts* k, int t, const float* transform, const float* imu) {
  if (t >= 18) return;
  const int j = t % 9;
  const float a = lo_to_end_angle(j, transform, imu);
  const float v = t < 9 ? dm::sin(a) : dm::cos(a);
  float* sv[9] = {&k->srx, &k->sry, &k->srz, &k->s[0], &k->s[1], &k->s[2], &k->s[3], &k->s[4], &k->s[5]};
  float* cv[9] = {&k->crx, &k->cry, &k->crz, &k->c[0], &k->c[1], &k->c[2], &k->c[3], &k->c[4], &k->c[5]};
  *(t < 9 ? sv[j] : cv[j]) = v;
}
// Same expression tree; the sines / cosines of the nine angles that do not depend on the point (the whole-sweep
// transform and the six IMU angles) are taken from `k` (lo_to_end_consts), where they were computed by the
// same functions on the same arguments -- identical bits, 18 fdlibm evaluations per point fewer.
__device__ __forceinline__ float4 lo_transform_to_end_pre(const float4 pi, const float* transform, const float* imu /*12*/, const ToEndConsts& k) {
  float s = 10 * (pi.w - int(pi.w));
  float rx = s * transform[0];
  float ry = s * transform[1];
  float rz = s * transform[2];
  float tx = s * transform[3];
  float ty = s * transform[4];
  float tz = s * transform[5];
  float x1 = dm::cos(rz) * (pi.x - tx) + dm::sin(rz) * (pi.y - ty);
  float y1 = -dm::sin(rz) * (pi.x - tx) + dm::cos(rz) * (pi.y - ty);
  float z1 = (pi.z - tz);
  float x2 = x1;
  float y2 = dm::cos(rx) * y1 + dm::sin(rx) * z1;
  float z2 = -dm::sin(rx) * y1 + dm::cos(rx) * z1;
  float x3 = dm::cos(ry) * x2 - dm::sin(ry) * z2;
  float y3 = y2;
  float z3 = dm::sin(ry) * x2 + dm::cos(ry) * z2;
  rx = transform[0];
  ry = transform[1];
  rz = transform[2];
  tx = transform[3];
  ty = transform[4];
  tz = transform[5];
  float x4 = k.cry * x3 + k.sry * z3;
  float y4 = y3;
  float z4 = -k.sry * x3 + k.cry * z3;
  float x5 = x4;
  float y5 = k.crx * y4 - k.srx * z4;
  float z5 = k.srx * y4 + k.crx * z4;
  float x6 = k.crz * x5 - k.srz * y5 + tx;
  float y6 = k.srz * x5 + k.crz * y5 + ty;
  float z6 = z5 + tz;
  const float imuPitchStart = imu[0], imuYawStart = imu[1], imuRollStart = imu[2];
  const float imuPitchLast = imu[3], imuYawLast = imu[4], imuRollLast = imu[5];
  const float imuShiftFromStartX = imu[6], imuShiftFromStartY = imu[7], imuShiftFromStartZ = imu[8];
  float x7 = k.c[0] * (x6 - imuShiftFromStartX) - k.s[0] * (y6 - imuShiftFromStartY);
  float y7 = k.s[0] * (x6 - imuShiftFromStartX) + k.c[0] * (y6 - imuShiftFromStartY);
  float z7 = z6 - imuShiftFromStartZ;
  float x8 = x7;
  float y8 = k.c[1] * y7 - k.s[1] * z7;
  float z8 = k.s[1] * y7 + k.c[1] * z7;
  float x9 = k.c[2] * x8 + k.s[2] * z8;
  float y9 = y8;
  float z9 = -k.s[2] * x8 + k.c[2] * z8;
  float x10 = k.c[3] * x9 - k.s[3] * z9;
  float y10 = y9;
  float z10 = k.s[3] * x9 + k.c[3] * z9;
  float x11 = x10;
  float y11 = k.c[4] * y10 + k.s[4] * z10;
  float z11 = -k.s[4] * y10 + k.c[4] * z10;
  float4 po;
  po.x = k.c[5] * x11 + k.s[5] * y11;
  po.y = -k.s[5] * x11 + k.c[5] * y11;
  po.z = z11;
  po.w = int(pi.w);
  return po;
}

// imuTrans as laserOdometry receives it (:337-351) from scanRegistration's state (:614-629)
__device__ __forceinline__ void lo_load_imu(const ImuState* s, float* imu /*12*/) {
  imu[0] = s->rpy_start[1]; imu[1] = s->rpy_start[2]; imu[2] = s->rpy_start[0];
  imu[3] = s->rpy_cur[1]; imu[4] = s->rpy_cur[2]; imu[5] = s->rpy_cur[0];
  for (int a = 0; a < 3; a++) { imu[6 + a] = s->shift_from_start[a]; imu[9 + a] = s->velo_from_start[a]; }
}

__device__ __forceinline__ float lo_sqdist(const float4 pj, const float4 sel) {
  return (pj.x - sel.x) * (pj.x - sel.x) + (pj.y - sel.y) * (pj.y - sel.y) + (pj.z - sel.z) * (pj.z - sel.z);
}

// warp-wide arg-min over (d, ord): the sequential loops update on strict '<', i.e. the earliest
// evaluated candidate wins ties; `ord` is the evaluation order.
__device__ __forceinline__ void warp_argmin(float& d, int& ord, int& idx) {
  for (int o = 16; o > 0; o >>= 1) {
    float od = __shfl_xor_sync(0xffffffffu, d, o);
    int oo = __shfl_xor_sync(0xffffffffu, ord, o);
    int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (od < d || (od == d && oo < ord)) { d = od; ord = oo; idx = oi; }
  }
}

// ring of position i in a ring-major cloud: largest r with roff[r] <= i (roff has R+1 entries).
// The "last" clouds are ring-major by construction (scanRegistration emits ring -> sector -> pick
// order / ring -> voxel order) and int(intensity) of every point equals its ring, so this is the
// same value as the reference's int(laserCloud*Last->points[i].intensity) without a dependent load.
__device__ __forceinline__ int lo_ring_of(const int* roff, int R, int i) {
  int lo = 0, hi = R - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (roff[mid] <= i) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// corner correspondence :478-527 for one feature (warp cooperative).  The reference's two linear
// scans stop at the first point whose ring leaves (c-2.5, c+2.5); with the ring offset table the
// window is known up front, so every load below is independent (no load -> ballot -> load chain).
__device__ void lo_corner_search(const LOArgs& A, const int* roff, int R, const float4 sel, int n_sharp, int n_last, int& ind1, int& ind2) {
  const int lane = lane_id();
  int bi; float bd;
  grid_nn1_warp(A.corner_grid, sel.x, sel.y, sel.z, 25.f, bi, bd);
  if (PHASE_ON && threadIdx.x == 0) g_dbg_clk[448 + blockIdx.x] = clock64();
  int closestPointInd = -1, minPointInd2 = -1;
  if (bi >= 0 && bd < 25) {
    closestPointInd = bi;
    const int c = lo_ring_of(roff, R, bi);   // closestPointScan
    float best = 25.f; int bord = 0x7fffffff, bidx = -1;
    const int jend = min(n_sharp, n_last);  // forward bound is cornerPointsSharpNum (:486), kept
    // forward: j > closest, ring in (c, c+2.5]  <=>  roff[c+1] <= j < roff[c+3]
    {
      const int a0 = max(bi + 1, roff[min(c + 1, R)]), b0 = min(jend, roff[min(c + 3, R)]);
      for (int j0 = a0 + lane; j0 < b0; j0 += 128) {
        float4 pc[4];
#pragma unroll
        for (int u = 0; u < 4; u++) pc[u] = __ldg(&A.corner_last[min(j0 + 32 * u, b0 - 1)]);
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int j = j0 + 32 * u;
          if (j >= b0) break;
          float dd = lo_sqdist(pc[u], sel);
          int ord = j - (bi + 1);
          if (dd < best || (dd == best && ord < bord)) { best = dd; bord = ord; bidx = j; }
        }
      }
    }
    // backward: j < closest, ring in [c-2.5, c)  <=>  roff[c-2] <= j < roff[c]
    {
      const int a0 = roff[max(c - 2, 0)], b0 = min(bi, roff[c]);
      for (int j0 = b0 - 1 - lane; j0 >= a0; j0 -= 128) {
        float4 pc[4];
#pragma unroll
        for (int u = 0; u < 4; u++) pc[u] = __ldg(&A.corner_last[max(j0 - 32 * u, a0)]);
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int j = j0 - 32 * u;
          if (j < a0) break;
          float dd = lo_sqdist(pc[u], sel);
          int ord = 0x40000000 + (bi - 1 - j);
          if (dd < best || (dd == best && ord < bord)) { best = dd; bord = ord; bidx = j; }
        }
      }
    }
    warp_argmin(best, bord, bidx);
    if (best < 25.f) minPointInd2 = bidx;
  }
  ind1 = closestPointInd;
  ind2 = minPointInd2;
}

// surf correspondence :590-650
__device__ void lo_surf_search(const LOArgs& A, const int* roff, int R, const float4 sel, int n_flat, int n_last, int& ind1, int& ind2, int& ind3) {
  const int lane = lane_id();
  int bi; float bd;
  grid_nn1_warp(A.surf_grid, sel.x, sel.y, sel.z, 25.f, bi, bd);
  if (PHASE_ON && threadIdx.x == 0) g_dbg_clk[448 + blockIdx.x] = clock64();
  int closestPointInd = -1, minPointInd2 = -1, minPointInd3 = -1;
  if (bi >= 0 && bd < 25) {
    closestPointInd = bi;
    const int c = lo_ring_of(roff, R, bi);
    float best2 = 25.f, best3 = 25.f;
    int ord2 = 0x7fffffff, ord3 = 0x7fffffff, idx2 = -1, idx3 = -1;
    const int jend = min(n_flat, n_last);  // forward bound is surfPointsFlatNum (:598), kept
    // forward: ring <= c (same ring) feeds minPointInd2, ring in (c, c+2.5] feeds minPointInd3
    {
      const int a0 = bi + 1, b0 = min(jend, roff[min(c + 3, R)]), split = roff[min(c + 1, R)];
      for (int j0 = a0 + lane; j0 < b0; j0 += 256) {
        float4 pc[8];
#pragma unroll
        for (int u = 0; u < 8; u++) pc[u] = __ldg(&A.surf_last[min(j0 + 32 * u, b0 - 1)]);
#pragma unroll
        for (int u = 0; u < 8; u++) {
        const int j = j0 + 32 * u;
        if (j >= b0) break;
        float dd = lo_sqdist(pc[u], sel);
        int ord = j - (bi + 1);
        if (j < split) {
          if (dd < best2 || (dd == best2 && ord < ord2)) { best2 = dd; ord2 = ord; idx2 = j; }
        } else {
          if (dd < best3 || (dd == best3 && ord < ord3)) { best3 = dd; ord3 = ord; idx3 = j; }
        }
        }
      }
    }
    // backward: ring >= c (same ring) feeds minPointInd2, ring in [c-2.5, c) feeds minPointInd3
    {
      const int a0 = roff[max(c - 2, 0)], split = roff[c];
      for (int j0 = bi - 1 - lane; j0 >= a0; j0 -= 256) {
        float4 pc[8];
#pragma unroll
        for (int u = 0; u < 8; u++) pc[u] = __ldg(&A.surf_last[max(j0 - 32 * u, a0)]);
#pragma unroll
        for (int u = 0; u < 8; u++) {
        const int j = j0 - 32 * u;
        if (j < a0) break;
        float dd = lo_sqdist(pc[u], sel);
        int ord = 0x40000000 + (bi - 1 - j);
        if (j >= split) {
          if (dd < best2 || (dd == best2 && ord < ord2)) { best2 = dd; ord2 = ord; idx2 = j; }
        } else {
          if (dd < best3 || (dd == best3 && ord < ord3)) { best3 = dd; ord3 = ord; idx3 = j; }
        }
        }
      }
    }
    warp_argmin(best2, ord2, idx2);
    warp_argmin(best3, ord3, idx3);
    if (best2 < 25.f) minPointInd2 = idx2;
    if (best3 < 25.f) minPointInd3 = idx3;
  }
  ind1 = closestPointInd; ind2 = minPointInd2; ind3 = minPointInd3;
}

// point-to-line residual :530-582.  returns validity
__device__ __forceinline__ bool lo_corner_residual(const float4 pointSel, const float4 tripod1, const float4 tripod2, int iterCount, float4& coeff) {
  float x0 = pointSel.x; float y0 = pointSel.y; float z0 = pointSel.z;
  float x1 = tripod1.x; float y1 = tripod1.y; float z1 = tripod1.z;
  float x2 = tripod2.x; float y2 = tripod2.y; float z2 = tripod2.z;
  float a012 = dm::sqrt(((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
             * ((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
             + ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
             * ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
             + ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))
             * ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1)));
  float l12 = dm::sqrt((x1 - x2)*(x1 - x2) + (y1 - y2)*(y1 - y2) + (z1 - z2)*(z1 - z2));
  float la = ((y1 - y2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
           + (z1 - z2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))) / a012 / l12;
  float lb = -((x1 - x2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
           - (z1 - z2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
  float lc = -((x1 - x2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
           + (y1 - y2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
  float ld2 = a012 / l12;
  float s = 1;
  if (iterCount >= 5) {
    s = 1 - 1.8 * dm::fabs(ld2);
  }
  coeff.x = s * la;
  coeff.y = s * lb;
  coeff.z = s * lc;
  coeff.w = s * ld2;
  return (s > 0.1 && ld2 != 0);
}

// point-to-plane residual :653-693
__device__ __forceinline__ bool lo_surf_residual(const float4 pointSel, const float4 tripod1, const float4 tripod2, const float4 tripod3,
                                                 int iterCount, float4& coeff) {
  float pa = (tripod2.y - tripod1.y) * (tripod3.z - tripod1.z) - (tripod3.y - tripod1.y) * (tripod2.z - tripod1.z);
  float pb = (tripod2.z - tripod1.z) * (tripod3.x - tripod1.x) - (tripod3.z - tripod1.z) * (tripod2.x - tripod1.x);
  float pc = (tripod2.x - tripod1.x) * (tripod3.y - tripod1.y) - (tripod3.x - tripod1.x) * (tripod2.y - tripod1.y);
  float pd = -(pa * tripod1.x + pb * tripod1.y + pc * tripod1.z);
  float ps = dm::sqrt(pa * pa + pb * pb + pc * pc);
  pa /= ps;
  pb /= ps;
  pc /= ps;
  pd /= ps;
  float pd2 = pa * pointSel.x + pb * pointSel.y + pc * pointSel.z + pd;
  float s = 1;
  if (iterCount >= 5) {
    s = 1 - 1.8 * dm::fabs(pd2) / dm::sqrt(dm::sqrt(pointSel.x * pointSel.x + pointSel.y * pointSel.y + pointSel.z * pointSel.z));
  }
  coeff.x = s * pa;
  coeff.y = s * pb;
  coeff.z = s * pc;
  coeff.w = s * pd2;
  return (s > 0.1 && pd2 != 0);
}

// one Jacobian row :709-763 -> a[6], b
__device__ __forceinline__ void lo_row(const float4 pointOri, const float4 coeffv, const float* tr6 /*srx,crx,sry,cry,srz,crz*/,
                                       const float* transform, float* a, float& b) {
  float s = 1;
  float srx = tr6[0], crx = tr6[1], sry = tr6[2], cry = tr6[3], srz = tr6[4], crz = tr6[5];
  float tx = s * transform[3];
  float ty = s * transform[4];
  float tz = s * transform[5];
  struct { float x, y, z; } coeff = {coeffv.x, coeffv.y, coeffv.z};
  float arx = (-s*crx*sry*srz*pointOri.x + s*crx*crz*sry*pointOri.y + s*srx*sry*pointOri.z
            + s*tx*crx*sry*srz - s*ty*crx*crz*sry - s*tz*srx*sry) * coeff.x
            + (s*srx*srz*pointOri.x - s*crz*srx*pointOri.y + s*crx*pointOri.z
            + s*ty*crz*srx - s*tz*crx - s*tx*srx*srz) * coeff.y
            + (s*crx*cry*srz*pointOri.x - s*crx*cry*crz*pointOri.y - s*cry*srx*pointOri.z
            + s*tz*cry*srx + s*ty*crx*cry*crz - s*tx*crx*cry*srz) * coeff.z;
  float ary = ((-s*crz*sry - s*cry*srx*srz)*pointOri.x
            + (s*cry*crz*srx - s*sry*srz)*pointOri.y - s*crx*cry*pointOri.z
            + tx*(s*crz*sry + s*cry*srx*srz) + ty*(s*sry*srz - s*cry*crz*srx)
            + s*tz*crx*cry) * coeff.x
            + ((s*cry*crz - s*srx*sry*srz)*pointOri.x
            + (s*cry*srz + s*crz*srx*sry)*pointOri.y - s*crx*sry*pointOri.z
            + s*tz*crx*sry - ty*(s*cry*srz + s*crz*srx*sry)
            - tx*(s*cry*crz - s*srx*sry*srz)) * coeff.z;
  float arz = ((-s*cry*srz - s*crz*srx*sry)*pointOri.x + (s*cry*crz - s*srx*sry*srz)*pointOri.y
            + tx*(s*cry*srz + s*crz*srx*sry) - ty*(s*cry*crz - s*srx*sry*srz)) * coeff.x
            + (-s*crx*crz*pointOri.x - s*crx*srz*pointOri.y
            + s*ty*crx*srz + s*tx*crx*crz) * coeff.y
            + ((s*cry*crz*srx - s*sry*srz)*pointOri.x + (s*crz*sry + s*cry*srx*srz)*pointOri.y
            + tx*(s*sry*srz - s*cry*crz*srx) - ty*(s*crz*sry + s*cry*srx*srz)) * coeff.z;
  float atx = -s*(cry*crz - srx*sry*srz) * coeff.x + s*crx*srz * coeff.y
            - s*(crz*sry + cry*srx*srz) * coeff.z;
  float aty = -s*(cry*srz + crz*srx*sry) * coeff.x - s*crx*crz * coeff.y
            - s*(sry*srz - cry*crz*srx) * coeff.z;
  float atz = s*crx*sry * coeff.x - s*srx * coeff.y - s*crx*cry * coeff.z;
  float d2 = coeffv.w;
  b = -0.05 * d2;
  a[0] = arx; a[1] = ary; a[2] = arz; a[3] = atx; a[4] = aty; a[5] = atz;
}

}  // namespace loamgpu
