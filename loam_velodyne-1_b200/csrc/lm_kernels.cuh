// lm_kernels.cuh -- laserMapping on sm_100a, part 1 (SURVEY.md §8a rows LM-1 .. LM-11):
// pose prediction, cube-map bookkeeping, local-map assembly and the sweep-to-map LM loop.
//
// Map layout in HBM: the 21x11x21 cube array of the reference (laserMapping.cpp:64-70,88-91)
// becomes a table of (offset, count) descriptors into a log-structured arena of float4 points
// per feature type.  Cubes are immutable segments; every mapped sweep appends one new block with
// the refiltered contents of the touched cubes (part 2), so shifting the window (:454-614) moves
// 8-byte descriptors instead of point clouds and the local map is one gather of <= 125 segments.
// The scalar expressions restated in this file (cited below) come from loam_velodyne:
//   Copyright 2013, Ji Zhang, Carnegie Mellon University
//   Further contributions copyright (c) 2016, Southwest Research Institute.  All rights reserved.
// BSD 3-clause; the full notice, conditions and disclaimer are retained in NOTICE.md at the repository root.
#pragma once
#include <cooperative_groups.h>
#include "grid_index.cuh"
#include "lm_cube_index.cuh"
#include "linalg.cuh"
#include "pose_math.cuh"
#include "lo_kernels2.cuh"

namespace loamgpu {

constexpr int PROC_MAX = 250;                            // cubes rewritten per mapped sweep
#ifndef LOAM_LM_THREADS
#define LOAM_LM_THREADS 256
#endif
constexpr int LM_THREADS = LOAM_LM_THREADS;   // 256 = half an SM: lo_optimize and the other streams run beside it
constexpr int LM_THREADS_ALONE = 512;       // mapping on its own (synchronous pipeline, mapping-only node): twice the warps hide the 5-NN latency

struct LMImu {   // host-maintained ring buffer, laserMapping.cpp:102-108,323-335
  double time[IMU_Q];
  float roll[IMU_Q], pitch[IMU_Q];
  int pointer_last;
  int pad;
};

struct LMState {
  float transformSum[6], transformIncre[6], transformTobeMapped[6], transformBefMapped[6], transformAftMapped[6];  // :96-100
  int cenW, cenH, cenD;                 // laserCloudCenWidth/Height/Depth :64-66
  int isDegenerate;                     // :385
  float matP[36];                       // :386
  int imuPointerFront;                  // :102
  int validNum, surroundNum;            // :616-617
  int validInd[125], surroundInd[125];  // :72-73
  int gatherOff[2][126];                // start of each valid cube inside the local map
  int gatherSrc[2][125];                // where each valid cube lay in the arena when the pass began (local-map download)
  int verNext;                          // last cube version handed out (CubeDesc::ver, lm_cube_index.cuh)
  int mapN[2];                          // laserCloudCornerFromMapNum / SurfFromMapNum :680-681
  int stackN2[2];                       // laserCloudCornerStackNum2 / Surf :683,688
  int stackN[2];                        // laserCloudCornerStackNum / Surf after the filters :696,701
  int optimize_enable;                  // :706
  int iters, rows;
  float predTrig[6];                    // sin/cos of the predicted transformTobeMapped (rx,ry,rz)
  float predT[6];                       // the predicted transformTobeMapped
  float pointOnYAxis[3];                // :440-444
  int bump[2];                          // arena append pointers
  int arenaCap;
  int error;                            // bit0 arena full, bit1 too many cubes, bit2 voxel key range, bit3 local map cap
  // refilter plan (part 2)
  int nproc[2];
  int procCube[2][PROC_MAX];
  int procMode[2][PROC_MAX];            // 1 = valid cube (refiltered), 0 = only appended to
  int procNewCnt[2][PROC_MAX];          // new points of this sweep per processed cube
  int procNewOff[2][PROC_MAX + 1];      // their start inside the sorted new-point list
  int newSel[2];                        // which ping-pong buffer holds the sorted new-point list
  int scratchBump[2];                   // general-path scratch allocation of this sweep
  int rfStat[2][8];                     // last sweep: [0] merge cubes [1] general cubes [2] append cubes [3] max n_old
                                        // [4] max n_new [5] points written [6] new-only voxels [7] cubes gone dirty
  int surroundN;                        // surround cloud size (LM-14)
  int surroundIn;                       // size of the gathered (unfiltered) surround cloud
  int surroundOff[250];                 // start of every (cube, type) segment inside it
  int registeredN;                      // size of the registered full-resolution cloud
  int vxOverflow[4];
  // snapshot taken by lm_insert_ids at the end of a pass: the map update, the registered cloud and the
  // pose record of pass k run on the side stream while the front of pass k+1 (pose prediction, incoming
  // filters) already overwrites stackN / transformTobeMapped on the main stream
  int insN[2];
  float insT[6];
  // the same snapshot, two deep, of what the pipeline's pose record of the pass reports: the record kernel runs on the
  // second side stream, off the map-update chain, and may be late by a pass
  struct Rec { float aft[6], bef[6]; int iters, rows, deg, mapN[2], stackN[2], pad; } rec[2];
};

struct LMTrig { float srx, crx, sry, cry, srz, crz; };

__device__ __forceinline__ LMTrig lm_trig(const float* t) {
  LMTrig g;
  g.srx = dm::sin(t[0]); g.crx = dm::cos(t[0]);
  g.sry = dm::sin(t[1]); g.cry = dm::cos(t[1]);
  g.srz = dm::sin(t[2]); g.crz = dm::cos(t[2]);
  return g;
}

// pointAssociateToMap :234-252 with cos/sin(transformTobeMapped[k]) supplied (same values)
__device__ __forceinline__ float4 lm_to_map(const float4 pi, const LMTrig& g, const float* t) {
  float x1 = g.crz * pi.x - g.srz * pi.y;
  float y1 = g.srz * pi.x + g.crz * pi.y;
  float z1 = pi.z;
  float x2 = x1;
  float y2 = g.crx * y1 - g.srx * z1;
  float z2 = g.srx * y1 + g.crx * z1;
  float4 po;
  po.x = g.cry * x2 + g.sry * z2 + t[3];
  po.y = y2 + t[4];
  po.z = -g.sry * x2 + g.cry * z2 + t[5];
  po.w = pi.w;
  return po;
}

// pointAssociateTobeMapped :254-272
__device__ __forceinline__ float4 lm_to_sensor(const float4 pi, const LMTrig& g, const float* t) {
  float x1 = g.cry * (pi.x - t[3]) - g.sry * (pi.z - t[5]);
  float y1 = pi.y - t[4];
  float z1 = g.sry * (pi.x - t[3]) + g.cry * (pi.z - t[5]);
  float x2 = x1;
  float y2 = g.crx * y1 + g.srx * z1;
  float z2 = -g.srx * y1 + g.crx * z1;
  float4 po;
  po.x = g.crz * x2 + g.srz * y2;
  po.y = -g.srz * x2 + g.crz * y2;
  po.z = z2;
  po.w = pi.w;
  return po;
}

// cube index of a map coordinate :446-452 / :983-989 (double arithmetic, truncation + fix-up)
__device__ __forceinline__ int lm_cube_coord(float v, int cen) {
  int c = int((v + 25.0) / 50.0) + cen;
  if (v + 25.0 < 0) c--;
  return c;
}

// topic hand-over laserOdometry -> laserMapping (see LMHandle::Input): copies the three clouds and
// the odometry state of this sweep into an input bank
struct LMHandoffArgs {
  const LOState* lo;
  const float4* src[3];
  float4* dst[3];
  LOState* lo_out;
  int cap;
};
__global__ void __launch_bounds__(256) lm_handoff(LMHandoffArgs A) {
  const int n0 = min(A.lo->n_corner_last, A.cap), n1 = min(A.lo->n_surf_last, A.cap), n2 = min(A.lo->n_full_end, A.cap);
  const int total = n0 + n1 + n2;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    if (i < n0) A.dst[0][i] = A.src[0][i];
    else if (i < n0 + n1) A.dst[1][i - n0] = A.src[1][i - n0];
    else A.dst[2][i - n0 - n1] = A.src[2][i - n0 - n1];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *A.lo_out = *A.lo;
}

struct LMSetupArgs {
  LMState* st;
  const LOState* lo;        // transformSum of this sweep, cloud sizes
  CubeDesc* cubes[2];
  int phase;                // 1: pose prediction only (nothing of the cube map is touched); 2: the rest; 0: both
  CIState* ci;              // per-cube index (lm_cube_index.cuh); nullptr = the index is rebuilt over the gathered local map
};

// K1: one block.  /laser_odom_to_init hand-off (:304-318), transformAssociateToMap (:422),
// FOV probe point, centre cube + window shifts (:446-614), cube selection (:616-672) and the
// offsets of the local-map gather (:674-681).
// Two kernels from one body: the pose prediction (phase 1, one warp, main stream) and the cube bookkeeping (phase 2,
// side stream) are launched separately anyway, and a kernel that carries only its own phase's code misses the
// instruction cache far less (both ran at 3 % issue utilisation, most stalls "no instruction").
template <int PHASE>
__device__ __forceinline__ void lm_setup_body(const LMSetupArgs& A) {
  LMState* st = A.st;
  __shared__ int s_shift[3], s_center[3];
  __shared__ unsigned char s_fov[125], s_in[125];
  __shared__ int s_ind[125];
  if (PHASE != 2 && threadIdx.x < 32) {
    // pose prediction on one warp: the sines / cosines of the nine input angles of transformAssociateToMap are
    // evaluated by 18 lanes at once, lane 0 runs the closed forms against the table (pm::transformAssociateToMapV:
    // same expression trees, same bits)
    __shared__ float s_ang[16], s_sn[16], s_cs[16];
    const int lane = threadIdx.x;
    if (lane == 0) {
      float sum[6];
      if (A.lo->sum_decoded) { for (int i = 0; i < 6; i++) sum[i] = A.lo->transformSum[i]; }
      else pm::odometry_msg_roundtrip(A.lo->transformSum, sum);
      for (int i = 0; i < 6; i++) st->transformSum[i] = sum[i];
      for (int i = 0; i < 3; i++) { s_ang[i] = sum[i]; s_ang[3 + i] = st->transformBefMapped[i]; s_ang[6 + i] = st->transformAftMapped[i]; }
    }
    __syncwarp();
    {
      const int j = lane & 15;
      if (j < 9) {
        const float a = s_ang[j];
        if (lane < 16) s_sn[j] = dm::sin(a); else s_cs[j] = dm::cos(a);
      }
      __syncwarp();
    }
    if (lane == 0) {
      pm::transformAssociateToMapV(s_sn, s_cs, st->transformSum, st->transformBefMapped, st->transformAftMapped, st->transformIncre, st->transformTobeMapped);
      const float* T = st->transformTobeMapped;
      LMTrig g = lm_trig(T);
      st->predTrig[0] = g.srx; st->predTrig[1] = g.crx; st->predTrig[2] = g.sry; st->predTrig[3] = g.cry; st->predTrig[4] = g.srz; st->predTrig[5] = g.crz;
      for (int i = 0; i < 6; i++) st->predT[i] = T[i];
      float4 py = lm_to_map(make_float4(0.0f, 10.0f, 0.0f, 0.f), g, T);
      st->pointOnYAxis[0] = py.x; st->pointOnYAxis[1] = py.y; st->pointOnYAxis[2] = py.z;
      st->stackN2[0] = A.lo->n_corner_last; st->stackN2[1] = A.lo->n_surf_last;
    }
  }
  if (PHASE == 1) return;
  if (threadIdx.x == 0) {
    const float* T = st->transformTobeMapped;
    int cI = lm_cube_coord(T[3], st->cenW), cJ = lm_cube_coord(T[4], st->cenH), cK = lm_cube_coord(T[5], st->cenD);
    // net shifts per axis: +1 per `while (center < 3)`, -1 per `while (center >= dim - 3)`
    int dims[3] = {CUBE_W, CUBE_H, CUBE_D};
    int c[3] = {cI, cJ, cK};
    for (int a = 0; a < 3; a++) {
      int sh = 0;
      while (c[a] < 3) { c[a]++; sh++; }
      while (c[a] >= dims[a] - 3) { c[a]--; sh--; }
      s_shift[a] = sh;
      s_center[a] = c[a];
    }
    st->cenW += s_shift[0]; st->cenH += s_shift[1]; st->cenD += s_shift[2];
  }
  __syncthreads();
  // apply the shifts to the descriptor tables, one (line, type) per thread
  for (int a = 0; a < 3; a++) {
    int sh = s_shift[a];
    if (sh == 0) continue;
    int dims[3] = {CUBE_W, CUBE_H, CUBE_D};
    int n = dims[a], n1 = dims[(a + 1) % 3], n2 = dims[(a + 2) % 3];
    for (int line = threadIdx.x; line < n1 * n2 * 2; line += blockDim.x) {
      int T = line / (n1 * n2), l = line % (n1 * n2);
      int u = l / n2, v = l % n2;
      CubeDesc* cubes = A.cubes[T];
      int steps = sh > 0 ? sh : -sh;
      for (int sidx = 0; sidx < steps; sidx++) {
        if (sh > 0) {
          for (int t = n - 1; t >= 1; t--) {
            int cc[3], cp[3];
            cc[a] = t; cc[(a + 1) % 3] = u; cc[(a + 2) % 3] = v;
            cp[a] = t - 1; cp[(a + 1) % 3] = u; cp[(a + 2) % 3] = v;
            cubes[cc[0] + CUBE_W * cc[1] + CUBE_W * CUBE_H * cc[2]] = cubes[cp[0] + CUBE_W * cp[1] + CUBE_W * CUBE_H * cp[2]];
          }
          int c0[3];
          c0[a] = 0; c0[(a + 1) % 3] = u; c0[(a + 2) % 3] = v;
          { CubeDesc& cd = cubes[c0[0] + CUBE_W * c0[1] + CUBE_W * CUBE_H * c0[2]]; cd.cnt = 0; cd.dirty = 0; }
        } else {
          for (int t = 0; t < n - 1; t++) {
            int cc[3], cp[3];
            cc[a] = t; cc[(a + 1) % 3] = u; cc[(a + 2) % 3] = v;
            cp[a] = t + 1; cp[(a + 1) % 3] = u; cp[(a + 2) % 3] = v;
            cubes[cc[0] + CUBE_W * cc[1] + CUBE_W * CUBE_H * cc[2]] = cubes[cp[0] + CUBE_W * cp[1] + CUBE_W * CUBE_H * cp[2]];
          }
          int c0[3];
          c0[a] = n - 1; c0[(a + 1) % 3] = u; c0[(a + 2) % 3] = v;
          { CubeDesc& cd = cubes[c0[0] + CUBE_W * c0[1] + CUBE_W * CUBE_H * c0[2]]; cd.cnt = 0; cd.dirty = 0; }
        }
      }
    }
    __syncthreads();
  }
  // 5x5x5 neighbourhood, FOV test on the 8 corners (:618-672)
  if (threadIdx.x < 125) {
    int di = threadIdx.x / 25, dj = (threadIdx.x / 5) % 5, dk = threadIdx.x % 5;
    int i = s_center[0] - 2 + di, j = s_center[1] - 2 + dj, k = s_center[2] - 2 + dk;
    bool in = i >= 0 && i < CUBE_W && j >= 0 && j < CUBE_H && k >= 0 && k < CUBE_D;
    bool isInLaserFOV = false;
    if (in) {
      const float* T = st->transformTobeMapped;
      const float* P = st->pointOnYAxis;
      float centerX = 50.0 * (i - st->cenW);
      float centerY = 50.0 * (j - st->cenH);
      float centerZ = 50.0 * (k - st->cenD);
      for (int ii = -1; ii <= 1; ii += 2) {
        for (int jj = -1; jj <= 1; jj += 2) {
          for (int kk = -1; kk <= 1; kk += 2) {
            float cornerX = centerX + 25.0 * ii;
            float cornerY = centerY + 25.0 * jj;
            float cornerZ = centerZ + 25.0 * kk;
            float squaredSide1 = (T[3] - cornerX) * (T[3] - cornerX) + (T[4] - cornerY) * (T[4] - cornerY) + (T[5] - cornerZ) * (T[5] - cornerZ);
            float squaredSide2 = (P[0] - cornerX) * (P[0] - cornerX) + (P[1] - cornerY) * (P[1] - cornerY) + (P[2] - cornerZ) * (P[2] - cornerZ);
            float check1 = 100.0 + squaredSide1 - squaredSide2 - 10.0 * dm::dsqrt(3.0) * dm::sqrt(squaredSide1);
            float check2 = 100.0 + squaredSide1 - squaredSide2 + 10.0 * dm::dsqrt(3.0) * dm::sqrt(squaredSide1);
            if (check1 < 0 && check2 > 0) isInLaserFOV = true;
          }
        }
      }
    }
    s_in[threadIdx.x] = in ? 1 : 0;
    s_fov[threadIdx.x] = isInLaserFOV ? 1 : 0;
    s_ind[threadIdx.x] = in ? i + CUBE_W * j + CUBE_W * CUBE_H * k : -1;
  }
  __syncthreads();
  __shared__ int s_vlist[125], s_nv, s_rankq[125];
  __shared__ int s_cnt[2][128];
  if (threadIdx.x == 0) {
    int nv = 0, ns = 0;
    for (int q = 0; q < 125; q++) {  // (i, j, k) loop order == q order
      s_rankq[q] = -1;
      if (!s_in[q]) continue;
      if (s_fov[q]) { s_rankq[q] = nv; s_vlist[nv] = s_ind[q]; st->validInd[nv++] = s_ind[q]; }
      st->surroundInd[ns++] = s_ind[q];
    }
    st->validNum = nv; st->surroundNum = ns;
    s_nv = nv;
    st->iters = 0; st->rows = 0;
  }
  __syncthreads();
  // offsets of the valid cubes inside the local map: the cube sizes are fetched in parallel
  // (a serial loop here is 250 dependent global loads on the critical path of every mapped sweep)
  const int nv = s_nv;
  if (threadIdx.x < 250) {
    const int T = threadIdx.x / 125, v = threadIdx.x % 125;
    CubeDesc d;
    d.off = 0; d.cnt = 0;
    if (v < nv) d = A.cubes[T][s_vlist[v]];
    s_cnt[T][v] = d.cnt;
    st->gatherSrc[T][v] = d.off;
  }
  __syncthreads();
  if (threadIdx.x < 64) {   // warp T scans its 125 counts
    const int T = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int run = 0;
    for (int v0 = 0; v0 < 128; v0 += 32) {
      const int v = v0 + lane;
      const int c = v < 125 ? s_cnt[T][v] : 0;
      int x = c;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      if (v < nv) st->gatherOff[T][v] = run + x - c;
      run += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) { st->gatherOff[T][nv] = run; st->mapN[T] = run; s_cnt[T][127] = run; }
  }
  __syncthreads();
  if (threadIdx.x == 0) st->optimize_enable = (s_cnt[0][127] > 10 && s_cnt[1][127] > 100) ? 1 : 0;  // :706
  if (A.ci == nullptr) return;
  // ---- per-cube index (lm_cube_index.cuh): which slots the queries of this pass search, which of them are stale ----
  CIState* ci = A.ci;
  __shared__ int s_need[2], s_cur[2], s_reset[2], s_nb, s_npts, s_bcnt[256];
  s_bcnt[threadIdx.x] = -1;
  if (threadIdx.x < 2) { s_need[threadIdx.x] = 0; s_cur[threadIdx.x] = ci->cursor[threadIdx.x]; s_reset[threadIdx.x] = 0; }
  if (threadIdx.x == 0) { s_nb = 0; s_npts = 0; }
  __syncthreads();
  const int T = threadIdx.x / 125, q = threadIdx.x % 125;
  const bool mine = threadIdx.x < 250;
  int slot = 0, wi = 0, wj = 0, wk = 0, cnt = 0, ver = 0, need = 0;
  bool present = false, stale = false;
  CISlot S;
  if (mine) {
    wi = s_center[0] - 2 + q / 25 - st->cenW; wj = s_center[1] - 2 + (q / 5) % 5 - st->cenH; wk = s_center[2] - 2 + q % 5 - st->cenD;
    slot = ci_pmod5(wi) + 5 * ci_pmod5(wj) + 25 * ci_pmod5(wk);
    S = ci->slot[T][slot];
    if (s_rankq[q] >= 0) {
      const CubeDesc d = A.cubes[T][s_ind[q]];
      cnt = d.cnt; ver = d.ver;
      present = cnt > 0;
    }
    if (present) {
      stale = !(S.wi == wi && S.wj == wj && S.wk == wk && S.ver == ver && S.cnt == cnt && S.room >= cnt);
      if (stale && S.room < cnt) { need = cnt + cnt / 8 + 256; atomicAdd(&s_need[T], need); }
      if (cnt >= (1 << 24)) atomicOr(&st->error, 32);   // the position inside a cube is a 24-bit field of the candidate id
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 && s_cur[threadIdx.x] + s_need[threadIdx.x] > ci->cap) { s_reset[threadIdx.x] = 1; s_cur[threadIdx.x] = 0; s_need[threadIdx.x] = 0; }
  __syncthreads();
  if (mine && s_reset[T]) {
    // the index arena of this type is exhausted: every region is dropped and the local map is indexed from scratch
    S.room = 0; S.ver = -1;
    ci->slot[T][slot] = S;
    need = 0;
    if (present) { stale = true; need = cnt + cnt / 8 + 256; atomicAdd(&s_need[T], need); }
  }
  __syncthreads();
  if (threadIdx.x < 2 && s_reset[threadIdx.x]) {
    atomicAdd(&ci->stat[2], 1);
    if (s_need[threadIdx.x] > ci->cap) { atomicOr(&st->error, 32); s_reset[threadIdx.x] = 2; }
  }
  __syncthreads();
  if (mine) {
    const bool fatal = s_reset[T] == 2;   // the local map does not fit the index arena: nothing is searched, the error flag is up
    if (stale && !fatal) {
      if (need) { S.base = atomicAdd(&s_cur[T], need); S.room = need; }
      S.wi = wi; S.wj = wj; S.wk = wk; S.ver = ver; S.cnt = cnt; S.fill = 0;
      ci->slot[T][slot] = S;
      atomicAdd(&s_nb, 1);
      atomicAdd(&s_npts, cnt);
      s_bcnt[threadIdx.x] = cnt;
    }
    ci->win_slot[T][q] = (present && !fatal) ? slot : -1;
    ci->win_cube[T][q] = s_ind[q];
    if (T == 0) ci->win_rank[q] = s_rankq[q];
  }
  __syncthreads();
  // build list, largest cube first: the clusters of ci_build walk it with a fixed stride, so the big cubes start at once
  // and the small ones fill in behind them
  if (mine && s_bcnt[threadIdx.x] >= 0) {
    const int c = s_bcnt[threadIdx.x];
    int rank = 0;
    for (int j = 0; j < 250; j++) {
      const int o = s_bcnt[j];
      rank += (o > c || (o == c && j < (int)threadIdx.x)) ? 1 : 0;
    }
    ci->build[rank] = T * 125 + q;
  }
  if (threadIdx.x == 0) {
    ci->nbuild = s_nb;
    ci->cursor[0] = s_cur[0]; ci->cursor[1] = s_cur[1];
    ci->org[0] = 50.0f * (float)(s_center[0] - 2 - st->cenW) - 25.0f;
    ci->org[1] = 50.0f * (float)(s_center[1] - 2 - st->cenH) - 25.0f;
    ci->org[2] = 50.0f * (float)(s_center[2] - 2 - st->cenD) - 25.0f;
    ci->stat[0] += s_nb; ci->stat[1] += s_npts; ci->stat[3] = s_nb; ci->stat[4] = s_npts;
  }
}

__global__ void __launch_bounds__(32) lm_setup(LMSetupArgs A) { lm_setup_body<1>(A); }
__global__ void __launch_bounds__(256) lm_setup_cubes(LMSetupArgs A) { lm_setup_body<2>(A); }

// K2: incoming clouds to the map with the predicted pose and back (:424-434, :683-691)
struct LMStackArgs {
  const LMState* st;
  const float4* in[2];
  float4* out[2];
};
__global__ void __launch_bounds__(256) lm_stack_transform(LMStackArgs A) {
  const LMState* st = A.st;
  LMTrig g;
  g.srx = st->predTrig[0]; g.crx = st->predTrig[1]; g.sry = st->predTrig[2]; g.cry = st->predTrig[3]; g.srz = st->predTrig[4]; g.crz = st->predTrig[5];
  float T[6];
  for (int i = 0; i < 6; i++) T[i] = st->predT[i];
  const int n0 = st->stackN2[0], n1 = st->stackN2[1];
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n0 + n1; i += gridDim.x * 256) {
    int t = i < n0 ? 0 : 1;
    int k = t ? i - n0 : i;
    A.out[t][k] = lm_to_sensor(lm_to_map(A.in[t][k], g, T), g, T);
  }
}

// K4: local map = concat of the valid cubes in (i,j,k) loop order (:674-679)
struct LMGatherArgs {
  LMState* st;
  const CubeDesc* cubes[2];
  const float4* arena[2];
  float4* E[2];
  int ecap;
};
// Flat form: thread i copies local-map point i; its cube is found by a binary search over the
// <= 126 segment starts staged in shared memory.  The work per CTA is the same whatever the cube
// sizes are (a dense map keeps most of its points in a handful of cubes), and every load is
// independent -- four in flight per thread.
__global__ void __launch_bounds__(256) lm_gather_map(LMGatherArgs A) {
  const LMState* st = A.st;
  const int T = blockIdx.y;
  __shared__ int s_off[128], s_src[128];
  const int nv = st->validNum;
  for (int v = threadIdx.x; v <= nv; v += 256) {
    s_off[v] = st->gatherOff[T][v];
    s_src[v] = v < nv ? st->gatherSrc[T][v] : 0;
  }
  __syncthreads();
  const int total = s_off[nv];
  if (total > A.ecap) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(&A.st->error, 8); return; }
  const float4* __restrict__ arena = A.arena[T];
  float4* __restrict__ E = A.E[T];
  for (int base = blockIdx.x * 1024; base < total; base += gridDim.x * 1024) {
    float4 p[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = min(base + u * 256 + (int)threadIdx.x, total - 1);
      int lo = 0, hi = nv;          // largest v with s_off[v] <= i (empty cubes share a start: take the last)
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_off[mid] <= i) lo = mid; else hi = mid;
      }
      p[u] = __ldg(&arena[s_src[lo] + (i - s_off[lo])]);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = base + u * 256 + (int)threadIdx.x;
      if (i < total) E[i] = p[u];
    }
  }
}

// ---- exact 5-NN inside radius 1 (the only case the reference accepts, :719,:826) ------------
// A group of W lanes (half a warp) cooperates on one query.  Returns true when the 5th neighbour
// has d2 < 1.0; neighbours ascending by (d2, index).  out[r] = (x, y, z, original index bits) of
// neighbour r, written by the lane that owns it (out is shared memory).  One batch of cell-header
// loads (grid_rows9), then every candidate load is independent: the dependent-load depth per
// query is 2, not 3 x 27.  The 5-way merge uses the warp-reduce unit (redux.sync) on the distance
// bits (non-negative floats order like unsigned ints), ties resolved by a second reduce on the index.
// have_prev: out[0..4] still holds the five neighbours this query had in the PREVIOUS LM iteration
// (same cloud, the query has moved by one small pose update).  Their largest distance from the new
// position bounds the new fifth-neighbour distance from above, so only the cells that ball reaches
// are scanned (grid_rows9_bounded): typically a third of the candidates, same exact answer.
template <int W>
__device__ __forceinline__ bool lm_knn5_group(const GridView& g, const GridDesc& d, float qx, float qy, float qz, float4* out, bool have_prev = false) {
  const int lane = lane_id();
  const int gl = lane & (W - 1);
  const unsigned int gmask = (W == 32) ? 0xffffffffu : (((1u << W) - 1u) << (lane & ~(W - 1)));
  float ld[5]; int li[5]; float lx[5], ly[5], lz[5];
#pragma unroll
  for (int k = 0; k < 5; k++) { ld[k] = 3.0e38f; li[k] = 0x7fffffff; lx[k] = ly[k] = lz[k] = 0.f; }
  int total = 0;
  if (d.valid) {
    int cx = (int)floorf((qx - d.ox) * d.inv_h), cy = (int)floorf((qy - d.oy) * d.inv_h), cz = (int)floorf((qz - d.oz) * d.inv_h);
    // cell size >= 1.02 m: every point with d2 < 1 lies in the 3x3x3 block around the query cell
    GridRows9 R;
    if (have_prev) {   // uniform inside the group
      float r2 = 0.f;
#pragma unroll
      for (int k = 0; k < 5; k++) {
        const float4 p = out[k];
        float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
        float dd = 0.f;
        dd += dx * dx; dd += dy * dy; dd += dz * dz;
        r2 = fmaxf(r2, dd);
      }
      __syncwarp(gmask);   // every lane has read the old neighbours before any lane stores new ones
      if constexpr (W >= 16) grid_rows9_bounded<W>(g, d, cx, cy, cz, qx, qy, qz, r2, R);
      else grid_rows9<W>(g, d, cx, cy, cz, R);   // the bounded form needs a lane per row
    } else {
      grid_rows9<W>(g, d, cx, cy, cz, R);
    }
    total = R.pre[9];
    if (total >= 5) {
      // four independent candidate loads in flight per lane before the first one is consumed
      for (int t0 = gl; t0 < total; t0 += 4 * W) {
        float4 pc[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int t = t0 + u * W;
          pc[u] = g.sorted_pts[grid_rows9_pos(R, t < total ? t : 0)];
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          if (t0 + u * W >= total) break;
          const float4 p = pc[u];
          float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
          float dd = 0.f;
          dd += dx * dx; dd += dy * dy; dd += dz * dz;
          int id = __float_as_int(p.w);
          if (nn_less(dd, id, ld[4], li[4])) {
            ld[4] = dd; li[4] = id; lx[4] = p.x; ly[4] = p.y; lz[4] = p.z;
#pragma unroll
            for (int k = 4; k > 0; k--) {
              if (nn_less(ld[k], li[k], ld[k - 1], li[k - 1])) {
                float td = ld[k]; ld[k] = ld[k - 1]; ld[k - 1] = td;
                int ti = li[k]; li[k] = li[k - 1]; li[k - 1] = ti;
                float tx = lx[k]; lx[k] = lx[k - 1]; lx[k - 1] = tx;
                float ty = ly[k]; ly[k] = ly[k - 1]; ly[k - 1] = ty;
                float tz = lz[k]; lz[k] = lz[k - 1]; lz[k - 1] = tz;
              }
            }
          }
        }
      }
    }
  }
  if (total < 5) return false;   // uniform inside the group
  // merge the W sorted lists: five rounds, the winner pops its head
  unsigned int bdu = 0; int bi = 0;
#pragma unroll
  for (int r = 0; r < 5; r++) {
    bdu = __reduce_min_sync(gmask, __float_as_uint(ld[0]));
    const bool tie = __float_as_uint(ld[0]) == bdu;
    bi = (int)__reduce_min_sync(gmask, tie ? (unsigned int)li[0] : 0x7fffffffu);
    if (tie && li[0] == bi && bi != 0x7fffffff) {   // indices are unique: exactly one owner
      out[r] = make_float4(lx[0], ly[0], lz[0], __int_as_float(bi));
#pragma unroll
      for (int k = 0; k < 4; k++) { ld[k] = ld[k + 1]; li[k] = li[k + 1]; lx[k] = lx[k + 1]; ly[k] = ly[k + 1]; lz[k] = lz[k + 1]; }
      ld[4] = 3.0e38f; li[4] = 0x7fffffff;
    }
  }
  return bi != 0x7fffffff && __uint_as_float(bdu) < 1.0;
}

}  // namespace loamgpu
