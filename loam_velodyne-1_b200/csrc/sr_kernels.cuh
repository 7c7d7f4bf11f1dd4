// sr_kernels.cuh -- scanRegistration on sm_100a (SURVEY.md §8a rows SR-1 .. SR-6).
//
// The reference's laserCloudHandler (scanRegistration.cpp:211-636) is a chain of sequential
// loops with hidden state (halfPassed latch, stable ring bucketing, suppression flags spilling
// between sectors).  Here each loop is re-expressed as a data-parallel kernel with the state made
// explicit:
//   sr_ring_ori     ring id + azimuth per raw point, per-block ring histogram, latch index (min)
//   sr_ring_scan    exclusive scan ring-major -> stable scatter bases, scanStartInd/scanEndInd
//   sr_scatter      relTime/intensity, stable ring-major scatter into the float4 cloud
//   sr_imu_deskew   (IMU only) prefix-max IMU pointer, interpolation, shift/rotate to sweep start
//   sr_curv_reject  11-tap curvature + occlusion/parallel-beam rejection, halo staged in smem
//   sr_sort_sectors per (ring, sector) bitonic sort of (curvature, index) keys in smem
//   sr_select       one warp per ring: greedy sharp/less-sharp/flat pick with ballot scans
//   sr_lessflat     one block per ring: label<=0 compaction + PCL-identical 0.2 m voxel filter
//   sr_compact      emits the four feature clouds in ring -> sector -> pick order
// All float expressions mirror the reference's expression trees (compiled with -fmad=false).
#pragma once
#include "common.cuh"

namespace loamgpu {

constexpr int SR_BLK = 256;          // points per block in ring/scatter kernels
constexpr int SR_MAX_RINGS = 128;
constexpr int SR_SECTOR_CAP = 1024;  // max points per (ring, sector) for the smem sort
constexpr int SR_RING_CAP = 4096;    // max less-flat candidates per ring for the smem voxel filter
constexpr int SR_SHARP_PER_RING = 12, SR_LESS_SHARP_PER_RING = 120, SR_FLAT_PER_RING = 24;
constexpr int IMU_Q = 200;           // imuQueLength, scanRegistration.cpp:70

struct SRCounts {
  int n_full;         // cloudSize after ring filtering
  int h_min;          // raw index of the point that sets halfPassed (INT_MAX if none)
  int n_sharp, n_less_sharp, n_flat, n_less_flat;
  int error;          // bit 0: sector cap, bit 1: ring cap, bit 2: max_points
  int last_valid;     // raw index of the last point that survives NaN + ring filtering (-1 if none)
};

// IMU ring buffer snapshot uploaded per sweep (host maintains it, scanRegistration.cpp:638-660)
struct ImuSnapshot {
  double time[IMU_Q];
  float roll[IMU_Q], pitch[IMU_Q], yaw[IMU_Q];
  float velo[3][IMU_Q], shift[3][IMU_Q];
  int pointer_last;
  int pad;
};

// IMU state that persists across sweeps on the device (double buffered by sweep parity)
struct ImuState {
  int pointer_front;
  float rpy_start[3];     // roll, pitch, yaw
  float velo_start[3], shift_start[3];
  float rpy_cur[3];
  float velo_cur[3], shift_cur[3];
  float shift_from_start[3], velo_from_start[3];
};

struct SRParamsDev {
  int n_scans, ring_model;
  float ring_lower, ring_upper;
  int max_points;
};

// ---- ring id, scanRegistration.cpp:248-260 (+ bk variant :268-274 for ring_model 1) ----
__device__ __forceinline__ int sr_ring_id(float px, float py, float pz, const SRParamsDev& P) {
  float angle = dm::atan(py / dm::sqrt(px * px + pz * pz)) * 180 / M_PI;
  int scanID;
  if (P.ring_model == 0) {
    int roundedAngle = int(angle + (angle < 0.0 ? -0.5 : +0.5));
    if (roundedAngle > 0) scanID = roundedAngle;
    else scanID = roundedAngle + (P.n_scans - 1);
  } else {
    const float angleLowerBoundDeg = P.ring_lower;
    const float angleUpperBoundDeg = P.ring_upper;
    const float angleSpan = angleUpperBoundDeg - angleLowerBoundDeg;
    const float angleStep = angleSpan / (P.n_scans - 1);
    float angleID = (angle - angleLowerBoundDeg) / angleStep;
    scanID = int(angleID + 0.5f);
  }
  if (scanID > (P.n_scans - 1) || scanID < 0) return -1;
  return scanID;
}

// branch taken before halfPassed latches, :263-272.  returns ori, sets *latch
__device__ __forceinline__ float sr_ori_branch_a(float ori, float startOri, bool* latch) {
  if (ori < startOri - M_PI / 2) {
    ori += 2 * M_PI;
  } else if (ori > startOri + M_PI * 3 / 2) {
    ori -= 2 * M_PI;
  }
  *latch = (ori - startOri > M_PI);
  return ori;
}
// branch taken after the latch, :273-281
__device__ __forceinline__ float sr_ori_branch_b(float ori, float endOri) {
  ori += 2 * M_PI;
  if (ori < endOri - M_PI * 3 / 2) {
    ori += 2 * M_PI;
  } else if (ori > endOri + M_PI / 2) {
    ori -= 2 * M_PI;
  }
  return ori;
}

// K-A: per raw point ring id + raw azimuth; per-block ring histogram; latch index by atomicMin
__device__ __forceinline__ void sr_ring_ori_body(const float* __restrict__ raw, int n, int stride_floats,
                                                      SRParamsDev P, float startOri, short* __restrict__ ring,
                                                      float* __restrict__ ori_raw, int* __restrict__ hist, int nblk,
                                                      SRCounts* cnt) {
  __shared__ int sh[SR_MAX_RINGS];
  for (int r = threadIdx.x; r < P.n_scans; r += SR_BLK) sh[r] = 0;
  __syncthreads();
  int i = blockIdx.x * SR_BLK + threadIdx.x;
  if (i < n) {
    const float* p = raw + (size_t)i * stride_floats;
    float vx = p[0], vy = p[1], vz = p[2];
    int rid = -1;
    float ori = 0.f;
    if (isfinite(vx) && isfinite(vy) && isfinite(vz)) {  // removeNaNFromPointCloud :228
      float px = vy, py = vz, pz = vx;                   // axis swap :244-246
      rid = sr_ring_id(px, py, pz, P);
      if (rid >= 0) {
        ori = -dm::atan2(px, pz);  // :262
        bool latch;
        sr_ori_branch_a(ori, startOri, &latch);
        if (latch) atomicMin(&cnt->h_min, i);
        atomicMax(&cnt->last_valid, i);
        atomicAdd(&sh[rid], 1);
      }
    }
    ring[i] = (short)rid;
    ori_raw[i] = ori;
  }
  __syncthreads();
  for (int r = threadIdx.x; r < P.n_scans; r += SR_BLK) hist[r * nblk + blockIdx.x] = sh[r];
}

// K-B: single block.  Stable ring bucketing (:350-357): one warp per ring scans that ring's
// per-block counts (exclusive, relative to the ring), the ring totals give ring_start[] and
// cloudSize; scanStartInd/scanEndInd exactly as the loop :359-393 derives them from ring changes
// inside [5, cloudSize-6].  sr_scatter adds ring_start[ring] to the relative offsets.
// One CTA per ring scans that ring's row of block counts; the last CTA to finish (ticket counter)
// turns the ring totals into ring_start / scanStartInd / scanEndInd.
constexpr int RSC_THREADS = 256;
__device__ __forceinline__ void sr_ring_scan_body(int* __restrict__ hist, int nblk, int n_scans,
                                                            int* __restrict__ ring_start, int* __restrict__ scan_start,
                                                            int* __restrict__ scan_end, SRCounts* cnt, int max_points,
                                                            int* ring_tot, unsigned int* ticket) {
  __shared__ int s_w[RSC_THREADS / 32];
  __shared__ int s_run, s_last;
  __shared__ int s_tot[SR_MAX_RINGS];
  __shared__ int s_start[SR_MAX_RINGS + 1];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  {
    const int r = blockIdx.x;
    int* row = hist + (size_t)r * nblk;
    if (threadIdx.x == 0) s_run = 0;
    __syncthreads();
    for (int b0 = 0; b0 < nblk; b0 += RSC_THREADS) {
      const int i = b0 + threadIdx.x;
      const int v = i < nblk ? row[i] : 0;
      int x = v;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      if (lane == 31) s_w[warp] = x;
      __syncthreads();
      int off = s_run;
      for (int w = 0; w < warp; w++) off += s_w[w];
      if (i < nblk) row[i] = off + x - v;
      __syncthreads();
      if (threadIdx.x == RSC_THREADS - 1) s_run = off + x;
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      ring_tot[r] = s_run;
      __threadfence();
      const unsigned int t = atomicAdd(ticket, 1u);
      s_last = (t == (unsigned int)n_scans - 1u) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) *ticket = 0u;   // re-arm for the next sweep (stream order)
    __threadfence();
  }
  for (int r = threadIdx.x; r < n_scans; r += RSC_THREADS) s_tot[r] = *((volatile int*)&ring_tot[r]);
  __syncthreads();
  if (warp == 0) {
    int run = 0;
    for (int r0 = 0; r0 < n_scans; r0 += 32) {
      int r = r0 + lane;
      int v = r < n_scans ? s_tot[r] : 0;
      int x = v;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      if (r < n_scans) { s_start[r] = run + x - v; ring_start[r] = run + x - v; }
      run += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) { s_start[n_scans] = run; ring_start[n_scans] = run; }
  }
  __syncthreads();
  const int cloudSize = s_start[n_scans];
  for (int r = threadIdx.x; r < n_scans; r += RSC_THREADS) { scan_start[r] = 0; scan_end[r] = 0; }
  __syncthreads();
  // every ring r > 0 that owns a position inside [5, cloudSize-6] starts a scan (:383-390) and
  // closes the scan of ring r-1 (the ring id the reference tracks is simply "previous ring + 1")
  for (int r = threadIdx.x; r < n_scans; r += RSC_THREADS) {
    int f = s_start[r], c = s_start[r + 1] - f;
    if (c <= 0 || r == 0) continue;
    int lo = f > 5 ? f : 5;
    int hi = f + c - 1 < cloudSize - 6 ? f + c - 1 : cloudSize - 6;
    if (lo <= hi) {
      scan_start[r] = lo + 5;
      scan_end[r - 1] = lo - 5;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    cnt->n_full = cloudSize;
    if (cloudSize > max_points) cnt->error |= 4;
    scan_start[0] = 5;                     // :392
    scan_end[n_scans - 1] = cloudSize - 5; // :393
  }
}

// K-C: final azimuth -> relTime -> intensity (:262-284) and the stable ring-major scatter
__device__ __forceinline__ void sr_scatter_body(const float* __restrict__ raw, int n, int stride_floats,
                                                     SRParamsDev P, float startOri, float endOri,
                                                     const short* __restrict__ ring, const float* __restrict__ ori_raw,
                                                     const int* __restrict__ base, const int* __restrict__ ring_start, int nblk,
                                                     const SRCounts* cnt, float4* __restrict__ full, int* __restrict__ pos_out,
                                                     float* __restrict__ reltime_out, int meta = 0) {
  __shared__ unsigned short wcount[SR_BLK / 32][SR_MAX_RINGS];
  for (int k = threadIdx.x; k < (SR_BLK / 32) * SR_MAX_RINGS; k += SR_BLK) (&wcount[0][0])[k] = 0;
  __syncthreads();
  int i = blockIdx.x * SR_BLK + threadIdx.x;
  int rid = i < n ? (int)ring[i] : -1;
  int key = rid < 0 ? SR_MAX_RINGS : rid;
  unsigned int m = __match_any_sync(0xffffffffu, key);
  int rank_in_warp = __popc(m & ((1u << lane_id()) - 1u));
  int warp = threadIdx.x >> 5;
  if (rid >= 0 && rank_in_warp == 0) wcount[warp][rid] = (unsigned short)__popc(m);
  __syncthreads();
  if (rid >= 0) {
    int off = 0;
    for (int w = 0; w < warp; w++) off += wcount[w][rid];
    int pos = ring_start[rid] + base[rid * nblk + blockIdx.x] + off + rank_in_warp;
    const float* p = raw + (size_t)i * stride_floats;
    float px = p[1], py = p[2], pz = p[0];
    float relTime;
    if (meta) {   // packet metadata front end (pk_ring_meta): ori_raw already holds relTime
      relTime = ori_raw[i];
    } else {
      float ori = ori_raw[i];
      if (i <= cnt->h_min) {
        bool latch;
        ori = sr_ori_branch_a(ori, startOri, &latch);
      } else {
        ori = sr_ori_branch_b(ori, endOri);
      }
      relTime = (ori - startOri) / (endOri - startOri);
    }
    const double scanPeriod = 0.1;  // :55
    float intensity = rid + scanPeriod * relTime;
    if (pos < P.max_points) full[pos] = make_float4(px, py, pz, intensity);
    pos_out[i] = pos;
    reltime_out[i] = relTime;
  } else if (i < n) {
    pos_out[i] = -1;
  }
}

// ---- IMU deskew (:286-349 and :108-171) -------------------------------------------------
// f(t) = number of ring-buffer steps the sequential `while` loop (:288-293) would advance from
// the sweep-start pointer if it saw only this point; the sequential pointer is the running max.
__device__ __forceinline__ int sr_imu_steps(double t, const double* s_time, int front0, int L) {
  int p = 0;
  while (p < L) {
    if (t < s_time[(front0 + p) % IMU_Q]) break;
    p++;
  }
  return p;
}

struct ImuCur { float roll, pitch, yaw, vx, vy, vz, sx, sy, sz; };

// interpolation :295-331 at ring-buffer slot `front`
__device__ __forceinline__ ImuCur sr_imu_interp(const ImuSnapshot* S, int front, double timeScanCur, float pointTime) {
  ImuCur c;
  if (timeScanCur + pointTime > S->time[front]) {
    c.roll = S->roll[front]; c.pitch = S->pitch[front]; c.yaw = S->yaw[front];
    c.vx = S->velo[0][front]; c.vy = S->velo[1][front]; c.vz = S->velo[2][front];
    c.sx = S->shift[0][front]; c.sy = S->shift[1][front]; c.sz = S->shift[2][front];
  } else {
    int back = (front + IMU_Q - 1) % IMU_Q;
    float ratioFront = (timeScanCur + pointTime - S->time[back]) / (S->time[front] - S->time[back]);
    float ratioBack = (S->time[front] - timeScanCur - pointTime) / (S->time[front] - S->time[back]);
    c.roll = S->roll[front] * ratioFront + S->roll[back] * ratioBack;
    c.pitch = S->pitch[front] * ratioFront + S->pitch[back] * ratioBack;
    if (S->yaw[front] - S->yaw[back] > M_PI) {
      c.yaw = S->yaw[front] * ratioFront + (S->yaw[back] + 2 * M_PI) * ratioBack;
    } else if (S->yaw[front] - S->yaw[back] < -M_PI) {
      c.yaw = S->yaw[front] * ratioFront + (S->yaw[back] - 2 * M_PI) * ratioBack;
    } else {
      c.yaw = S->yaw[front] * ratioFront + S->yaw[back] * ratioBack;
    }
    c.vx = S->velo[0][front] * ratioFront + S->velo[0][back] * ratioBack;
    c.vy = S->velo[1][front] * ratioFront + S->velo[1][back] * ratioBack;
    c.vz = S->velo[2][front] * ratioFront + S->velo[2][back] * ratioBack;
    c.sx = S->shift[0][front] * ratioFront + S->shift[0][back] * ratioBack;
    c.sy = S->shift[1][front] * ratioFront + S->shift[1][back] * ratioBack;
    c.sz = S->shift[2][front] * ratioFront + S->shift[2][back] * ratioBack;
  }
  return c;
}

// K-D1: per-block max of the per-point pointer advance
__global__ void __launch_bounds__(SR_BLK) sr_imu_steps_kernel(int n, const short* __restrict__ ring,
                                                              const float* __restrict__ reltime, double timeScanCur,
                                                              const ImuSnapshot* __restrict__ S, const ImuState* __restrict__ st_in,
                                                              unsigned char* __restrict__ steps, int* __restrict__ blockmax) {
  __shared__ double s_time[IMU_Q];
  __shared__ int s_max;
  for (int k = threadIdx.x; k < IMU_Q; k += SR_BLK) s_time[k] = S->time[k];
  if (threadIdx.x == 0) s_max = 0;
  __syncthreads();
  int front0 = st_in->pointer_front;
  int L = (S->pointer_last - front0 + IMU_Q) % IMU_Q;
  int i = blockIdx.x * SR_BLK + threadIdx.x;
  int f = 0;
  if (i < n && ring[i] >= 0) {
    const double scanPeriod = 0.1;
    float pointTime = reltime[i] * scanPeriod;
    f = sr_imu_steps(timeScanCur + pointTime, s_time, front0, L);
  }
  if (i < n) steps[i] = (unsigned char)f;
  int wm = __reduce_max_sync(0xffffffffu, f);
  if (lane_id() == 0) atomicMax(&s_max, wm);
  __syncthreads();
  if (threadIdx.x == 0) blockmax[blockIdx.x] = s_max;
}

// K-D2: running max -> slot, interpolate, latch *Start on raw point `first` (i == 0 after NaN
// removal, :332-343), otherwise ShiftToStartIMU / VeloToStartIMU / TransformToStartIMU.
__global__ void __launch_bounds__(SR_BLK) sr_imu_deskew(int n, int first, const short* __restrict__ ring,
                                                        const float* __restrict__ reltime, const int* __restrict__ pos,
                                                        double timeScanCur, const ImuSnapshot* __restrict__ S,
                                                        const ImuState* __restrict__ st_in, ImuState* __restrict__ st_out,
                                                        const unsigned char* __restrict__ steps, const int* __restrict__ blockmax,
                                                        int nblk, float4* __restrict__ full, int max_points,
                                                        const SRCounts* __restrict__ cnt) {
  __shared__ int s_base;
  __shared__ int s_wmax[SR_BLK / 32];
  __shared__ ImuCur s_start;
  const double scanPeriod = 0.1;
  int front0 = st_in->pointer_front;
  if (threadIdx.x < 32) {
    int m = 0;
    for (int b = threadIdx.x; b < blockIdx.x; b += 32) m = max(m, blockmax[b]);
    m = __reduce_max_sync(0xffffffffu, m);
    if (threadIdx.x == 0) s_base = m;
  }
  if (threadIdx.x == 32) {
    // *Start values: latched from raw point `first` when it survives ring filtering, else kept
    if (first < n && ring[first] >= 0) {
      float pt = reltime[first] * scanPeriod;
      int slot = (front0 + (int)steps[first]) % IMU_Q;
      s_start = sr_imu_interp(S, slot, timeScanCur, pt);
    } else {
      ImuCur c;
      c.roll = st_in->rpy_start[0]; c.pitch = st_in->rpy_start[1]; c.yaw = st_in->rpy_start[2];
      c.vx = st_in->velo_start[0]; c.vy = st_in->velo_start[1]; c.vz = st_in->velo_start[2];
      c.sx = st_in->shift_start[0]; c.sy = st_in->shift_start[1]; c.sz = st_in->shift_start[2];
      s_start = c;
    }
  }
  __syncthreads();
  int i = blockIdx.x * SR_BLK + threadIdx.x;
  bool valid = i < n && ring[i] >= 0;
  int f = valid ? (int)steps[i] : 0;
  int x = f;
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane_id() >= o) x = max(x, y);
  }
  if (lane_id() == 31) s_wmax[threadIdx.x >> 5] = x;
  __syncthreads();
  int run = s_base;
  for (int w = 0; w < (int)(threadIdx.x >> 5); w++) run = max(run, s_wmax[w]);
  run = max(run, x);
  ImuCur cur;
  float sfs[3] = {0, 0, 0}, vfs[3] = {0, 0, 0};
  if (valid) {
    int slot = (front0 + run) % IMU_Q;
    float pointTime = reltime[i] * scanPeriod;
    cur = sr_imu_interp(S, slot, timeScanCur, pointTime);
    if (i != first) {
      const ImuCur st = s_start;
      float imuRollStart = st.roll, imuPitchStart = st.pitch, imuYawStart = st.yaw;
      // ShiftToStartIMU :108-125
      float sX = cur.sx - st.sx - st.vx * pointTime;
      float sY = cur.sy - st.sy - st.vy * pointTime;
      float sZ = cur.sz - st.sz - st.vz * pointTime;
      float x1 = dm::cos(imuYawStart) * sX - dm::sin(imuYawStart) * sZ;
      float y1 = sY;
      float z1 = dm::sin(imuYawStart) * sX + dm::cos(imuYawStart) * sZ;
      float x2 = x1;
      float y2 = dm::cos(imuPitchStart) * y1 + dm::sin(imuPitchStart) * z1;
      float z2 = -dm::sin(imuPitchStart) * y1 + dm::cos(imuPitchStart) * z1;
      sfs[0] = dm::cos(imuRollStart) * x2 + dm::sin(imuRollStart) * y2;
      sfs[1] = -dm::sin(imuRollStart) * x2 + dm::cos(imuRollStart) * y2;
      sfs[2] = z2;
      // VeloToStartIMU :127-144
      float vX = cur.vx - st.vx, vY = cur.vy - st.vy, vZ = cur.vz - st.vz;
      x1 = dm::cos(imuYawStart) * vX - dm::sin(imuYawStart) * vZ;
      y1 = vY;
      z1 = dm::sin(imuYawStart) * vX + dm::cos(imuYawStart) * vZ;
      x2 = x1;
      y2 = dm::cos(imuPitchStart) * y1 + dm::sin(imuPitchStart) * z1;
      z2 = -dm::sin(imuPitchStart) * y1 + dm::cos(imuPitchStart) * z1;
      vfs[0] = dm::cos(imuRollStart) * x2 + dm::sin(imuRollStart) * y2;
      vfs[1] = -dm::sin(imuRollStart) * x2 + dm::cos(imuRollStart) * y2;
      vfs[2] = z2;
      // TransformToStartIMU :146-171
      int ps = pos[i];
      if (ps >= 0 && ps < max_points) {
        float4 p = full[ps];
        float imuRollCur = cur.roll, imuPitchCur = cur.pitch, imuYawCur = cur.yaw;
        float a1 = dm::cos(imuRollCur) * p.x - dm::sin(imuRollCur) * p.y;
        float b1 = dm::sin(imuRollCur) * p.x + dm::cos(imuRollCur) * p.y;
        float c1 = p.z;
        float a2 = a1;
        float b2 = dm::cos(imuPitchCur) * b1 - dm::sin(imuPitchCur) * c1;
        float c2 = dm::sin(imuPitchCur) * b1 + dm::cos(imuPitchCur) * c1;
        float a3 = dm::cos(imuYawCur) * a2 + dm::sin(imuYawCur) * c2;
        float b3 = b2;
        float c3 = -dm::sin(imuYawCur) * a2 + dm::cos(imuYawCur) * c2;
        float a4 = dm::cos(imuYawStart) * a3 - dm::sin(imuYawStart) * c3;
        float b4 = b3;
        float c4 = dm::sin(imuYawStart) * a3 + dm::cos(imuYawStart) * c3;
        float a5 = a4;
        float b5 = dm::cos(imuPitchStart) * b4 + dm::sin(imuPitchStart) * c4;
        float c5 = -dm::sin(imuPitchStart) * b4 + dm::cos(imuPitchStart) * c4;
        p.x = dm::cos(imuRollStart) * a5 + dm::sin(imuRollStart) * b5 + sfs[0];
        p.y = -dm::sin(imuRollStart) * a5 + dm::cos(imuRollStart) * b5 + sfs[1];
        p.z = c5 + sfs[2];
        full[ps] = p;
      }
    }
  }
  // the globals exported through imuTrans (:614-629) are those left by the LAST surviving raw
  // point (st_out was pre-initialised with *st_in by the host)
  if (valid && i == cnt->last_valid) {
    {
      ImuState o = *st_in;
      o.pointer_front = (front0 + run) % IMU_Q;
      const ImuCur st = s_start;
      o.rpy_start[0] = st.roll; o.rpy_start[1] = st.pitch; o.rpy_start[2] = st.yaw;
      o.velo_start[0] = st.vx; o.velo_start[1] = st.vy; o.velo_start[2] = st.vz;
      o.shift_start[0] = st.sx; o.shift_start[1] = st.sy; o.shift_start[2] = st.sz;
      o.rpy_cur[0] = cur.roll; o.rpy_cur[1] = cur.pitch; o.rpy_cur[2] = cur.yaw;
      o.velo_cur[0] = cur.vx; o.velo_cur[1] = cur.vy; o.velo_cur[2] = cur.vz;
      o.shift_cur[0] = cur.sx; o.shift_cur[1] = cur.sy; o.shift_cur[2] = cur.sz;
      if (i != first) {
        for (int a = 0; a < 3; a++) { o.shift_from_start[a] = sfs[a]; o.velo_from_start[a] = vfs[a]; }
      }
      *st_out = o;
    }
  }
}

}  // namespace loamgpu
