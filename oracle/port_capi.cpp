// ORACLE (test infrastructure).  C interface ("checker API") over the CPU restatement in
// loam_port.hpp so that tests/ and bench.py's cpu_baseline leg can drive it through ctypes.
// oracle/_ref (the reference's own sources compiled from /root/reference) exports the SAME
// symbols, so either can serve as the checker.
#include <chrono>
#include <cstdio>
#include "loam_port.hpp"
#include "packet_port.hpp"
#include "checker_api.h"

using namespace oracle;

struct chk_handle {
  loam_params P;
  ScanRegistration sr;
  LaserOdometry lo;
  LaserMapping lm;
  TransformMaintenance tm;
  SRFeatures feat;
  LOOutput odom;
  loam_sweep_result res;
  double busy[4];
  chk_handle(const loam_params& p) : P(p), sr(p), lo(p), lm(p) { memset(&res, 0, sizeof(res)); busy[0] = busy[1] = busy[2] = busy[3] = 0; }
};

static int copy_cloud(const Cloud& c, float* out, int cap) {
  int n = (int)c.points.size();
  for (int i = 0; i < n && i < cap; i++) {
    out[4 * i] = c.points[i].x; out[4 * i + 1] = c.points[i].y; out[4 * i + 2] = c.points[i].z; out[4 * i + 3] = c.points[i].intensity;
  }
  return n;
}
static int copy_ints(const std::vector<int>& v, int* out, int cap) {
  for (int i = 0; i < (int)v.size() && i < cap; i++) out[i] = v[i];
  return (int)v.size();
}
static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

extern "C" {

const char* chk_kind(void) {
#ifdef ORACLE_LIBM
  return "port-libm";
#else
  return "port";
#endif
}

void* chk_create(const loam_params* p) { return new chk_handle(*p); }
void chk_destroy(void* hv) { delete (chk_handle*)hv; }

int chk_imu(void* hv, double stamp, const double quat[4], const float acc[3]) {
  chk_handle* h = (chk_handle*)hv;
  h->sr.imu(stamp, quat, acc);
  h->lm.imu(stamp, quat);
  return 0;
}

// laserOdometry -> transformMaintenance -> (laserMapping -> transformMaintenance) on h->feat
static void run_from_features(chk_handle* h, double t1) {
  loam_sweep_result& r = h->res;
  {
    r.sr_status = LOAM_OK;
    r.n_full = (int)h->feat.full.size(); r.n_sharp = (int)h->feat.sharp.size(); r.n_less_sharp = (int)h->feat.lessSharp.size();
    r.n_flat = (int)h->feat.flat.size(); r.n_less_flat = (int)h->feat.lessFlat.size();
    memcpy(r.imu_trans, h->feat.imuTrans, sizeof(r.imu_trans));
    bool lok = h->lo.process(h->feat, h->odom);
    double t2 = now_s();
    h->busy[1] += t2 - t1;
    r.lo_iters = h->lo.lastIters; r.lo_rows = h->lo.lastRows; r.lo_degenerate = h->lo.isDegenerate;
    memcpy(r.transform, h->lo.transform, sizeof(r.transform));
    memcpy(r.transform_sum, h->lo.transformSum, sizeof(r.transform_sum));
    if (lok) {
      r.lo_status = LOAM_OK;
      float sumMsg[6];
      odometry_msg_roundtrip(h->odom.transformSum, sumMsg);
      h->tm.onOdometry(sumMsg, r.transform_mapped);
      double t3 = now_s();
      h->busy[3] += t3 - t2;
      if (h->odom.hasMappingInput) {
        bool mok = h->lm.process(h->odom, sumMsg);
        double t4 = now_s();
        h->busy[2] += t4 - t3;
        if (mok) {
          r.lm_status = LOAM_OK;
          r.lm_iters = h->lm.lastIters; r.lm_rows = h->lm.lastRows; r.lm_degenerate = h->lm.isDegenerate;
          r.n_map_corner = h->lm.lastCornerFromMapNum; r.n_map_surf = h->lm.lastSurfFromMapNum;
          r.n_stack_corner = h->lm.lastCornerStackNum; r.n_stack_surf = h->lm.lastSurfStackNum;
          memcpy(r.transform_aft_mapped, h->lm.transformAftMapped, sizeof(float) * 6);
          memcpy(r.transform_bef_mapped, h->lm.transformBefMapped, sizeof(float) * 6);
          float aftMsg[6];
          aft_mapped_msg_roundtrip(h->lm.transformAftMapped, aftMsg);
          h->tm.onMapping(aftMsg, h->lm.transformBefMapped);
        }
      }
    }
  }
}

int chk_push_sweep(void* hv, const float* xyz, int n, double stamp, loam_sweep_result* out) {
  chk_handle* h = (chk_handle*)hv;
  loam_sweep_result& r = h->res;
  r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
  double t0 = now_s();
  bool ok = h->sr.process(xyz, n, 3, stamp, h->feat);
  double t1 = now_s();
  h->busy[0] += t1 - t0;
  if (ok) run_from_features(h, t1);
  if (out) *out = r;
  return r.sr_status;
}

// the same chain fed with HOST copies of scanRegistration's six topics (x,y,z,intensity per point;
// imu_trans = the 12 floats of /imu_trans, NULL = zeros): what laserOdometry.cpp:281-381 receives
int chk_lo_push(void* hv, const float* full, int n_full, const float* sharp, int n_sharp, const float* less_sharp, int n_less_sharp,
                const float* flat, int n_flat, const float* less_flat, int n_less_flat, const float imu_trans[12], double stamp,
                loam_sweep_result* out) {
  chk_handle* h = (chk_handle*)hv;
  loam_sweep_result& r = h->res;
  r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
  auto fill = [](Cloud& c, const float* p, int n) {
    c.points.resize(n);
    for (int i = 0; i < n; i++) { c.points[i].x = p[4 * i]; c.points[i].y = p[4 * i + 1]; c.points[i].z = p[4 * i + 2]; c.points[i].intensity = p[4 * i + 3]; }
    c.width = n; c.height = 1;
  };
  SRFeatures& f = h->feat;
  fill(f.full, full, n_full); fill(f.sharp, sharp, n_sharp); fill(f.lessSharp, less_sharp, n_less_sharp);
  fill(f.flat, flat, n_flat); fill(f.lessFlat, less_flat, n_less_flat);
  for (int i = 0; i < 12; i++) f.imuTrans[i] = imu_trans ? imu_trans[i] : 0.f;
  f.sharpIdx.clear(); f.lessSharpIdx.clear(); f.flatIdx.clear();
  f.stamp = stamp;
  run_from_features(h, now_s());
  if (out) *out = r;
  return r.sr_status;
}


int chk_get_cloud(void* hv, int which, float* out, int cap) {
  chk_handle* h = (chk_handle*)hv;
  switch (which) {
    case LOAM_CLOUD_FULL: return copy_cloud(h->feat.full, out, cap);
    case LOAM_CLOUD_SHARP: return copy_cloud(h->feat.sharp, out, cap);
    case LOAM_CLOUD_LESS_SHARP: return copy_cloud(h->feat.lessSharp, out, cap);
    case LOAM_CLOUD_FLAT: return copy_cloud(h->feat.flat, out, cap);
    case LOAM_CLOUD_LESS_FLAT: return copy_cloud(h->feat.lessFlat, out, cap);
    case LOAM_CLOUD_CORNER_LAST: return copy_cloud(*h->lo.laserCloudCornerLast, out, cap);
    case LOAM_CLOUD_SURF_LAST: return copy_cloud(*h->lo.laserCloudSurfLast, out, cap);
    case LOAM_CLOUD_FULL_END: return copy_cloud(h->odom.fullRes, out, cap);
    case LOAM_CLOUD_REGISTERED: return copy_cloud(*h->lm.laserCloudFullRes, out, cap);
    case LOAM_CLOUD_SURROUND: return copy_cloud(*h->lm.laserCloudSurround, out, cap);
    case LOAM_CLOUD_MAP_CORNER: return copy_cloud(*h->lm.laserCloudCornerFromMap, out, cap);
    case LOAM_CLOUD_MAP_SURF: return copy_cloud(*h->lm.laserCloudSurfFromMap, out, cap);
    case LOAM_CLOUD_STACK_CORNER: return copy_cloud(*h->lm.laserCloudCornerStack, out, cap);
    case LOAM_CLOUD_STACK_SURF: return copy_cloud(*h->lm.laserCloudSurfStack, out, cap);
  }
  return -1;
}

int chk_get_ints(void* hv, int which, int* out, int cap) {
  chk_handle* h = (chk_handle*)hv;
  int nfull = (int)h->feat.full.size();
  auto arr = [&](const std::vector<int>& v, int n) { for (int i = 0; i < n && i < cap; i++) out[i] = v[i]; return n; };
  auto farr = [&](const std::vector<float>& v, int n) { for (int i = 0; i < n && i < cap; i++) out[i] = (int)v[i]; return n; };
  switch (which) {
    case LOAM_INTS_LABEL: return arr(h->sr.cloudLabel, nfull);
    case LOAM_INTS_PICKED: return arr(h->sr.cloudNeighborPicked, nfull);
    case LOAM_INTS_SORT_IND: return arr(h->sr.cloudSortInd, nfull);
    case LOAM_INTS_SHARP_IDX: return copy_ints(h->feat.sharpIdx, out, cap);
    case LOAM_INTS_LESS_SHARP_IDX: return copy_ints(h->feat.lessSharpIdx, out, cap);
    case LOAM_INTS_FLAT_IDX: return copy_ints(h->feat.flatIdx, out, cap);
    case LOAM_INTS_CORNER_IND1: return farr(h->lo.pointSearchCornerInd1, (int)h->feat.sharp.size());
    case LOAM_INTS_CORNER_IND2: return farr(h->lo.pointSearchCornerInd2, (int)h->feat.sharp.size());
    case LOAM_INTS_SURF_IND1: return farr(h->lo.pointSearchSurfInd1, (int)h->feat.flat.size());
    case LOAM_INTS_SURF_IND2: return farr(h->lo.pointSearchSurfInd2, (int)h->feat.flat.size());
    case LOAM_INTS_SURF_IND3: return farr(h->lo.pointSearchSurfInd3, (int)h->feat.flat.size());
    case LOAM_INTS_SCAN_START: return arr(h->sr.scanStartInd, (int)h->sr.scanStartInd.size());
    case LOAM_INTS_SCAN_END: return arr(h->sr.scanEndInd, (int)h->sr.scanEndInd.size());
    case 100: return copy_ints(h->lm.cornerNN, out, cap);  // LM 5-NN indices of the last iteration
    case 101: return copy_ints(h->lm.surfNN, out, cap);
  }
  return -1;
}

int chk_get_floats(void* hv, int which, float* out, int cap) {
  chk_handle* h = (chk_handle*)hv;
  if (which == LOAM_FLOATS_CURVATURE) {
    int n = (int)h->feat.full.size();
    for (int i = 0; i < n && i < cap; i++) out[i] = h->sr.cloudCurvature[i];
    return n;
  }
  if (which == 1 || which == 2) {  // matP of laserOdometry (:405) / laserMapping (:386)
    const cv::Mat& P = which == 1 ? h->lo.matP : h->lm.matP;
    for (int i = 0; i < 36 && i < cap; i++) out[i] = P.at<float>(i / 6, i % 6);
    return 36;
  }
  return -1;
}

// per-node busy seconds accumulated so far: sr, lo, lm, tm
void chk_busy_seconds(void* hv, double out[4]) { memcpy(out, ((chk_handle*)hv)->busy, sizeof(double) * 4); }

// teacher forcing (parity tests): overwrite laserOdometry's transform / transformSum
void chk_lo_set_state(void* hv, const float transform[6], const float transformSum[6]) {
  chk_handle* h = (chk_handle*)hv;
  memcpy(h->lo.transform, transform, sizeof(float) * 6);
  memcpy(h->lo.transformSum, transformSum, sizeof(float) * 6);
}


// the solve step of one LM iteration on given normal equations: cv::solve(QR), and -- as at
// iterCount == 0 -- cv::eigen, the threshold scan and matP = matV.inv() * matV2, then the projection
// (laserOdometry.cpp:765-797 with thr 10, laserMapping.cpp:922-954 with thr 100).  `batch` systems:
// AtA [batch][36], AtB [batch][6] -> isDegenerate [batch], matP [batch][36], X [batch][6]
void chk_solve_step(const float* AtA, const float* AtB, int batch, float thr, int* isDegenerate, float* matP36, float* X6) {
  for (int t = 0; t < batch; t++) {
    cv::Mat matAtA(6, 6, CV_32F, cv::Scalar::all(0)), matAtB(6, 1, CV_32F, cv::Scalar::all(0)), matX(6, 1, CV_32F, cv::Scalar::all(0));
    cv::Mat matP(6, 6, CV_32F, cv::Scalar::all(0));
    for (int i = 0; i < 6; i++) {
      for (int j = 0; j < 6; j++) matAtA.at<float>(i, j) = AtA[t * 36 + i * 6 + j];
      matAtB.at<float>(i, 0) = AtB[t * 6 + i];
    }
    cv::solve(matAtA, matAtB, matX, cv::DECOMP_QR);
    cv::Mat matE(1, 6, CV_32F, cv::Scalar::all(0));
    cv::Mat matV(6, 6, CV_32F, cv::Scalar::all(0));
    cv::Mat matV2(6, 6, CV_32F, cv::Scalar::all(0));
    cv::eigen(matAtA, matE, matV);
    matV.copyTo(matV2);
    bool deg = false;
    for (int i = 5; i >= 0; i--) {
      if (matE.at<float>(0, i) < thr) {
        for (int j = 0; j < 6; j++) matV2.at<float>(i, j) = 0;
        deg = true;
      } else {
        break;
      }
    }
    matP = matV.inv() * matV2;
    if (deg) {
      cv::Mat matX2(6, 1, CV_32F, cv::Scalar::all(0));
      matX.copyTo(matX2);
      matX = matP * matX2;
    }
    isDegenerate[t] = deg ? 1 : 0;
    for (int i = 0; i < 6; i++) {
      for (int j = 0; j < 6; j++) matP36[t * 36 + i * 6 + j] = matP.at<float>(i, j);
      X6[t * 6 + i] = matX.at<float>(i, 0);
    }
  }
}

// ---- laserMapping alone, teacher-forced (BASELINE.json configs[2]) and its map file ----------
// one loop-body pass of laserMapping.cpp:408-1096 fed with recorded copies of the four topics;
// transformSum is laserOdometry's value BEFORE the message encoding (the round trip of :304-318 is applied here)
int chk_lm_push(void* hv, const float* corner, int nc, const float* surf, int ns, const float* full, int nf,
                const float transformSum[6], double stamp, loam_sweep_result* out) {
  chk_handle* h = (chk_handle*)hv;
  loam_sweep_result& r = h->res;
  memset(&r, 0, sizeof(r));
  r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
  LOOutput& o = h->odom;
  auto fill = [](Cloud& c, const float* p, int n) {
    c.points.resize(n);
    for (int i = 0; i < n; i++) { c.points[i].x = p[4 * i]; c.points[i].y = p[4 * i + 1]; c.points[i].z = p[4 * i + 2]; c.points[i].intensity = p[4 * i + 3]; }
    c.width = n; c.height = 1;
  };
  fill(o.cornerLast, corner, nc); fill(o.surfLast, surf, ns); fill(o.fullRes, full, nf);
  memcpy(o.transformSum, transformSum, sizeof(float) * 6);
  o.hasMappingInput = true;
  o.stamp = stamp;
  float sumMsg[6];
  odometry_msg_roundtrip(o.transformSum, sumMsg);
  double t0 = now_s();
  bool mok = h->lm.process(o, sumMsg);
  h->busy[2] += now_s() - t0;
  memcpy(r.transform_sum, transformSum, sizeof(float) * 6);
  if (mok) {
    r.lm_status = LOAM_OK;
    r.lm_iters = h->lm.lastIters; r.lm_rows = h->lm.lastRows; r.lm_degenerate = h->lm.isDegenerate;
    r.n_map_corner = h->lm.lastCornerFromMapNum; r.n_map_surf = h->lm.lastSurfFromMapNum;
    r.n_stack_corner = h->lm.lastCornerStackNum; r.n_stack_surf = h->lm.lastSurfStackNum;
    memcpy(r.transform_aft_mapped, h->lm.transformAftMapped, sizeof(float) * 6);
    memcpy(r.transform_bef_mapped, h->lm.transformBefMapped, sizeof(float) * 6);
  }
  if (out) *out = r;
  return r.lm_status;
}

// "LOAMMAP1" (layout stated in loam_velodyne-1_b200/csrc/lm_persist.cuh; this is an independent
// reader / writer over the oracle's own containers).  The oracle makes no claim about voxel order,
// so it writes dirty = 1 for every non-empty cube and ignores the flag on load.
namespace {
struct MapHdr {
  char magic[8]; int version; int dims[3]; int cen[3]; int frameCount, mapFrameCount; int isDegenerate;
  float tf[5][6]; int imuPointerFront, imuPointerLast, hasImu, imuQueLength; long long total[2];
};
}
int chk_lm_save_map(void* hv, const char* path) {
  LaserMapping& m = ((chk_handle*)hv)->lm;
  MapHdr H;
  memset(&H, 0, sizeof(H));
  memcpy(H.magic, "LOAMMAP1", 8);
  H.version = 1;
  H.dims[0] = LaserMapping::laserCloudWidth; H.dims[1] = LaserMapping::laserCloudHeight; H.dims[2] = LaserMapping::laserCloudDepth;
  H.cen[0] = m.laserCloudCenWidth; H.cen[1] = m.laserCloudCenHeight; H.cen[2] = m.laserCloudCenDepth;
  H.frameCount = m.frameCount; H.mapFrameCount = m.mapFrameCount; H.isDegenerate = m.isDegenerate ? 1 : 0;
  const float* tfs[5] = {m.transformSum, m.transformIncre, m.transformTobeMapped, m.transformBefMapped, m.transformAftMapped};
  for (int k = 0; k < 5; k++) memcpy(H.tf[k], tfs[k], sizeof(float) * 6);
  H.imuPointerFront = m.imuPointerFront; H.imuPointerLast = m.imuPointerLast; H.hasImu = m.imuPointerLast >= 0; H.imuQueLength = LaserMapping::imuQueLength;
  const std::vector<Cloud::Ptr>* arr[2] = {&m.laserCloudCornerArray, &m.laserCloudSurfArray};
  for (int T = 0; T < 2; T++)
    for (int c = 0; c < LaserMapping::laserCloudNum; c++) H.total[T] += (long long)(*arr[T])[c]->points.size();
  FILE* f = fopen(path, "wb");
  if (!f) return -1;
  fwrite(&H, sizeof(H), 1, f);
  fwrite(m.imuTime, sizeof(double), LaserMapping::imuQueLength, f);
  fwrite(m.imuRoll, sizeof(float), LaserMapping::imuQueLength, f);
  fwrite(m.imuPitch, sizeof(float), LaserMapping::imuQueLength, f);
  for (int T = 0; T < 2; T++) {
    std::vector<int> cnt(LaserMapping::laserCloudNum);
    std::vector<unsigned char> dirty(LaserMapping::laserCloudNum);
    for (int c = 0; c < LaserMapping::laserCloudNum; c++) { cnt[c] = (int)(*arr[T])[c]->points.size(); dirty[c] = cnt[c] > 0; }
    fwrite(cnt.data(), sizeof(int), cnt.size(), f);
    fwrite(dirty.data(), 1, dirty.size(), f);
    for (int c = 0; c < LaserMapping::laserCloudNum; c++)
      for (const PointType& p : (*arr[T])[c]->points) { float v[4] = {p.x, p.y, p.z, p.intensity}; fwrite(v, sizeof(float), 4, f); }
  }
  return fclose(f) == 0 ? 0 : -1;
}
int chk_lm_load_map(void* hv, const char* path) {
  LaserMapping& m = ((chk_handle*)hv)->lm;
  FILE* f = fopen(path, "rb");
  if (!f) return -1;
  MapHdr H;
  if (fread(&H, sizeof(H), 1, f) != 1 || memcmp(H.magic, "LOAMMAP1", 8) != 0 || H.version != 1 || H.dims[0] != LaserMapping::laserCloudWidth ||
      H.dims[1] != LaserMapping::laserCloudHeight || H.dims[2] != LaserMapping::laserCloudDepth || H.imuQueLength != LaserMapping::imuQueLength) {
    fclose(f);
    return -2;
  }
  bool ok = fread(m.imuTime, sizeof(double), LaserMapping::imuQueLength, f) == (size_t)LaserMapping::imuQueLength;
  ok = ok && fread(m.imuRoll, sizeof(float), LaserMapping::imuQueLength, f) == (size_t)LaserMapping::imuQueLength;
  ok = ok && fread(m.imuPitch, sizeof(float), LaserMapping::imuQueLength, f) == (size_t)LaserMapping::imuQueLength;
  m.laserCloudCenWidth = H.cen[0]; m.laserCloudCenHeight = H.cen[1]; m.laserCloudCenDepth = H.cen[2];
  m.frameCount = H.frameCount; m.mapFrameCount = H.mapFrameCount; m.isDegenerate = H.isDegenerate != 0;
  float* tfs[5] = {m.transformSum, m.transformIncre, m.transformTobeMapped, m.transformBefMapped, m.transformAftMapped};
  for (int k = 0; k < 5; k++) memcpy(tfs[k], H.tf[k], sizeof(float) * 6);
  m.imuPointerFront = H.imuPointerFront; m.imuPointerLast = H.hasImu ? H.imuPointerLast : -1;
  std::vector<Cloud::Ptr>* arr[2] = {&m.laserCloudCornerArray, &m.laserCloudSurfArray};
  for (int T = 0; T < 2 && ok; T++) {
    std::vector<int> cnt(LaserMapping::laserCloudNum);
    std::vector<unsigned char> dirty(LaserMapping::laserCloudNum);
    ok = ok && fread(cnt.data(), sizeof(int), cnt.size(), f) == cnt.size();
    ok = ok && fread(dirty.data(), 1, dirty.size(), f) == dirty.size();
    for (int c = 0; c < LaserMapping::laserCloudNum && ok; c++) {
      Cloud& cl = *(*arr[T])[c];
      cl.points.resize(cnt[c] > 0 ? cnt[c] : 0);
      for (int i = 0; i < cnt[c] && ok; i++) {
        float v[4];
        ok = fread(v, sizeof(float), 4, f) == 4;
        cl.points[i].x = v[0]; cl.points[i].y = v[1]; cl.points[i].z = v[2]; cl.points[i].intensity = v[3];
      }
      cl.width = (uint32_t)cl.points.size(); cl.height = 1;
    }
  }
  fclose(f);
  return ok ? 0 : -3;
}


// The reference runs its nodes as separate processes joined by topic queues: sweep k+1 is registered
// while sweep k is in the odometry and an earlier one in the mapping.  Same structure with one thread
// per node and queues of depth 2 (the reference's subscriber queue size) between them; a full queue
// blocks the producer instead of dropping the message, so every sweep is processed.  IMU-free
// sequences only.  xyz: n_sweeps arrays back to back, counts[k] points each.  poses: n_sweeps x 6
// (/integrated_to_init).  Returns the wall-clock seconds of the run.
}  // extern "C"
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
namespace {
template <class T> struct BoundedQueue {
  std::deque<T> q; std::mutex m; std::condition_variable cv; size_t cap = 2; bool closed = false;
  void push(T&& v) { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return q.size() < cap; }); q.push_back(std::move(v)); cv.notify_all(); }
  bool pop(T& v) { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return !q.empty() || closed; }); if (q.empty()) return false; v = std::move(q.front()); q.pop_front(); cv.notify_all(); return true; }
  void close() { std::unique_lock<std::mutex> l(m); closed = true; cv.notify_all(); }
};
struct FeatMsg { int k; bool ok; SRFeatures f; };
struct OdomMsg { int k; bool ok; LOOutput o; };
}
extern "C" {
double chk_run_pipelined(void* hv, const float* xyz, const int* counts, const double* stamps, int n_sweeps, float* poses) {
  chk_handle* h = (chk_handle*)hv;
  BoundedQueue<FeatMsg> q1;
  BoundedQueue<OdomMsg> q2;
  std::mutex tm_mu;   // transformMaintenance receives from both nodes (its two callbacks never overlap in the reference either)
  auto t0 = std::chrono::steady_clock::now();
  std::thread t_sr([&] {
    size_t off = 0;
    for (int k = 0; k < n_sweeps; k++) {
      FeatMsg m; m.k = k;
      m.ok = h->sr.process(xyz + off * 3, counts[k], 3, stamps[k], m.f);
      off += (size_t)counts[k];
      q1.push(std::move(m));
    }
    q1.close();
  });
  std::thread t_lo([&] {
    FeatMsg m;
    while (q1.pop(m)) {
      if (!m.ok) continue;
      OdomMsg o; o.k = m.k;
      o.ok = h->lo.process(m.f, o.o);
      if (!o.ok) continue;
      float sumMsg[6];
      odometry_msg_roundtrip(o.o.transformSum, sumMsg);
      { std::lock_guard<std::mutex> l(tm_mu); h->tm.onOdometry(sumMsg, poses + 6 * (size_t)m.k); }
      if (o.o.hasMappingInput) q2.push(std::move(o));
    }
    q2.close();
  });
  std::thread t_lm([&] {
    OdomMsg o;
    while (q2.pop(o)) {
      float sumMsg[6];
      odometry_msg_roundtrip(o.o.transformSum, sumMsg);
      if (h->lm.process(o.o, sumMsg)) {
        float aftMsg[6];
        aft_mapped_msg_roundtrip(h->lm.transformAftMapped, aftMsg);
        std::lock_guard<std::mutex> l(tm_mu);
        h->tm.onMapping(aftMsg, h->lm.transformBefMapped);
      }
    }
  });
  t_sr.join(); t_lo.join(); t_lm.join();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// the packet front end of the sensor p describes (VLP-16: 16 rings / ring_model 0, HDL-64E: 64 / 1): decoded xyz and,
// when ring / reltime are given, the packet-metadata SR-1 quantities (packet_port.hpp)
int chk_packet_front_end(const loam_params* p, const unsigned char* packets, int n_packets, float* xyz, int* ring, float* reltime) {
  int kind;
  if (p->n_scans == 16 && p->ring_model == 0) kind = 0;
  else if (p->n_scans == 64 && p->ring_model == 1) kind = 1;
  else return -1;
  if (ring && reltime) {
    sr1_meta_from_packets(packets, n_packets, kind, p->ring_lower_deg, p->ring_upper_deg, xyz, ring, reltime);
  } else {
    for (int q = 0; q < n_packets; q++) {
      if (kind == 0) decode_vlp16_packet(packets + (size_t)1206 * q, xyz + (size_t)384 * 3 * q);
      else decode_hdl64_packet(packets + (size_t)1206 * q, p->ring_lower_deg, p->ring_upper_deg, xyz + (size_t)384 * 3 * q);
    }
  }
  return 0;
}
// one revolution of data packets through the whole chain; packet_metadata as loam_params.packet_metadata
int chk_push_packets(void* hv, const unsigned char* packets, int n_packets, int packet_metadata, double stamp, loam_sweep_result* out) {
  chk_handle* h = (chk_handle*)hv;
  const int n = n_packets * 384;
  std::vector<float> xyz((size_t)n * 3), rt;
  std::vector<int> ring;
  if (packet_metadata) { ring.resize(n); rt.resize(n); }
  if (chk_packet_front_end(&h->P, packets, n_packets, xyz.data(), packet_metadata ? ring.data() : nullptr, packet_metadata ? rt.data() : nullptr) != 0) return -1;
  loam_sweep_result& r = h->res;
  r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
  double t0 = now_s();
  bool ok = h->sr.process(xyz.data(), n, 3, stamp, h->feat, packet_metadata ? ring.data() : nullptr, packet_metadata ? rt.data() : nullptr);
  double t1 = now_s();
  h->busy[0] += t1 - t0;
  if (ok) run_from_features(h, t1);
  if (out) *out = r;
  return r.sr_status;
}
// raw-packet front end: n_packets VLP-16 data packets -> n_packets * 384 points (NaN = no return)
void chk_decode_packets(const unsigned char* packets, int n_packets, float* xyz) {
  for (int p = 0; p < n_packets; p++) decode_vlp16_packet(packets + (size_t)1206 * p, xyz + (size_t)384 * 3 * p);
}

}  // extern "C"

// ---- building blocks exposed for unit tests (oracle/defs/*) ----
extern "C" {
// fn: 0 sin 1 cos 2 atan 3 atan2(y,x) 4 asin 5 sqrt 6 x/y 7 x*y+x ; the library's math (detmath or glibc)
void chk_math(int fn, const float* x, const float* y, float* out, int n) {
  for (int i = 0; i < n; i++) {
    float r = 0.f;
    switch (fn) {
      case 0: r = omath::sin(x[i]); break;
      case 1: r = omath::cos(x[i]); break;
      case 2: r = omath::atan(x[i]); break;
      case 3: r = omath::atan2(y[i], x[i]); break;
      case 4: r = omath::asin(x[i]); break;
      case 5: r = omath::sqrt(x[i]); break;
      case 6: r = x[i] / y[i]; break;
      case 7: r = x[i] * y[i] + x[i]; break;
    }
    out[i] = r;
  }
}
void chk_linalg(int op, const float* A, const float* b, float* out, int m, int n, int batch) {
  for (int t = 0; t < batch; t++) {
    float a[36], bb[8], w[8], v[36];
    for (int i = 0; i < m * n; i++) a[i] = A[(size_t)t * m * n + i];
    if (op == 0) {
      for (int i = 0; i < m; i++) bb[i] = b[(size_t)t * m + i];
      int ok = cvl::qr_solve(a, m, n, bb, 1);
      for (int i = 0; i < n; i++) out[(size_t)t * n + i] = ok ? bb[i] : 0.f;
    } else if (op == 1) {
      cvl::jacobi_eigen(a, w, v, n);
      float* o = out + (size_t)t * (n + n * n);
      for (int i = 0; i < n; i++) o[i] = w[i];
      for (int i = 0; i < n * n; i++) o[n + i] = v[i];
    } else if (op == 2) {
      cvl::lu_inverse(a, v, n);
      for (int i = 0; i < n * n; i++) out[(size_t)t * n * n + i] = v[i];
    }
  }
}
void chk_gemm(const float* A, const float* B, float* C, int m, int k, int n) { cvl::gemm(A, B, C, m, k, n); }
int chk_voxel_filter(const float* xyzi, int n, float leaf, float* out, int cap) {
  Cloud::Ptr in(new Cloud());
  in->points.resize(n);
  for (int i = 0; i < n; i++) { in->points[i].x = xyzi[4 * i]; in->points[i].y = xyzi[4 * i + 1]; in->points[i].z = xyzi[4 * i + 2]; in->points[i].intensity = xyzi[4 * i + 3]; }
  pcl::VoxelGrid<PointType> f;
  f.setLeafSize(leaf, leaf, leaf);
  f.setInputCloud(in);
  Cloud o;
  f.filter(o);
  return copy_cloud(o, out, cap);
}
// k-NN of each query in the cloud through the kd-tree: idx[q*k..], d2[q*k..]
void chk_knn(const float* xyzi, int n, const float* queries, int nq, int k, int* idx, float* d2) {
  Cloud::Ptr in(new Cloud());
  in->points.resize(n);
  for (int i = 0; i < n; i++) { in->points[i].x = xyzi[4 * i]; in->points[i].y = xyzi[4 * i + 1]; in->points[i].z = xyzi[4 * i + 2]; in->points[i].intensity = xyzi[4 * i + 3]; }
  pcl::KdTreeFLANN<PointType> tree;
  tree.setInputCloud(in);
  std::vector<int> ii;
  std::vector<float> dd;
  for (int q = 0; q < nq; q++) {
    PointType p;
    p.x = queries[3 * q]; p.y = queries[3 * q + 1]; p.z = queries[3 * q + 2];
    int got = tree.nearestKSearch(p, k, ii, dd);
    for (int j = 0; j < k; j++) { idx[q * k + j] = j < got ? ii[j] : -1; d2[q * k + j] = j < got ? dd[j] : -1.f; }
  }
}
}
