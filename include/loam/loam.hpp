// loam.hpp -- header-only C++ facade over the C ABI (include/loam_b200.h): the node-level operator
// surface of Maofei/loam_velodyne-1 with ROS stripped.  One class per reference node:
//   loam::ScanRegistration     src/scanRegistration.cpp   (imuHandler :638-660, laserCloudHandler :211-636)
//   loam::LaserOdometry        src/laserOdometry.cpp      (main loop body :410-931)
//   loam::LaserMapping         src/laserMapping.cpp       (imuHandler :323-335, main loop body :408-1096)
//   loam::TransformMaintenance src/transformMaintenance.cpp (handlers :147-203), host only
//   loam::Pipeline             the four wired as the reference's topic graph on one CUDA stream
// Sweeps come in as pcl::PointCloud<pcl::PointXYZI> (or XYZ) in the sensor frame, firing order;
// feature clouds and poses come out with the reference's conventions (intensity = ring +
// 0.1 * relTime; transform[0..2] = rx, ry, rz, [3..5] = tx, ty, tz).
// Return values: LOAM_OK (0), LOAM_NOT_READY (1: the reference would return / continue without
// publishing).  Errors throw loam::Error (nothing throws across the C ABI itself).
#pragma once
#include <stdexcept>
#include <string>
#include <vector>
#include "../loam_b200.h"
#include "pcl_min.h"

namespace loam {

typedef pcl::PointXYZI PointType;                 // include/loam_velodyne/common.h:37
typedef pcl::PointCloud<PointType> Cloud;

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};
inline int check(int rc) {
  if (rc < 0) throw Error(rc, std::string("loam_b200: ") + loam_last_error());
  return rc;
}

struct Params : loam_params {
  // 1 = VLP-16 (the reference's literals), 2 = HDL-64E, 4 = 128-ring (BASELINE.json configs)
  static Params preset(int which, int system_delay = 20) {
    Params p;
    check(loam_params_preset(&p, which));
    p.system_delay = system_delay;
    return p;
  }
};

typedef loam_sweep_result SweepResult;
struct Pose6 { float v[6]; };

namespace detail {
// packed float4 (device layout) -> pcl cloud
inline void unpack(const std::vector<float>& xyzi, int n, Cloud* out) {
  out->points.resize((size_t)n);
  for (int i = 0; i < n; i++) {
    PointType& p = out->points[(size_t)i];
    p.x = xyzi[4 * (size_t)i]; p.y = xyzi[4 * (size_t)i + 1]; p.z = xyzi[4 * (size_t)i + 2];
    p.intensity = xyzi[4 * (size_t)i + 3];
  }
  out->width = (uint32_t)n; out->height = 1; out->is_dense = true;
}
// pcl cloud (32-byte points) -> packed float4 x,y,z,intensity
inline void pack(const Cloud& in, std::vector<float>* xyzi) {
  xyzi->resize(4 * in.points.size() + 4);
  for (size_t i = 0; i < in.points.size(); i++) {
    const PointType& p = in.points[i];
    (*xyzi)[4 * i] = p.x; (*xyzi)[4 * i + 1] = p.y; (*xyzi)[4 * i + 2] = p.z; (*xyzi)[4 * i + 3] = p.intensity;
  }
}
template <typename H, typename F>
inline void download(H* h, F fn, int which, Cloud* out) {
  int n = 0;
  check(fn(h, which, nullptr, 0, &n));
  std::vector<float> buf((size_t)(n > 0 ? n : 1) * 4);
  check(fn(h, which, buf.data(), n, &n));
  unpack(buf, n, out);
}
}  // namespace detail

// the five clouds + imuTrans that scanRegistration publishes (:584-634)
struct Features {
  Cloud full, sharp, less_sharp, flat, less_flat;
  float imu_trans[12];
};

class ScanRegistration {
 public:
  explicit ScanRegistration(const Params& p, int device = 0) : h_(nullptr), owned_(true) { check(loam_sr_create(&p, device, &h_)); }
  ScanRegistration(loam_sr* borrowed) : h_(borrowed), owned_(false) {}
  ~ScanRegistration() { if (owned_ && h_) loam_sr_destroy(h_); }
  ScanRegistration(const ScanRegistration&) = delete;
  ScanRegistration& operator=(const ScanRegistration&) = delete;
  // imuHandler: orientation quaternion (x, y, z, w), linear acceleration
  void imu(double stamp, const double quat[4], const float acc[3]) { check(loam_sr_imu(h_, stamp, quat, acc)); }
  // laserCloudHandler.  Works for any PCL point type that starts with float x, y, z.
  template <typename PointT>
  int process(const pcl::PointCloud<PointT>& sweep, double stamp) {
    if (sweep.points.empty()) return LOAM_NOT_READY;
    return check(loam_sr_process(h_, &sweep.points[0].x, (int)sweep.points.size(), (int)sizeof(PointT), stamp));
  }
  template <typename PointT>
  int process(const pcl::PointCloud<PointT>& sweep, double stamp, Features* out) {
    int rc = process(sweep, stamp);
    if (rc == LOAM_OK && out) features(out);
    return rc;
  }
  // the same handler fed with the raw VLP-16 data packets of one revolution (n_packets x 1206 bytes)
  int processPackets(const unsigned char* packets, int n_packets, double stamp, Features* out = nullptr) {
    int rc = check(loam_sr_process_packets(h_, packets, n_packets, stamp));
    if (rc == LOAM_OK && out) features(out);
    return rc;
  }
  void cloud(int which, Cloud* out) { detail::download(h_, loam_sr_download, which, out); }
  void features(Features* out) {
    cloud(LOAM_CLOUD_FULL, &out->full); cloud(LOAM_CLOUD_SHARP, &out->sharp); cloud(LOAM_CLOUD_LESS_SHARP, &out->less_sharp);
    cloud(LOAM_CLOUD_FLAT, &out->flat); cloud(LOAM_CLOUD_LESS_FLAT, &out->less_flat);
    check(loam_sr_imu_trans(h_, out->imu_trans));
  }
  loam_sr* handle() { return h_; }
 private:
  loam_sr* h_;
  bool owned_;
};

struct OdomResult {
  float transform[6];       // laserOdometry.cpp:93
  float transform_sum[6];   // :94, /laser_odom_to_init
  int iterations, rows, degenerate;
  bool has_mapping_input;   // clouds published to laserMapping this sweep (:910-930)
};

class LaserOdometry {
 public:
  explicit LaserOdometry(const Params& p, int device = 0) : h_(nullptr), owned_(true) { check(loam_lo_create(&p, device, &h_)); }
  LaserOdometry(loam_lo* borrowed) : h_(borrowed), owned_(false) {}
  ~LaserOdometry() { if (owned_ && h_) loam_lo_destroy(h_); }
  LaserOdometry(const LaserOdometry&) = delete;
  LaserOdometry& operator=(const LaserOdometry&) = delete;
  // consumes the features the given ScanRegistration produced last (same device)
  int process(ScanRegistration& sr, OdomResult* out) {
    int pub = 0;
    int rc = check(loam_lo_process(h_, sr.handle(), &pub));
    if (out) {
      int st[4] = {0, 0, 0, 0};
      check(loam_lo_state(h_, out->transform, out->transform_sum, st));
      out->iterations = st[0]; out->rows = st[1]; out->degenerate = st[2];
      out->has_mapping_input = pub != 0;
    }
    return rc;
  }
  // the same pass on features held by the caller (LaserOdometry::process(const Features&, OdomResult*))
  int process(const Features& f, OdomResult* out) {
    std::vector<float> c[5];
    detail::pack(f.full, &c[0]); detail::pack(f.sharp, &c[1]); detail::pack(f.less_sharp, &c[2]); detail::pack(f.flat, &c[3]); detail::pack(f.less_flat, &c[4]);
    int pub = 0;
    int rc = check(loam_lo_process_host(h_, c[0].data(), (int)f.full.points.size(), c[1].data(), (int)f.sharp.points.size(),
                                        c[2].data(), (int)f.less_sharp.points.size(), c[3].data(), (int)f.flat.points.size(),
                                        c[4].data(), (int)f.less_flat.points.size(), f.imu_trans, &pub));
    if (out) {
      int st[4] = {0, 0, 0, 0};
      check(loam_lo_state(h_, out->transform, out->transform_sum, st));
      out->iterations = st[0]; out->rows = st[1]; out->degenerate = st[2];
      out->has_mapping_input = pub != 0;
    }
    return rc;
  }
  void cloud(int which, Cloud* out) { detail::download(h_, loam_lo_download, which, out); }
  loam_lo* handle() { return h_; }
 private:
  loam_lo* h_;
  bool owned_;
};

struct MapResult {
  float transform_aft_mapped[6];   // laserMapping.cpp:100, /aft_mapped_to_init
  float transform_bef_mapped[6];   // :99
  int iterations, rows, degenerate;
};

class LaserMapping {
 public:
  explicit LaserMapping(const Params& p, int device = 0) : h_(nullptr), owned_(true) { check(loam_lm_create(&p, device, &h_)); }
  LaserMapping(loam_lm* borrowed) : h_(borrowed), owned_(false) {}
  ~LaserMapping() { if (owned_ && h_) loam_lm_destroy(h_); }
  LaserMapping(const LaserMapping&) = delete;
  LaserMapping& operator=(const LaserMapping&) = delete;
  void imu(double stamp, const double quat[4]) { check(loam_lm_imu(h_, stamp, quat)); }
  int process(LaserOdometry& lo, double stamp, MapResult* out) {
    int rc = check(loam_lm_process(h_, lo.handle(), stamp));
    if (rc == LOAM_OK && out) {
      float tobe[6];
      int st[8];
      check(loam_lm_state(h_, out->transform_aft_mapped, out->transform_bef_mapped, tobe, st));
      out->iterations = st[0]; out->rows = st[1]; out->degenerate = st[2];
    }
    return rc;
  }
  // the same pass on laserOdometry output held by the caller (recorded / replayed / produced elsewhere):
  // the three clouds of /laser_cloud_corner_last, /laser_cloud_surf_last, /velodyne_cloud_3 and transformSum
  int process(const Cloud& corner_last, const Cloud& surf_last, const Cloud& full_res, const float transform_sum[6], double stamp, MapResult* out) {
    std::vector<float> c, s, f;
    detail::pack(corner_last, &c); detail::pack(surf_last, &s); detail::pack(full_res, &f);
    int rc = check(loam_lm_process_host(h_, c.data(), (int)corner_last.points.size(), s.data(), (int)surf_last.points.size(),
                                        f.data(), (int)full_res.points.size(), transform_sum, stamp));
    if (rc == LOAM_OK && out) {
      float tobe[6];
      int st[8];
      check(loam_lm_state(h_, out->transform_aft_mapped, out->transform_bef_mapped, tobe, st));
      out->iterations = st[0]; out->rows = st[1]; out->degenerate = st[2];
    }
    return rc;
  }
  // map persistence: checkpoint / resume of the cube map, window and poses; PCD export of a cloud
  // (which = LOAM_CLOUD_* of this node, or 20 / 21 = the whole corner / surf map)
  void saveMap(const std::string& path) { check(loam_lm_save_map(h_, path.c_str())); }
  void loadMap(const std::string& path) { check(loam_lm_load_map(h_, path.c_str())); }
  void exportPCD(int which, const std::string& path) { check(loam_lm_export_pcd(h_, which, path.c_str())); }
  void cloud(int which, Cloud* out) { detail::download(h_, loam_lm_download, which, out); }
  loam_lm* handle() { return h_; }
 private:
  loam_lm* h_;
  bool owned_;
};

class TransformMaintenance {
 public:
  TransformMaintenance() : h_(nullptr) { check(loam_tm_create(&h_)); }
  ~TransformMaintenance() { if (h_) loam_tm_destroy(h_); }
  TransformMaintenance(const TransformMaintenance&) = delete;
  TransformMaintenance& operator=(const TransformMaintenance&) = delete;
  void onMapping(const float aft_mapped[6], const float bef_mapped[6]) { check(loam_tm_on_mapping(h_, aft_mapped, bef_mapped)); }
  Pose6 onOdometry(const float transform_sum[6]) {
    Pose6 p;
    check(loam_tm_on_odometry(h_, transform_sum, p.v));
    return p;
  }
 private:
  loam_tm* h_;
};

// sr -> lo -> tm -> (lm -> tm) on one device / one CUDA stream; push() is the synchronous form,
// pushAsync() + collect() keep the device queue full (no host round trip per sweep).
class Pipeline {
 public:
  explicit Pipeline(const Params& p, int device = 0) : h_(nullptr) { check(loam_pipeline_create(&p, device, &h_)); }
  ~Pipeline() { if (h_) loam_pipeline_destroy(h_); }
  Pipeline(const Pipeline&) = delete;
  Pipeline& operator=(const Pipeline&) = delete;
  void imu(double stamp, const double quat[4], const float acc[3]) { check(loam_pipeline_imu(h_, stamp, quat, acc)); }
  template <typename PointT>
  int push(const pcl::PointCloud<PointT>& sweep, double stamp, SweepResult* res) {
    if (sweep.points.empty()) return LOAM_NOT_READY;
    return check(loam_pipeline_push(h_, &sweep.points[0].x, (int)sweep.points.size(), (int)sizeof(PointT), stamp, res));
  }
  int push(const float* xyz, size_t n, size_t stride_bytes, double stamp, SweepResult* res) {
    return check(loam_pipeline_push(h_, xyz, (int)n, (int)stride_bytes, stamp, res));
  }
  template <typename PointT>
  int pushAsync(const pcl::PointCloud<PointT>& sweep, double stamp) {
    if (sweep.points.empty()) return LOAM_NOT_READY;
    return check(loam_pipeline_push_async(h_, &sweep.points[0].x, (int)sweep.points.size(), (int)sizeof(PointT), stamp));
  }
  int pushPackets(const unsigned char* packets, int n_packets, double stamp, SweepResult* res) {
    return check(loam_pipeline_push_packets(h_, packets, n_packets, stamp, res));
  }
  int pushPacketsAsync(const unsigned char* packets, int n_packets, double stamp) {
    return check(loam_pipeline_push_packets_async(h_, packets, n_packets, stamp));
  }
  std::vector<SweepResult> collect() {
    std::vector<SweepResult> out(1024);
    int n = 0;
    check(loam_pipeline_collect(h_, out.data(), (int)out.size(), &n));
    out.resize((size_t)(n < 1024 ? n : 1024));
    return out;
  }
  ScanRegistration registration() { return ScanRegistration(loam_pipeline_sr(h_)); }
  LaserOdometry odometry() { return LaserOdometry(loam_pipeline_lo(h_)); }
  LaserMapping mapping() { return LaserMapping(loam_pipeline_lm(h_)); }
 private:
  loam_pipeline* h_;
};

}  // namespace loam
