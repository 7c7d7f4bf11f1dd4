// pcl_min.h -- the two PCL types the LOAM node interfaces exchange, for builds without PCL.
// Layout-compatible with PCL 1.7+: pcl::PointXYZ is 16 bytes {x, y, z, pad}, pcl::PointXYZI is
// 32 bytes {x, y, z, pad, intensity, pad[3]}, both 16-byte aligned; pcl::PointCloud<T> exposes
// points / width / height / is_dense.  When the real PCL headers are present they are used instead.
#pragma once
#if defined(__has_include)
#if __has_include(<pcl/point_cloud.h>) && __has_include(<pcl/point_types.h>) && !defined(LOAM_FORCE_PCL_MIN)
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#define LOAM_HAVE_PCL 1
#endif
#endif

#ifndef LOAM_HAVE_PCL
#include <stddef.h>
#include <stdint.h>
#include <vector>

namespace pcl {

struct alignas(16) PointXYZ {
  float x, y, z, data_pad;
  PointXYZ() : x(0), y(0), z(0), data_pad(1.f) {}
  PointXYZ(float x_, float y_, float z_) : x(x_), y(y_), z(z_), data_pad(1.f) {}
};

struct alignas(16) PointXYZI {
  float x, y, z, data_pad;
  float intensity, pad_[3];
  PointXYZI() : x(0), y(0), z(0), data_pad(1.f), intensity(0), pad_{0, 0, 0} {}
};

template <typename PointT>
class PointCloud {
 public:
  std::vector<PointT> points;
  uint32_t width = 0, height = 0;
  bool is_dense = true;
  size_t size() const { return points.size(); }
  bool empty() const { return points.empty(); }
  void clear() { points.clear(); width = height = 0; }
  void resize(size_t n) { points.resize(n); width = (uint32_t)n; height = 1; }
  void push_back(const PointT& p) { points.push_back(p); width = (uint32_t)points.size(); height = 1; }
  PointT& operator[](size_t i) { return points[i]; }
  const PointT& operator[](size_t i) const { return points[i]; }
};

}  // namespace pcl
#endif
