// ORACLE (test infrastructure, never linked into the product).
// pcl_lite.h -- the subset of PCL 1.7 semantics the reference's hot path relies on, restated.
// PCL is NOT vendored under /root/reference (CMakeLists.txt:14 `find_package(PCL REQUIRED)`,
// version unpinned; the only version evidence in the tree is pcl-1.7-1.7.1), so
// these are the published algorithms of pcl::PointCloud, pcl::VoxelGrid::applyFilter,
// pcl::KdTreeFLANN::nearestKSearch and pcl::removeNaNFromPointCloud as adopted in SURVEY.md
// App. B.1/B.2.  Call sites they serve: scanRegistration.cpp:228,575-579;
// laserOdometry.cpp:436-437,478,590,905-906; laserMapping.cpp:388-395,694-700,707-708,717,824,
// 1022-1027,1050-1051.
#pragma once
#include <math.h>
#include <stdint.h>
#include <algorithm>
#include <limits>
#include <memory>
#include <vector>

namespace pcl {

struct alignas(16) PointXYZ {
  float x, y, z, pad_;
  PointXYZ() : x(0), y(0), z(0), pad_(1.f) {}
  PointXYZ(float a, float b, float c) : x(a), y(b), z(c), pad_(1.f) {}
};

struct alignas(16) PointXYZI {
  float x, y, z, pad_;
  float intensity;
  float pad2_[3];
  PointXYZI() : x(0), y(0), z(0), pad_(1.f), intensity(0) { pad2_[0] = pad2_[1] = pad2_[2] = 0; }
};

template <typename PointT>
class PointCloud {
 public:
  typedef std::shared_ptr<PointCloud<PointT> > Ptr;
  typedef std::shared_ptr<const PointCloud<PointT> > ConstPtr;
  std::vector<PointT> points;
  uint32_t width, height;
  bool is_dense;
  PointCloud() : width(0), height(0), is_dense(true) {}
  PointCloud(uint32_t w, uint32_t h) : points(size_t(w) * h), width(w), height(h), is_dense(true) {}
  void clear() { points.clear(); width = 0; height = 0; }
  size_t size() const { return points.size(); }
  void push_back(const PointT& p) {
    points.push_back(p);
    width = (uint32_t)points.size();
    height = 1;
  }
  PointCloud& operator+=(const PointCloud& rhs) {
    points.insert(points.end(), rhs.points.begin(), rhs.points.end());
    width = (uint32_t)points.size();
    height = 1;
    is_dense = is_dense && rhs.is_dense;
    return *this;
  }
};

// pcl::removeNaNFromPointCloud: keeps points whose x,y,z are all finite, preserving order.
template <typename PointT>
void removeNaNFromPointCloud(const PointCloud<PointT>& in, PointCloud<PointT>& out,
                             std::vector<int>& index) {
  if (&in != &out) out.points.resize(in.points.size());
  index.resize(in.points.size());
  size_t j = 0;
  for (size_t i = 0; i < in.points.size(); ++i) {
    const PointT& p = in.points[i];
    if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
    out.points[j] = in.points[i];
    index[j] = (int)i;
    j++;
  }
  out.points.resize(j);
  index.resize(j);
  out.height = 1;
  out.width = (uint32_t)j;
  out.is_dense = true;
}

namespace detail {
inline void field4(const PointXYZI& p, float f[4]) { f[0] = p.x; f[1] = p.y; f[2] = p.z; f[3] = p.intensity; }
inline void setfield4(PointXYZI& p, const float f[4]) { p.x = f[0]; p.y = f[1]; p.z = f[2]; p.intensity = f[3]; }
}  // namespace detail

// pcl::VoxelGrid<PointXYZI>, defaults (downsample_all_data=true, min_points_per_voxel=0).
// App. B.1: cell = (int)(floor(p*inv_leaf) - (float)min_b); idx = i + j*dx + k*dx*dy; one
// output per non-empty voxel in ascending idx; centroid of ALL FOUR fields = fp32 running sum
// (starting from 0) in ascending input index, divided by (float)count.
template <typename PointT>
class VoxelGrid {
 public:
  VoxelGrid() { leaf_[0] = leaf_[1] = leaf_[2] = 0; inv_[0] = inv_[1] = inv_[2] = 0; }
  void setLeafSize(float lx, float ly, float lz) {
    leaf_[0] = lx; leaf_[1] = ly; leaf_[2] = lz;
    for (int a = 0; a < 3; a++) inv_[a] = 1.0f / leaf_[a];
  }
  void setInputCloud(const typename PointCloud<PointT>::Ptr& c) { input_ = c; }
  void setInputCloud(const typename PointCloud<PointT>::ConstPtr& c) { input_ = c; }
  void filter(PointCloud<PointT>& out) {
    const std::vector<PointT>& in = input_->points;
    if (&out == input_.get()) {  // PCL filters into a temporary when in == out
      PointCloud<PointT> tmp;
      filter_impl(in, tmp);
      out = tmp;
      return;
    }
    filter_impl(in, out);
  }

 private:
  struct Key { unsigned int idx; unsigned int pt; };
  void filter_impl(const std::vector<PointT>& in, PointCloud<PointT>& out) {
    out.points.clear();
    out.height = 1;
    out.is_dense = true;
    if (in.empty()) { out.width = 0; return; }
    float mn[3] = {std::numeric_limits<float>::max(), std::numeric_limits<float>::max(),
                   std::numeric_limits<float>::max()};
    float mx[3] = {-mn[0], -mn[1], -mn[2]};
    size_t nfinite = 0;
    for (size_t i = 0; i < in.size(); i++) {
      const PointT& p = in[i];
      if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
      nfinite++;
      mn[0] = std::min(mn[0], p.x); mx[0] = std::max(mx[0], p.x);
      mn[1] = std::min(mn[1], p.y); mx[1] = std::max(mx[1], p.y);
      mn[2] = std::min(mn[2], p.z); mx[2] = std::max(mx[2], p.z);
    }
    if (nfinite == 0) { out.width = 0; return; }
    int64_t dx = (int64_t)((mx[0] - mn[0]) * inv_[0]) + 1;
    int64_t dy = (int64_t)((mx[1] - mn[1]) * inv_[1]) + 1;
    int64_t dz = (int64_t)((mx[2] - mn[2]) * inv_[2]) + 1;
    if (dx * dy * dz > (int64_t)std::numeric_limits<int32_t>::max()) {
      out.points = in;  // PCL warns and returns the input unfiltered
      out.width = (uint32_t)out.points.size();
      return;
    }
    int minb[3], maxb[3], divb[3];
    for (int a = 0; a < 3; a++) {
      minb[a] = (int)floorf(mn[a] * inv_[a]);
      maxb[a] = (int)floorf(mx[a] * inv_[a]);
      divb[a] = maxb[a] - minb[a] + 1;
    }
    int mul[3] = {1, divb[0], divb[0] * divb[1]};
    std::vector<Key> keys;
    keys.reserve(nfinite);
    for (size_t i = 0; i < in.size(); i++) {
      const PointT& p = in[i];
      if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
      int i0 = (int)(floorf(p.x * inv_[0]) - (float)minb[0]);
      int i1 = (int)(floorf(p.y * inv_[1]) - (float)minb[1]);
      int i2 = (int)(floorf(p.z * inv_[2]) - (float)minb[2]);
      Key k;
      k.idx = (unsigned int)(i0 * mul[0] + i1 * mul[1] + i2 * mul[2]);
      k.pt = (unsigned int)i;
      keys.push_back(k);
    }
    // adopted definition (App. B.1 §5): stable, i.e. ascending (idx, input index)
    std::sort(keys.begin(), keys.end(), [](const Key& a, const Key& b) {
      return a.idx < b.idx || (a.idx == b.idx && a.pt < b.pt);
    });
    size_t first = 0;
    while (first < keys.size()) {
      size_t last = first + 1;
      while (last < keys.size() && keys[last].idx == keys[first].idx) last++;
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      for (size_t q = first; q < last; q++) {
        float f[4];
        detail::field4(in[keys[q].pt], f);
        c[0] += f[0]; c[1] += f[1]; c[2] += f[2]; c[3] += f[3];
      }
      float cnt = (float)(last - first);
      c[0] /= cnt; c[1] /= cnt; c[2] /= cnt; c[3] /= cnt;
      PointT o;
      detail::setfield4(o, c);
      out.points.push_back(o);
      first = last;
    }
    out.width = (uint32_t)out.points.size();
  }
  float leaf_[3], inv_[3];
  typename PointCloud<PointT>::ConstPtr input_;
};

// pcl::KdTreeFLANN<PointXYZI>: exact k-NN over x,y,z, distance = flann::L2_Simple<float>
// (((dx*dx)+dy*dy)+dz*dz in fp32), results ascending by (d2, index)  -- App. B.2.
// Implemented as a median-split kd-tree with leaf size 15 (FLANN KDTreeSingleIndex's default in
// PCL); only the RESULT is part of the contract, the traversal is not.
template <typename PointT>
class KdTreeFLANN {
 public:
  typedef std::shared_ptr<KdTreeFLANN<PointT> > Ptr;
  void setInputCloud(const typename PointCloud<PointT>::ConstPtr& cloud) { build(cloud->points); }
  void setInputCloud(const typename PointCloud<PointT>::Ptr& cloud) { build(cloud->points); }
  int nearestKSearch(const PointT& q, int k, std::vector<int>& idx, std::vector<float>& d2) const {
    int n = (int)pts_.size() / 3;
    if (k > n) k = n;
    idx.resize(k);
    d2.resize(k);
    if (k == 0) return 0;
    Heap h;
    h.k = k; h.n = 0;
    h.d.resize(k); h.i.resize(k);
    float qq[3] = {q.x, q.y, q.z};
    float off[3] = {0, 0, 0};
    float rootd = 0;
    for (int a = 0; a < 3; a++) {
      if (qq[a] < bbmin_[a]) { off[a] = (qq[a] - bbmin_[a]) * (qq[a] - bbmin_[a]); }
      else if (qq[a] > bbmax_[a]) { off[a] = (qq[a] - bbmax_[a]) * (qq[a] - bbmax_[a]); }
      rootd += off[a];
    }
    search(0, qq, h, rootd, off);
    for (int j = 0; j < k; j++) { idx[j] = map_[h.i[j]]; d2[j] = h.d[j]; }
    return k;
  }

 private:
  struct Node { int lo, hi; int axis; float split_lo, split_hi; int left, right; };
  struct Heap {  // sorted insertion array, ascending (d, i); k is tiny (1 or 5)
    int k, n;
    std::vector<float> d;
    std::vector<int> i;
    float worst() const { return n < k ? std::numeric_limits<float>::infinity() : d[k - 1]; }
    void push(float dd, int ii, const std::vector<int>& map) {
      if (n == k) {
        if (dd > d[k - 1] || (dd == d[k - 1] && map[ii] > map[i[k - 1]])) return;
      }
      int pos = n < k ? n : k - 1;
      while (pos > 0 && (d[pos - 1] > dd || (d[pos - 1] == dd && map[i[pos - 1]] > map[ii]))) {
        d[pos] = d[pos - 1]; i[pos] = i[pos - 1];
        pos--;
      }
      d[pos] = dd; i[pos] = ii;
      if (n < k) n++;
    }
  };
  void build(const std::vector<PointT>& in) {
    pts_.clear(); map_.clear(); nodes_.clear();
    for (size_t i = 0; i < in.size(); i++) {
      const PointT& p = in[i];
      if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
      pts_.push_back(p.x); pts_.push_back(p.y); pts_.push_back(p.z);
      map_.push_back((int)i);
    }
    int n = (int)map_.size();
    order_.resize(n);
    for (int i = 0; i < n; i++) order_[i] = i;
    for (int a = 0; a < 3; a++) { bbmin_[a] = std::numeric_limits<float>::max(); bbmax_[a] = -bbmin_[a]; }
    for (int i = 0; i < n; i++)
      for (int a = 0; a < 3; a++) {
        bbmin_[a] = std::min(bbmin_[a], pts_[3 * i + a]);
        bbmax_[a] = std::max(bbmax_[a], pts_[3 * i + a]);
      }
    if (n > 0) { nodes_.reserve(2 * n / 8 + 4); split(0, n); }
    // reorder points for locality
    std::vector<float> p2(pts_.size());
    std::vector<int> m2(map_.size());
    for (int i = 0; i < n; i++) {
      int o = order_[i];
      p2[3 * i] = pts_[3 * o]; p2[3 * i + 1] = pts_[3 * o + 1]; p2[3 * i + 2] = pts_[3 * o + 2];
      m2[i] = map_[o];
    }
    pts_.swap(p2);
    map_.swap(m2);
  }
  int split(int lo, int hi) {
    int id = (int)nodes_.size();
    nodes_.push_back(Node());
    Node nd;
    nd.lo = lo; nd.hi = hi; nd.left = nd.right = -1; nd.axis = -1; nd.split_lo = nd.split_hi = 0;
    if (hi - lo > 15) {
      float mn[3], mx[3];
      for (int a = 0; a < 3; a++) { mn[a] = std::numeric_limits<float>::max(); mx[a] = -mn[a]; }
      for (int i = lo; i < hi; i++)
        for (int a = 0; a < 3; a++) {
          float v = pts_[3 * order_[i] + a];
          mn[a] = std::min(mn[a], v); mx[a] = std::max(mx[a], v);
        }
      int ax = 0;
      if (mx[1] - mn[1] > mx[ax] - mn[ax]) ax = 1;
      if (mx[2] - mn[2] > mx[ax] - mn[ax]) ax = 2;
      int mid = (lo + hi) / 2;
      std::nth_element(order_.begin() + lo, order_.begin() + mid, order_.begin() + hi,
                       [&](int a, int b) { return pts_[3 * a + ax] < pts_[3 * b + ax]; });
      float slo = -std::numeric_limits<float>::max(), shi = std::numeric_limits<float>::max();
      for (int i = lo; i < mid; i++) slo = std::max(slo, pts_[3 * order_[i] + ax]);
      for (int i = mid; i < hi; i++) shi = std::min(shi, pts_[3 * order_[i] + ax]);
      nd.axis = ax; nd.split_lo = slo; nd.split_hi = shi;
      nd.left = split(lo, mid);
      nd.right = split(mid, hi);
    }
    nodes_[id] = nd;
    return id;
  }
  void search(int id, const float* q, Heap& h, float mind, float* off) const {
    const Node& nd = nodes_[id];
    if (nd.axis < 0) {
      for (int i = nd.lo; i < nd.hi; i++) {
        float dx = q[0] - pts_[3 * i], dy = q[1] - pts_[3 * i + 1], dz = q[2] - pts_[3 * i + 2];
        float d = 0.f;
        d += dx * dx; d += dy * dy; d += dz * dz;  // L2_Simple accumulation order
        h.push(d, i, map_);
      }
      return;
    }
    int ax = nd.axis;
    float v = q[ax];
    float d1 = v - nd.split_lo, d2 = v - nd.split_hi;
    int near, far;
    float cut;
    if (d1 + d2 < 0) { near = nd.left; far = nd.right; cut = d2 * d2; }
    else { near = nd.right; far = nd.left; cut = d1 * d1; }
    search(near, q, h, mind, off);
    float old = off[ax];
    float nmind = mind - old + cut;
    // conservative prune: bound is a lower bound up to fp32 rounding, keep ties (<=)
    if (nmind * 0.9999f <= h.worst()) {
      off[ax] = cut;
      search(far, q, h, nmind, off);
      off[ax] = old;
    }
  }
  std::vector<float> pts_;
  std::vector<int> map_, order_;
  std::vector<Node> nodes_;
  float bbmin_[3], bbmax_[3];
};

}  // namespace pcl
