// ORACLE (test infrastructure).  In-process stand-ins for the ROS / tf / pcl_conversions surface
// used by the reference's four nodes, so that /root/reference/src/*.cpp compile unmodified
// (except the scripted capacity / N_SCANS / ring-formula patch in oracle/build_ref.sh) and run
// inside one process.  Numeric third-party semantics come from oracle/defs/{pcl,cv,tf}_lite.h.
//
// Execution model: each node's main() runs on its own thread as a coroutine.  ros::spinOnce() /
// ros::spin() hand control back to the driver ("idle"), the driver posts messages and calls
// cycle(), the node thread delivers them to the subscribed callbacks and runs one loop body.
#pragma once
#include <math.h>
#include <cmath>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include <map>
#include <memory>
#include <functional>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <chrono>
#include "defs/oracle_math.h"
#include "defs/pcl_lite.h"
#include "defs/cv_lite.h"
#include "defs/tf_lite.h"

namespace ros {
struct Time {
  double t;
  Time() : t(0) {}
  Time& fromSec(double s) { t = s; return *this; }
  double toSec() const { return t; }
};
}  // namespace ros

namespace std_msgs { struct Header { ros::Time stamp; std::string frame_id; }; }
namespace geometry_msgs {
struct Quaternion { double x, y, z, w; Quaternion() : x(0), y(0), z(0), w(1) {} };
struct Vector3 { double x, y, z; Vector3() : x(0), y(0), z(0) {} };
struct Point { double x, y, z; Point() : x(0), y(0), z(0) {} };
struct Pose { Point position; Quaternion orientation; };
struct PoseWithCovariance { Pose pose; };
struct Twist { Vector3 linear, angular; };
struct TwistWithCovariance { Twist twist; };
}  // namespace geometry_msgs
namespace sensor_msgs {
struct PointCloud2 {
  std_msgs::Header header;
  int nfields;               // 3 (xyz) or 4 (xyzi)
  std::vector<float> data;   // nfields floats per point
  PointCloud2() : nfields(4) {}
  typedef std::shared_ptr<PointCloud2> Ptr;
  typedef std::shared_ptr<const PointCloud2> ConstPtr;
};
typedef std::shared_ptr<const PointCloud2> PointCloud2ConstPtr;
struct Imu {
  std_msgs::Header header;
  geometry_msgs::Quaternion orientation;
  geometry_msgs::Vector3 angular_velocity, linear_acceleration;
  typedef std::shared_ptr<const Imu> ConstPtr;
};
}  // namespace sensor_msgs
namespace nav_msgs {
struct Odometry {
  std_msgs::Header header;
  std::string child_frame_id;
  geometry_msgs::PoseWithCovariance pose;
  geometry_msgs::TwistWithCovariance twist;
  typedef std::shared_ptr<const Odometry> ConstPtr;
};
}  // namespace nav_msgs

namespace pcl {
inline void fromROSMsg(const sensor_msgs::PointCloud2& m, PointCloud<PointXYZ>& c) {
  size_t n = m.data.size() / m.nfields;
  c.points.resize(n);
  for (size_t i = 0; i < n; i++) {
    c.points[i].x = m.data[i * m.nfields]; c.points[i].y = m.data[i * m.nfields + 1]; c.points[i].z = m.data[i * m.nfields + 2];
  }
  c.width = (uint32_t)n; c.height = 1; c.is_dense = false;
}
inline void fromROSMsg(const sensor_msgs::PointCloud2& m, PointCloud<PointXYZI>& c) {
  size_t n = m.data.size() / m.nfields;
  c.points.resize(n);
  for (size_t i = 0; i < n; i++) {
    c.points[i].x = m.data[i * m.nfields]; c.points[i].y = m.data[i * m.nfields + 1]; c.points[i].z = m.data[i * m.nfields + 2];
    c.points[i].intensity = m.nfields > 3 ? m.data[i * m.nfields + 3] : 0.f;
  }
  c.width = (uint32_t)n; c.height = 1; c.is_dense = false;
}
inline void toROSMsg(const PointCloud<PointXYZ>& c, sensor_msgs::PointCloud2& m) {
  m.nfields = 3;
  m.data.resize(c.points.size() * 3);
  for (size_t i = 0; i < c.points.size(); i++) { m.data[3 * i] = c.points[i].x; m.data[3 * i + 1] = c.points[i].y; m.data[3 * i + 2] = c.points[i].z; }
}
inline void toROSMsg(const PointCloud<PointXYZI>& c, sensor_msgs::PointCloud2& m) {
  m.nfields = 4;
  m.data.resize(c.points.size() * 4);
  for (size_t i = 0; i < c.points.size(); i++) {
    m.data[4 * i] = c.points[i].x; m.data[4 * i + 1] = c.points[i].y; m.data[4 * i + 2] = c.points[i].z; m.data[4 * i + 3] = c.points[i].intensity;
  }
}
}  // namespace pcl

namespace tf {
struct Quaternion {
  double x_, y_, z_, w_;
  Quaternion() : x_(0), y_(0), z_(0), w_(1) {}
  Quaternion(double x, double y, double z, double w) : x_(x), y_(y), z_(z), w_(w) {}
};
struct Vector3 { double x, y, z; Vector3() : x(0), y(0), z(0) {} Vector3(double a, double b, double c) : x(a), y(b), z(c) {} };
struct Matrix3x3 {
  tfl::Quat q;
  explicit Matrix3x3(const Quaternion& a) { q.x = a.x_; q.y = a.y_; q.z = a.z_; q.w = a.w_; }
  void getRPY(double& roll, double& pitch, double& yaw) const { tfl::rpy_from_quat(q, roll, pitch, yaw); }
};
inline void quaternionMsgToTF(const geometry_msgs::Quaternion& m, Quaternion& q) { q = Quaternion(m.x, m.y, m.z, m.w); }
inline geometry_msgs::Quaternion createQuaternionMsgFromRollPitchYaw(double roll, double pitch, double yaw) {
  tfl::Quat q = tfl::quat_from_rpy(roll, pitch, yaw);
  geometry_msgs::Quaternion m;
  m.x = q.x; m.y = q.y; m.z = q.z; m.w = q.w;
  return m;
}
struct StampedTransform {
  ros::Time stamp_;
  std::string frame_id_, child_frame_id_;
  void setRotation(const Quaternion&) {}
  void setOrigin(const Vector3&) {}
};
struct TransformBroadcaster { void sendTransform(const StampedTransform&) {} };
}  // namespace tf

namespace ros {

// ---- per-node bus (everything static: each node is its own shared object) ----
namespace shim {
struct Bus {
  std::map<std::string, std::function<void(const std::shared_ptr<void>&)> > subs;
  std::vector<std::pair<std::string, std::shared_ptr<void> > > inbox;
  std::map<std::string, std::shared_ptr<void> > outbox;   // last message per topic
  std::map<std::string, int> outcount;                    // publishes since last clear
  std::mutex mu;
  std::condition_variable cv;
  bool go = false, idle = false, shutdown = false, started = false;
  std::thread th;
  double busy = 0;  // seconds spent between wake-up and next idle (node busy time)
};
inline Bus& bus() { static Bus b; return b; }

// node side: called from ros::spinOnce()/spin(): become idle, wait for go, deliver inbox
inline void node_wait_and_deliver(std::chrono::steady_clock::time_point& woke) {
  Bus& b = bus();
  std::vector<std::pair<std::string, std::shared_ptr<void> > > msgs;
  {
    std::unique_lock<std::mutex> lk(b.mu);
    if (b.started) b.busy += std::chrono::duration<double>(std::chrono::steady_clock::now() - woke).count();
    b.started = true;
    b.idle = true;
    b.cv.notify_all();
    b.cv.wait(lk, [&] { return b.go || b.shutdown; });
    b.go = false;
    b.idle = false;
    msgs.swap(b.inbox);
  }
  woke = std::chrono::steady_clock::now();
  for (auto& m : msgs) {
    auto it = b.subs.find(m.first);
    if (it != b.subs.end()) it->second(m.second);
  }
}
inline std::chrono::steady_clock::time_point& woke_tp() { static std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now(); return t; }

// driver side
inline void wait_idle() {
  Bus& b = bus();
  std::unique_lock<std::mutex> lk(b.mu);
  b.cv.wait(lk, [&] { return b.idle; });
}
inline void cycle() {  // run one loop iteration of the node with whatever was posted
  Bus& b = bus();
  {
    std::unique_lock<std::mutex> lk(b.mu);
    b.cv.wait(lk, [&] { return b.idle; });
    b.idle = false;
    b.go = true;
    b.cv.notify_all();
  }
  wait_idle();
}
template <class M>
inline void post(const std::string& topic, const std::shared_ptr<M>& m) {
  Bus& b = bus();
  std::unique_lock<std::mutex> lk(b.mu);
  b.inbox.push_back(std::make_pair(topic, std::static_pointer_cast<void>(m)));
}
}  // namespace shim

inline void init(int&, char**, const char*) {}
inline bool ok() { return !shim::bus().shutdown; }
inline void spinOnce() { shim::node_wait_and_deliver(shim::woke_tp()); }
inline void spin() { while (ok()) spinOnce(); }
struct Rate { explicit Rate(double) {} void sleep() {} };

class Publisher {
 public:
  std::string topic;
  template <class M>
  void publish(const M& m) const {
    shim::Bus& b = shim::bus();
    b.outbox[topic] = std::static_pointer_cast<void>(std::make_shared<M>(m));
    b.outcount[topic]++;
  }
};
class Subscriber {};
class NodeHandle {
 public:
  template <class M>
  Subscriber subscribe(const std::string& topic, uint32_t, void (*cb)(const std::shared_ptr<M const>&)) {
    shim::bus().subs[topic] = [cb](const std::shared_ptr<void>& p) { cb(std::static_pointer_cast<const M>(p)); };
    return Subscriber();
  }
  template <class M>
  Publisher advertise(const std::string& topic, uint32_t) {
    Publisher p;
    p.topic = topic;
    return p;
  }
};
}  // namespace ros

// ---- math: route the reference's unqualified / std:: libm calls to the oracle's definition ----
#ifndef ORACLE_LIBM
inline float ref_sin(float x) { return omath::sin(x); }
inline double ref_sin(double x) { return omath::sin(x); }
inline float ref_cos(float x) { return omath::cos(x); }
inline double ref_cos(double x) { return omath::cos(x); }
inline float ref_atan(float x) { return omath::atan(x); }
inline double ref_atan(double x) { return omath::atan(x); }
inline float ref_atan2(float y, float x) { return omath::atan2(y, x); }
inline double ref_atan2(double y, double x) { return omath::atan2(y, x); }
inline float ref_asin(float x) { return omath::asin(x); }
inline double ref_asin(double x) { return omath::asin(x); }
inline double ref_pow(double x, int e) { return omath::pow(x, e); }
namespace std { using ::ref_sin; using ::ref_cos; using ::ref_atan; using ::ref_atan2; using ::ref_asin; using ::ref_pow; }
#define sin ref_sin
#define cos ref_cos
#define atan ref_atan
#define atan2 ref_atan2
#define asin ref_asin
#define pow ref_pow
#endif
