// adapter_common.h -- shared by the four ROS nodes of adapters/ros/: message <-> packed-array
// conversions and the pose <-> nav_msgs/Odometry encodings of the reference's topic graph
// (SURVEY.md §3.0).  Compiles against real ROS (roscpp, sensor_msgs, nav_msgs, tf,
// pcl_conversions) and, in this repository's tests, against the in-process stand-ins of
// oracle/shims/ (ROS is not installed here).
#pragma once
#include <stdlib.h>
#include <math.h>
#include <vector>
#include <ros/ros.h>
#include <nav_msgs/Odometry.h>
#include <sensor_msgs/Imu.h>
#include <sensor_msgs/PointCloud2.h>
#include <tf/transform_datatypes.h>
#include <tf/transform_broadcaster.h>
#include <pcl_conversions/pcl_conversions.h>
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#include "loam_b200.h"

#ifndef ROS_FATAL_STREAM_OR_PRINT
#ifdef ROS_FATAL_STREAM
#define ROS_FATAL_STREAM_OR_PRINT(what, why) ROS_FATAL_STREAM(what << ": " << why)
#else
#include <stdio.h>
#define ROS_FATAL_STREAM_OR_PRINT(what, why) fprintf(stderr, "%s: %s\n", what, why)
#endif
#endif

namespace loam_ros {

typedef pcl::PointXYZI PointType;   // include/loam_velodyne/common.h:37

// BASELINE.json presets (1 = VLP-16, the reference's literals; 2 = HDL-64E; 4 = 128 rings); a launch
// file sets LOAM_B200_PRESET / LOAM_B200_DEVICE in the node's environment
inline int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}
#ifndef LOAM_ADAPTER_PRESET
#define LOAM_ADAPTER_PRESET 1
#endif
inline loam_params params_from_env(int system_delay) {
  loam_params p;
  loam_params_preset(&p, env_int("LOAM_B200_PRESET", LOAM_ADAPTER_PRESET));
  p.system_delay = system_delay;
  return p;
}
inline void die_on_error(int rc, const char* what) {
  if (rc < 0) { ROS_FATAL_STREAM_OR_PRINT(what, loam_last_error()); exit(1); }
}

// sensor_msgs/PointCloud2 -> packed x,y,z,intensity floats; NaN points dropped as
// pcl::removeNaNFromPointCloud does in the reference's handlers (laserOdometry.cpp:282, ...)
inline void unpack_msg(const sensor_msgs::PointCloud2& msg, std::vector<float>* xyzi, bool drop_nan) {
  pcl::PointCloud<PointType> c;
  pcl::fromROSMsg(msg, c);
  xyzi->clear();
  xyzi->reserve(4 * c.points.size() + 4);
  for (size_t i = 0; i < c.points.size(); i++) {
    const PointType& q = c.points[i];
    if (drop_nan && !(std::isfinite(q.x) && std::isfinite(q.y) && std::isfinite(q.z))) continue;
    xyzi->push_back(q.x); xyzi->push_back(q.y); xyzi->push_back(q.z); xyzi->push_back(q.intensity);
  }
}
inline void publish_cloud(const ros::Publisher& pub, const std::vector<float>& xyzi, int n, double stamp, const char* frame) {
  pcl::PointCloud<PointType> c;
  c.points.resize((size_t)n);
  for (int i = 0; i < n; i++) {
    PointType& q = c.points[(size_t)i];
    q.x = xyzi[4 * (size_t)i]; q.y = xyzi[4 * (size_t)i + 1]; q.z = xyzi[4 * (size_t)i + 2]; q.intensity = xyzi[4 * (size_t)i + 3];
  }
  c.width = (uint32_t)n; c.height = 1;
  sensor_msgs::PointCloud2 msg;
  pcl::toROSMsg(c, msg);
  msg.header.stamp = ros::Time().fromSec(stamp);
  msg.header.frame_id = frame;
  pub.publish(msg);
}

// (rx, ry, rz, tx, ty, tz) -> pose of a nav_msgs/Odometry, the encoding of laserOdometry.cpp:858-869
inline void pose_to_msg(const float t[6], nav_msgs::Odometry* m) {
  geometry_msgs::Quaternion q = tf::createQuaternionMsgFromRollPitchYaw(t[2], -t[0], -t[1]);
  m->pose.pose.orientation.x = -q.y;
  m->pose.pose.orientation.y = -q.z;
  m->pose.pose.orientation.z = q.x;
  m->pose.pose.orientation.w = q.w;
  m->pose.pose.position.x = t[3];
  m->pose.pose.position.y = t[4];
  m->pose.pose.position.z = t[5];
}
// ... and the decoding of laserMapping.cpp:304-318 / transformMaintenance.cpp:149-161
inline void pose_from_msg(const nav_msgs::Odometry& m, float t[6]) {
  double roll, pitch, yaw;
  const geometry_msgs::Quaternion& q = m.pose.pose.orientation;
  tf::Matrix3x3(tf::Quaternion(q.z, -q.x, -q.y, q.w)).getRPY(roll, pitch, yaw);
  t[0] = -pitch; t[1] = -yaw; t[2] = roll;
  t[3] = m.pose.pose.position.x; t[4] = m.pose.pose.position.y; t[5] = m.pose.pose.position.z;
}
inline void send_tf(tf::TransformBroadcaster& br, tf::StampedTransform& tr, const nav_msgs::Odometry& m, double stamp) {
  const geometry_msgs::Quaternion& q = m.pose.pose.orientation;
  tr.stamp_ = ros::Time().fromSec(stamp);
  tr.setRotation(tf::Quaternion(q.x, q.y, q.z, q.w));
  tr.setOrigin(tf::Vector3(m.pose.pose.position.x, m.pose.pose.position.y, m.pose.pose.position.z));
  br.sendTransform(tr);
}

}  // namespace loam_ros

