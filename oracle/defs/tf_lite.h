// ORACLE (test infrastructure, never linked into the product).
// tf_lite.h -- the tf (bullet LinearMath) quaternion <-> roll/pitch/yaw algebra the reference uses
// for the IMU orientation (scanRegistration.cpp:640-643, laserMapping.cpp:325-328) and for the
// pose hand-offs between nodes (laserOdometry.cpp:858-864 -> laserMapping.cpp:304-318 ->
// transformMaintenance.cpp:147-203).  tf is not vendored under /root/reference; this restates
// tf::Quaternion::setRPY, tf::Matrix3x3::setRotation and getEulerYPR in double.
// The trig goes through ORACLE_MATH (detmath by default, glibc with -DORACLE_LIBM).
#pragma once
#include "oracle_math.h"

namespace tfl {

struct Quat { double x, y, z, w; };

inline Quat quat_from_rpy(double roll, double pitch, double yaw) {
  double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
  double cosYaw = omath::cos(halfYaw), sinYaw = omath::sin(halfYaw);
  double cosPitch = omath::cos(halfPitch), sinPitch = omath::sin(halfPitch);
  double cosRoll = omath::cos(halfRoll), sinRoll = omath::sin(halfRoll);
  Quat q;
  q.x = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
  q.y = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
  q.z = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
  q.w = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  return q;
}

inline void rpy_from_quat(const Quat& q, double& roll, double& pitch, double& yaw) {
  double d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
  double s = 2.0 / d;
  double xs = q.x * s, ys = q.y * s, zs = q.z * s;
  double wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  double xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
  double yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  double m00 = 1.0 - (yy + zz);
  double m10 = xy + wz;
  double m20 = xz - wy, m21 = yz + wx, m22 = 1.0 - (xx + yy);
  if (omath::fabs(m20) >= 1) {
    yaw = 0;
    double m01 = xy - wz, m02 = xz + wy;
    if (m20 < 0) {
      double delta = omath::atan2(m01, m02);
      pitch = 3.14159265358979323846 / 2.0;
      roll = delta;
    } else {
      double delta = omath::atan2(-m01, -m02);
      pitch = -3.14159265358979323846 / 2.0;
      roll = delta;
    }
  } else {
    pitch = -omath::asin(m20);
    roll = omath::atan2(m21 / omath::cos(pitch), m22 / omath::cos(pitch));
    yaw = omath::atan2(m10 / omath::cos(pitch), m00 / omath::cos(pitch));
  }
}

}  // namespace tfl
