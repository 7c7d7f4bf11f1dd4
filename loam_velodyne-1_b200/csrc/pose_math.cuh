// pose_math.cuh -- scalar pose algebra of the reference, callable from host and device.
// Restates (same expression trees, fp32 unless the reference promotes to double):
//   AccumulateRotation        laserOdometry.cpp:256-273
//   PluginIMURotation         laserOdometry.cpp:196-254
//   transformAssociateToMap   laserMapping.cpp:110-197 == transformMaintenance.cpp:60-145
//   the nav_msgs/Odometry pose hand-off (quaternion encode/decode in double):
//     laserOdometry.cpp:858-864 -> laserMapping.cpp:304-318 / transformMaintenance.cpp:147-161
//   tf::Quaternion::setRPY / tf::Matrix3x3::getRPY (tf is not vendored by the reference)
// Trigonometry goes through loam/detmath.h so host and device agree bit for bit.
// The scalar expressions restated in this file (cited below) come from loam_velodyne:
//   Copyright 2013, Ji Zhang, Carnegie Mellon University
//   Further contributions copyright (c) 2016, Southwest Research Institute.  All rights reserved.
// BSD 3-clause; the full notice, conditions and disclaimer are retained in NOTICE.md at the repository root.
#pragma once
#include "loam/detmath.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace loamgpu {
namespace pm {
using loam::dm::sin; using loam::dm::cos; using loam::dm::atan2; using loam::dm::asin;
namespace dm = loam::dm;

struct Quat { double x, y, z, w; };

LOAM_HD Quat quat_from_rpy(double roll, double pitch, double yaw) {
  double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
  double cosYaw = dm::cos(halfYaw), sinYaw = dm::sin(halfYaw);
  double cosPitch = dm::cos(halfPitch), sinPitch = dm::sin(halfPitch);
  double cosRoll = dm::cos(halfRoll), sinRoll = dm::sin(halfRoll);
  Quat q;
  q.x = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
  q.y = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
  q.z = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
  q.w = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  return q;
}

LOAM_HD void rpy_from_quat(const Quat& q, double& roll, double& pitch, double& yaw) {
  double d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
  double s = 2.0 / d;
  double xs = q.x * s, ys = q.y * s, zs = q.z * s;
  double wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  double xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
  double yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  double m00 = 1.0 - (yy + zz);
  double m10 = xy + wz;
  double m20 = xz - wy, m21 = yz + wx, m22 = 1.0 - (xx + yy);
  if (dm::fabs(m20) >= 1) {
    yaw = 0;
    double m01 = xy - wz, m02 = xz + wy;
    if (m20 < 0) {
      double delta = dm::atan2(m01, m02);
      pitch = 3.14159265358979323846 / 2.0;
      roll = delta;
    } else {
      double delta = dm::atan2(-m01, -m02);
      pitch = -3.14159265358979323846 / 2.0;
      roll = delta;
    }
  } else {
    pitch = -dm::asin(m20);
    roll = dm::atan2(m21 / dm::cos(pitch), m22 / dm::cos(pitch));
    yaw = dm::atan2(m10 / dm::cos(pitch), m00 / dm::cos(pitch));
  }
}

LOAM_HD void rpy_from_quat(double x, double y, double z, double w, double& roll, double& pitch, double& yaw) {
  Quat q = {x, y, z, w};
  rpy_from_quat(q, roll, pitch, yaw);
}

// /laser_odom_to_init and /aft_mapped_to_init carry (rx,ry,rz) as a quaternion built from
// createQuaternionMsgFromRollPitchYaw(rz, -rx, -ry) with components shuffled to
// (-q.y, -q.z, q.x, q.w); receivers rebuild Quaternion(o.z, -o.x, -o.y, o.w) and take getRPY.
LOAM_HD void odometry_msg_roundtrip(const float in[6], float out[6]) {
  Quat g = quat_from_rpy(in[2], -in[0], -in[1]);
  double ox = -g.y, oy = -g.z, oz = g.x, ow = g.w;
  Quat r = {oz, -ox, -oy, ow};
  double roll, pitch, yaw;
  rpy_from_quat(r, roll, pitch, yaw);
  out[0] = -pitch; out[1] = -yaw; out[2] = roll;
  out[3] = in[3]; out[4] = in[4]; out[5] = in[5];
}

LOAM_HD void PluginIMURotation(float bcx, float bcy, float bcz, float blx, float bly, float blz,
                                float alx, float aly, float alz, float& acx, float& acy, float& acz) {
    float sbcx = sin(bcx); float cbcx = cos(bcx); float sbcy = sin(bcy); float cbcy = cos(bcy);
    float sbcz = sin(bcz); float cbcz = cos(bcz);
    float sblx = sin(blx); float cblx = cos(blx); float sbly = sin(bly); float cbly = cos(bly);
    float sblz = sin(blz); float cblz = cos(blz);
    float salx = sin(alx); float calx = cos(alx); float saly = sin(aly); float caly = cos(aly);
    float salz = sin(alz); float calz = cos(alz);
    float srx = -sbcx*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly)
              - cbcx*cbcz*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
              - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
              - cbcx*sbcz*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
              - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz);
    acx = -asin(srx);
    float srycrx = (cbcy*sbcz - cbcz*sbcx*sbcy)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
                 - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
                 - (cbcy*cbcz + sbcx*sbcy*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
                 - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
                 + cbcx*sbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
    float crycrx = (cbcz*sbcy - cbcy*sbcx*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
                 - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
                 - (sbcy*sbcz + cbcy*cbcz*sbcx)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
                 - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
                 + cbcx*cbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
    acy = atan2(srycrx / cos(acx), crycrx / cos(acx));
    float srzcrx = sbcx*(cblx*cbly*(calz*saly - caly*salx*salz)
                 - cblx*sbly*(caly*calz + salx*saly*salz) + calx*salz*sblx)
                 - cbcx*cbcz*((caly*calz + salx*saly*salz)*(cbly*sblz - cblz*sblx*sbly)
                 + (calz*saly - caly*salx*salz)*(sbly*sblz + cbly*cblz*sblx)
                 - calx*cblx*cblz*salz) + cbcx*sbcz*((caly*calz + salx*saly*salz)*(cbly*cblz
                 + sblx*sbly*sblz) + (calz*saly - caly*salx*salz)*(cblz*sbly - cbly*sblx*sblz)
                 + calx*cblx*salz*sblz);
    float crzcrx = sbcx*(cblx*sbly*(caly*salz - calz*salx*saly)
                 - cblx*cbly*(saly*salz + caly*calz*salx) + calx*calz*sblx)
                 + cbcx*cbcz*((saly*salz + caly*calz*salx)*(sbly*sblz + cbly*cblz*sblx)
                 + (caly*salz - calz*salx*saly)*(cbly*sblz - cblz*sblx*sbly)
                 + calx*calz*cblx*cblz) - cbcx*sbcz*((saly*salz + caly*calz*salx)*(cblz*sbly
                 - cbly*sblx*sblz) + (caly*salz - calz*salx*saly)*(cbly*cblz + sblx*sbly*sblz)
                 - calx*calz*cblx*sblz);
    acz = atan2(srzcrx / cos(acx), crzcrx / cos(acx));
  }

LOAM_HD void AccumulateRotation(float cx, float cy, float cz, float lx, float ly, float lz,
                                 float& ox, float& oy, float& oz) {
    float srx = cos(lx)*cos(cx)*sin(ly)*sin(cz) - cos(cx)*cos(cz)*sin(lx) - cos(lx)*cos(ly)*sin(cx);
    ox = -asin(srx);
    float srycrx = sin(lx)*(cos(cy)*sin(cz) - cos(cz)*sin(cx)*sin(cy)) + cos(lx)*sin(ly)*(cos(cy)*cos(cz)
                 + sin(cx)*sin(cy)*sin(cz)) + cos(lx)*cos(ly)*cos(cx)*sin(cy);
    float crycrx = cos(lx)*cos(ly)*cos(cx)*cos(cy) - cos(lx)*sin(ly)*(cos(cz)*sin(cy)
                 - cos(cy)*sin(cx)*sin(cz)) - sin(lx)*(sin(cy)*sin(cz) + cos(cy)*cos(cz)*sin(cx));
    oy = atan2(srycrx / cos(ox), crycrx / cos(ox));
    float srzcrx = sin(cx)*(cos(lz)*sin(ly) - cos(ly)*sin(lx)*sin(lz)) + cos(cx)*sin(cz)*(cos(ly)*cos(lz)
                 + sin(lx)*sin(ly)*sin(lz)) + cos(lx)*cos(cx)*cos(cz)*sin(lz);
    float crzcrx = cos(lx)*cos(lz)*cos(cx)*cos(cz) - cos(cx)*sin(cz)*(cos(ly)*sin(lz)
                 - cos(lz)*sin(lx)*sin(ly)) - sin(cx)*(sin(ly)*sin(lz) + cos(ly)*cos(lz)*sin(lx));
    oz = atan2(srzcrx / cos(ox), crzcrx / cos(ox));
  }

LOAM_HD void transformAssociateToMap(const float transformSum[6], const float transformBefMapped[6],
                                    const float transformAftMapped[6], float transformIncre[6],
                                    float transformTobeMapped[6]) {
  float x1 = cos(transformSum[1]) * (transformBefMapped[3] - transformSum[3])
           - sin(transformSum[1]) * (transformBefMapped[5] - transformSum[5]);
  float y1 = transformBefMapped[4] - transformSum[4];
  float z1 = sin(transformSum[1]) * (transformBefMapped[3] - transformSum[3])
           + cos(transformSum[1]) * (transformBefMapped[5] - transformSum[5]);
  float x2 = x1;
  float y2 = cos(transformSum[0]) * y1 + sin(transformSum[0]) * z1;
  float z2 = -sin(transformSum[0]) * y1 + cos(transformSum[0]) * z1;
  transformIncre[3] = cos(transformSum[2]) * x2 + sin(transformSum[2]) * y2;
  transformIncre[4] = -sin(transformSum[2]) * x2 + cos(transformSum[2]) * y2;
  transformIncre[5] = z2;
  float sbcx = sin(transformSum[0]); float cbcx = cos(transformSum[0]);
  float sbcy = sin(transformSum[1]); float cbcy = cos(transformSum[1]);
  float sbcz = sin(transformSum[2]); float cbcz = cos(transformSum[2]);
  float sblx = sin(transformBefMapped[0]); float cblx = cos(transformBefMapped[0]);
  float sbly = sin(transformBefMapped[1]); float cbly = cos(transformBefMapped[1]);
  float sblz = sin(transformBefMapped[2]); float cblz = cos(transformBefMapped[2]);
  float salx = sin(transformAftMapped[0]); float calx = cos(transformAftMapped[0]);
  float saly = sin(transformAftMapped[1]); float caly = cos(transformAftMapped[1]);
  float salz = sin(transformAftMapped[2]); float calz = cos(transformAftMapped[2]);
  float srx = -sbcx*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly)
            - cbcx*cbcz*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
            - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
            - cbcx*sbcz*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
            - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz);
  transformTobeMapped[0] = -asin(srx);
  float srycrx = (cbcy*sbcz - cbcz*sbcx*sbcy)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
               - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
               - (cbcy*cbcz + sbcx*sbcy*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
               - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
               + cbcx*sbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
  float crycrx = (cbcz*sbcy - cbcy*sbcx*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
               - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
               - (sbcy*sbcz + cbcy*cbcz*sbcx)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
               - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
               + cbcx*cbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
  transformTobeMapped[1] = atan2(srycrx / cos(transformTobeMapped[0]), crycrx / cos(transformTobeMapped[0]));
  float srzcrx = sbcx*(cblx*cbly*(calz*saly - caly*salx*salz)
               - cblx*sbly*(caly*calz + salx*saly*salz) + calx*salz*sblx)
               - cbcx*cbcz*((caly*calz + salx*saly*salz)*(cbly*sblz - cblz*sblx*sbly)
               + (calz*saly - caly*salx*salz)*(sbly*sblz + cbly*cblz*sblx)
               - calx*cblx*cblz*salz) + cbcx*sbcz*((caly*calz + salx*saly*salz)*(cbly*cblz
               + sblx*sbly*sblz) + (calz*saly - caly*salx*salz)*(cblz*sbly - cbly*sblx*sblz)
               + calx*cblx*salz*sblz);
  float crzcrx = sbcx*(cblx*sbly*(caly*salz - calz*salx*saly)
               - cblx*cbly*(saly*salz + caly*calz*salx) + calx*calz*sblx)
               + cbcx*cbcz*((saly*salz + caly*calz*salx)*(sbly*sblz + cbly*cblz*sblx)
               + (caly*salz - calz*salx*saly)*(cbly*sblz - cblz*sblx*sbly)
               + calx*calz*cblx*cblz) - cbcx*sbcz*((saly*salz + caly*calz*salx)*(cblz*sbly
               - cbly*sblx*sblz) + (caly*salz - calz*salx*saly)*(cbly*cblz + sblx*sbly*sblz)
               - calx*calz*cblx*sblz);
  transformTobeMapped[2] = atan2(srzcrx / cos(transformTobeMapped[0]), crzcrx / cos(transformTobeMapped[0]));
  x1 = cos(transformTobeMapped[2]) * transformIncre[3] - sin(transformTobeMapped[2]) * transformIncre[4];
  y1 = sin(transformTobeMapped[2]) * transformIncre[3] + cos(transformTobeMapped[2]) * transformIncre[4];
  z1 = transformIncre[5];
  x2 = x1;
  y2 = cos(transformTobeMapped[0]) * y1 - sin(transformTobeMapped[0]) * z1;
  z2 = sin(transformTobeMapped[0]) * y1 + cos(transformTobeMapped[0]) * z1;
  transformTobeMapped[3] = transformAftMapped[3] - (cos(transformTobeMapped[1]) * x2 + sin(transformTobeMapped[1]) * z2);
  transformTobeMapped[4] = transformAftMapped[4] - y2;
  transformTobeMapped[5] = transformAftMapped[5] - (-sin(transformTobeMapped[1]) * x2 + cos(transformTobeMapped[1]) * z2);
}



// ---- the same three functions with the sines / cosines of their INPUT angles supplied by the caller --------
// (identical expression trees, generated from the text above: sin(a) -> sn[i], cos(a) -> cs[i], i = position of
// the angle in the argument list).  On the device the table is filled by the lanes of a warp at once; each entry
// is what the same detmath function returns for the same argument, so the results are bit-identical.  Angles
// that are computed inside (asin / atan2 results) are still evaluated in place.
LOAM_HD void AccumulateRotationV(const float* sn, const float* cs, float cx, float cy, float cz, float lx, float ly, float lz,
                                 float& ox, float& oy, float& oz) {
    float srx = cs[3]*cs[0]*sn[4]*sn[2] - cs[0]*cs[2]*sn[3] - cs[3]*cs[4]*sn[0];
    ox = -asin(srx);
    float srycrx = sn[3]*(cs[1]*sn[2] - cs[2]*sn[0]*sn[1]) + cs[3]*sn[4]*(cs[1]*cs[2]
                 + sn[0]*sn[1]*sn[2]) + cs[3]*cs[4]*cs[0]*sn[1];
    float crycrx = cs[3]*cs[4]*cs[0]*cs[1] - cs[3]*sn[4]*(cs[2]*sn[1]
                 - cs[1]*sn[0]*sn[2]) - sn[3]*(sn[1]*sn[2] + cs[1]*cs[2]*sn[0]);
    oy = atan2(srycrx / cos(ox), crycrx / cos(ox));
    float srzcrx = sn[0]*(cs[5]*sn[4] - cs[4]*sn[3]*sn[5]) + cs[0]*sn[2]*(cs[4]*cs[5]
                 + sn[3]*sn[4]*sn[5]) + cs[3]*cs[0]*cs[2]*sn[5];
    float crzcrx = cs[3]*cs[5]*cs[0]*cs[2] - cs[0]*sn[2]*(cs[4]*sn[5]
                 - cs[5]*sn[3]*sn[4]) - sn[0]*(sn[4]*sn[5] + cs[4]*cs[5]*sn[3]);
    oz = atan2(srzcrx / cos(ox), crzcrx / cos(ox));
  }

LOAM_HD void PluginIMURotationV(const float* sn, const float* cs, float bcx, float bcy, float bcz, float blx, float bly, float blz,
                                float alx, float aly, float alz, float& acx, float& acy, float& acz) {
    float sbcx = sn[0]; float cbcx = cs[0]; float sbcy = sn[1]; float cbcy = cs[1];
    float sbcz = sn[2]; float cbcz = cs[2];
    float sblx = sn[3]; float cblx = cs[3]; float sbly = sn[4]; float cbly = cs[4];
    float sblz = sn[5]; float cblz = cs[5];
    float salx = sn[6]; float calx = cs[6]; float saly = sn[7]; float caly = cs[7];
    float salz = sn[8]; float calz = cs[8];
    float srx = -sbcx*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly)
              - cbcx*cbcz*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
              - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
              - cbcx*sbcz*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
              - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz);
    acx = -asin(srx);
    float srycrx = (cbcy*sbcz - cbcz*sbcx*sbcy)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
                 - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
                 - (cbcy*cbcz + sbcx*sbcy*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
                 - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
                 + cbcx*sbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
    float crycrx = (cbcz*sbcy - cbcy*sbcx*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
                 - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
                 - (sbcy*sbcz + cbcy*cbcz*sbcx)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
                 - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
                 + cbcx*cbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
    acy = atan2(srycrx / cos(acx), crycrx / cos(acx));
    float srzcrx = sbcx*(cblx*cbly*(calz*saly - caly*salx*salz)
                 - cblx*sbly*(caly*calz + salx*saly*salz) + calx*salz*sblx)
                 - cbcx*cbcz*((caly*calz + salx*saly*salz)*(cbly*sblz - cblz*sblx*sbly)
                 + (calz*saly - caly*salx*salz)*(sbly*sblz + cbly*cblz*sblx)
                 - calx*cblx*cblz*salz) + cbcx*sbcz*((caly*calz + salx*saly*salz)*(cbly*cblz
                 + sblx*sbly*sblz) + (calz*saly - caly*salx*salz)*(cblz*sbly - cbly*sblx*sblz)
                 + calx*cblx*salz*sblz);
    float crzcrx = sbcx*(cblx*sbly*(caly*salz - calz*salx*saly)
                 - cblx*cbly*(saly*salz + caly*calz*salx) + calx*calz*sblx)
                 + cbcx*cbcz*((saly*salz + caly*calz*salx)*(sbly*sblz + cbly*cblz*sblx)
                 + (caly*salz - calz*salx*saly)*(cbly*sblz - cblz*sblx*sbly)
                 + calx*calz*cblx*cblz) - cbcx*sbcz*((saly*salz + caly*calz*salx)*(cblz*sbly
                 - cbly*sblx*sblz) + (caly*salz - calz*salx*saly)*(cbly*cblz + sblx*sbly*sblz)
                 - calx*calz*cblx*sblz);
    acz = atan2(srzcrx / cos(acx), crzcrx / cos(acx));
  }

LOAM_HD void transformAssociateToMapV(const float* sn, const float* cs, const float transformSum[6], const float transformBefMapped[6],
                                    const float transformAftMapped[6], float transformIncre[6],
                                    float transformTobeMapped[6]) {
  float x1 = cs[1] * (transformBefMapped[3] - transformSum[3])
           - sn[1] * (transformBefMapped[5] - transformSum[5]);
  float y1 = transformBefMapped[4] - transformSum[4];
  float z1 = sn[1] * (transformBefMapped[3] - transformSum[3])
           + cs[1] * (transformBefMapped[5] - transformSum[5]);
  float x2 = x1;
  float y2 = cs[0] * y1 + sn[0] * z1;
  float z2 = -sn[0] * y1 + cs[0] * z1;
  transformIncre[3] = cs[2] * x2 + sn[2] * y2;
  transformIncre[4] = -sn[2] * x2 + cs[2] * y2;
  transformIncre[5] = z2;
  float sbcx = sn[0]; float cbcx = cs[0];
  float sbcy = sn[1]; float cbcy = cs[1];
  float sbcz = sn[2]; float cbcz = cs[2];
  float sblx = sn[3]; float cblx = cs[3];
  float sbly = sn[4]; float cbly = cs[4];
  float sblz = sn[5]; float cblz = cs[5];
  float salx = sn[6]; float calx = cs[6];
  float saly = sn[7]; float caly = cs[7];
  float salz = sn[8]; float calz = cs[8];
  float srx = -sbcx*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly)
            - cbcx*cbcz*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
            - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
            - cbcx*sbcz*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
            - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz);
  transformTobeMapped[0] = -asin(srx);
  float srycrx = (cbcy*sbcz - cbcz*sbcx*sbcy)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
               - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
               - (cbcy*cbcz + sbcx*sbcy*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
               - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
               + cbcx*sbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
  float crycrx = (cbcz*sbcy - cbcy*sbcx*sbcz)*(calx*caly*(cblz*sbly - cbly*sblx*sblz)
               - calx*saly*(cbly*cblz + sblx*sbly*sblz) + cblx*salx*sblz)
               - (sbcy*sbcz + cbcy*cbcz*sbcx)*(calx*saly*(cbly*sblz - cblz*sblx*sbly)
               - calx*caly*(sbly*sblz + cbly*cblz*sblx) + cblx*cblz*salx)
               + cbcx*cbcy*(salx*sblx + calx*caly*cblx*cbly + calx*cblx*saly*sbly);
  transformTobeMapped[1] = atan2(srycrx / cos(transformTobeMapped[0]), crycrx / cos(transformTobeMapped[0]));
  float srzcrx = sbcx*(cblx*cbly*(calz*saly - caly*salx*salz)
               - cblx*sbly*(caly*calz + salx*saly*salz) + calx*salz*sblx)
               - cbcx*cbcz*((caly*calz + salx*saly*salz)*(cbly*sblz - cblz*sblx*sbly)
               + (calz*saly - caly*salx*salz)*(sbly*sblz + cbly*cblz*sblx)
               - calx*cblx*cblz*salz) + cbcx*sbcz*((caly*calz + salx*saly*salz)*(cbly*cblz
               + sblx*sbly*sblz) + (calz*saly - caly*salx*salz)*(cblz*sbly - cbly*sblx*sblz)
               + calx*cblx*salz*sblz);
  float crzcrx = sbcx*(cblx*sbly*(caly*salz - calz*salx*saly)
               - cblx*cbly*(saly*salz + caly*calz*salx) + calx*calz*sblx)
               + cbcx*cbcz*((saly*salz + caly*calz*salx)*(sbly*sblz + cbly*cblz*sblx)
               + (caly*salz - calz*salx*saly)*(cbly*sblz - cblz*sblx*sbly)
               + calx*calz*cblx*cblz) - cbcx*sbcz*((saly*salz + caly*calz*salx)*(cblz*sbly
               - cbly*sblx*sblz) + (caly*salz - calz*salx*saly)*(cbly*cblz + sblx*sbly*sblz)
               - calx*calz*cblx*sblz);
  transformTobeMapped[2] = atan2(srzcrx / cos(transformTobeMapped[0]), crzcrx / cos(transformTobeMapped[0]));
  x1 = cos(transformTobeMapped[2]) * transformIncre[3] - sin(transformTobeMapped[2]) * transformIncre[4];
  y1 = sin(transformTobeMapped[2]) * transformIncre[3] + cos(transformTobeMapped[2]) * transformIncre[4];
  z1 = transformIncre[5];
  x2 = x1;
  y2 = cos(transformTobeMapped[0]) * y1 - sin(transformTobeMapped[0]) * z1;
  z2 = sin(transformTobeMapped[0]) * y1 + cos(transformTobeMapped[0]) * z1;
  transformTobeMapped[3] = transformAftMapped[3] - (cos(transformTobeMapped[1]) * x2 + sin(transformTobeMapped[1]) * z2);
  transformTobeMapped[4] = transformAftMapped[4] - y2;
  transformTobeMapped[5] = transformAftMapped[5] - (-sin(transformTobeMapped[1]) * x2 + cos(transformTobeMapped[1]) * z2);
}

}  // namespace pm
}  // namespace loamgpu
