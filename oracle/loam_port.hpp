// ORACLE (test infrastructure; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use it; never linked into or called by the product).
//
// loam_port.hpp -- CPU restatement of the reference's per-sweep numeric pipeline
// (Maofei/loam_velodyne-1 at /root/reference), ROS stripped, parameterised for N_SCANS / ring
// model / capacity.  Each function cites the reference lines it follows.  Expression trees,
// float/double promotion and evaluation order mirror the source text so that, compiled with
// `g++ -O2 -ffp-contract=off`, it produces the same bits as the reference compiled against the
// same third-party definitions (oracle/defs/*.h); tests/test_oracle_vs_ref.py checks exactly that
// against oracle/_ref (the reference's own .cpp files compiled from /root/reference).
// The scalar expressions restated in this file (cited below) come from loam_velodyne:
//   Copyright 2013, Ji Zhang, Carnegie Mellon University
//   Further contributions copyright (c) 2016, Southwest Research Institute.  All rights reserved.
// BSD 3-clause; the full notice, conditions and disclaimer are retained in NOTICE.md at the repository root.
#pragma once
#include <stdio.h>
#include <string.h>
#include <vector>
#include <algorithm>
#include "loam_b200.h"
#include "defs/oracle_math.h"
#include "defs/pcl_lite.h"
#include "defs/cv_lite.h"
#include "defs/tf_lite.h"

namespace oracle {

typedef pcl::PointXYZI PointType;
typedef pcl::PointCloud<PointType> Cloud;
using omath::sin; using omath::cos; using omath::atan; using omath::atan2; using omath::asin;
using omath::sqrt; using omath::fabs; using omath::pow;

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
inline double rad2deg(double radians) { return radians * 180.0 / M_PI; }  // common.h:39-42

// ------------------------------------------------------------------------------------------
// scanRegistration.cpp
// ------------------------------------------------------------------------------------------
struct SRFeatures {
  Cloud full, sharp, lessSharp, flat, lessFlat;
  float imuTrans[12];
  std::vector<int> sharpIdx, lessSharpIdx, flatIdx;  // indices into `full` (debug/parity)
  double stamp;
};

class ScanRegistration {
 public:
  explicit ScanRegistration(const loam_params& p) : P(p) {
    int cap = P.max_points;
    cloudCurvature.assign(cap, 0.f); cloudSortInd.assign(cap, 0);
    cloudNeighborPicked.assign(cap, 0); cloudLabel.assign(cap, 0);
    for (int i = 0; i < imuQueLength; i++) {
      imuTime[i] = 0; imuRoll[i] = imuPitch[i] = imuYaw[i] = 0; imuAccX[i] = imuAccY[i] = imuAccZ[i] = 0;
      imuVeloX[i] = imuVeloY[i] = imuVeloZ[i] = 0; imuShiftX[i] = imuShiftY[i] = imuShiftZ[i] = 0;
    }
  }
  static const int imuQueLength = 200;  // scanRegistration.cpp:70
  loam_params P;
  const double scanPeriod = 0.1;  // :55 (double in this node)
  int systemInitCount = 0; bool systemInited = false;  // :58-59
  std::vector<float> cloudCurvature; std::vector<int> cloudSortInd, cloudNeighborPicked, cloudLabel;  // :63-66
  std::vector<int> scanStartInd, scanEndInd;
  int imuPointerFront = 0, imuPointerLast = -1;  // :68-69
  float imuRollStart = 0, imuPitchStart = 0, imuYawStart = 0, imuRollCur = 0, imuPitchCur = 0, imuYawCur = 0;
  float imuVeloXStart = 0, imuVeloYStart = 0, imuVeloZStart = 0, imuShiftXStart = 0, imuShiftYStart = 0, imuShiftZStart = 0;
  float imuVeloXCur = 0, imuVeloYCur = 0, imuVeloZCur = 0, imuShiftXCur = 0, imuShiftYCur = 0, imuShiftZCur = 0;
  float imuShiftFromStartXCur = 0, imuShiftFromStartYCur = 0, imuShiftFromStartZCur = 0;
  float imuVeloFromStartXCur = 0, imuVeloFromStartYCur = 0, imuVeloFromStartZCur = 0;
  double imuTime[imuQueLength];
  float imuRoll[imuQueLength], imuPitch[imuQueLength], imuYaw[imuQueLength];
  float imuAccX[imuQueLength], imuAccY[imuQueLength], imuAccZ[imuQueLength];
  float imuVeloX[imuQueLength], imuVeloY[imuQueLength], imuVeloZ[imuQueLength];
  float imuShiftX[imuQueLength], imuShiftY[imuQueLength], imuShiftZ[imuQueLength];

  // :108-125
  void ShiftToStartIMU(float pointTime) {
    imuShiftFromStartXCur = imuShiftXCur - imuShiftXStart - imuVeloXStart * pointTime;
    imuShiftFromStartYCur = imuShiftYCur - imuShiftYStart - imuVeloYStart * pointTime;
    imuShiftFromStartZCur = imuShiftZCur - imuShiftZStart - imuVeloZStart * pointTime;
    float x1 = cos(imuYawStart) * imuShiftFromStartXCur - sin(imuYawStart) * imuShiftFromStartZCur;
    float y1 = imuShiftFromStartYCur;
    float z1 = sin(imuYawStart) * imuShiftFromStartXCur + cos(imuYawStart) * imuShiftFromStartZCur;
    float x2 = x1;
    float y2 = cos(imuPitchStart) * y1 + sin(imuPitchStart) * z1;
    float z2 = -sin(imuPitchStart) * y1 + cos(imuPitchStart) * z1;
    imuShiftFromStartXCur = cos(imuRollStart) * x2 + sin(imuRollStart) * y2;
    imuShiftFromStartYCur = -sin(imuRollStart) * x2 + cos(imuRollStart) * y2;
    imuShiftFromStartZCur = z2;
  }
  // :127-144
  void VeloToStartIMU() {
    imuVeloFromStartXCur = imuVeloXCur - imuVeloXStart;
    imuVeloFromStartYCur = imuVeloYCur - imuVeloYStart;
    imuVeloFromStartZCur = imuVeloZCur - imuVeloZStart;
    float x1 = cos(imuYawStart) * imuVeloFromStartXCur - sin(imuYawStart) * imuVeloFromStartZCur;
    float y1 = imuVeloFromStartYCur;
    float z1 = sin(imuYawStart) * imuVeloFromStartXCur + cos(imuYawStart) * imuVeloFromStartZCur;
    float x2 = x1;
    float y2 = cos(imuPitchStart) * y1 + sin(imuPitchStart) * z1;
    float z2 = -sin(imuPitchStart) * y1 + cos(imuPitchStart) * z1;
    imuVeloFromStartXCur = cos(imuRollStart) * x2 + sin(imuRollStart) * y2;
    imuVeloFromStartYCur = -sin(imuRollStart) * x2 + cos(imuRollStart) * y2;
    imuVeloFromStartZCur = z2;
  }
  // :146-171
  void TransformToStartIMU(PointType* p) {
    float x1 = cos(imuRollCur) * p->x - sin(imuRollCur) * p->y;
    float y1 = sin(imuRollCur) * p->x + cos(imuRollCur) * p->y;
    float z1 = p->z;
    float x2 = x1;
    float y2 = cos(imuPitchCur) * y1 - sin(imuPitchCur) * z1;
    float z2 = sin(imuPitchCur) * y1 + cos(imuPitchCur) * z1;
    float x3 = cos(imuYawCur) * x2 + sin(imuYawCur) * z2;
    float y3 = y2;
    float z3 = -sin(imuYawCur) * x2 + cos(imuYawCur) * z2;
    float x4 = cos(imuYawStart) * x3 - sin(imuYawStart) * z3;
    float y4 = y3;
    float z4 = sin(imuYawStart) * x3 + cos(imuYawStart) * z3;
    float x5 = x4;
    float y5 = cos(imuPitchStart) * y4 + sin(imuPitchStart) * z4;
    float z5 = -sin(imuPitchStart) * y4 + cos(imuPitchStart) * z4;
    p->x = cos(imuRollStart) * x5 + sin(imuRollStart) * y5 + imuShiftFromStartXCur;
    p->y = -sin(imuRollStart) * x5 + cos(imuRollStart) * y5 + imuShiftFromStartYCur;
    p->z = z5 + imuShiftFromStartZCur;
  }
  // :173-209
  void AccumulateIMUShift() {
    float roll = imuRoll[imuPointerLast];
    float pitch = imuPitch[imuPointerLast];
    float yaw = imuYaw[imuPointerLast];
    float accX = imuAccX[imuPointerLast];
    float accY = imuAccY[imuPointerLast];
    float accZ = imuAccZ[imuPointerLast];
    float x1 = cos(roll) * accX - sin(roll) * accY;
    float y1 = sin(roll) * accX + cos(roll) * accY;
    float z1 = accZ;
    float x2 = x1;
    float y2 = cos(pitch) * y1 - sin(pitch) * z1;
    float z2 = sin(pitch) * y1 + cos(pitch) * z1;
    accX = cos(yaw) * x2 + sin(yaw) * z2;
    accY = y2;
    accZ = -sin(yaw) * x2 + cos(yaw) * z2;
    int imuPointerBack = (imuPointerLast + imuQueLength - 1) % imuQueLength;
    double timeDiff = imuTime[imuPointerLast] - imuTime[imuPointerBack];
    if (timeDiff < scanPeriod) {
      imuShiftX[imuPointerLast] = imuShiftX[imuPointerBack] + imuVeloX[imuPointerBack] * timeDiff + accX * timeDiff * timeDiff / 2;
      imuShiftY[imuPointerLast] = imuShiftY[imuPointerBack] + imuVeloY[imuPointerBack] * timeDiff + accY * timeDiff * timeDiff / 2;
      imuShiftZ[imuPointerLast] = imuShiftZ[imuPointerBack] + imuVeloZ[imuPointerBack] * timeDiff + accZ * timeDiff * timeDiff / 2;
      imuVeloX[imuPointerLast] = imuVeloX[imuPointerBack] + accX * timeDiff;
      imuVeloY[imuPointerLast] = imuVeloY[imuPointerBack] + accY * timeDiff;
      imuVeloZ[imuPointerLast] = imuVeloZ[imuPointerBack] + accZ * timeDiff;
    }
  }
  // imuHandler :638-660
  void imu(double stamp, const double quat[4], const float acc[3]) {
    double roll, pitch, yaw;
    tfl::Quat q = {quat[0], quat[1], quat[2], quat[3]};
    tfl::rpy_from_quat(q, roll, pitch, yaw);
    // linear_acceleration is float64 in sensor_msgs/Imu
    double lax = acc[0], lay = acc[1], laz = acc[2];
    float accX = lay - sin(roll) * cos(pitch) * 9.81;
    float accY = laz - cos(roll) * cos(pitch) * 9.81;
    float accZ = lax + sin(pitch) * 9.81;
    imuPointerLast = (imuPointerLast + 1) % imuQueLength;
    imuTime[imuPointerLast] = stamp;
    imuRoll[imuPointerLast] = roll;
    imuPitch[imuPointerLast] = pitch;
    imuYaw[imuPointerLast] = yaw;
    imuAccX[imuPointerLast] = accX;
    imuAccY[imuPointerLast] = accY;
    imuAccZ[imuPointerLast] = accZ;
    AccumulateIMUShift();
  }

  // laserCloudHandler :211-636.  xyz = /velodyne_points payload (3 floats per point).
  // meta_ring / meta_reltime (both or neither; indexed like xyz): SR-1's ring id and relTime supplied by the
  // caller instead of being derived from the point with atan / atan2 / halfPassed (:248-284) -- the product's
  // packet-metadata front end (packet_port.hpp sr1_meta_from_packets); everything else is the same handler
  bool process(const float* xyz, int n, int stride_floats, double stamp, SRFeatures& out,
               const int* meta_ring = nullptr, const float* meta_reltime = nullptr) {
    if (!systemInited) {  // :213-219
      systemInitCount++;
      if (systemInitCount >= P.system_delay) systemInited = true;
      return false;  // note: with system_delay = 0 the first sweep is still swallowed, as in the reference
    }
    const int N_SCANS = P.n_scans;
    scanStartInd.assign(N_SCANS, 0);
    scanEndInd.assign(N_SCANS, 0);
    double timeScanCur = stamp;
    pcl::PointCloud<pcl::PointXYZ> laserCloudIn;
    laserCloudIn.points.reserve(n);
    std::vector<int> rawIndex;
    for (int i = 0; i < n; i++) {
      const float* p = xyz + (size_t)i * stride_floats;
      if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2])) continue;  // :228
      laserCloudIn.points.push_back(pcl::PointXYZ(p[0], p[1], p[2]));
      if (meta_ring) rawIndex.push_back(i);
    }
    int cloudSize = (int)laserCloudIn.points.size();
    if (cloudSize == 0) return false;
    float startOri = -atan2(laserCloudIn.points[0].y, laserCloudIn.points[0].x);
    float endOri = -atan2(laserCloudIn.points[cloudSize - 1].y, laserCloudIn.points[cloudSize - 1].x) + 2 * M_PI;
    if (endOri - startOri > 3 * M_PI) {
      endOri -= 2 * M_PI;
    } else if (endOri - startOri < M_PI) {
      endOri += 2 * M_PI;
    }
    bool halfPassed = false;
    int count = cloudSize;
    PointType point;
    std::vector<Cloud> laserCloudScans(N_SCANS);
    for (int i = 0; i < cloudSize; i++) {  // :243-351
      point.x = laserCloudIn.points[i].y;
      point.y = laserCloudIn.points[i].z;
      point.z = laserCloudIn.points[i].x;
      float angle = meta_ring ? 0.f : atan(point.y / sqrt(point.x * point.x + point.z * point.z)) * 180 / M_PI;
      int scanID;
      if (meta_ring) {
        scanID = meta_ring[rawIndex[i]];
      } else if (P.ring_model == 0) {  // :249-256
        int roundedAngle = int(angle + (angle < 0.0 ? -0.5 : +0.5));
        if (roundedAngle > 0) {
          scanID = roundedAngle;
        } else {
          scanID = roundedAngle + (N_SCANS - 1);
        }
      } else {  // bk variant scanRegistration.cpp:268-274 (HDL-64E block), bounds parameterised
        const float angleLowerBoundDeg = P.ring_lower_deg;
        const float angleUpperBoundDeg = P.ring_upper_deg;
        const float angleSpan = angleUpperBoundDeg - angleLowerBoundDeg;
        const float angleStep = angleSpan / (N_SCANS - 1);
        float angleID = (angle - angleLowerBoundDeg) / angleStep;
        scanID = int(angleID + 0.5f);
      }
      if (scanID > (N_SCANS - 1) || scanID < 0) {  // :257-260
        count--;
        continue;
      }
      float ori = meta_ring ? 0.f : -atan2(point.x, point.z);  // :262-281
      if (meta_ring) {
      } else if (!halfPassed) {
        if (ori < startOri - M_PI / 2) {
          ori += 2 * M_PI;
        } else if (ori > startOri + M_PI * 3 / 2) {
          ori -= 2 * M_PI;
        }
        if (ori - startOri > M_PI) {
          halfPassed = true;
        }
      } else {
        ori += 2 * M_PI;
        if (ori < endOri - M_PI * 3 / 2) {
          ori += 2 * M_PI;
        } else if (ori > endOri + M_PI / 2) {
          ori -= 2 * M_PI;
        }
      }
      float relTime = meta_ring ? meta_reltime[rawIndex[i]] : (ori - startOri) / (endOri - startOri);
      point.intensity = scanID + scanPeriod * relTime;  // :284

      if (imuPointerLast >= 0) {  // :286-349
        float pointTime = relTime * scanPeriod;
        while (imuPointerFront != imuPointerLast) {
          if (timeScanCur + pointTime < imuTime[imuPointerFront]) break;
          imuPointerFront = (imuPointerFront + 1) % imuQueLength;
        }
        if (timeScanCur + pointTime > imuTime[imuPointerFront]) {
          imuRollCur = imuRoll[imuPointerFront];
          imuPitchCur = imuPitch[imuPointerFront];
          imuYawCur = imuYaw[imuPointerFront];
          imuVeloXCur = imuVeloX[imuPointerFront];
          imuVeloYCur = imuVeloY[imuPointerFront];
          imuVeloZCur = imuVeloZ[imuPointerFront];
          imuShiftXCur = imuShiftX[imuPointerFront];
          imuShiftYCur = imuShiftY[imuPointerFront];
          imuShiftZCur = imuShiftZ[imuPointerFront];
        } else {
          int imuPointerBack = (imuPointerFront + imuQueLength - 1) % imuQueLength;
          float ratioFront = (timeScanCur + pointTime - imuTime[imuPointerBack]) / (imuTime[imuPointerFront] - imuTime[imuPointerBack]);
          float ratioBack = (imuTime[imuPointerFront] - timeScanCur - pointTime) / (imuTime[imuPointerFront] - imuTime[imuPointerBack]);
          imuRollCur = imuRoll[imuPointerFront] * ratioFront + imuRoll[imuPointerBack] * ratioBack;
          imuPitchCur = imuPitch[imuPointerFront] * ratioFront + imuPitch[imuPointerBack] * ratioBack;
          if (imuYaw[imuPointerFront] - imuYaw[imuPointerBack] > M_PI) {
            imuYawCur = imuYaw[imuPointerFront] * ratioFront + (imuYaw[imuPointerBack] + 2 * M_PI) * ratioBack;
          } else if (imuYaw[imuPointerFront] - imuYaw[imuPointerBack] < -M_PI) {
            imuYawCur = imuYaw[imuPointerFront] * ratioFront + (imuYaw[imuPointerBack] - 2 * M_PI) * ratioBack;
          } else {
            imuYawCur = imuYaw[imuPointerFront] * ratioFront + imuYaw[imuPointerBack] * ratioBack;
          }
          imuVeloXCur = imuVeloX[imuPointerFront] * ratioFront + imuVeloX[imuPointerBack] * ratioBack;
          imuVeloYCur = imuVeloY[imuPointerFront] * ratioFront + imuVeloY[imuPointerBack] * ratioBack;
          imuVeloZCur = imuVeloZ[imuPointerFront] * ratioFront + imuVeloZ[imuPointerBack] * ratioBack;
          imuShiftXCur = imuShiftX[imuPointerFront] * ratioFront + imuShiftX[imuPointerBack] * ratioBack;
          imuShiftYCur = imuShiftY[imuPointerFront] * ratioFront + imuShiftY[imuPointerBack] * ratioBack;
          imuShiftZCur = imuShiftZ[imuPointerFront] * ratioFront + imuShiftZ[imuPointerBack] * ratioBack;
        }
        if (i == 0) {
          imuRollStart = imuRollCur; imuPitchStart = imuPitchCur; imuYawStart = imuYawCur;
          imuVeloXStart = imuVeloXCur; imuVeloYStart = imuVeloYCur; imuVeloZStart = imuVeloZCur;
          imuShiftXStart = imuShiftXCur; imuShiftYStart = imuShiftYCur; imuShiftZStart = imuShiftZCur;
        } else {
          ShiftToStartIMU(pointTime);
          VeloToStartIMU();
          TransformToStartIMU(&point);
        }
      }
      laserCloudScans[scanID].push_back(point);
    }
    cloudSize = count;
    if (cloudSize > P.max_points) return false;  // the reference would overrun its arrays

    Cloud& laserCloud = out.full;  // :354-357
    laserCloud.clear();
    for (int i = 0; i < N_SCANS; i++) laserCloud += laserCloudScans[i];
    const std::vector<PointType>& pts = laserCloud.points;
    int scanCount = -1;
    for (int i = 5; i < cloudSize - 5; i++) {  // :359-391
      float diffX = pts[i - 5].x + pts[i - 4].x + pts[i - 3].x + pts[i - 2].x + pts[i - 1].x - 10 * pts[i].x
                  + pts[i + 1].x + pts[i + 2].x + pts[i + 3].x + pts[i + 4].x + pts[i + 5].x;
      float diffY = pts[i - 5].y + pts[i - 4].y + pts[i - 3].y + pts[i - 2].y + pts[i - 1].y - 10 * pts[i].y
                  + pts[i + 1].y + pts[i + 2].y + pts[i + 3].y + pts[i + 4].y + pts[i + 5].y;
      float diffZ = pts[i - 5].z + pts[i - 4].z + pts[i - 3].z + pts[i - 2].z + pts[i - 1].z - 10 * pts[i].z
                  + pts[i + 1].z + pts[i + 2].z + pts[i + 3].z + pts[i + 4].z + pts[i + 5].z;
      cloudCurvature[i] = diffX * diffX + diffY * diffY + diffZ * diffZ;
      cloudSortInd[i] = i;
      cloudNeighborPicked[i] = 0;
      cloudLabel[i] = 0;
      if (int(pts[i].intensity) != scanCount) {
        scanCount = int(pts[i].intensity);
        if (scanCount > 0 && scanCount < N_SCANS) {
          scanStartInd[scanCount] = i + 5;
          scanEndInd[scanCount - 1] = i - 5;
        }
      }
    }
    scanStartInd[0] = 5;  // :392-393
    scanEndInd.back() = cloudSize - 5;

    for (int i = 5; i < cloudSize - 6; i++) {  // :395-452
      float diffX = pts[i + 1].x - pts[i].x;
      float diffY = pts[i + 1].y - pts[i].y;
      float diffZ = pts[i + 1].z - pts[i].z;
      float diff = diffX * diffX + diffY * diffY + diffZ * diffZ;
      if (diff > 0.1) {
        float depth1 = sqrt(pts[i].x * pts[i].x + pts[i].y * pts[i].y + pts[i].z * pts[i].z);
        float depth2 = sqrt(pts[i + 1].x * pts[i + 1].x + pts[i + 1].y * pts[i + 1].y + pts[i + 1].z * pts[i + 1].z);
        if (depth1 > depth2) {
          diffX = pts[i + 1].x - pts[i].x * depth2 / depth1;
          diffY = pts[i + 1].y - pts[i].y * depth2 / depth1;
          diffZ = pts[i + 1].z - pts[i].z * depth2 / depth1;
          if (sqrt(diffX * diffX + diffY * diffY + diffZ * diffZ) / depth2 < 0.1) {
            cloudNeighborPicked[i - 5] = 1; cloudNeighborPicked[i - 4] = 1; cloudNeighborPicked[i - 3] = 1;
            cloudNeighborPicked[i - 2] = 1; cloudNeighborPicked[i - 1] = 1; cloudNeighborPicked[i] = 1;
          }
        } else {
          diffX = pts[i + 1].x * depth1 / depth2 - pts[i].x;
          diffY = pts[i + 1].y * depth1 / depth2 - pts[i].y;
          diffZ = pts[i + 1].z * depth1 / depth2 - pts[i].z;
          if (sqrt(diffX * diffX + diffY * diffY + diffZ * diffZ) / depth1 < 0.1) {
            cloudNeighborPicked[i + 1] = 1; cloudNeighborPicked[i + 2] = 1; cloudNeighborPicked[i + 3] = 1;
            cloudNeighborPicked[i + 4] = 1; cloudNeighborPicked[i + 5] = 1; cloudNeighborPicked[i + 6] = 1;
          }
        }
      }
      float diffX2 = pts[i].x - pts[i - 1].x;
      float diffY2 = pts[i].y - pts[i - 1].y;
      float diffZ2 = pts[i].z - pts[i - 1].z;
      float diff2 = diffX2 * diffX2 + diffY2 * diffY2 + diffZ2 * diffZ2;
      float dis = pts[i].x * pts[i].x + pts[i].y * pts[i].y + pts[i].z * pts[i].z;
      if (diff > 0.0002 * dis && diff2 > 0.0002 * dis) {
        cloudNeighborPicked[i] = 1;
      }
    }

    out.sharp.clear(); out.lessSharp.clear(); out.flat.clear(); out.lessFlat.clear();
    out.sharpIdx.clear(); out.lessSharpIdx.clear(); out.flatIdx.clear();
    for (int i = 0; i < N_SCANS; i++) {  // :460-582
      Cloud::Ptr surfPointsLessFlatScan(new Cloud());
      for (int j = 0; j < 6; j++) {
        int sp = (scanStartInd[i] * (6 - j) + scanEndInd[i] * j) / 6;
        int ep = (scanStartInd[i] * (5 - j) + scanEndInd[i] * (j + 1)) / 6 - 1;
        // :466-474 is a bubble sort that swaps only on strict '<': a stable ascending sort
        if (ep > sp) {
          std::stable_sort(cloudSortInd.begin() + sp, cloudSortInd.begin() + ep + 1,
                           [&](int a, int b) { return cloudCurvature[a] < cloudCurvature[b]; });
        }
        int largestPickedNum = 0;
        for (int k = ep; k >= sp; k--) {  // :476-522
          int ind = cloudSortInd[k];
          if (cloudNeighborPicked[ind] == 0 && cloudCurvature[ind] > 0.1) {
            largestPickedNum++;
            if (largestPickedNum <= 2) {
              cloudLabel[ind] = 2;
              out.sharp.push_back(pts[ind]); out.sharpIdx.push_back(ind);
              out.lessSharp.push_back(pts[ind]); out.lessSharpIdx.push_back(ind);
            } else if (largestPickedNum <= 20) {
              cloudLabel[ind] = 1;
              out.lessSharp.push_back(pts[ind]); out.lessSharpIdx.push_back(ind);
            } else {
              break;
            }
            cloudNeighborPicked[ind] = 1;
            for (int l = 1; l <= 5; l++) {
              float diffX = pts[ind + l].x - pts[ind + l - 1].x;
              float diffY = pts[ind + l].y - pts[ind + l - 1].y;
              float diffZ = pts[ind + l].z - pts[ind + l - 1].z;
              if (diffX * diffX + diffY * diffY + diffZ * diffZ > 0.05) break;
              cloudNeighborPicked[ind + l] = 1;
            }
            for (int l = -1; l >= -5; l--) {
              float diffX = pts[ind + l].x - pts[ind + l + 1].x;
              float diffY = pts[ind + l].y - pts[ind + l + 1].y;
              float diffZ = pts[ind + l].z - pts[ind + l + 1].z;
              if (diffX * diffX + diffY * diffY + diffZ * diffZ > 0.05) break;
              cloudNeighborPicked[ind + l] = 1;
            }
          }
        }
        int smallestPickedNum = 0;
        for (int k = sp; k <= ep; k++) {  // :524-566
          int ind = cloudSortInd[k];
          if (cloudNeighborPicked[ind] == 0 && cloudCurvature[ind] < 0.1) {
            cloudLabel[ind] = -1;
            out.flat.push_back(pts[ind]); out.flatIdx.push_back(ind);
            smallestPickedNum++;
            if (smallestPickedNum >= 4) break;
            cloudNeighborPicked[ind] = 1;
            for (int l = 1; l <= 5; l++) {
              float diffX = pts[ind + l].x - pts[ind + l - 1].x;
              float diffY = pts[ind + l].y - pts[ind + l - 1].y;
              float diffZ = pts[ind + l].z - pts[ind + l - 1].z;
              if (diffX * diffX + diffY * diffY + diffZ * diffZ > 0.05) break;
              cloudNeighborPicked[ind + l] = 1;
            }
            for (int l = -1; l >= -5; l--) {
              float diffX = pts[ind + l].x - pts[ind + l + 1].x;
              float diffY = pts[ind + l].y - pts[ind + l + 1].y;
              float diffZ = pts[ind + l].z - pts[ind + l + 1].z;
              if (diffX * diffX + diffY * diffY + diffZ * diffZ > 0.05) break;
              cloudNeighborPicked[ind + l] = 1;
            }
          }
        }
        for (int k = sp; k <= ep; k++) {  // :568-572
          if (cloudLabel[k] <= 0) surfPointsLessFlatScan->push_back(pts[k]);
        }
      }
      Cloud surfPointsLessFlatScanDS;  // :575-581
      pcl::VoxelGrid<PointType> downSizeFilter;
      downSizeFilter.setInputCloud(surfPointsLessFlatScan);
      downSizeFilter.setLeafSize(0.2, 0.2, 0.2);
      downSizeFilter.filter(surfPointsLessFlatScanDS);
      out.lessFlat += surfPointsLessFlatScanDS;
    }
    // :614-629
    out.imuTrans[0] = imuPitchStart; out.imuTrans[1] = imuYawStart; out.imuTrans[2] = imuRollStart;
    out.imuTrans[3] = imuPitchCur; out.imuTrans[4] = imuYawCur; out.imuTrans[5] = imuRollCur;
    out.imuTrans[6] = imuShiftFromStartXCur; out.imuTrans[7] = imuShiftFromStartYCur; out.imuTrans[8] = imuShiftFromStartZCur;
    out.imuTrans[9] = imuVeloFromStartXCur; out.imuTrans[10] = imuVeloFromStartYCur; out.imuTrans[11] = imuVeloFromStartZCur;
    out.stamp = stamp;
    return true;
  }
};

// ------------------------------------------------------------------------------------------
// laserOdometry.cpp
// ------------------------------------------------------------------------------------------
struct LOOutput {
  float transformSum[6];    // /laser_odom_to_init (every sweep)
  bool hasMappingInput;     // :910-930 publishes the three clouds
  Cloud cornerLast, surfLast, fullRes;
  double stamp;
};

class LaserOdometry {
 public:
  explicit LaserOdometry(const loam_params& p) : P(p), matP(6, 6, CV_32F, cv::Scalar::all(0)) {
    int cap = P.max_points;
    pointSearchCornerInd1.assign(cap, 0.f); pointSearchCornerInd2.assign(cap, 0.f);
    pointSearchSurfInd1.assign(cap, 0.f); pointSearchSurfInd2.assign(cap, 0.f); pointSearchSurfInd3.assign(cap, 0.f);
    for (int i = 0; i < 6; i++) transform[i] = transformSum[i] = 0;
    cornerPointsSharp.reset(new Cloud()); cornerPointsLessSharp.reset(new Cloud());
    surfPointsFlat.reset(new Cloud()); surfPointsLessFlat.reset(new Cloud());
    laserCloudCornerLast.reset(new Cloud()); laserCloudSurfLast.reset(new Cloud());
    laserCloudOri.reset(new Cloud()); coeffSel.reset(new Cloud()); laserCloudFullRes.reset(new Cloud());
    kdtreeCornerLast.reset(new pcl::KdTreeFLANN<PointType>()); kdtreeSurfLast.reset(new pcl::KdTreeFLANN<PointType>());
  }
  loam_params P;
  const float scanPeriod = 0.1;  // :49 (float in this node)
  static const int skipFrameNum = 1;  // :51
  bool systemInited = false;
  Cloud::Ptr cornerPointsSharp, cornerPointsLessSharp, surfPointsFlat, surfPointsLessFlat;
  Cloud::Ptr laserCloudCornerLast, laserCloudSurfLast, laserCloudOri, coeffSel, laserCloudFullRes;
  pcl::KdTreeFLANN<PointType>::Ptr kdtreeCornerLast, kdtreeSurfLast;
  int laserCloudCornerLastNum = 0, laserCloudSurfLastNum = 0;  // :81-82 (zero-initialised globals)
  std::vector<float> pointSearchCornerInd1, pointSearchCornerInd2;  // :85-86 (indices stored as float)
  std::vector<float> pointSearchSurfInd1, pointSearchSurfInd2, pointSearchSurfInd3;  // :89-91
  float transform[6], transformSum[6];  // :93-94
  float imuRollStart = 0, imuPitchStart = 0, imuYawStart = 0, imuRollLast = 0, imuPitchLast = 0, imuYawLast = 0;
  float imuShiftFromStartX = 0, imuShiftFromStartY = 0, imuShiftFromStartZ = 0;
  float imuVeloFromStartX = 0, imuVeloFromStartY = 0, imuVeloFromStartZ = 0;
  bool isDegenerate = false;  // :404
  cv::Mat matP;               // :405
  int frameCount = skipFrameNum;  // :407
  // stats of the last sweep
  int lastIters = 0, lastRows = 0;

  // :101-124
  void TransformToStart(PointType const* const pi, PointType* const po) {
    float s = 10 * (pi->intensity - int(pi->intensity));
    float rx = s * transform[0];
    float ry = s * transform[1];
    float rz = s * transform[2];
    float tx = s * transform[3];
    float ty = s * transform[4];
    float tz = s * transform[5];
    float x1 = cos(rz) * (pi->x - tx) + sin(rz) * (pi->y - ty);
    float y1 = -sin(rz) * (pi->x - tx) + cos(rz) * (pi->y - ty);
    float z1 = (pi->z - tz);
    float x2 = x1;
    float y2 = cos(rx) * y1 + sin(rx) * z1;
    float z2 = -sin(rx) * y1 + cos(rx) * z1;
    po->x = cos(ry) * x2 - sin(ry) * z2;
    po->y = y2;
    po->z = sin(ry) * x2 + cos(ry) * z2;
    po->intensity = pi->intensity;
  }
  // :126-194
  void TransformToEnd(PointType const* const pi, PointType* const po) {
    float s = 10 * (pi->intensity - int(pi->intensity));
    float rx = s * transform[0];
    float ry = s * transform[1];
    float rz = s * transform[2];
    float tx = s * transform[3];
    float ty = s * transform[4];
    float tz = s * transform[5];
    float x1 = cos(rz) * (pi->x - tx) + sin(rz) * (pi->y - ty);
    float y1 = -sin(rz) * (pi->x - tx) + cos(rz) * (pi->y - ty);
    float z1 = (pi->z - tz);
    float x2 = x1;
    float y2 = cos(rx) * y1 + sin(rx) * z1;
    float z2 = -sin(rx) * y1 + cos(rx) * z1;
    float x3 = cos(ry) * x2 - sin(ry) * z2;
    float y3 = y2;
    float z3 = sin(ry) * x2 + cos(ry) * z2;
    rx = transform[0];
    ry = transform[1];
    rz = transform[2];
    tx = transform[3];
    ty = transform[4];
    tz = transform[5];
    float x4 = cos(ry) * x3 + sin(ry) * z3;
    float y4 = y3;
    float z4 = -sin(ry) * x3 + cos(ry) * z3;
    float x5 = x4;
    float y5 = cos(rx) * y4 - sin(rx) * z4;
    float z5 = sin(rx) * y4 + cos(rx) * z4;
    float x6 = cos(rz) * x5 - sin(rz) * y5 + tx;
    float y6 = sin(rz) * x5 + cos(rz) * y5 + ty;
    float z6 = z5 + tz;
    float x7 = cos(imuRollStart) * (x6 - imuShiftFromStartX) - sin(imuRollStart) * (y6 - imuShiftFromStartY);
    float y7 = sin(imuRollStart) * (x6 - imuShiftFromStartX) + cos(imuRollStart) * (y6 - imuShiftFromStartY);
    float z7 = z6 - imuShiftFromStartZ;
    float x8 = x7;
    float y8 = cos(imuPitchStart) * y7 - sin(imuPitchStart) * z7;
    float z8 = sin(imuPitchStart) * y7 + cos(imuPitchStart) * z7;
    float x9 = cos(imuYawStart) * x8 + sin(imuYawStart) * z8;
    float y9 = y8;
    float z9 = -sin(imuYawStart) * x8 + cos(imuYawStart) * z8;
    float x10 = cos(imuYawLast) * x9 - sin(imuYawLast) * z9;
    float y10 = y9;
    float z10 = sin(imuYawLast) * x9 + cos(imuYawLast) * z9;
    float x11 = x10;
    float y11 = cos(imuPitchLast) * y10 + sin(imuPitchLast) * z10;
    float z11 = -sin(imuPitchLast) * y10 + cos(imuPitchLast) * z10;
    po->x = cos(imuRollLast) * x11 + sin(imuRollLast) * y11;
    po->y = -sin(imuRollLast) * x11 + cos(imuRollLast) * y11;
    po->z = z11;
    po->intensity = int(pi->intensity);
  }
  // :196-254
  static void PluginIMURotation(float bcx, float bcy, float bcz, float blx, float bly, float blz,
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
  // :256-273
  static void AccumulateRotation(float cx, float cy, float cz, float lx, float ly, float lz,
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

  // one pass of the main loop body :413-931 for one synchronised set of the six topics.
  // returns false on the first sweep (:427-456, nothing but the init publish happens).
  bool process(const SRFeatures& f, LOOutput& out) {
    *cornerPointsSharp = f.sharp; *cornerPointsLessSharp = f.lessSharp;
    *surfPointsFlat = f.flat; *surfPointsLessFlat = f.lessFlat; *laserCloudFullRes = f.full;
    // imuTransHandler :337-351
    imuPitchStart = f.imuTrans[0]; imuYawStart = f.imuTrans[1]; imuRollStart = f.imuTrans[2];
    imuPitchLast = f.imuTrans[3]; imuYawLast = f.imuTrans[4]; imuRollLast = f.imuTrans[5];
    imuShiftFromStartX = f.imuTrans[6]; imuShiftFromStartY = f.imuTrans[7]; imuShiftFromStartZ = f.imuTrans[8];
    imuVeloFromStartX = f.imuTrans[9]; imuVeloFromStartY = f.imuTrans[10]; imuVeloFromStartZ = f.imuTrans[11];
    out.stamp = f.stamp;
    out.hasMappingInput = false;
    lastIters = 0; lastRows = 0;

    if (!systemInited) {  // :427-456
      std::swap(cornerPointsLessSharp, laserCloudCornerLast);
      std::swap(surfPointsLessFlat, laserCloudSurfLast);
      kdtreeCornerLast->setInputCloud(laserCloudCornerLast);
      kdtreeSurfLast->setInputCloud(laserCloudSurfLast);
      transformSum[0] += imuPitchStart;
      transformSum[2] += imuRollStart;
      systemInited = true;
      return false;
    }
    laserCloudOri->clear();  // :458-459 (once per sweep, NOT per iteration)
    coeffSel->clear();
    transform[3] -= imuVeloFromStartX * scanPeriod;  // :461-463
    transform[4] -= imuVeloFromStartY * scanPeriod;
    transform[5] -= imuVeloFromStartZ * scanPeriod;

    std::vector<int> pointSearchInd;
    std::vector<float> pointSearchSqDis;
    PointType pointOri, pointSel, tripod1, tripod2, tripod3, coeff;

    if (laserCloudCornerLastNum > 10 && laserCloudSurfLastNum > 100) {  // :465
      int cornerPointsSharpNum = (int)cornerPointsSharp->points.size();
      int surfPointsFlatNum = (int)surfPointsFlat->points.size();
      const int lastCornerSize = (int)laserCloudCornerLast->points.size();
      const int lastSurfSize = (int)laserCloudSurfLast->points.size();
      for (int iterCount = 0; iterCount < P.lo_max_iter; iterCount++) {  // :470
        lastIters = iterCount + 1;
        for (int i = 0; i < cornerPointsSharpNum; i++) {  // :471-584
          TransformToStart(&cornerPointsSharp->points[i], &pointSel);
          if (iterCount % 5 == 0) {
            kdtreeCornerLast->nearestKSearch(pointSel, 1, pointSearchInd, pointSearchSqDis);
            int closestPointInd = -1, minPointInd2 = -1;
            if (pointSearchSqDis[0] < 25) {
              closestPointInd = pointSearchInd[0];
              int closestPointScan = int(laserCloudCornerLast->points[closestPointInd].intensity);
              float pointSqDis, minPointSqDis2 = 25;
              // forward bound is cornerPointsSharpNum (:486) -- reference behaviour, kept; the
              // additional size guard only prevents reading past the cloud (UB in the reference)
              for (int j = closestPointInd + 1; j < cornerPointsSharpNum && j < lastCornerSize; j++) {
                if (int(laserCloudCornerLast->points[j].intensity) > closestPointScan + 2.5) break;
                pointSqDis = (laserCloudCornerLast->points[j].x - pointSel.x) * (laserCloudCornerLast->points[j].x - pointSel.x) +
                             (laserCloudCornerLast->points[j].y - pointSel.y) * (laserCloudCornerLast->points[j].y - pointSel.y) +
                             (laserCloudCornerLast->points[j].z - pointSel.z) * (laserCloudCornerLast->points[j].z - pointSel.z);
                if (int(laserCloudCornerLast->points[j].intensity) > closestPointScan) {
                  if (pointSqDis < minPointSqDis2) { minPointSqDis2 = pointSqDis; minPointInd2 = j; }
                }
              }
              for (int j = closestPointInd - 1; j >= 0; j--) {
                if (int(laserCloudCornerLast->points[j].intensity) < closestPointScan - 2.5) break;
                pointSqDis = (laserCloudCornerLast->points[j].x - pointSel.x) * (laserCloudCornerLast->points[j].x - pointSel.x) +
                             (laserCloudCornerLast->points[j].y - pointSel.y) * (laserCloudCornerLast->points[j].y - pointSel.y) +
                             (laserCloudCornerLast->points[j].z - pointSel.z) * (laserCloudCornerLast->points[j].z - pointSel.z);
                if (int(laserCloudCornerLast->points[j].intensity) < closestPointScan) {
                  if (pointSqDis < minPointSqDis2) { minPointSqDis2 = pointSqDis; minPointInd2 = j; }
                }
              }
            }
            pointSearchCornerInd1[i] = closestPointInd;
            pointSearchCornerInd2[i] = minPointInd2;
          }
          if (pointSearchCornerInd2[i] >= 0) {  // :530-583
            tripod1 = laserCloudCornerLast->points[(int)pointSearchCornerInd1[i]];
            tripod2 = laserCloudCornerLast->points[(int)pointSearchCornerInd2[i]];
            float x0 = pointSel.x; float y0 = pointSel.y; float z0 = pointSel.z;
            float x1 = tripod1.x; float y1 = tripod1.y; float z1 = tripod1.z;
            float x2 = tripod2.x; float y2 = tripod2.y; float z2 = tripod2.z;
            float a012 = sqrt(((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                       * ((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                       + ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
                       * ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
                       + ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))
                       * ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1)));
            float l12 = sqrt((x1 - x2)*(x1 - x2) + (y1 - y2)*(y1 - y2) + (z1 - z2)*(z1 - z2));
            float la = ((y1 - y2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                     + (z1 - z2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))) / a012 / l12;
            float lb = -((x1 - x2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                     - (z1 - z2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
            float lc = -((x1 - x2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
                     + (y1 - y2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
            float ld2 = a012 / l12;
            float s = 1;
            if (iterCount >= 5) s = 1 - 1.8 * fabs(ld2);
            coeff.x = s * la; coeff.y = s * lb; coeff.z = s * lc; coeff.intensity = s * ld2;
            if (s > 0.1 && ld2 != 0) {
              laserCloudOri->push_back(cornerPointsSharp->points[i]);
              coeffSel->push_back(coeff);
            }
          }
        }
        for (int i = 0; i < surfPointsFlatNum; i++) {  // :586-695
          TransformToStart(&surfPointsFlat->points[i], &pointSel);
          if (iterCount % 5 == 0) {
            kdtreeSurfLast->nearestKSearch(pointSel, 1, pointSearchInd, pointSearchSqDis);
            int closestPointInd = -1, minPointInd2 = -1, minPointInd3 = -1;
            if (pointSearchSqDis[0] < 25) {
              closestPointInd = pointSearchInd[0];
              int closestPointScan = int(laserCloudSurfLast->points[closestPointInd].intensity);
              float pointSqDis, minPointSqDis2 = 25, minPointSqDis3 = 25;
              for (int j = closestPointInd + 1; j < surfPointsFlatNum && j < lastSurfSize; j++) {  // :598 bound kept
                if (int(laserCloudSurfLast->points[j].intensity) > closestPointScan + 2.5) break;
                pointSqDis = (laserCloudSurfLast->points[j].x - pointSel.x) * (laserCloudSurfLast->points[j].x - pointSel.x) +
                             (laserCloudSurfLast->points[j].y - pointSel.y) * (laserCloudSurfLast->points[j].y - pointSel.y) +
                             (laserCloudSurfLast->points[j].z - pointSel.z) * (laserCloudSurfLast->points[j].z - pointSel.z);
                if (int(laserCloudSurfLast->points[j].intensity) <= closestPointScan) {
                  if (pointSqDis < minPointSqDis2) { minPointSqDis2 = pointSqDis; minPointInd2 = j; }
                } else {
                  if (pointSqDis < minPointSqDis3) { minPointSqDis3 = pointSqDis; minPointInd3 = j; }
                }
              }
              for (int j = closestPointInd - 1; j >= 0; j--) {
                if (int(laserCloudSurfLast->points[j].intensity) < closestPointScan - 2.5) break;
                pointSqDis = (laserCloudSurfLast->points[j].x - pointSel.x) * (laserCloudSurfLast->points[j].x - pointSel.x) +
                             (laserCloudSurfLast->points[j].y - pointSel.y) * (laserCloudSurfLast->points[j].y - pointSel.y) +
                             (laserCloudSurfLast->points[j].z - pointSel.z) * (laserCloudSurfLast->points[j].z - pointSel.z);
                if (int(laserCloudSurfLast->points[j].intensity) >= closestPointScan) {
                  if (pointSqDis < minPointSqDis2) { minPointSqDis2 = pointSqDis; minPointInd2 = j; }
                } else {
                  if (pointSqDis < minPointSqDis3) { minPointSqDis3 = pointSqDis; minPointInd3 = j; }
                }
              }
            }
            pointSearchSurfInd1[i] = closestPointInd;
            pointSearchSurfInd2[i] = minPointInd2;
            pointSearchSurfInd3[i] = minPointInd3;
          }
          if (pointSearchSurfInd2[i] >= 0 && pointSearchSurfInd3[i] >= 0) {  // :653-694
            tripod1 = laserCloudSurfLast->points[(int)pointSearchSurfInd1[i]];
            tripod2 = laserCloudSurfLast->points[(int)pointSearchSurfInd2[i]];
            tripod3 = laserCloudSurfLast->points[(int)pointSearchSurfInd3[i]];
            float pa = (tripod2.y - tripod1.y) * (tripod3.z - tripod1.z) - (tripod3.y - tripod1.y) * (tripod2.z - tripod1.z);
            float pb = (tripod2.z - tripod1.z) * (tripod3.x - tripod1.x) - (tripod3.z - tripod1.z) * (tripod2.x - tripod1.x);
            float pc = (tripod2.x - tripod1.x) * (tripod3.y - tripod1.y) - (tripod3.x - tripod1.x) * (tripod2.y - tripod1.y);
            float pd = -(pa * tripod1.x + pb * tripod1.y + pc * tripod1.z);
            float ps = sqrt(pa * pa + pb * pb + pc * pc);
            pa /= ps; pb /= ps; pc /= ps; pd /= ps;
            float pd2 = pa * pointSel.x + pb * pointSel.y + pc * pointSel.z + pd;
            float s = 1;
            if (iterCount >= 5) {
              s = 1 - 1.8 * fabs(pd2) / sqrt(sqrt(pointSel.x * pointSel.x + pointSel.y * pointSel.y + pointSel.z * pointSel.z));
            }
            coeff.x = s * pa; coeff.y = s * pb; coeff.z = s * pc; coeff.intensity = s * pd2;
            if (s > 0.1 && pd2 != 0) {
              laserCloudOri->push_back(surfPointsFlat->points[i]);
              coeffSel->push_back(coeff);
            }
          }
        }
        int pointSelNum = (int)laserCloudOri->points.size();  // :697-700
        lastRows = pointSelNum;
        if (pointSelNum < 10) continue;

        cv::Mat matA(pointSelNum, 6, CV_32F, cv::Scalar::all(0));
        cv::Mat matAt(6, pointSelNum, CV_32F, cv::Scalar::all(0));
        cv::Mat matAtA(6, 6, CV_32F, cv::Scalar::all(0));
        cv::Mat matB(pointSelNum, 1, CV_32F, cv::Scalar::all(0));
        cv::Mat matAtB(6, 1, CV_32F, cv::Scalar::all(0));
        cv::Mat matX(6, 1, CV_32F, cv::Scalar::all(0));
        for (int i = 0; i < pointSelNum; i++) {  // :708-764
          pointOri = laserCloudOri->points[i];
          coeff = coeffSel->points[i];
          float s = 1;
          float srx = sin(s * transform[0]); float crx = cos(s * transform[0]);
          float sry = sin(s * transform[1]); float cry = cos(s * transform[1]);
          float srz = sin(s * transform[2]); float crz = cos(s * transform[2]);
          float tx = s * transform[3]; float ty = s * transform[4]; float tz = s * transform[5];
          float arx = (-s*crx*sry*srz*pointOri.x + s*crx*crz*sry*pointOri.y + s*srx*sry*pointOri.z
                    + s*tx*crx*sry*srz - s*ty*crx*crz*sry - s*tz*srx*sry) * coeff.x
                    + (s*srx*srz*pointOri.x - s*crz*srx*pointOri.y + s*crx*pointOri.z
                    + s*ty*crz*srx - s*tz*crx - s*tx*srx*srz) * coeff.y
                    + (s*crx*cry*srz*pointOri.x - s*crx*cry*crz*pointOri.y - s*cry*srx*pointOri.z
                    + s*tz*cry*srx + s*ty*crx*cry*crz - s*tx*crx*cry*srz) * coeff.z;
          float ary = ((-s*crz*sry - s*cry*srx*srz)*pointOri.x
                    + (s*cry*crz*srx - s*sry*srz)*pointOri.y - s*crx*cry*pointOri.z
                    + tx*(s*crz*sry + s*cry*srx*srz) + ty*(s*sry*srz - s*cry*crz*srx)
                    + s*tz*crx*cry) * coeff.x
                    + ((s*cry*crz - s*srx*sry*srz)*pointOri.x
                    + (s*cry*srz + s*crz*srx*sry)*pointOri.y - s*crx*sry*pointOri.z
                    + s*tz*crx*sry - ty*(s*cry*srz + s*crz*srx*sry)
                    - tx*(s*cry*crz - s*srx*sry*srz)) * coeff.z;
          float arz = ((-s*cry*srz - s*crz*srx*sry)*pointOri.x + (s*cry*crz - s*srx*sry*srz)*pointOri.y
                    + tx*(s*cry*srz + s*crz*srx*sry) - ty*(s*cry*crz - s*srx*sry*srz)) * coeff.x
                    + (-s*crx*crz*pointOri.x - s*crx*srz*pointOri.y
                    + s*ty*crx*srz + s*tx*crx*crz) * coeff.y
                    + ((s*cry*crz*srx - s*sry*srz)*pointOri.x + (s*crz*sry + s*cry*srx*srz)*pointOri.y
                    + tx*(s*sry*srz - s*cry*crz*srx) - ty*(s*crz*sry + s*cry*srx*srz)) * coeff.z;
          float atx = -s*(cry*crz - srx*sry*srz) * coeff.x + s*crx*srz * coeff.y
                    - s*(crz*sry + cry*srx*srz) * coeff.z;
          float aty = -s*(cry*srz + crz*srx*sry) * coeff.x - s*crx*crz * coeff.y
                    - s*(sry*srz - cry*crz*srx) * coeff.z;
          float atz = s*crx*sry * coeff.x - s*srx * coeff.y - s*crx*cry * coeff.z;
          float d2 = coeff.intensity;
          matA.at<float>(i, 0) = arx; matA.at<float>(i, 1) = ary; matA.at<float>(i, 2) = arz;
          matA.at<float>(i, 3) = atx; matA.at<float>(i, 4) = aty; matA.at<float>(i, 5) = atz;
          matB.at<float>(i, 0) = -0.05 * d2;
        }
        cv::transpose(matA, matAt);  // :765-768
        matAtA = matAt * matA;
        matAtB = matAt * matB;
        cv::solve(matAtA, matAtB, matX, cv::DECOMP_QR);

        if (iterCount == 0) {  // :770-791
          cv::Mat matE(1, 6, CV_32F, cv::Scalar::all(0));
          cv::Mat matV(6, 6, CV_32F, cv::Scalar::all(0));
          cv::Mat matV2(6, 6, CV_32F, cv::Scalar::all(0));
          cv::eigen(matAtA, matE, matV);
          matV.copyTo(matV2);
          isDegenerate = false;
          float eignThre[6] = {10, 10, 10, 10, 10, 10};
          for (int i = 5; i >= 0; i--) {
            if (matE.at<float>(0, i) < eignThre[i]) {
              for (int j = 0; j < 6; j++) matV2.at<float>(i, j) = 0;
              isDegenerate = true;
            } else {
              break;
            }
          }
          matP = matV.inv() * matV2;
        }
        if (isDegenerate) {  // :793-797
          cv::Mat matX2(6, 1, CV_32F, cv::Scalar::all(0));
          matX.copyTo(matX2);
          matX = matP * matX2;
        }
        // :800-811 NaN guard
        if (std::isnan(matX.at<float>(0, 0)) || std::isnan(matX.at<float>(1, 0)) || std::isnan(matX.at<float>(2, 0)) ||
            std::isnan(matX.at<float>(3, 0)) || std::isnan(matX.at<float>(4, 0)) || std::isnan(matX.at<float>(5, 0))) {
        } else {
          transform[0] += matX.at<float>(0, 0); transform[1] += matX.at<float>(1, 0); transform[2] += matX.at<float>(2, 0);
          transform[3] += matX.at<float>(3, 0); transform[4] += matX.at<float>(4, 0); transform[5] += matX.at<float>(5, 0);
        }
        float deltaR = sqrt(pow(rad2deg(matX.at<float>(0, 0)), 2) + pow(rad2deg(matX.at<float>(1, 0)), 2) +
                            pow(rad2deg(matX.at<float>(2, 0)), 2));
        float deltaT = sqrt(pow(matX.at<float>(3, 0) * 100, 2) + pow(matX.at<float>(4, 0) * 100, 2) +
                            pow(matX.at<float>(5, 0) * 100, 2));
        if (deltaR < 0.1 && deltaT < 0.1) break;
      }
    }

    float rx, ry, rz, tx, ty, tz;  // :830-856
    AccumulateRotation(transformSum[0], transformSum[1], transformSum[2],
                       -transform[0], -transform[1] * 1.05, -transform[2], rx, ry, rz);
    float x1 = cos(rz) * (transform[3] - imuShiftFromStartX) - sin(rz) * (transform[4] - imuShiftFromStartY);
    float y1 = sin(rz) * (transform[3] - imuShiftFromStartX) + cos(rz) * (transform[4] - imuShiftFromStartY);
    float z1 = transform[5] * 1.05 - imuShiftFromStartZ;
    float x2 = x1;
    float y2 = cos(rx) * y1 - sin(rx) * z1;
    float z2 = sin(rx) * y1 + cos(rx) * z1;
    tx = transformSum[3] - (cos(ry) * x2 + sin(ry) * z2);
    ty = transformSum[4] - y2;
    tz = transformSum[5] - (-sin(ry) * x2 + cos(ry) * z2);
    PluginIMURotation(rx, ry, rz, imuPitchStart, imuYawStart, imuRollStart,
                      imuPitchLast, imuYawLast, imuRollLast, rx, ry, rz);
    transformSum[0] = rx; transformSum[1] = ry; transformSum[2] = rz;
    transformSum[3] = tx; transformSum[4] = ty; transformSum[5] = tz;
    for (int i = 0; i < 6; i++) out.transformSum[i] = transformSum[i];

    int cornerPointsLessSharpNum = (int)cornerPointsLessSharp->points.size();  // :875-891
    for (int i = 0; i < cornerPointsLessSharpNum; i++)
      TransformToEnd(&cornerPointsLessSharp->points[i], &cornerPointsLessSharp->points[i]);
    int surfPointsLessFlatNum = (int)surfPointsLessFlat->points.size();
    for (int i = 0; i < surfPointsLessFlatNum; i++)
      TransformToEnd(&surfPointsLessFlat->points[i], &surfPointsLessFlat->points[i]);
    frameCount++;
    if (frameCount >= skipFrameNum + 1) {
      int laserCloudFullResNum = (int)laserCloudFullRes->points.size();
      for (int i = 0; i < laserCloudFullResNum; i++)
        TransformToEnd(&laserCloudFullRes->points[i], &laserCloudFullRes->points[i]);
    }
    std::swap(cornerPointsLessSharp, laserCloudCornerLast);  // :893-899
    std::swap(surfPointsLessFlat, laserCloudSurfLast);
    laserCloudCornerLastNum = (int)laserCloudCornerLast->points.size();
    laserCloudSurfLastNum = (int)laserCloudSurfLast->points.size();
    if (laserCloudCornerLastNum > 10 && laserCloudSurfLastNum > 100) {
      kdtreeCornerLast->setInputCloud(laserCloudCornerLast);
      kdtreeSurfLast->setInputCloud(laserCloudSurfLast);
    }
    if (frameCount >= skipFrameNum + 1) {  // :910-930
      frameCount = 0;
      out.hasMappingInput = true;
      out.cornerLast = *laserCloudCornerLast;
      out.surfLast = *laserCloudSurfLast;
      out.fullRes = *laserCloudFullRes;
    }
    return true;
  }
};

// pose hand-off /laser_odom_to_init: laserOdometry.cpp:858-867 -> laserMapping.cpp:304-318 (and
// transformMaintenance.cpp:147-161): RPY -> quaternion (double) -> RPY -> float
inline void odometry_msg_roundtrip(const float in[6], float out[6]) {
  tfl::Quat g = tfl::quat_from_rpy(in[2], -in[0], -in[1]);  // createQuaternionMsgFromRollPitchYaw(rz,-rx,-ry)
  // message orientation = (-g.y, -g.z, g.x, g.w); receiver builds Quaternion(o.z, -o.x, -o.y, o.w)
  double ox = -g.y, oy = -g.z, oz = g.x, ow = g.w;
  tfl::Quat r = {oz, -ox, -oy, ow};
  double roll, pitch, yaw;
  tfl::rpy_from_quat(r, roll, pitch, yaw);
  out[0] = -pitch; out[1] = -yaw; out[2] = roll;
  out[3] = in[3]; out[4] = in[4]; out[5] = in[5];  // float -> float64 -> float is exact
}

// transformAssociateToMap, identical text in laserMapping.cpp:110-197 and
// transformMaintenance.cpp:60-145
inline void transformAssociateToMap(const float transformSum[6], const float transformBefMapped[6],
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

// ------------------------------------------------------------------------------------------
// laserMapping.cpp
// ------------------------------------------------------------------------------------------
class LaserMapping {
 public:
  static const int laserCloudWidth = 21, laserCloudHeight = 11, laserCloudDepth = 21;  // :67-69
  static const int laserCloudNum = laserCloudWidth * laserCloudHeight * laserCloudDepth;
  static const int imuQueLength = 200;  // :104
  static const int stackFrameNum = 1, mapFrameNum = 5;  // :51-52
  explicit LaserMapping(const loam_params& p)
      : P(p), matA0(5, 3, CV_32F, cv::Scalar::all(0)), matB0(5, 1, CV_32F, cv::Scalar::all(-1)),
        matX0(3, 1, CV_32F, cv::Scalar::all(0)), matA1(3, 3, CV_32F, cv::Scalar::all(0)),
        matD1(1, 3, CV_32F, cv::Scalar::all(0)), matV1(3, 3, CV_32F, cv::Scalar::all(0)),
        matP(6, 6, CV_32F, cv::Scalar::all(0)) {
    for (int i = 0; i < 6; i++) transformSum[i] = transformIncre[i] = transformTobeMapped[i] = transformBefMapped[i] = transformAftMapped[i] = 0;
    for (int i = 0; i < imuQueLength; i++) { imuTime[i] = 0; imuRoll[i] = imuPitch[i] = 0; }
    laserCloudCornerStack.reset(new Cloud()); laserCloudSurfStack.reset(new Cloud());
    laserCloudCornerStack2.reset(new Cloud()); laserCloudSurfStack2.reset(new Cloud());
    laserCloudOri.reset(new Cloud()); coeffSel.reset(new Cloud());
    laserCloudSurround.reset(new Cloud()); laserCloudSurround2.reset(new Cloud());
    laserCloudCornerFromMap.reset(new Cloud()); laserCloudSurfFromMap.reset(new Cloud());
    laserCloudFullRes.reset(new Cloud());
    laserCloudCornerArray.resize(laserCloudNum); laserCloudSurfArray.resize(laserCloudNum);
    laserCloudCornerArray2.resize(laserCloudNum); laserCloudSurfArray2.resize(laserCloudNum);
    for (int i = 0; i < laserCloudNum; i++) {  // :397-402
      laserCloudCornerArray[i].reset(new Cloud()); laserCloudSurfArray[i].reset(new Cloud());
      laserCloudCornerArray2[i].reset(new Cloud()); laserCloudSurfArray2[i].reset(new Cloud());
    }
    kdtreeCornerFromMap.reset(new pcl::KdTreeFLANN<PointType>()); kdtreeSurfFromMap.reset(new pcl::KdTreeFLANN<PointType>());
    downSizeFilterCorner.setLeafSize(0.2, 0.2, 0.2);  // :388-392
    downSizeFilterSurf.setLeafSize(0.4, 0.4, 0.4);
  }
  loam_params P;
  const float scanPeriod = 0.1;  // :49
  double timeLaserOdometry = 0;
  int laserCloudCenWidth = 10, laserCloudCenHeight = 5, laserCloudCenDepth = 10;  // :64-66
  int laserCloudValidInd[125], laserCloudSurroundInd[125];
  Cloud::Ptr laserCloudCornerStack, laserCloudSurfStack, laserCloudCornerStack2, laserCloudSurfStack2;
  Cloud::Ptr laserCloudOri, coeffSel, laserCloudSurround, laserCloudSurround2;
  Cloud::Ptr laserCloudCornerFromMap, laserCloudSurfFromMap, laserCloudFullRes;
  std::vector<Cloud::Ptr> laserCloudCornerArray, laserCloudSurfArray, laserCloudCornerArray2, laserCloudSurfArray2;
  pcl::KdTreeFLANN<PointType>::Ptr kdtreeCornerFromMap, kdtreeSurfFromMap;
  float transformSum[6], transformIncre[6], transformTobeMapped[6], transformBefMapped[6], transformAftMapped[6];
  int imuPointerFront = 0, imuPointerLast = -1;
  double imuTime[imuQueLength];
  float imuRoll[imuQueLength], imuPitch[imuQueLength];
  cv::Mat matA0, matB0, matX0, matA1, matD1, matV1;
  bool isDegenerate = false;
  cv::Mat matP;
  pcl::VoxelGrid<PointType> downSizeFilterCorner, downSizeFilterSurf;
  int frameCount = stackFrameNum - 1, mapFrameCount = mapFrameNum - 1;  // :404-405
  // outputs / stats
  bool surroundPublished = false;
  int lastIters = 0, lastRows = 0, lastValidNum = 0;
  int lastCornerFromMapNum = 0, lastSurfFromMapNum = 0, lastCornerStackNum = 0, lastSurfStackNum = 0;
  // NN correspondences of the LAST LM iteration (parity): 5 indices per query or -1
  std::vector<int> cornerNN, surfNN;

  // :199-232
  void transformUpdate() {
    if (imuPointerLast >= 0) {
      float imuRollLast = 0, imuPitchLast = 0;
      while (imuPointerFront != imuPointerLast) {
        if (timeLaserOdometry + scanPeriod < imuTime[imuPointerFront]) break;
        imuPointerFront = (imuPointerFront + 1) % imuQueLength;
      }
      if (timeLaserOdometry + scanPeriod > imuTime[imuPointerFront]) {
        imuRollLast = imuRoll[imuPointerFront];
        imuPitchLast = imuPitch[imuPointerFront];
      } else {
        int imuPointerBack = (imuPointerFront + imuQueLength - 1) % imuQueLength;
        float ratioFront = (timeLaserOdometry + scanPeriod - imuTime[imuPointerBack]) / (imuTime[imuPointerFront] - imuTime[imuPointerBack]);
        float ratioBack = (imuTime[imuPointerFront] - timeLaserOdometry - scanPeriod) / (imuTime[imuPointerFront] - imuTime[imuPointerBack]);
        imuRollLast = imuRoll[imuPointerFront] * ratioFront + imuRoll[imuPointerBack] * ratioBack;
        imuPitchLast = imuPitch[imuPointerFront] * ratioFront + imuPitch[imuPointerBack] * ratioBack;
      }
      transformTobeMapped[0] = 0.998 * transformTobeMapped[0] + 0.002 * imuPitchLast;
      transformTobeMapped[2] = 0.998 * transformTobeMapped[2] + 0.002 * imuRollLast;
    }
    for (int i = 0; i < 6; i++) {
      transformBefMapped[i] = transformSum[i];
      transformAftMapped[i] = transformTobeMapped[i];
    }
  }
  // :234-252
  void pointAssociateToMap(PointType const* const pi, PointType* const po) {
    float x1 = cos(transformTobeMapped[2]) * pi->x - sin(transformTobeMapped[2]) * pi->y;
    float y1 = sin(transformTobeMapped[2]) * pi->x + cos(transformTobeMapped[2]) * pi->y;
    float z1 = pi->z;
    float x2 = x1;
    float y2 = cos(transformTobeMapped[0]) * y1 - sin(transformTobeMapped[0]) * z1;
    float z2 = sin(transformTobeMapped[0]) * y1 + cos(transformTobeMapped[0]) * z1;
    po->x = cos(transformTobeMapped[1]) * x2 + sin(transformTobeMapped[1]) * z2 + transformTobeMapped[3];
    po->y = y2 + transformTobeMapped[4];
    po->z = -sin(transformTobeMapped[1]) * x2 + cos(transformTobeMapped[1]) * z2 + transformTobeMapped[5];
    po->intensity = pi->intensity;
  }
  // :254-272
  void pointAssociateTobeMapped(PointType const* const pi, PointType* const po) {
    float x1 = cos(transformTobeMapped[1]) * (pi->x - transformTobeMapped[3]) - sin(transformTobeMapped[1]) * (pi->z - transformTobeMapped[5]);
    float y1 = pi->y - transformTobeMapped[4];
    float z1 = sin(transformTobeMapped[1]) * (pi->x - transformTobeMapped[3]) + cos(transformTobeMapped[1]) * (pi->z - transformTobeMapped[5]);
    float x2 = x1;
    float y2 = cos(transformTobeMapped[0]) * y1 + sin(transformTobeMapped[0]) * z1;
    float z2 = -sin(transformTobeMapped[0]) * y1 + cos(transformTobeMapped[0]) * z1;
    po->x = cos(transformTobeMapped[2]) * x2 + sin(transformTobeMapped[2]) * y2;
    po->y = -sin(transformTobeMapped[2]) * x2 + cos(transformTobeMapped[2]) * y2;
    po->z = z2;
    po->intensity = pi->intensity;
  }
  // imuHandler :323-335
  void imu(double stamp, const double quat[4]) {
    double roll, pitch, yaw;
    tfl::Quat q = {quat[0], quat[1], quat[2], quat[3]};
    tfl::rpy_from_quat(q, roll, pitch, yaw);
    imuPointerLast = (imuPointerLast + 1) % imuQueLength;
    imuTime[imuPointerLast] = stamp;
    imuRoll[imuPointerLast] = roll;
    imuPitch[imuPointerLast] = pitch;
  }

  // shift helpers for :454-614.  dir=+1 moves cube contents toward higher index along `axis`.
  void shiftCubes(int axis, int dir) {
    const int W = laserCloudWidth, H = laserCloudHeight, D = laserCloudDepth;
    int dims[3] = {W, H, D};
    int n = dims[axis];
    for (int a = 0; a < dims[(axis + 1) % 3]; a++)
      for (int b = 0; b < dims[(axis + 2) % 3]; b++) {
        auto idx = [&](int t) {
          int c[3];
          c[axis] = t; c[(axis + 1) % 3] = a; c[(axis + 2) % 3] = b;
          return c[0] + W * c[1] + W * H * c[2];
        };
        if (dir > 0) {
          Cloud::Ptr cc = laserCloudCornerArray[idx(n - 1)], cs = laserCloudSurfArray[idx(n - 1)];
          for (int t = n - 1; t >= 1; t--) {
            laserCloudCornerArray[idx(t)] = laserCloudCornerArray[idx(t - 1)];
            laserCloudSurfArray[idx(t)] = laserCloudSurfArray[idx(t - 1)];
          }
          laserCloudCornerArray[idx(0)] = cc; laserCloudSurfArray[idx(0)] = cs;
          cc->clear(); cs->clear();
        } else {
          Cloud::Ptr cc = laserCloudCornerArray[idx(0)], cs = laserCloudSurfArray[idx(0)];
          for (int t = 0; t < n - 1; t++) {
            laserCloudCornerArray[idx(t)] = laserCloudCornerArray[idx(t + 1)];
            laserCloudSurfArray[idx(t)] = laserCloudSurfArray[idx(t + 1)];
          }
          laserCloudCornerArray[idx(n - 1)] = cc; laserCloudSurfArray[idx(n - 1)] = cs;
          cc->clear(); cs->clear();
        }
      }
  }

  // one pass of the main loop body :411-1096.  transformSumIn = /laser_odom_to_init after the
  // message round trip (:304-318).
  bool process(const LOOutput& in, const float transformSumIn[6]) {
    *laserCloudFullRes = in.fullRes;
    timeLaserOdometry = in.stamp;
    for (int i = 0; i < 6; i++) transformSum[i] = transformSumIn[i];
    const Cloud& laserCloudCornerLast = in.cornerLast;
    const Cloud& laserCloudSurfLast = in.surfLast;
    PointType pointOri, pointSel, coeff;
    std::vector<int> pointSearchInd;
    std::vector<float> pointSearchSqDis;
    surroundPublished = false;
    lastIters = 0; lastRows = 0;

    frameCount++;
    if (frameCount < stackFrameNum) return false;
    transformAssociateToMap(transformSum, transformBefMapped, transformAftMapped, transformIncre, transformTobeMapped);  // :422
    int laserCloudCornerLastNum = (int)laserCloudCornerLast.points.size();
    for (int i = 0; i < laserCloudCornerLastNum; i++) {
      pointAssociateToMap(&laserCloudCornerLast.points[i], &pointSel);
      laserCloudCornerStack2->push_back(pointSel);
    }
    int laserCloudSurfLastNum = (int)laserCloudSurfLast.points.size();
    for (int i = 0; i < laserCloudSurfLastNum; i++) {
      pointAssociateToMap(&laserCloudSurfLast.points[i], &pointSel);
      laserCloudSurfStack2->push_back(pointSel);
    }
    frameCount = 0;

    PointType pointOnYAxis;  // :440-444
    pointOnYAxis.x = 0.0; pointOnYAxis.y = 10.0; pointOnYAxis.z = 0.0;
    pointAssociateToMap(&pointOnYAxis, &pointOnYAxis);

    int centerCubeI = int((transformTobeMapped[3] + 25.0) / 50.0) + laserCloudCenWidth;  // :446-452
    int centerCubeJ = int((transformTobeMapped[4] + 25.0) / 50.0) + laserCloudCenHeight;
    int centerCubeK = int((transformTobeMapped[5] + 25.0) / 50.0) + laserCloudCenDepth;
    if (transformTobeMapped[3] + 25.0 < 0) centerCubeI--;
    if (transformTobeMapped[4] + 25.0 < 0) centerCubeJ--;
    if (transformTobeMapped[5] + 25.0 < 0) centerCubeK--;

    while (centerCubeI < 3) { shiftCubes(0, +1); centerCubeI++; laserCloudCenWidth++; }                      // :454-479
    while (centerCubeI >= laserCloudWidth - 3) { shiftCubes(0, -1); centerCubeI--; laserCloudCenWidth--; }   // :481-506
    while (centerCubeJ < 3) { shiftCubes(1, +1); centerCubeJ++; laserCloudCenHeight++; }                     // :508-533
    while (centerCubeJ >= laserCloudHeight - 3) { shiftCubes(1, -1); centerCubeJ--; laserCloudCenHeight--; } // :535-560
    while (centerCubeK < 3) { shiftCubes(2, +1); centerCubeK++; laserCloudCenDepth++; }                      // :562-587
    while (centerCubeK >= laserCloudDepth - 3) { shiftCubes(2, -1); centerCubeK--; laserCloudCenDepth--; }   // :589-614

    int laserCloudValidNum = 0, laserCloudSurroundNum = 0;  // :616-672
    for (int i = centerCubeI - 2; i <= centerCubeI + 2; i++) {
      for (int j = centerCubeJ - 2; j <= centerCubeJ + 2; j++) {
        for (int k = centerCubeK - 2; k <= centerCubeK + 2; k++) {
          if (i >= 0 && i < laserCloudWidth && j >= 0 && j < laserCloudHeight && k >= 0 && k < laserCloudDepth) {
            float centerX = 50.0 * (i - laserCloudCenWidth);
            float centerY = 50.0 * (j - laserCloudCenHeight);
            float centerZ = 50.0 * (k - laserCloudCenDepth);
            bool isInLaserFOV = false;
            for (int ii = -1; ii <= 1; ii += 2) {
              for (int jj = -1; jj <= 1; jj += 2) {
                for (int kk = -1; kk <= 1; kk += 2) {
                  float cornerX = centerX + 25.0 * ii;
                  float cornerY = centerY + 25.0 * jj;
                  float cornerZ = centerZ + 25.0 * kk;
                  float squaredSide1 = (transformTobeMapped[3] - cornerX) * (transformTobeMapped[3] - cornerX)
                                     + (transformTobeMapped[4] - cornerY) * (transformTobeMapped[4] - cornerY)
                                     + (transformTobeMapped[5] - cornerZ) * (transformTobeMapped[5] - cornerZ);
                  float squaredSide2 = (pointOnYAxis.x - cornerX) * (pointOnYAxis.x - cornerX)
                                     + (pointOnYAxis.y - cornerY) * (pointOnYAxis.y - cornerY)
                                     + (pointOnYAxis.z - cornerZ) * (pointOnYAxis.z - cornerZ);
                  float check1 = 100.0 + squaredSide1 - squaredSide2 - 10.0 * sqrt(3.0) * sqrt(squaredSide1);
                  float check2 = 100.0 + squaredSide1 - squaredSide2 + 10.0 * sqrt(3.0) * sqrt(squaredSide1);
                  if (check1 < 0 && check2 > 0) isInLaserFOV = true;
                }
              }
            }
            if (isInLaserFOV) {
              laserCloudValidInd[laserCloudValidNum] = i + laserCloudWidth * j + laserCloudWidth * laserCloudHeight * k;
              laserCloudValidNum++;
            }
            laserCloudSurroundInd[laserCloudSurroundNum] = i + laserCloudWidth * j + laserCloudWidth * laserCloudHeight * k;
            laserCloudSurroundNum++;
          }
        }
      }
    }
    lastValidNum = laserCloudValidNum;

    laserCloudCornerFromMap->clear();  // :674-681
    laserCloudSurfFromMap->clear();
    for (int i = 0; i < laserCloudValidNum; i++) {
      *laserCloudCornerFromMap += *laserCloudCornerArray[laserCloudValidInd[i]];
      *laserCloudSurfFromMap += *laserCloudSurfArray[laserCloudValidInd[i]];
    }
    int laserCloudCornerFromMapNum = (int)laserCloudCornerFromMap->points.size();
    int laserCloudSurfFromMapNum = (int)laserCloudSurfFromMap->points.size();

    int laserCloudCornerStackNum2 = (int)laserCloudCornerStack2->points.size();  // :683-691
    for (int i = 0; i < laserCloudCornerStackNum2; i++)
      pointAssociateTobeMapped(&laserCloudCornerStack2->points[i], &laserCloudCornerStack2->points[i]);
    int laserCloudSurfStackNum2 = (int)laserCloudSurfStack2->points.size();
    for (int i = 0; i < laserCloudSurfStackNum2; i++)
      pointAssociateTobeMapped(&laserCloudSurfStack2->points[i], &laserCloudSurfStack2->points[i]);

    laserCloudCornerStack->clear();  // :693-704
    downSizeFilterCorner.setInputCloud(laserCloudCornerStack2);
    downSizeFilterCorner.filter(*laserCloudCornerStack);
    int laserCloudCornerStackNum = (int)laserCloudCornerStack->points.size();
    laserCloudSurfStack->clear();
    downSizeFilterSurf.setInputCloud(laserCloudSurfStack2);
    downSizeFilterSurf.filter(*laserCloudSurfStack);
    int laserCloudSurfStackNum = (int)laserCloudSurfStack->points.size();
    laserCloudCornerStack2->clear();
    laserCloudSurfStack2->clear();
    lastCornerFromMapNum = laserCloudCornerFromMapNum; lastSurfFromMapNum = laserCloudSurfFromMapNum;
    lastCornerStackNum = laserCloudCornerStackNum; lastSurfStackNum = laserCloudSurfStackNum;
    cornerNN.assign((size_t)laserCloudCornerStackNum * 5, -1);
    surfNN.assign((size_t)laserCloudSurfStackNum * 5, -1);

    if (laserCloudCornerFromMapNum > 10 && laserCloudSurfFromMapNum > 100) {  // :706
      kdtreeCornerFromMap->setInputCloud(laserCloudCornerFromMap);
      kdtreeSurfFromMap->setInputCloud(laserCloudSurfFromMap);
      for (int iterCount = 0; iterCount < P.lm_max_iter; iterCount++) {  // :710
        lastIters = iterCount + 1;
        laserCloudOri->clear();
        coeffSel->clear();
        for (int i = 0; i < laserCloudCornerStackNum; i++) {  // :714-819
          pointOri = laserCloudCornerStack->points[i];
          pointAssociateToMap(&pointOri, &pointSel);
          kdtreeCornerFromMap->nearestKSearch(pointSel, 5, pointSearchInd, pointSearchSqDis);
          for (int j = 0; j < 5; j++) cornerNN[(size_t)i * 5 + j] = -1;
          if (pointSearchSqDis[4] < 1.0) {
            for (int j = 0; j < 5; j++) cornerNN[(size_t)i * 5 + j] = pointSearchInd[j];
            float cx = 0; float cy = 0; float cz = 0;
            for (int j = 0; j < 5; j++) {
              cx += laserCloudCornerFromMap->points[pointSearchInd[j]].x;
              cy += laserCloudCornerFromMap->points[pointSearchInd[j]].y;
              cz += laserCloudCornerFromMap->points[pointSearchInd[j]].z;
            }
            cx /= 5; cy /= 5; cz /= 5;
            float a11 = 0; float a12 = 0; float a13 = 0; float a22 = 0; float a23 = 0; float a33 = 0;
            for (int j = 0; j < 5; j++) {
              float ax = laserCloudCornerFromMap->points[pointSearchInd[j]].x - cx;
              float ay = laserCloudCornerFromMap->points[pointSearchInd[j]].y - cy;
              float az = laserCloudCornerFromMap->points[pointSearchInd[j]].z - cz;
              a11 += ax * ax; a12 += ax * ay; a13 += ax * az;
              a22 += ay * ay; a23 += ay * az; a33 += az * az;
            }
            a11 /= 5; a12 /= 5; a13 /= 5; a22 /= 5; a23 /= 5; a33 /= 5;
            matA1.at<float>(0, 0) = a11; matA1.at<float>(0, 1) = a12; matA1.at<float>(0, 2) = a13;
            matA1.at<float>(1, 0) = a12; matA1.at<float>(1, 1) = a22; matA1.at<float>(1, 2) = a23;
            matA1.at<float>(2, 0) = a13; matA1.at<float>(2, 1) = a23; matA1.at<float>(2, 2) = a33;
            cv::eigen(matA1, matD1, matV1);
            if (matD1.at<float>(0, 0) > 3 * matD1.at<float>(0, 1)) {
              float x0 = pointSel.x; float y0 = pointSel.y; float z0 = pointSel.z;
              float x1 = cx + 0.1 * matV1.at<float>(0, 0);
              float y1 = cy + 0.1 * matV1.at<float>(0, 1);
              float z1 = cz + 0.1 * matV1.at<float>(0, 2);
              float x2 = cx - 0.1 * matV1.at<float>(0, 0);
              float y2 = cy - 0.1 * matV1.at<float>(0, 1);
              float z2 = cz - 0.1 * matV1.at<float>(0, 2);
              float a012 = sqrt(((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                         * ((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                         + ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
                         * ((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
                         + ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))
                         * ((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1)));
              float l12 = sqrt((x1 - x2)*(x1 - x2) + (y1 - y2)*(y1 - y2) + (z1 - z2)*(z1 - z2));
              float la = ((y1 - y2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                       + (z1 - z2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))) / a012 / l12;
              float lb = -((x1 - x2)*((x0 - x1)*(y0 - y2) - (x0 - x2)*(y0 - y1))
                       - (z1 - z2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
              float lc = -((x1 - x2)*((x0 - x1)*(z0 - z2) - (x0 - x2)*(z0 - z1))
                       + (y1 - y2)*((y0 - y1)*(z0 - z2) - (y0 - y2)*(z0 - z1))) / a012 / l12;
              float ld2 = a012 / l12;
              float s = 1 - 0.9 * fabs(ld2);
              coeff.x = s * la; coeff.y = s * lb; coeff.z = s * lc; coeff.intensity = s * ld2;
              if (s > 0.1) {
                laserCloudOri->push_back(pointOri);
                coeffSel->push_back(coeff);
              }
            }
          }
        }
        for (int i = 0; i < laserCloudSurfStackNum; i++) {  // :821-877
          pointOri = laserCloudSurfStack->points[i];
          pointAssociateToMap(&pointOri, &pointSel);
          kdtreeSurfFromMap->nearestKSearch(pointSel, 5, pointSearchInd, pointSearchSqDis);
          for (int j = 0; j < 5; j++) surfNN[(size_t)i * 5 + j] = -1;
          if (pointSearchSqDis[4] < 1.0) {
            for (int j = 0; j < 5; j++) surfNN[(size_t)i * 5 + j] = pointSearchInd[j];
            for (int j = 0; j < 5; j++) {
              matA0.at<float>(j, 0) = laserCloudSurfFromMap->points[pointSearchInd[j]].x;
              matA0.at<float>(j, 1) = laserCloudSurfFromMap->points[pointSearchInd[j]].y;
              matA0.at<float>(j, 2) = laserCloudSurfFromMap->points[pointSearchInd[j]].z;
            }
            cv::solve(matA0, matB0, matX0, cv::DECOMP_QR);
            float pa = matX0.at<float>(0, 0);
            float pb = matX0.at<float>(1, 0);
            float pc = matX0.at<float>(2, 0);
            float pd = 1;
            float ps = sqrt(pa * pa + pb * pb + pc * pc);
            pa /= ps; pb /= ps; pc /= ps; pd /= ps;
            bool planeValid = true;
            for (int j = 0; j < 5; j++) {
              if (fabs(pa * laserCloudSurfFromMap->points[pointSearchInd[j]].x +
                       pb * laserCloudSurfFromMap->points[pointSearchInd[j]].y +
                       pc * laserCloudSurfFromMap->points[pointSearchInd[j]].z + pd) > 0.2) {
                planeValid = false;
                break;
              }
            }
            if (planeValid) {
              float pd2 = pa * pointSel.x + pb * pointSel.y + pc * pointSel.z + pd;
              float s = 1 - 0.9 * fabs(pd2) / sqrt(sqrt(pointSel.x * pointSel.x + pointSel.y * pointSel.y + pointSel.z * pointSel.z));
              coeff.x = s * pa; coeff.y = s * pb; coeff.z = s * pc; coeff.intensity = s * pd2;
              if (s > 0.1) {
                laserCloudOri->push_back(pointOri);
                coeffSel->push_back(coeff);
              }
            }
          }
        }
        float srx = sin(transformTobeMapped[0]); float crx = cos(transformTobeMapped[0]);  // :879-884
        float sry = sin(transformTobeMapped[1]); float cry = cos(transformTobeMapped[1]);
        float srz = sin(transformTobeMapped[2]); float crz = cos(transformTobeMapped[2]);
        int laserCloudSelNum = (int)laserCloudOri->points.size();
        lastRows = laserCloudSelNum;
        if (laserCloudSelNum < 50) continue;  // :887

        cv::Mat matA(laserCloudSelNum, 6, CV_32F, cv::Scalar::all(0));
        cv::Mat matAt(6, laserCloudSelNum, CV_32F, cv::Scalar::all(0));
        cv::Mat matAtA(6, 6, CV_32F, cv::Scalar::all(0));
        cv::Mat matB(laserCloudSelNum, 1, CV_32F, cv::Scalar::all(0));
        cv::Mat matAtB(6, 1, CV_32F, cv::Scalar::all(0));
        cv::Mat matX(6, 1, CV_32F, cv::Scalar::all(0));
        for (int i = 0; i < laserCloudSelNum; i++) {  // :897-921
          pointOri = laserCloudOri->points[i];
          coeff = coeffSel->points[i];
          float arx = (crx*sry*srz*pointOri.x + crx*crz*sry*pointOri.y - srx*sry*pointOri.z) * coeff.x
                    + (-srx*srz*pointOri.x - crz*srx*pointOri.y - crx*pointOri.z) * coeff.y
                    + (crx*cry*srz*pointOri.x + crx*cry*crz*pointOri.y - cry*srx*pointOri.z) * coeff.z;
          float ary = ((cry*srx*srz - crz*sry)*pointOri.x
                    + (sry*srz + cry*crz*srx)*pointOri.y + crx*cry*pointOri.z) * coeff.x
                    + ((-cry*crz - srx*sry*srz)*pointOri.x
                    + (cry*srz - crz*srx*sry)*pointOri.y - crx*sry*pointOri.z) * coeff.z;
          float arz = ((crz*srx*sry - cry*srz)*pointOri.x + (-cry*crz-srx*sry*srz)*pointOri.y)*coeff.x
                    + (crx*crz*pointOri.x - crx*srz*pointOri.y) * coeff.y
                    + ((sry*srz + cry*crz*srx)*pointOri.x + (crz*sry-cry*srx*srz)*pointOri.y)*coeff.z;
          matA.at<float>(i, 0) = arx; matA.at<float>(i, 1) = ary; matA.at<float>(i, 2) = arz;
          matA.at<float>(i, 3) = coeff.x; matA.at<float>(i, 4) = coeff.y; matA.at<float>(i, 5) = coeff.z;
          matB.at<float>(i, 0) = -coeff.intensity;
        }
        cv::transpose(matA, matAt);  // :922-925
        matAtA = matAt * matA;
        matAtB = matAt * matB;
        cv::solve(matAtA, matAtB, matX, cv::DECOMP_QR);
        if (iterCount == 0) {  // :927-948
          cv::Mat matE(1, 6, CV_32F, cv::Scalar::all(0));
          cv::Mat matV(6, 6, CV_32F, cv::Scalar::all(0));
          cv::Mat matV2(6, 6, CV_32F, cv::Scalar::all(0));
          cv::eigen(matAtA, matE, matV);
          matV.copyTo(matV2);
          isDegenerate = false;
          float eignThre[6] = {100, 100, 100, 100, 100, 100};
          for (int i = 5; i >= 0; i--) {
            if (matE.at<float>(0, i) < eignThre[i]) {
              for (int j = 0; j < 6; j++) matV2.at<float>(i, j) = 0;
              isDegenerate = true;
            } else {
              break;
            }
          }
          matP = matV.inv() * matV2;
        }
        if (isDegenerate) {  // :950-954
          cv::Mat matX2(6, 1, CV_32F, cv::Scalar::all(0));
          matX.copyTo(matX2);
          matX = matP * matX2;
        }
        transformTobeMapped[0] += matX.at<float>(0, 0); transformTobeMapped[1] += matX.at<float>(1, 0);  // :956-961
        transformTobeMapped[2] += matX.at<float>(2, 0); transformTobeMapped[3] += matX.at<float>(3, 0);
        transformTobeMapped[4] += matX.at<float>(4, 0); transformTobeMapped[5] += matX.at<float>(5, 0);
        float deltaR = sqrt(pow(rad2deg(matX.at<float>(0, 0)), 2) + pow(rad2deg(matX.at<float>(1, 0)), 2) +
                            pow(rad2deg(matX.at<float>(2, 0)), 2));
        float deltaT = sqrt(pow(matX.at<float>(3, 0) * 100, 2) + pow(matX.at<float>(4, 0) * 100, 2) +
                            pow(matX.at<float>(5, 0) * 100, 2));
        if (deltaR < 0.05 && deltaT < 0.05) break;
      }
      transformUpdate();  // :977
    }

    for (int i = 0; i < laserCloudCornerStackNum; i++) {  // :980-997
      pointAssociateToMap(&laserCloudCornerStack->points[i], &pointSel);
      int cubeI = int((pointSel.x + 25.0) / 50.0) + laserCloudCenWidth;
      int cubeJ = int((pointSel.y + 25.0) / 50.0) + laserCloudCenHeight;
      int cubeK = int((pointSel.z + 25.0) / 50.0) + laserCloudCenDepth;
      if (pointSel.x + 25.0 < 0) cubeI--;
      if (pointSel.y + 25.0 < 0) cubeJ--;
      if (pointSel.z + 25.0 < 0) cubeK--;
      if (cubeI >= 0 && cubeI < laserCloudWidth && cubeJ >= 0 && cubeJ < laserCloudHeight && cubeK >= 0 && cubeK < laserCloudDepth) {
        int cubeInd = cubeI + laserCloudWidth * cubeJ + laserCloudWidth * laserCloudHeight * cubeK;
        laserCloudCornerArray[cubeInd]->push_back(pointSel);
      }
    }
    for (int i = 0; i < laserCloudSurfStackNum; i++) {  // :999-1016
      pointAssociateToMap(&laserCloudSurfStack->points[i], &pointSel);
      int cubeI = int((pointSel.x + 25.0) / 50.0) + laserCloudCenWidth;
      int cubeJ = int((pointSel.y + 25.0) / 50.0) + laserCloudCenHeight;
      int cubeK = int((pointSel.z + 25.0) / 50.0) + laserCloudCenDepth;
      if (pointSel.x + 25.0 < 0) cubeI--;
      if (pointSel.y + 25.0 < 0) cubeJ--;
      if (pointSel.z + 25.0 < 0) cubeK--;
      if (cubeI >= 0 && cubeI < laserCloudWidth && cubeJ >= 0 && cubeJ < laserCloudHeight && cubeK >= 0 && cubeK < laserCloudDepth) {
        int cubeInd = cubeI + laserCloudWidth * cubeJ + laserCloudWidth * laserCloudHeight * cubeK;
        laserCloudSurfArray[cubeInd]->push_back(pointSel);
      }
    }
    for (int i = 0; i < laserCloudValidNum; i++) {  // :1018-1036
      int ind = laserCloudValidInd[i];
      laserCloudCornerArray2[ind]->clear();
      downSizeFilterCorner.setInputCloud(laserCloudCornerArray[ind]);
      downSizeFilterCorner.filter(*laserCloudCornerArray2[ind]);
      laserCloudSurfArray2[ind]->clear();
      downSizeFilterSurf.setInputCloud(laserCloudSurfArray[ind]);
      downSizeFilterSurf.filter(*laserCloudSurfArray2[ind]);
      std::swap(laserCloudCornerArray[ind], laserCloudCornerArray2[ind]);
      std::swap(laserCloudSurfArray[ind], laserCloudSurfArray2[ind]);
    }
    mapFrameCount++;  // :1038-1058
    if (mapFrameCount >= mapFrameNum) {
      mapFrameCount = 0;
      laserCloudSurround2->clear();
      for (int i = 0; i < laserCloudSurroundNum; i++) {
        int ind = laserCloudSurroundInd[i];
        *laserCloudSurround2 += *laserCloudCornerArray[ind];
        *laserCloudSurround2 += *laserCloudSurfArray[ind];
      }
      laserCloudSurround->clear();
      downSizeFilterCorner.setInputCloud(laserCloudSurround2);
      downSizeFilterCorner.filter(*laserCloudSurround);
      surroundPublished = true;
    }
    int laserCloudFullResNum = (int)laserCloudFullRes->points.size();  // :1060-1063
    for (int i = 0; i < laserCloudFullResNum; i++)
      pointAssociateToMap(&laserCloudFullRes->points[i], &laserCloudFullRes->points[i]);
    return true;
  }
};

// /aft_mapped_to_init hand-off: laserMapping.cpp:1071-1087 -> transformMaintenance.cpp:182-203
inline void aft_mapped_msg_roundtrip(const float aft[6], float out[6]) { odometry_msg_roundtrip(aft, out); }

// ------------------------------------------------------------------------------------------
// transformMaintenance.cpp
// ------------------------------------------------------------------------------------------
class TransformMaintenance {
 public:
  float transformSum[6], transformIncre[6], transformMapped[6], transformBefMapped[6], transformAftMapped[6];  // :49-53
  TransformMaintenance() {
    for (int i = 0; i < 6; i++) transformSum[i] = transformIncre[i] = transformMapped[i] = transformBefMapped[i] = transformAftMapped[i] = 0;
  }
  // odomAftMappedHandler :182-203 (aft after the quaternion round trip; bef travels in twist as float64)
  void onMapping(const float aftMappedMsg[6], const float befMapped[6]) {
    for (int i = 0; i < 6; i++) { transformAftMapped[i] = aftMappedMsg[i]; transformBefMapped[i] = befMapped[i]; }
  }
  // laserOdometryHandler :147-180
  void onOdometry(const float transformSumMsg[6], float integrated[6]) {
    for (int i = 0; i < 6; i++) transformSum[i] = transformSumMsg[i];
    transformAssociateToMap(transformSum, transformBefMapped, transformAftMapped, transformIncre, transformMapped);
    for (int i = 0; i < 6; i++) integrated[i] = transformMapped[i];
  }
};

}  // namespace oracle
