// ORACLE (test infrastructure).  CPU restatement of the Velodyne VLP-16 data-packet decoder that
// produces /velodyne_points for the reference (scanRegistration.cpp:675 subscribes to it).  The
// decoder itself is NOT under /root/reference: it is the un-vendored ROS package
// velodyne_pointcloud (ros-drivers/velodyne, rawdata.cc unpack_vlp16; version unpinned by the
// reference's package.xml) -- "parity unpinned" for this row.  What is restated is the published
// packet format (VLP-16 User Manual ch. 9: 12 blocks x [0xFFEE, azimuth, 2 x 16 x (distance, reflectivity)],
// 2 mm distance units, 0.01 deg azimuth units, 55.296 us per firing sequence, 2.304 us per laser)
// and the driver's azimuth interpolation, with an ideal calibration (vertical angles only).
// Written independently of loam_velodyne-1_b200/csrc/pk_decode.cuh: loops over the packet
// structure instead of decoding a flat point index.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include "defs/oracle_math.h"

namespace oracle {

inline void decode_vlp16_packet(const uint8_t* pkt, float* xyz /* 384 x 3 */) {
  static const int laser_deg[16] = {-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15};
  const float nanv = std::numeric_limits<float>::quiet_NaN();
  int last_diff = 0;
  for (int block = 0; block < 12; block++) {
    const uint8_t* b = pkt + 100 * block;
    const unsigned flag = b[0] | (b[1] << 8);
    const int azimuth = b[2] | (b[3] << 8);
    int azimuth_diff;
    if (block < 11) {
      const uint8_t* nb = b + 100;
      azimuth_diff = (36000 + (nb[2] | (nb[3] << 8)) - azimuth) % 36000;
      last_diff = azimuth_diff;
    } else {
      azimuth_diff = last_diff;
    }
    for (int firing = 0, k = 0; firing < 2; firing++) {
      for (int dsr = 0; dsr < 16; dsr++, k++) {
        float* out = xyz + 3 * (32 * block + k);
        const unsigned raw = b[4 + 3 * k] | (b[5 + 3 * k] << 8);
        if (flag != 0xEEFFu || raw == 0) { out[0] = out[1] = out[2] = nanv; continue; }
        float azimuth_corrected_f = (float)azimuth + ((float)azimuth_diff * ((float)dsr * 2.304f + (float)firing * 55.296f)) / 110.592f;
        int azimuth_corrected = ((int)std::round(azimuth_corrected_f)) % 36000;
        if (azimuth_corrected < 0) azimuth_corrected += 36000;
        float distance = (float)raw * 0.002f;
        float rotation = (float)((double)azimuth_corrected * (M_PI / 18000.0));
        float vert = (float)((double)laser_deg[dsr] * (M_PI / 180.0));
        float cos_vert = omath::cos(vert), sin_vert = omath::sin(vert);
        float cos_rot = omath::cos(rotation), sin_rot = omath::sin(rotation);
        float xy_distance = distance * cos_vert;
        float x = xy_distance * sin_rot;
        float y = xy_distance * cos_rot;
        float z = distance * sin_vert;
        out[0] = y;     // ROS axes: x forward, y left, z up
        out[1] = -x;
        out[2] = z;
      }
    }
  }
}

}  // namespace oracle
