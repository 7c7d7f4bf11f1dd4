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
#include <vector>
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

// HDL-64E data packet (HDL-64E S2 manual app. E), ideal calibration: 12 blocks, flag 0xEEFF = upper block
// (lasers 32..63), 0xDDFF = lower block (lasers 0..31), one azimuth per block, laser k at elevation
// lower + (upper - lower) * k / 63 degrees (the ring model of BASELINE.json configs[1]); distances in 2 mm units
inline void decode_hdl64_packet(const uint8_t* pkt, float lower_deg, float upper_deg, float* xyz /* 384 x 3 */) {
  const float nanv = std::numeric_limits<float>::quiet_NaN();
  for (int block = 0; block < 12; block++) {
    const uint8_t* b = pkt + 100 * block;
    const unsigned flag = b[0] | (b[1] << 8);
    const int azimuth = (b[2] | (b[3] << 8)) % 36000;
    const bool upper = flag == 0xEEFFu, lower = flag == 0xDDFFu;
    for (int k = 0; k < 32; k++) {
      float* out = xyz + 3 * (32 * block + k);
      const unsigned raw = b[4 + 3 * k] | (b[5 + 3 * k] << 8);
      if ((!upper && !lower) || raw == 0) { out[0] = out[1] = out[2] = nanv; continue; }
      const int laser = upper ? 32 + k : k;
      double elev = (double)lower_deg + ((double)upper_deg - (double)lower_deg) * (double)laser / 63.0;
      float distance = (float)raw * 0.002f;
      float rotation = (float)((double)azimuth * (M_PI / 18000.0));
      float vert = (float)(elev * (M_PI / 180.0));
      float cos_vert = omath::cos(vert), sin_vert = omath::sin(vert);
      float cos_rot = omath::cos(rotation), sin_rot = omath::sin(rotation);
      float xy_distance = distance * cos_vert;
      out[0] = xy_distance * cos_rot;     // ROS axes: x forward, y left, z up
      out[1] = -(xy_distance * sin_rot);
      out[2] = distance * sin_vert;
    }
  }
}

// SR-1 from packet metadata (the product's loam_params.packet_metadata = 1; there is no such mode in the
// reference, which only ever sees the decoded cloud): per slot of a revolution of `n_packets` packets
//   ring    = position of the slot's laser in elevation order (VLP-16: nominal elevation w deg -> w > 0 ? w : w + 15,
//             what scanRegistration.cpp:249-256 yields for exactly that elevation; HDL-64E: the laser number)
//   relTime = unwrapped azimuth offset from the first return / that of the last return, azimuths in the
//             packet's 0.01 deg units: offset = (az - az_first + 36000) % 36000, and + 36000 when the slot lies in
//             the second half of the slot range [first return, last return] while its offset is below 18000
// and the decoded xyz (NaN for a slot without return, ring -1).  kind: 0 VLP-16, 1 HDL-64E.
inline void sr1_meta_from_packets(const uint8_t* packets, int n_packets, int kind, float lower_deg, float upper_deg,
                                  float* xyz, int* ring, float* reltime) {
  static const int vlp_deg[16] = {-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15};
  const int n = n_packets * 384;
  std::vector<int> az(n, 0);
  for (int p = 0; p < n_packets; p++) {
    const uint8_t* pkt = packets + (size_t)1206 * p;
    if (kind == 0) decode_vlp16_packet(pkt, xyz + (size_t)384 * 3 * p);
    else decode_hdl64_packet(pkt, lower_deg, upper_deg, xyz + (size_t)384 * 3 * p);
    int last_diff = 0;
    for (int block = 0; block < 12; block++) {
      const uint8_t* b = pkt + 100 * block;
      const unsigned flag = b[0] | (b[1] << 8);
      const int azimuth = b[2] | (b[3] << 8);
      int diff = last_diff;
      if (block < 11) { diff = (36000 + (b[102] | (b[103] << 8)) - azimuth) % 36000; last_diff = diff; }
      for (int k = 0; k < 32; k++) {
        const int s = 384 * p + 32 * block + k;
        if (kind == 0) {
          const int firing = k / 16, dsr = k % 16;
          float f = (float)azimuth + ((float)diff * ((float)dsr * 2.304f + (float)firing * 55.296f)) / 110.592f;
          int a = ((int)std::round(f)) % 36000;
          if (a < 0) a += 36000;
          az[s] = a;
          ring[s] = vlp_deg[dsr] > 0 ? vlp_deg[dsr] : vlp_deg[dsr] + 15;
        } else {
          az[s] = azimuth % 36000;
          ring[s] = flag == 0xEEFFu ? 32 + k : k;
        }
      }
    }
  }
  int first = -1, last = -1;
  for (int s = 0; s < n; s++) {
    const bool valid = std::isfinite(xyz[3 * s]);
    if (!valid) { ring[s] = -1; continue; }
    if (first < 0) first = s;
    last = s;
  }
  for (int s = 0; s < n; s++) reltime[s] = 0.f;
  if (first < 0) return;
  auto offset = [&](int s) {
    int d = (az[s] - az[first] + 36000) % 36000;
    if ((long long)(s - first) * 2 > (long long)(last - first) && d < 18000) d += 36000;
    return d;
  };
  const int span = offset(last);
  for (int s = first; s <= last; s++)
    if (ring[s] >= 0 && span > 0) reltime[s] = (float)offset(s) / (float)span;
}

}  // namespace oracle
