// pk_decode.cuh -- raw-packet front end: Velodyne VLP-16 data packets ->
// the XYZ points /velodyne_points carries (what scanRegistration.cpp:675 subscribes to), decoded
// on the device so that a sweep costs 3.1 bytes per point of host->device traffic instead of 12.
//
// The decoder the reference depends on lives in the un-vendored ROS package velodyne_pointcloud
// (ros-drivers/velodyne, rawdata.cc `unpack_vlp16`, version unpinned); its arithmetic is restated
// here from the published packet format (VLP-16 User Manual, ch. 9) with an ideal calibration
// (vertical angles only, every other correction zero):
//   packet  = 12 blocks x 100 B + 4 B timestamp + 2 B factory        = 1206 B
//   block   = flag 0xFFEE (2 B, little endian 0xEEFF) + azimuth (u16, 0.01 deg) + 32 x {distance u16 (2 mm), reflectivity u8}
//             = two firing sequences of the 16 lasers, laser l at elevation (l odd ? l : l - 15) deg
//   azimuth of (block b, firing f, laser l) = az_b + diff_b * (2.304 l + 55.296 f) / 110.592, rounded to 0.01 deg,
//             diff_b = (36000 + az_{b+1} - az_b) % 36000, the last block of a packet re-uses diff of the one before
//   point   = d cos(w) cos(a), -d cos(w) sin(a), d sin(w)   (ROS axes: x forward, y left, z up; a clockwise)
//   distance 0 = no return: the slot is written as NaN and dropped by scanRegistration's NaN filter (:228)
// Every floating-point step is written once, in pk::decode_point, and used by the kernel and by the
// host (which needs the first and last valid point of the sweep, scanRegistration.cpp:230-232).
#pragma once
#include <math.h>
#include <stdint.h>
#include "common.cuh"

namespace loamgpu {
namespace pk {

constexpr int PACKET_BYTES = 1206, BLOCKS = 12, BLOCK_BYTES = 100, POINTS_PER_PACKET = 384;

__host__ __device__ __forceinline__ unsigned int u16le(const unsigned char* p) { return (unsigned int)p[0] | ((unsigned int)p[1] << 8); }

// point `idx` (0..383: block, firing, laser order) of one packet; false (out = NaN) for a no-return slot
__host__ __device__ inline bool decode_point(const unsigned char* pkt, int idx, float out[3]) {
  const int b = idx >> 5, f = (idx >> 4) & 1, l = idx & 15;
  const unsigned char* blk = pkt + b * BLOCK_BYTES;
  const unsigned int raw = u16le(blk + 4 + 3 * (idx & 31));
  if (raw == 0 || u16le(blk) != 0xEEFFu) {
#ifdef __CUDA_ARCH__
    out[0] = out[1] = out[2] = __int_as_float(0x7fc00000);
#else
    out[0] = out[1] = out[2] = NAN;
#endif
    return false;
  }
  const int az = (int)u16le(blk + 2);
  int diff;
  if (b < BLOCKS - 1) diff = (36000 + (int)u16le(blk + BLOCK_BYTES + 2) - az) % 36000;
  else diff = (36000 + az - (int)u16le(blk - BLOCK_BYTES + 2)) % 36000;
  const float t_off = (float)l * 2.304f + (float)f * 55.296f;
  const float az_f = (float)az + ((float)diff * t_off) / 110.592f;
  int az_i = (int)roundf(az_f) % 36000;
  if (az_i < 0) az_i += 36000;
  const float d = (float)raw * 0.002f;
  const float rot = (float)((double)az_i * (3.14159265358979323846 / 18000.0));
  const int w_deg = (l & 1) ? l : l - 15;
  const float vert = (float)((double)w_deg * (3.14159265358979323846 / 180.0));
  const float cos_v = dm::cos(vert), sin_v = dm::sin(vert);
  const float cos_r = dm::cos(rot), sin_r = dm::sin(rot);
  const float xy = d * cos_v;
  const float xv = xy * sin_r, yv = xy * cos_r;
  out[0] = yv;
  out[1] = -xv;
  out[2] = d * sin_v;
  return true;
}

}  // namespace pk

// one thread per point slot; the 1206-byte packet is staged in shared memory by the CTA (one
// packet per CTA of 384 threads: coalesced byte loads in, coalesced 12-byte stores out)
__global__ void __launch_bounds__(384) pk_decode_vlp16(const unsigned char* __restrict__ packets, int n_packets, float* __restrict__ xyz) {
  __shared__ unsigned char s_pkt[pk::PACKET_BYTES + 2];
  for (int p = blockIdx.x; p < n_packets; p += gridDim.x) {
    const unsigned char* src = packets + (size_t)p * pk::PACKET_BYTES;
    for (int i = threadIdx.x; i < pk::PACKET_BYTES; i += 384) s_pkt[i] = src[i];
    __syncthreads();
    float o[3];
    pk::decode_point(s_pkt, threadIdx.x, o);
    float* dst = xyz + ((size_t)p * pk::POINTS_PER_PACKET + threadIdx.x) * 3;
    dst[0] = o[0]; dst[1] = o[1]; dst[2] = o[2];
    __syncthreads();
  }
}

}  // namespace loamgpu
