// pk_decode.cuh -- raw-packet front end: Velodyne data packets (VLP-16 and
// HDL-64E) -> what scanRegistration needs, decoded on the device so that a sweep costs 3.1 bytes per
// point of host->device traffic instead of 12.
//
// The decoder the reference depends on lives in the un-vendored ROS package velodyne_pointcloud
// (ros-drivers/velodyne, rawdata.cc, version unpinned); its arithmetic is restated here from the
// published packet formats (VLP-16 User Manual ch. 9; HDL-64E S2 manual app. E) with an ideal
// calibration (vertical angles only, every other correction zero):
//   packet  = 12 blocks x 100 B + 4 B timestamp + 2 B factory        = 1206 B
//   block   = flag (u16) + azimuth (u16, 0.01 deg) + 32 x {distance u16 (2 mm), reflectivity u8}
//   VLP-16:  flag 0xEEFF; a block = two firing sequences of the 16 lasers, laser l at elevation
//            (l odd ? l : l - 15) deg; azimuth of (block b, firing f, laser l) =
//            az_b + diff_b * (2.304 l + 55.296 f) / 110.592, rounded to 0.01 deg,
//            diff_b = (36000 + az_{b+1} - az_b) % 36000, the last block of a packet re-uses the diff before
//   HDL-64E: flag 0xEEFF = upper block (lasers 32..63), 0xDDFF = lower block (lasers 0..31); an upper and
//            a lower block with the same azimuth make one firing of the 64 lasers; laser k at elevation
//            ring_lower + (ring_upper - ring_lower) * k / 63 (loam_params), no azimuth interpolation
//   point   = d cos(w) cos(a), -d cos(w) sin(a), d sin(w)   (ROS axes: x forward, y left, z up; a clockwise)
//   distance 0 = no return: the slot is written as NaN and dropped by scanRegistration's NaN filter (:228)
//
// loam_params.packet_metadata = 1 ("SR-1 from packet metadata", pk_ring_meta below): instead of recovering
// the ring id with atan(y / sqrt(x^2 + z^2)) and the sweep phase with atan2 and the halfPassed latch from
// the decoded point (scanRegistration.cpp:248-284), both are taken from what the packet already says:
//   ring     = rank of the slot's laser in elevation order -- VLP-16: w > 0 ? w : w + 15 with w the laser's
//              nominal elevation in degrees (the value the reference's rounding yields for it); HDL-64E: k
//   relTime  = daz(slot) / daz(last valid slot), daz = (az - az_first + 36000) % 36000 in 0.01 deg units,
//              plus 36000 for a slot in the second half of the revolution's slots whose daz is below 18000
//              (the revolution ran past its starting azimuth) -- integer arithmetic, one float division
// DIFFERENCE to the reference: the ring ids are the same for an ideally calibrated sensor (the nominal
// elevation is far from the +-0.5-step rounding boundaries); relTime agrees with the atan2-derived value
// to ~1e-5 (0.01 deg azimuth quantisation against a float atan2), i.e. the intensity channel carries
// slightly different fractions and everything downstream (deskew, TransformToStart) follows them.  The
// oracle restates this variant separately (oracle/packet_port.hpp sr1_meta_from_packets) and the GPU path
// is bit-equal to THAT; with packet_metadata = 0 it is bit-equal to the reference fed with the decoded cloud.
// Every floating-point step is written once, in pk::decode_slot, and used by the kernels and by the host
// (which needs the first and last valid point of the sweep, scanRegistration.cpp:230-232).
#pragma once
#include <math.h>
#include <stdint.h>
#include "common.cuh"
#include "sr_kernels.cuh"

namespace loamgpu {
namespace pk {

constexpr int PACKET_BYTES = 1206, BLOCKS = 12, BLOCK_BYTES = 100, POINTS_PER_PACKET = 384;
enum Sensor { VLP16 = 0, HDL64E = 1 };
struct SensorDesc { int kind; int n_scans; float lower_deg, upper_deg; };

__host__ __device__ __forceinline__ unsigned int u16le(const unsigned char* p) { return (unsigned int)p[0] | ((unsigned int)p[1] << 8); }

struct Slot { float xyz[3]; int az; int ring; bool valid; };

// slot `idx` (0..383: block-major order) of one packet
__host__ __device__ inline Slot decode_slot(const unsigned char* pkt, int idx, const SensorDesc& S) {
  Slot o;
  const int b = idx >> 5, k = idx & 31;
  const unsigned char* blk = pkt + b * BLOCK_BYTES;
  const unsigned int raw = u16le(blk + 4 + 3 * k);
  const unsigned int flag = u16le(blk);
  const int az = (int)u16le(blk + 2);
  o.az = 0; o.ring = -1;
  bool ok = raw != 0;
  float vert;
  int az_i;
  if (S.kind == VLP16) {
    ok = ok && flag == 0xEEFFu;
    const int f = k >> 4, l = k & 15;
    int diff;
    if (b < BLOCKS - 1) diff = (36000 + (int)u16le(blk + BLOCK_BYTES + 2) - az) % 36000;
    else diff = (36000 + az - (int)u16le(blk - BLOCK_BYTES + 2)) % 36000;
    const float t_off = (float)l * 2.304f + (float)f * 55.296f;
    const float az_f = (float)az + ((float)diff * t_off) / 110.592f;
    az_i = (int)roundf(az_f) % 36000;
    if (az_i < 0) az_i += 36000;
    const int w_deg = (l & 1) ? l : l - 15;
    vert = (float)((double)w_deg * (3.14159265358979323846 / 180.0));
    o.ring = w_deg > 0 ? w_deg : w_deg + 15;
  } else {
    const bool upper = flag == 0xEEFFu, lower = flag == 0xDDFFu;
    ok = ok && (upper || lower);
    const int laser = upper ? 32 + k : k;
    az_i = az % 36000;
    const double e = (double)S.lower_deg + ((double)S.upper_deg - (double)S.lower_deg) * (double)laser / (double)(S.n_scans - 1);
    vert = (float)(e * (3.14159265358979323846 / 180.0));
    o.ring = laser;
  }
  o.valid = ok;
  o.az = az_i;
  if (!ok) {
#ifdef __CUDA_ARCH__
    o.xyz[0] = o.xyz[1] = o.xyz[2] = __int_as_float(0x7fc00000);
#else
    o.xyz[0] = o.xyz[1] = o.xyz[2] = NAN;
#endif
    o.ring = -1;
    return o;
  }
  const float d = (float)raw * 0.002f;
  const float rot = (float)((double)az_i * (3.14159265358979323846 / 18000.0));
  const float cos_v = dm::cos(vert), sin_v = dm::sin(vert);
  const float cos_r = dm::cos(rot), sin_r = dm::sin(rot);
  const float xy = d * cos_v;
  const float xv = xy * sin_r, yv = xy * cos_r;
  o.xyz[0] = yv;
  o.xyz[1] = -xv;
  o.xyz[2] = d * sin_v;
  return o;
}

// sweep phase of a slot from its azimuth (0.01 deg) -- see the header comment
struct MetaSpan { int first, last, az0, span; };
__host__ __device__ __forceinline__ int meta_daz(int slot, int az, const MetaSpan& M) {
  int daz = (az - M.az0 + 36000) % 36000;
  if ((long long)(slot - M.first) * 2 > (long long)(M.last - M.first) && daz < 18000) daz += 36000;
  return daz;
}
__host__ __device__ __forceinline__ float meta_reltime(int slot, int az, const MetaSpan& M) {
  if (M.span <= 0) return 0.f;
  return (float)meta_daz(slot, az, M) / (float)M.span;
}

}  // namespace pk

// one thread per slot: decoded xyz only (packet_metadata = 0: the unchanged SR-1 follows)
__global__ void __launch_bounds__(256) pk_decode(const unsigned char* __restrict__ packets, int n_packets, pk::SensorDesc S, float* __restrict__ xyz) {
  const int n = n_packets * pk::POINTS_PER_PACKET;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    const int p = i / pk::POINTS_PER_PACKET, idx = i - p * pk::POINTS_PER_PACKET;
    const pk::Slot o = pk::decode_slot(packets + (size_t)p * pk::PACKET_BYTES, idx, S);
    float* dst = xyz + (size_t)i * 3;
    dst[0] = o.xyz[0]; dst[1] = o.xyz[1]; dst[2] = o.xyz[2];
  }
}

// packet_metadata = 1: decoder fused with SR-1's first kernel (replaces pk_decode + sr_ring_ori): per slot the
// decoded xyz, the ring id from the laser id, relTime from the azimuth field, and the per-block ring histogram
// that sr_ring_scan / sr_scatter need for the stable ring-major order.  No atan / atan2, no halfPassed latch.
__global__ void __launch_bounds__(SR_BLK) pk_ring_meta(const unsigned char* __restrict__ packets, int n_packets, pk::SensorDesc S, pk::MetaSpan M,
                                                       float* __restrict__ xyz, short* __restrict__ ring, float* __restrict__ reltime,
                                                       int* __restrict__ hist, int nblk, SRCounts* cnt) {
  __shared__ int sh[SR_MAX_RINGS];
  for (int r = threadIdx.x; r < S.n_scans; r += SR_BLK) sh[r] = 0;
  __syncthreads();
  const int n = n_packets * pk::POINTS_PER_PACKET;
  const int i = blockIdx.x * SR_BLK + threadIdx.x;
  if (i < n) {
    const int p = i / pk::POINTS_PER_PACKET, idx = i - p * pk::POINTS_PER_PACKET;
    const pk::Slot o = pk::decode_slot(packets + (size_t)p * pk::PACKET_BYTES, idx, S);
    float* dst = xyz + (size_t)i * 3;
    dst[0] = o.xyz[0]; dst[1] = o.xyz[1]; dst[2] = o.xyz[2];
    int rid = -1;
    float rt = 0.f;
    if (o.valid && o.ring >= 0 && o.ring < S.n_scans) {
      rid = o.ring;
      rt = pk::meta_reltime(i, o.az, M);
      atomicMax(&cnt->last_valid, i);
      atomicAdd(&sh[rid], 1);
    }
    ring[i] = (short)rid;
    reltime[i] = rt;
  }
  __syncthreads();
  for (int r = threadIdx.x; r < S.n_scans; r += SR_BLK) hist[r * nblk + blockIdx.x] = sh[r];
}

}  // namespace loamgpu
