// sr_batch.cuh -- the scanRegistration kernels as launchable entry points: each kernel of sr_kernels.cuh /
// sr_kernels2.cuh is a __device__ body (NAME_body) with two __global__ wrappers,
//   NAME(...)              one sweep of one sequence (the node-level API), and
//   NAME_b(items, ...)     the same step for a BATCH of sequences in one launch: blockIdx.z selects the sequence,
//                          its buffers and per-sweep scalars come from a device array of SRDev descriptors.
// BASELINE.json configs[4] (64 independent VLP-16 sequences per box) issues one launch chain for all sequences
// of a GPU this way: nine launches per step instead of 8 x 64, and every streaming kernel moves 64 sweeps
// (tens of MB) per launch instead of 0.35 MB -- the size at which an HBM roofline means something.
#pragma once
#include "sr_kernels2.cuh"

namespace loamgpu {

// everything one sequence's scanRegistration pass needs on the device (filled by SRHandle::fill_dev)
struct SRDev {
  const float* raw; int n, nblk; float startOri, endOri; int meta;
  short* ring; float* ori; int* hist; int* ring_start; int* scan_start; int* scan_end;
  SRCounts* cnt; int* ring_tot; unsigned int* ticket;
  float4* full; int* pos; float* reltime; float* curv; int* sort_ind;
  unsigned char* picked; signed char* label; unsigned char* gap;
  int* sharp_ring; int* less_sharp_ring; int* flat_ring; int* ring_counts;
  float4* lessflat_tmp; float4* sharp; float4* less_sharp; float4* flat; float4* less_flat;
  int* sharp_idx; int* less_sharp_idx; int* flat_idx;
};

// counters of the sweep start from a known state (single mode: uploaded from pinned memory instead)
__global__ void sr_batch_init_b(const SRDev* items) {
  if (threadIdx.x != 0) return;
  SRCounts c;
  c.n_full = 0; c.h_min = INT_MAX; c.n_sharp = c.n_less_sharp = c.n_flat = c.n_less_flat = 0; c.error = 0; c.last_valid = -1;
  *items[blockIdx.x].cnt = c;
}

__global__ void __launch_bounds__(SR_BLK) sr_ring_ori(const float* __restrict__ raw, int n, int stride_floats,
                                                      SRParamsDev P, float startOri, short* __restrict__ ring,
                                                      float* __restrict__ ori_raw, int* __restrict__ hist, int nblk,
                                                      SRCounts* cnt) {
  sr_ring_ori_body(raw, n, stride_floats, P, startOri, ring, ori_raw, hist, nblk, cnt);
}

__global__ void __launch_bounds__(SR_BLK) sr_ring_ori_b(const SRDev* __restrict__ items, SRParamsDev P) {
  const SRDev D = items[blockIdx.z];
  if ((int)blockIdx.x >= D.nblk) return;
  sr_ring_ori_body(D.raw, D.n, 3, P, D.startOri, D.ring, D.ori, D.hist, D.nblk, D.cnt);
}

__global__ void __launch_bounds__(RSC_THREADS) sr_ring_scan(int* __restrict__ hist, int nblk, int n_scans,
                                                            int* __restrict__ ring_start, int* __restrict__ scan_start,
                                                            int* __restrict__ scan_end, SRCounts* cnt, int max_points,
                                                            int* ring_tot, unsigned int* ticket) {
  sr_ring_scan_body(hist, nblk, n_scans, ring_start, scan_start, scan_end, cnt, max_points, ring_tot, ticket);
}

__global__ void __launch_bounds__(RSC_THREADS) sr_ring_scan_b(const SRDev* __restrict__ items, int n_scans, int max_points) {
  const SRDev D = items[blockIdx.z];
  sr_ring_scan_body(D.hist, D.nblk, n_scans, D.ring_start, D.scan_start, D.scan_end, D.cnt, max_points, D.ring_tot, D.ticket);
}

__global__ void __launch_bounds__(SR_BLK) sr_scatter(const float* __restrict__ raw, int n, int stride_floats,
                                                     SRParamsDev P, float startOri, float endOri,
                                                     const short* __restrict__ ring, const float* __restrict__ ori_raw,
                                                     const int* __restrict__ base, const int* __restrict__ ring_start, int nblk,
                                                     const SRCounts* cnt, float4* __restrict__ full, int* __restrict__ pos_out,
                                                     float* __restrict__ reltime_out, int meta = 0) {
  sr_scatter_body(raw, n, stride_floats, P, startOri, endOri, ring, ori_raw, base, ring_start, nblk, cnt, full, pos_out, reltime_out, meta);
}

__global__ void __launch_bounds__(SR_BLK) sr_scatter_b(const SRDev* __restrict__ items, SRParamsDev P) {
  const SRDev D = items[blockIdx.z];
  if ((int)blockIdx.x >= D.nblk) return;
  sr_scatter_body(D.raw, D.n, 3, P, D.startOri, D.endOri, D.ring, D.ori, D.hist, D.ring_start, D.nblk, D.cnt, D.full, D.pos, D.reltime, D.meta);
}

__global__ void __launch_bounds__(CR_TILE) sr_curv_reject(const float4* __restrict__ full, const SRCounts* __restrict__ cnt,
                                                          float* __restrict__ curvature, int* __restrict__ sort_ind,
                                                          unsigned char* __restrict__ picked, signed char* __restrict__ label,
                                                          unsigned char* __restrict__ gapflag) {
  sr_curv_reject_body(full, cnt, curvature, sort_ind, picked, label, gapflag);
}

__global__ void __launch_bounds__(CR_TILE) sr_curv_reject_b(const SRDev* __restrict__ items) {
  const SRDev D = items[blockIdx.z];
  sr_curv_reject_body(D.full, D.cnt, D.curv, D.sort_ind, D.picked, D.label, D.gap);
}

__global__ void __launch_bounds__(512) sr_sort_sectors(const float* __restrict__ curvature, int* __restrict__ sort_ind,
                                                       const int* __restrict__ scan_start, const int* __restrict__ scan_end,
                                                       SRCounts* cnt) {
  sr_sort_sectors_body(curvature, sort_ind, scan_start, scan_end, cnt);
}

__global__ void __launch_bounds__(512) sr_sort_sectors_b(const SRDev* __restrict__ items) {
  const SRDev D = items[blockIdx.z];
  sr_sort_sectors_body(D.curv, D.sort_ind, D.scan_start, D.scan_end, D.cnt);
}

__global__ void __launch_bounds__(SEL_THREADS) sr_select(const float* __restrict__ curvature, const int* __restrict__ sort_ind,
                                                         const int* __restrict__ scan_start, const int* __restrict__ scan_end,
                                                         const int* __restrict__ ring_start, const unsigned char* __restrict__ gapflag,
                                                         unsigned char* __restrict__ picked, signed char* __restrict__ label,
                                                         int* __restrict__ sharp_idx, int* __restrict__ less_sharp_idx,
                                                         int* __restrict__ flat_idx, int* __restrict__ ring_counts,
                                                         SRCounts* __restrict__ cnt) {
  sr_select_body(curvature, sort_ind, scan_start, scan_end, ring_start, gapflag, picked, label, sharp_idx, less_sharp_idx, flat_idx, ring_counts, cnt);
}

__global__ void __launch_bounds__(SEL_THREADS) sr_select_b(const SRDev* __restrict__ items) {
  const SRDev D = items[blockIdx.z];
  sr_select_body(D.curv, D.sort_ind, D.scan_start, D.scan_end, D.ring_start, D.gap, D.picked, D.label, D.sharp_ring, D.less_sharp_ring, D.flat_ring, D.ring_counts, D.cnt);
}

__global__ void __launch_bounds__(1024, 2) sr_lessflat(const float4* __restrict__ full, const signed char* __restrict__ label,
                                                    const int* __restrict__ scan_start, const int* __restrict__ scan_end,
                                                    float4* __restrict__ out_tmp, int* __restrict__ ring_counts,
                                                    SRCounts* cnt) {
  sr_lessflat_body(full, label, scan_start, scan_end, out_tmp, ring_counts, cnt);
}

__global__ void __launch_bounds__(1024, 2) sr_lessflat_b(const SRDev* __restrict__ items) {
  const SRDev D = items[blockIdx.z];
  sr_lessflat_body(D.full, D.label, D.scan_start, D.scan_end, D.lessflat_tmp, D.ring_counts, D.cnt);
}

__global__ void __launch_bounds__(128) sr_compact(const float4* __restrict__ full, const int* __restrict__ ring_counts,
                                                  int n_scans, const int* __restrict__ sharp_idx,
                                                  const int* __restrict__ less_sharp_idx, const int* __restrict__ flat_idx,
                                                  const float4* __restrict__ lessflat_tmp, float4* __restrict__ sharp,
                                                  float4* __restrict__ less_sharp, float4* __restrict__ flat,
                                                  float4* __restrict__ less_flat, int* __restrict__ sharp_out_idx,
                                                  int* __restrict__ less_sharp_out_idx, int* __restrict__ flat_out_idx,
                                                  SRCounts* cnt, int max_points) {
  sr_compact_body(full, ring_counts, n_scans, sharp_idx, less_sharp_idx, flat_idx, lessflat_tmp, sharp, less_sharp, flat, less_flat, sharp_out_idx, less_sharp_out_idx, flat_out_idx, cnt, max_points);
}

__global__ void __launch_bounds__(128) sr_compact_b(const SRDev* __restrict__ items, int n_scans, int max_points) {
  const SRDev D = items[blockIdx.z];
  sr_compact_body(D.full, D.ring_counts, n_scans, D.sharp_ring, D.less_sharp_ring, D.flat_ring, D.lessflat_tmp, D.sharp, D.less_sharp, D.flat, D.less_flat, D.sharp_idx, D.less_sharp_idx, D.flat_idx, D.cnt, max_points);
}

}  // namespace loamgpu
