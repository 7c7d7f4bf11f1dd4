// cta_radix.cuh -- stable LSD radix sort of (32-bit key, 32-bit payload) pairs executed by ONE
// CTA, data in global memory (L2-resident at the sizes it is used for: 1e3 .. 1e5 elements).
//
// Why a single CTA: the mapping node sorts a few thousand to a few ten-thousand elements several
// times per sweep (incoming-sweep voxel filters laserMapping.cpp:693-701, new points per cube
// :980-1016).  A device-wide radix sort needs 3 launches per 8-bit pass and is launch/latency
// bound at these sizes; here a whole sort is part of one kernel and independent sorts run in
// different CTAs.  Each warp owns a contiguous chunk and walks it in order, ranking equal digits
// with match.any, so the sort is stable (payload order inside equal keys = input order).
#pragma once
#include "common.cuh"

namespace loamgpu {

constexpr int CR_MAX_WARPS = 32;

struct CtaRadixSmem {
  int cnt[CR_MAX_WARPS][256];   // per-warp digit counts, then running output cursors
  int tot[256];
  int wsum[CR_MAX_WARPS];
};

// one 8-bit pass.  All threads of the CTA must call it (contains __syncthreads).
__device__ __forceinline__ void cta_radix_pass(const unsigned int* __restrict__ kin, const unsigned int* __restrict__ vin,
                                               unsigned int* __restrict__ kout, unsigned int* __restrict__ vout,
                                               int n, int shift, CtaRadixSmem& S) {
  const int nwarps = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int L = ((n + nwarps * 32 - 1) / (nwarps * 32)) * 32;   // chunk per warp, multiple of 32
  const int c0 = warp * L, c1 = min(n, c0 + L);
  for (int k = threadIdx.x; k < nwarps * 256; k += blockDim.x) (&S.cnt[0][0])[k] = 0;
  __syncthreads();
  if (shift == 0) DBG_STAMP(8);
  // walk 1: digit counts of this warp's chunk
  for (int b = c0; b < c1; b += 128) {
    unsigned int key[4];
#pragma unroll
    for (int u = 0; u < 4; u++) { int i = b + u * 32 + lane; key[u] = i < c1 ? kin[i] : 0u; }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int i = b + u * 32 + lane;
      bool in = i < c1;
      int d = (int)((key[u] >> shift) & 255u);
      unsigned int m = warp_match8(d, in);
      if (in && (m & ((1u << lane) - 1u)) == 0) S.cnt[warp][d] += __popc(m);
      __syncwarp();
    }
  }
  __syncthreads();
  if (shift == 0) DBG_STAMP(9);
  // digit totals, per-warp exclusive offsets inside a digit
  for (int d = threadIdx.x; d < 256; d += blockDim.x) {
    int run = 0;
    for (int w = 0; w < nwarps; w++) { int c = S.cnt[w][d]; S.cnt[w][d] = run; run += c; }
    S.tot[d] = run;
  }
  __syncthreads();
  // exclusive scan of the 256 digit totals (warps 0..7, 32 digits each)
  {
    const bool act = threadIdx.x < 256;   // blockDim.x >= 256
    int v = act ? S.tot[threadIdx.x] : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (act && lane == 31) S.wsum[warp] = x;
    __syncthreads();
    int off = 0;
    for (int w = 0; w < warp && w < 8; w++) off += S.wsum[w];
    if (act) S.tot[threadIdx.x] = off + x - v;
  }
  __syncthreads();
  if (shift == 0) DBG_STAMP(10);
  // walk 2: stable scatter
  for (int b = c0; b < c1; b += 128) {
    unsigned int key[4], val[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int i = b + u * 32 + lane;
      key[u] = i < c1 ? kin[i] : 0u;
      val[u] = i < c1 ? vin[i] : 0u;
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int i = b + u * 32 + lane;
      bool in = i < c1;
      int d = (int)((key[u] >> shift) & 255u);
      unsigned int m = warp_match8(d, in);
      int r = __popc(m & ((1u << lane) - 1u));
      int dst = 0;
      if (in) dst = S.tot[d] + S.cnt[warp][d] + r;
      __syncwarp();
      if (in && r == 0) S.cnt[warp][d] += __popc(m);
      __syncwarp();
      if (in) { kout[dst] = key[u]; vout[dst] = val[u]; }
    }
  }
  __threadfence_block();
  __syncthreads();
  if (shift == 0) DBG_STAMP(11);
}

// The same pass for n <= (blockDim.x / 32) * 32 * ITEMS: every lane keeps its ITEMS elements in
// registers, so the chunk is read from global memory once (one load latency per pass instead of
// one per group of steps and per walk).
template <int ITEMS>
__device__ __forceinline__ void cta_radix_pass_reg(const unsigned int* __restrict__ kin, const unsigned int* __restrict__ vin,
                                                   unsigned int* __restrict__ kout, unsigned int* __restrict__ vout,
                                                   int n, int shift, CtaRadixSmem& S) {
  const int nwarps = blockDim.x >> 5;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int L = ((n + nwarps * 32 - 1) / (nwarps * 32)) * 32;   // chunk per warp, multiple of 32, <= 32 * ITEMS
  const int c0 = warp * L, c1 = min(n, c0 + L);
  unsigned int key[ITEMS], val[ITEMS];
#pragma unroll
  for (int u = 0; u < ITEMS; u++) {
    const int i = c0 + u * 32 + lane;
    key[u] = i < c1 ? kin[i] : 0u;
    val[u] = i < c1 ? vin[i] : 0u;
  }
  for (int k = threadIdx.x; k < nwarps * 256; k += blockDim.x) (&S.cnt[0][0])[k] = 0;
  __syncthreads();
  // walk the chunk once: rank of every element among the equal digits of this warp seen so far
  // (stable), per-warp digit counts as a by-product
  int rk[ITEMS];
#pragma unroll
  for (int u = 0; u < ITEMS; u++) {
    const bool in = c0 + u * 32 + lane < c1;
    const int d = (int)((key[u] >> shift) & 255u);
    const unsigned int m = warp_match8(d, in);
    const int r = __popc(m & ((1u << lane) - 1u));
    rk[u] = in ? S.cnt[warp][d] + r : -1;
    __syncwarp();
    if (in && r == 0) S.cnt[warp][d] += __popc(m);
    __syncwarp();
  }
  __syncthreads();
  for (int d = threadIdx.x; d < 256; d += blockDim.x) {
    int run = 0;
    for (int w = 0; w < nwarps; w++) { int c = S.cnt[w][d]; S.cnt[w][d] = run; run += c; }
    S.tot[d] = run;
  }
  __syncthreads();
  {
    const bool act = threadIdx.x < 256;
    int v = act ? S.tot[threadIdx.x] : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (act && lane == 31) S.wsum[warp] = x;
    __syncthreads();
    int off = 0;
    for (int w = 0; w < warp && w < 8; w++) off += S.wsum[w];
    if (act) S.tot[threadIdx.x] = off + x - v;
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < ITEMS; u++) {
    if (rk[u] >= 0) {
      const int d = (int)((key[u] >> shift) & 255u);
      const int dst = S.tot[d] + S.cnt[warp][d] + rk[u];
      kout[dst] = key[u]; vout[dst] = val[u];
    }
  }
  __threadfence_block();
  __syncthreads();
}

// sorts by key bits [0, 8*passes).  Buffers ping-pong; returns 0 if the result is in (k0, v0),
// 1 if it is in (k1, v1).
__device__ __forceinline__ int cta_radix_sort(unsigned int* k0, unsigned int* v0, unsigned int* k1, unsigned int* v1,
                                              int n, int passes, CtaRadixSmem& S) {
  int cur = 0;
  const bool small = n <= (int)(blockDim.x >> 5) * 32 * 8;
  for (int p = 0; p < passes; p++) {
    unsigned int* ki = cur ? k1 : k0; unsigned int* vi = cur ? v1 : v0;
    unsigned int* ko = cur ? k0 : k1; unsigned int* vo = cur ? v0 : v1;
    if (small) cta_radix_pass_reg<8>(ki, vi, ko, vo, n, 8 * p, S);
    else cta_radix_pass(ki, vi, ko, vo, n, 8 * p, S);
    cur ^= 1;
  }
  return cur;
}

}  // namespace loamgpu
