// vx_coop.cuh -- pcl::VoxelGrid of a whole cloud as ONE kernel:
// bounding box -> voxel keys -> stable LSD radix sort (only as many 8-bit passes as the key range
// needs) -> segment heads -> centroids.  One thread-block CLUSTER of VC_CLUSTER CTAs per job; the
// phases are separated by cluster barriers (~0.2 us, hardware) instead of kernel boundaries or
// grid-wide barriers (~2.5 us each, nine of them per filter).  Used for the incoming-sweep filters of laserMapping (laserMapping.cpp:693-701,
// downSizeFilterCorner 0.2 m / downSizeFilterSurf 0.4 m): two jobs of 5e3 .. 5e4 points per
// mapped sweep, for which the 18-launch device-wide path (sort_voxel.cuh) is launch/latency bound.
// Semantics are identical to sort_voxel.cuh: cell = (int)(floor(p * inv_leaf) - (float)min_b),
// idx = i + j*dx + k*dx*dy, output ascending idx, centroid of all four fields = fp32 running sum
// in input order / (float)count; non-finite points are dropped; an index range beyond INT_MAX
// returns the input unfiltered (PCL prints a warning and does the same).
#pragma once
#include <cooperative_groups.h>
#include "common.cuh"

namespace loamgpu {
namespace cg = cooperative_groups;

constexpr int VC_THREADS = 256;
constexpr int VC_ITEMS = 8;
constexpr int VC_TILE = VC_THREADS * VC_ITEMS;   // 2048 elements per tile
constexpr int VC_WARPS = VC_THREADS / 32;
constexpr int VC_CLUSTER = 16;                   // CTAs per job (non-portable cluster size: cudaFuncAttributeNonPortableClusterSizeAllowed)

struct VxCoopJob {
  const float4* pts;
  const int* n_ptr;
  unsigned int* bbox;        // 6 flipped floats, armed (min = 0xffffffff, max = 0); re-armed by the kernel
  unsigned int* keys[2];     // ping-pong
  unsigned int* vals[2];
  int* hist;                 // [256][tile_cap] per-tile digit counts
  int* heads;                // [tile_cap] segment heads per tile
  float4* sorted;            // points in sorted order (gathered by the last scatter pass)
  float4* out;
  int* out_count;
  int* overflow;
  float leaf;
  int tile_cap;
};
struct VxCoopArgs { VxCoopJob j[2]; };

__device__ __forceinline__ int vc_block_excl_scan256(int v, int* s_w) {   // 256 threads, returns exclusive prefix
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  int x = v;
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();
  if (lane == 31) s_w[warp] = x;
  __syncthreads();
  int off = 0;
#pragma unroll
  for (int w = 0; w < VC_WARPS; w++) if (w < warp) off += s_w[w];
  return off + x - v;
}

struct VcKeyParams { int minb[3], divb[3]; int overflow; int passes; };
__device__ __forceinline__ VcKeyParams vc_key_params(const VxCoopJob& J, int n) {
  VcKeyParams K;
  const float inv = 1.0f / J.leaf;
  float mn[3], mx[3];
  for (int a = 0; a < 3; a++) { mn[a] = float_unflip(__ldcg(&J.bbox[a])); mx[a] = float_unflip(__ldcg(&J.bbox[3 + a])); }
  long long dx = (long long)((mx[0] - mn[0]) * inv) + 1, dy = (long long)((mx[1] - mn[1]) * inv) + 1, dz = (long long)((mx[2] - mn[2]) * inv) + 1;
  K.overflow = ((n > 0) && (dx * dy * dz > 2147483647LL)) ? 1 : 0;
  for (int a = 0; a < 3; a++) {
    K.minb[a] = (int)floorf(mn[a] * inv);
    int maxb = (int)floorf(mx[a] * inv);
    K.divb[a] = maxb - K.minb[a] + 1;
  }
  K.passes = 4;
  if (!K.overflow && n > 0) {
    // every key is < ncells; non-finite points carry 0xffffffff and must still sort behind all of
    // them, so a pass count p needs ncells < 2^(8p)
    long long ncells = (long long)K.divb[0] * K.divb[1] * K.divb[2];
    K.passes = 1;
    while (K.passes < 4 && ncells >= (1LL << (8 * K.passes))) K.passes++;
  }
  if (n <= 0) K.passes = 0;
  return K;
}

__global__ void __cluster_dims__(VC_CLUSTER, 1, 1) __launch_bounds__(VC_THREADS) vx_filter_coop(VxCoopArgs A) {
  cg::cluster_group grid = cg::this_cluster();   // gridDim.x == VC_CLUSTER: the cluster is the job's whole "grid"
  const VxCoopJob J = A.j[blockIdx.y];
  const int G = gridDim.x;
  int n = *J.n_ptr;
  if ((n + VC_TILE - 1) / VC_TILE > J.tile_cap) n = 0;   // capacity is enforced by the host-side max_points check
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  __shared__ int s_hist[256];
  __shared__ int s_cnt[VC_WARPS][256];
  __shared__ int s_dbase[256];
  __shared__ int s_w[VC_WARPS];
  __shared__ int s_base;
  // ---- bounding box of the finite points ----
  {
    float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int i = blockIdx.x * VC_THREADS + threadIdx.x; i < n; i += G * VC_THREADS) {
      float4 p = J.pts[i];
      if (!isfinite(p.x) || !isfinite(p.y) || !isfinite(p.z)) continue;
      mn[0] = fminf(mn[0], p.x); mx[0] = fmaxf(mx[0], p.x);
      mn[1] = fminf(mn[1], p.y); mx[1] = fmaxf(mx[1], p.y);
      mn[2] = fminf(mn[2], p.z); mx[2] = fmaxf(mx[2], p.z);
    }
    for (int a = 0; a < 3; a++)
      for (int o = 16; o > 0; o >>= 1) {
        mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
        mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
      }
    if (lane == 0 && mn[0] <= mx[0])
      for (int a = 0; a < 3; a++) { atomicMin(&J.bbox[a], float_flip(mn[a])); atomicMax(&J.bbox[3 + a], float_flip(mx[a])); }
  }
  __threadfence();
  grid.sync();
  // ---- voxel keys ----
  const VcKeyParams K = vc_key_params(J, n);
  const int passes_all = K.passes;   // the jobs run in separate clusters: no common barrier count needed
  {
    const float inv = 1.0f / J.leaf;
    if (blockIdx.x == 0 && threadIdx.x == 0) *J.overflow = K.overflow;
    for (int i = blockIdx.x * VC_THREADS + threadIdx.x; i < n; i += G * VC_THREADS) {
      float4 p = J.pts[i];
      unsigned int key;
      if (K.overflow) key = (unsigned int)i;  // every point its own voxel, input order kept
      else if (!isfinite(p.x) || !isfinite(p.y) || !isfinite(p.z)) key = 0xffffffffu;
      else {
        int i0 = (int)(floorf(p.x * inv) - (float)K.minb[0]);
        int i1 = (int)(floorf(p.y * inv) - (float)K.minb[1]);
        int i2 = (int)(floorf(p.z * inv) - (float)K.minb[2]);
        key = (unsigned int)(i0 + i1 * K.divb[0] + i2 * K.divb[0] * K.divb[1]);
      }
      J.keys[0][i] = key;
      J.vals[0][i] = (unsigned int)i;
    }
  }
  __threadfence();
  grid.sync();
  if (blockIdx.x == 0 && threadIdx.x < 6) J.bbox[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;   // re-arm for the next use
  const int ntile = (n + VC_TILE - 1) / VC_TILE;
  const int tc = J.tile_cap;
  // ---- stable LSD radix sort, 8 bits per pass, tiles of VC_TILE elements ----
  for (int pass = 0; pass < passes_all; pass++) {
    const bool active = pass < K.passes;
    const unsigned int* kin = J.keys[pass & 1];
    const unsigned int* vin = J.vals[pass & 1];
    unsigned int* kout = J.keys[(pass & 1) ^ 1];
    unsigned int* vout = J.vals[(pass & 1) ^ 1];
    const int shift = 8 * pass;
    const bool last = pass == K.passes - 1;
    if (active) {
      for (int t = blockIdx.x; t < ntile; t += G) {
        s_hist[threadIdx.x] = 0;
        __syncthreads();
        const int base = t * VC_TILE;
#pragma unroll
        for (int k = 0; k < VC_ITEMS; k++) {
          int i = base + k * VC_THREADS + threadIdx.x;
          if (i < n) atomicAdd(&s_hist[(kin[i] >> shift) & 255u], 1);
        }
        __syncthreads();
        J.hist[threadIdx.x * tc + t] = s_hist[threadIdx.x];
        __syncthreads();
      }
    }
    __threadfence();
    grid.sync();
    if (active) {
      for (int t = blockIdx.x; t < ntile; t += G) {
        // digit d: elements of smaller digits anywhere + elements of digit d in earlier tiles
        {
          const int d = threadIdx.x;
          int tot = 0, pre = 0;
          for (int u = 0; u < ntile; u++) {
            int c = __ldcg(&J.hist[d * tc + u]);
            if (u < t) pre += c;
            tot += c;
          }
          int excl = vc_block_excl_scan256(tot, s_w);
          s_dbase[d] = excl + pre;
        }
        for (int k = threadIdx.x; k < VC_WARPS * 256; k += VC_THREADS) (&s_cnt[0][0])[k] = 0;
        __syncthreads();
        // each warp owns a contiguous chunk of the tile and walks it in order
        unsigned int key[VC_ITEMS], val[VC_ITEMS];
        int rank[VC_ITEMS];
        const int chunk0 = t * VC_TILE + warp * (VC_TILE / VC_WARPS);
#pragma unroll
        for (int k = 0; k < VC_ITEMS; k++) {
          int i = chunk0 + k * 32 + lane;
          bool in = i < n;
          key[k] = in ? kin[i] : 0xffffffffu;
          val[k] = in ? vin[i] : 0u;
        }
#pragma unroll
        for (int k = 0; k < VC_ITEMS; k++) {
          int i = chunk0 + k * 32 + lane;
          bool in = i < n;
          int d = (int)((key[k] >> shift) & 255u);
          unsigned int m = warp_match8(d, in);
          int r = __popc(m & ((1u << lane) - 1u));
          rank[k] = in ? s_cnt[warp][d] + r : -1;
          __syncwarp();
          if (in && r == 0) s_cnt[warp][d] += __popc(m);
          __syncwarp();
        }
        __syncthreads();
        {
          const int d = threadIdx.x;
          int run = 0;
#pragma unroll
          for (int w = 0; w < VC_WARPS; w++) { int c = s_cnt[w][d]; s_cnt[w][d] = run; run += c; }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < VC_ITEMS; k++) {
          if (rank[k] >= 0) {
            int d = (int)((key[k] >> shift) & 255u);
            int dst = s_dbase[d] + s_cnt[warp][d] + rank[k];
            kout[dst] = key[k];
            vout[dst] = val[k];
            if (last) J.sorted[dst] = J.pts[val[k]];   // payload gather rides on the last pass
          }
        }
        __syncthreads();
      }
    }
    __threadfence();
    grid.sync();
  }
  // ---- segment heads per tile ----
  const unsigned int* sk = J.keys[K.passes & 1];
  for (int t = blockIdx.x; t < ntile; t += G) {
    int c = 0;
    const int base = t * VC_TILE;
#pragma unroll
    for (int k = 0; k < VC_ITEMS; k++) {
      int i = base + k * VC_THREADS + threadIdx.x;
      if (i < n) {
        unsigned int key = sk[i];
        if (key != 0xffffffffu && (i == 0 || key != sk[i - 1])) c++;
      }
    }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    __syncthreads();
    if (lane == 0) s_w[warp] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
      for (int w = 0; w < VC_WARPS; w++) tot += s_w[w];
      J.heads[t] = tot;
    }
  }
  __threadfence();
  grid.sync();
  // ---- centroids: thread owns VC_ITEMS consecutive sorted elements, slots follow sorted order ----
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    int tot = 0;
    for (int u = 0; u < ntile; u++) tot += __ldcg(&J.heads[u]);
    *J.out_count = tot;
  }
  for (int t = blockIdx.x; t < ntile; t += G) {
    __syncthreads();
    if (threadIdx.x == 0) {
      int b = 0;
      for (int u = 0; u < t; u++) b += __ldcg(&J.heads[u]);
      s_base = b;
    }
    const int first = t * VC_TILE + threadIdx.x * VC_ITEMS;
    unsigned int key[VC_ITEMS + 1];
    float4 pt[VC_ITEMS];
    key[0] = (first > 0 && first <= n) ? sk[first - 1] : 0xffffffffu;
#pragma unroll
    for (int k = 0; k < VC_ITEMS; k++) {
      int i = first + k;
      key[k + 1] = i < n ? sk[i] : 0xffffffffu;
      pt[k] = i < n ? J.sorted[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    int hc = 0;
#pragma unroll
    for (int k = 0; k < VC_ITEMS; k++) {
      int i = first + k;
      bool head = (i < n) && key[k + 1] != 0xffffffffu && (i == 0 || key[k + 1] != key[k]);
      hc += head ? 1 : 0;
    }
    int slot = vc_block_excl_scan256(hc, s_w);
    __syncthreads();
    slot += s_base;
#pragma unroll
    for (int k = 0; k < VC_ITEMS; k++) {
      int i = first + k;
      bool head = (i < n) && key[k + 1] != 0xffffffffu && (i == 0 || key[k + 1] != key[k]);
      if (!head) continue;
      const unsigned int hk = key[k + 1];
      float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
      int cnt = 0;
#pragma unroll
      for (int q = k; q < VC_ITEMS; q++) {          // the part of the segment this thread already holds
        if (key[q + 1] != hk || first + q >= n) break;
        c0 += pt[q].x; c1 += pt[q].y; c2 += pt[q].z; c3 += pt[q].w;
        cnt++;
      }
      if (k + cnt == VC_ITEMS) {                     // it may run on into the next thread's elements
        for (int q = first + VC_ITEMS; q < n && sk[q] == hk; q++) {
          float4 p = J.sorted[q];
          c0 += p.x; c1 += p.y; c2 += p.z; c3 += p.w;
          cnt++;
        }
      }
      float cntf = (float)cnt;
      J.out[slot] = make_float4(c0 / cntf, c1 / cntf, c2 / cntf, c3 / cntf);
      slot++;
    }
  }
}

}  // namespace loamgpu
