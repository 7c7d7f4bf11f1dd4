// sort_voxel.cuh -- device-wide stable LSD radix sort (32-bit key, 32-bit payload) whose element
// count lives in DEVICE memory (no host round trip), and on top of it the pcl::VoxelGrid
// semantics of SURVEY.md App. B.1 for clouds too large for one block's shared memory:
//   laserMapping.cpp:693-701   downSizeFilterCorner/Surf on the incoming sweep   (LM-6)
//   laserMapping.cpp:1018-1036 per-cube refilter of the map                       (LM-13)
//   laserMapping.cpp:1049-1051 surround map                                       (LM-14)
// Every kernel handles up to two independent jobs (blockIdx.y): corner and surf.
#pragma once
#include "common.cuh"

namespace loamgpu {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 elements per block
constexpr int RS_WARPS = RS_THREADS / 32;

struct SortJob {
  unsigned int* keys[2];   // ping-pong
  unsigned int* vals[2];
  int* hist;               // [256][nblk_cap] digit-major per-tile counts (scanned in place)
  int* digit_total;        // [256]
  const int* n_ptr;
  int nblk_cap;
};
struct SortPair { SortJob j[2]; int njobs; };

__global__ void __launch_bounds__(RS_THREADS) rs_hist(SortPair sp, int pass) {
  const SortJob J = sp.j[blockIdx.y];
  const int n = *J.n_ptr;
  const int nblk = (n + RS_TILE - 1) / RS_TILE;
  __shared__ int sh[256];
  sh[threadIdx.x] = 0;
  __syncthreads();
  if ((int)blockIdx.x < nblk) {
    const unsigned int* keys = J.keys[pass & 1];
    int base = blockIdx.x * RS_TILE;
    for (int k = threadIdx.x; k < RS_TILE; k += RS_THREADS) {
      int i = base + k;
      if (i < n) atomicAdd(&sh[(keys[i] >> (8 * pass)) & 255], 1);
    }
  }
  __syncthreads();
  if ((int)blockIdx.x < J.nblk_cap) J.hist[threadIdx.x * J.nblk_cap + blockIdx.x] = ((int)blockIdx.x < nblk) ? sh[threadIdx.x] : 0;
}

// one block per digit: exclusive scan of that digit's per-tile counts + the digit's total
__global__ void __launch_bounds__(1024) rs_scan(SortPair sp) {
  const SortJob J = sp.j[blockIdx.y];
  const int n = *J.n_ptr;
  const int nblk = (n + RS_TILE - 1) / RS_TILE;
  int* row = J.hist + (size_t)blockIdx.x * J.nblk_cap;
  __shared__ int wt[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    int i = base + threadIdx.x;
    int v = i < nblk ? row[i] : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane_id() >= o) x += y;
    }
    if (lane_id() == 31) wt[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = wt[threadIdx.x], s = w;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane_id() >= o) s += y;
      }
      wt[threadIdx.x] = s - w;
    }
    __syncthreads();
    int excl = carry + wt[threadIdx.x >> 5] + x - v;
    if (i < nblk) row[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) J.digit_total[blockIdx.x] = carry;
}

// stable scatter: each warp owns a contiguous 512-element chunk of the tile and walks it in
// order, ranking equal digits with match_any; chunk bases come from a per-digit prefix over warps
__global__ void __launch_bounds__(RS_THREADS) rs_scatter(SortPair sp, int pass) {
  const SortJob J = sp.j[blockIdx.y];
  const int n = *J.n_ptr;
  const int nblk = (n + RS_TILE - 1) / RS_TILE;
  if ((int)blockIdx.x >= nblk) return;
  __shared__ int cnt[RS_WARPS][256];
  __shared__ int dbase[256];
  const unsigned int* kin = J.keys[pass & 1];
  const unsigned int* vin = J.vals[pass & 1];
  unsigned int* kout = J.keys[(pass & 1) ^ 1];
  unsigned int* vout = J.vals[(pass & 1) ^ 1];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  for (int k = threadIdx.x; k < RS_WARPS * 256; k += RS_THREADS) (&cnt[0][0])[k] = 0;
  {  // exclusive prefix of the digit totals
    int v = J.digit_total[threadIdx.x];
    __shared__ int wt[RS_WARPS];
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wt[warp] = x;
    __syncthreads();
    int off = 0;
    for (int w = 0; w < warp; w++) off += wt[w];
    dbase[threadIdx.x] = off + x - v + J.hist[(size_t)threadIdx.x * J.nblk_cap + blockIdx.x];
  }
  __syncthreads();
  unsigned int key[RS_ITEMS], val[RS_ITEMS];
  int rank[RS_ITEMS];
  const int chunk0 = blockIdx.x * RS_TILE + warp * (RS_TILE / RS_WARPS);
#pragma unroll
  for (int s = 0; s < RS_ITEMS; s++) {
    int i = chunk0 + s * 32 + lane;
    bool in = i < n;
    key[s] = in ? kin[i] : 0xffffffffu;
    val[s] = in ? vin[i] : 0u;
    int d = (int)((key[s] >> (8 * pass)) & 255);
    unsigned int m = warp_match8(d, in);
    int r = __popc(m & ((1u << lane) - 1u));
    if (in) {
      rank[s] = cnt[warp][d] + r;
    } else {
      rank[s] = -1;
    }
    __syncwarp();
    if (in && r == 0) cnt[warp][d] += __popc(m);
    __syncwarp();
  }
  __syncthreads();
  // per-digit exclusive prefix over warps (in place)
  {
    int d = threadIdx.x;
    int run = 0;
    for (int w = 0; w < RS_WARPS; w++) { int c = cnt[w][d]; cnt[w][d] = run; run += c; }
  }
  __syncthreads();
#pragma unroll
  for (int s = 0; s < RS_ITEMS; s++) {
    if (rank[s] >= 0) {
      int d = (key[s] >> (8 * pass)) & 255;
      int dst = dbase[d] + cnt[warp][d] + rank[s];
      kout[dst] = key[s];
      vout[dst] = val[s];
    }
  }
}

// sorts keys[0]/vals[0] (4 passes of 8 bits -> result back in keys[0]/vals[0])
inline int radix_sort_pairs(const SortPair& sp, int max_n_cap, cudaStream_t stream) {
  int nblk = (max_n_cap + RS_TILE - 1) / RS_TILE;
  if (nblk < 1) nblk = 1;
  for (int pass = 0; pass < 4; pass++) {
    LAUNCH("rs_hist", rs_hist, dim3(nblk, sp.njobs), RS_THREADS, 0, stream, sp, pass);
    LAUNCH("rs_scan", rs_scan, dim3(256, sp.njobs), 1024, 0, stream, sp);
    LAUNCH("rs_scatter", rs_scatter, dim3(nblk, sp.njobs), RS_THREADS, 0, stream, sp, pass);
  }
  CUDA_OK(cudaGetLastError());
  return LOAM_OK;
}

// ---- voxel grid: segment heads of the sorted keys -> one centroid per head --------------------
constexpr int VX_THREADS = 256;
constexpr int VX_ITEMS = 8;
constexpr int VX_TILE = VX_THREADS * VX_ITEMS;

struct VoxelJob {
  const unsigned int* keys;   // sorted
  const unsigned int* vals;   // sorted payload = index into pts
  const float4* pts;
  const int* n_ptr;
  int* block_heads;           // [nblk_cap] heads per tile (scanned in place)
  float4* out;                // output points (+ *out_base_ptr if given)
  const int* out_base_ptr;    // optional device offset into `out`
  int* out_count;             // number of output points
  int* seg_count;             // optional: per-segment output counts (segment = key >> seg_shift)
  int seg_shift;
  int nblk_cap;
};
struct VoxelPair { VoxelJob j[2]; int njobs; };

__device__ __forceinline__ bool vx_is_head(const unsigned int* keys, int i) {
  return keys[i] != 0xffffffffu && (i == 0 || keys[i] != keys[i - 1]);
}

__global__ void __launch_bounds__(VX_THREADS) vx_count_heads(VoxelPair vp) {
  const VoxelJob J = vp.j[blockIdx.y];
  const int n = *J.n_ptr;
  int c = 0;
  int base = blockIdx.x * VX_TILE;
  for (int k = threadIdx.x; k < VX_TILE; k += VX_THREADS) {
    int i = base + k;
    if (i < n && vx_is_head(J.keys, i)) c++;
  }
  __shared__ int ws[VX_THREADS / 32];
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane_id() == 0) ws[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < VX_THREADS / 32; w++) t += ws[w];
    J.block_heads[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(1024) vx_scan_blocks(VoxelPair vp) {
  const VoxelJob J = vp.j[blockIdx.x];
  const int n = *J.n_ptr;
  const int nblk = min((n + VX_TILE - 1) / VX_TILE, J.nblk_cap);
  __shared__ int wt[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    int i = base + threadIdx.x;
    int v = i < nblk ? J.block_heads[i] : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane_id() >= o) x += y;
    }
    if (lane_id() == 31) wt[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = wt[threadIdx.x], s = w;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane_id() >= o) s += y;
      }
      wt[threadIdx.x] = s - w;
    }
    __syncthreads();
    int excl = carry + wt[threadIdx.x >> 5] + x - v;
    if (i < nblk) J.block_heads[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *J.out_count = carry;
}

// centroid per head: fp32 running sum from 0 in sorted order (= ascending input index inside a
// voxel because the sort is stable), divided by (float)count -- PCL 1.7 VoxelGrid::applyFilter
__global__ void __launch_bounds__(VX_THREADS) vx_reduce(VoxelPair vp) {
  const VoxelJob J = vp.j[blockIdx.y];
  const int n = *J.n_ptr;
  const int base = blockIdx.x * VX_TILE;
  if (base >= n) return;
  __shared__ int ws[VX_THREADS / 32];
  const int out_base = (J.out_base_ptr ? *J.out_base_ptr : 0) + J.block_heads[blockIdx.x];
  // thread t owns items [t*VX_ITEMS, (t+1)*VX_ITEMS) of the tile so slots follow sorted order
  int first = base + threadIdx.x * VX_ITEMS;
  int hc = 0;
  bool head[VX_ITEMS];
#pragma unroll
  for (int s = 0; s < VX_ITEMS; s++) {
    int i = first + s;
    head[s] = (i < n) && vx_is_head(J.keys, i);
    hc += head[s] ? 1 : 0;
  }
  int x = hc;
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane_id() >= o) x += y;
  }
  if (lane_id() == 31) ws[threadIdx.x >> 5] = x;
  __syncthreads();
  int off = 0;
  for (int w = 0; w < (int)(threadIdx.x >> 5); w++) off += ws[w];
  int slot = out_base + off + x - hc;
#pragma unroll
  for (int s = 0; s < VX_ITEMS; s++) {
    if (!head[s]) continue;
    int i = first + s;
    unsigned int hk = J.keys[i];
    float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
    int q = i;
    while (q < n && J.keys[q] == hk) {
      float4 p = J.pts[J.vals[q]];
      c0 += p.x; c1 += p.y; c2 += p.z; c3 += p.w;
      q++;
    }
    float cntf = (float)(q - i);
    c0 /= cntf; c1 /= cntf; c2 /= cntf; c3 /= cntf;
    J.out[slot] = make_float4(c0, c1, c2, c3);
    if (J.seg_count) atomicAdd(&J.seg_count[hk >> J.seg_shift], 1);
    slot++;
  }
}

inline int voxel_reduce_sorted(const VoxelPair& vp, int max_n_cap, cudaStream_t stream) {
  int nblk = (max_n_cap + VX_TILE - 1) / VX_TILE;
  if (nblk < 1) nblk = 1;
  LAUNCH("vx_count_heads", vx_count_heads, dim3(nblk, vp.njobs), VX_THREADS, 0, stream, vp);
  LAUNCH("vx_scan_blocks", vx_scan_blocks, vp.njobs, 1024, 0, stream, vp);
  LAUNCH("vx_reduce", vx_reduce, dim3(nblk, vp.njobs), VX_THREADS, 0, stream, vp);
  CUDA_OK(cudaGetLastError());
  return LOAM_OK;
}

// ---- whole-cloud pcl::VoxelGrid key: idx = i + j*dx + k*dx*dy relative to the bounding box ----
struct CloudKeyJob {
  const float4* pts;
  const int* n_ptr;
  unsigned int* bbox;     // 6 flipped floats, armed (min=0xffffffff, max=0)
  unsigned int* keys;
  unsigned int* vals;
  int* overflow;          // set when dx*dy*dz exceeds INT32_MAX (PCL returns the input unfiltered)
  float leaf;
};
struct CloudKeyPair { CloudKeyJob j[2]; int njobs; };

__global__ void __launch_bounds__(256) vx_bbox(CloudKeyPair cp) {
  const CloudKeyJob J = cp.j[blockIdx.y];
  const int n = *J.n_ptr;
  float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
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
  if (lane_id() == 0 && mn[0] <= mx[0])
    for (int a = 0; a < 3; a++) { atomicMin(&J.bbox[a], float_flip(mn[a])); atomicMax(&J.bbox[3 + a], float_flip(mx[a])); }
}

__global__ void __launch_bounds__(256) vx_cloud_keys(CloudKeyPair cp) {
  const CloudKeyJob J = cp.j[blockIdx.y];
  const int n = *J.n_ptr;
  const float inv = 1.0f / J.leaf;
  float mn[3], mx[3];
  for (int a = 0; a < 3; a++) { mn[a] = float_unflip(J.bbox[a]); mx[a] = float_unflip(J.bbox[3 + a]); }
  long long dx = (long long)((mx[0] - mn[0]) * inv) + 1, dy = (long long)((mx[1] - mn[1]) * inv) + 1, dz = (long long)((mx[2] - mn[2]) * inv) + 1;
  bool overflow = (n > 0) && (dx * dy * dz > 2147483647LL);
  int minb[3], divb[3];
  for (int a = 0; a < 3; a++) {
    minb[a] = (int)floorf(mn[a] * inv);
    int maxb = (int)floorf(mx[a] * inv);
    divb[a] = maxb - minb[a] + 1;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *J.overflow = overflow ? 1 : 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    float4 p = J.pts[i];
    unsigned int key;
    if (overflow) key = (unsigned int)i;  // every point its own voxel, input order kept
    else if (!isfinite(p.x) || !isfinite(p.y) || !isfinite(p.z)) key = 0xffffffffu;
    else {
      int i0 = (int)(floorf(p.x * inv) - (float)minb[0]);
      int i1 = (int)(floorf(p.y * inv) - (float)minb[1]);
      int i2 = (int)(floorf(p.z * inv) - (float)minb[2]);
      key = (unsigned int)(i0 + i1 * divb[0] + i2 * divb[0] * divb[1]);
    }
    J.keys[i] = key;
    J.vals[i] = (unsigned int)i;
  }
}

__global__ void vx_rearm_bbox(CloudKeyPair cp) {
  if (threadIdx.x < 6) {
    for (int j = 0; j < cp.njobs; j++) cp.j[j].bbox[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;
  }
}

}  // namespace loamgpu
