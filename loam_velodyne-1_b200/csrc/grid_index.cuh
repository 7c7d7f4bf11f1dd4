// grid_index.cuh -- uniform-grid spatial index replacing pcl::KdTreeFLANN.
//
// Only the RESULT of nearestKSearch is part of the contract: exact k-NN under
// d2 = ((dx*dx)+dy*dy)+dz*dz in fp32 (flann::L2_Simple), ascending (d2, index).  Any exact index
// is admissible, so instead of a kd-tree (pointer chasing, sequential build) the cloud is
// counting-sorted into cells of a dense grid over its bounding box: build is 5 streaming
// kernels, queries are warp-cooperative scans of neighbouring cells with conservative pruning.
//   laserOdometry.cpp:436-437,905-906  kdtree*Last->setInputCloud      -> grid build, h ~ 1 m
//   laserMapping.cpp:707-708           kdtree*FromMap->setInputCloud   -> grid build, h = 1.02 m
#pragma once
#include <cooperative_groups.h>
#include "common.cuh"

namespace loamgpu {

struct GridDesc {
  float ox, oy, oz, h, inv_h;
  int nx, ny, nz, ncells;
  int n_points;
  int valid;
};

struct GridBuf {         // passed by value to kernels
  GridDesc* desc;        // device
  unsigned int* bbox;    // 6 flipped floats: min xyz, max xyz
  int2* cell;            // max_cells: {start, count}; the cells of one x-row are contiguous in sorted_pts
  int* cursor;           // 1
  int* pt_cell;          // cap
  int* pt_rank;          // cap
  float4* sorted_pts;    // cap: (x, y, z, original index as int bits)
  const float4* pts;     // source cloud
  const int* n_ptr;      // device count of the source cloud
  const int* enable;     // optional device flag (nullptr = always build)
  int max_cells;
  int cap;
  float h0;
};

struct GridView {        // what queries need
  const GridDesc* desc;
  const int2* cell;
  const float4* sorted_pts;
};

struct GridStorage {
  DevBuf<GridDesc> desc;
  DevBuf<unsigned int> bbox;
  DevBuf<int2> cell;
  DevBuf<int> cursor, pt_cell, pt_rank;
  DevBuf<float4> sorted_pts;
  int max_cells = 0, cap = 0;
  int alloc(int cap_points, int max_cells_, cudaStream_t s) {
    cap = cap_points; max_cells = max_cells_;
    int r = LOAM_OK;
    r |= desc.alloc_zero(1, s); r |= bbox.alloc(6); r |= cell.alloc_zero(max_cells, s);
    r |= cursor.alloc_zero(1, s); r |= pt_cell.alloc(cap); r |= pt_rank.alloc(cap); r |= sorted_pts.alloc(cap);
    if (r == LOAM_OK) {
      unsigned int init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
      if (cudaMemcpyAsync(bbox.p, init, sizeof(init), cudaMemcpyHostToDevice, s) != cudaSuccess) r = LOAM_ERR_CUDA;
      cudaStreamSynchronize(s);
    }
    return r;
  }
  GridBuf buf(const float4* pts, const int* n_ptr, const int* enable, float h0) {
    GridBuf b;
    b.desc = desc.p; b.bbox = bbox.p; b.cell = cell.p; b.cursor = cursor.p;
    b.pt_cell = pt_cell.p; b.pt_rank = pt_rank.p; b.sorted_pts = sorted_pts.p;
    b.pts = pts; b.n_ptr = n_ptr; b.enable = enable; b.max_cells = max_cells; b.cap = cap; b.h0 = h0;
    return b;
  }
  GridView view() const {
    GridView v;
    v.desc = desc.p; v.cell = cell.p; v.sorted_pts = sorted_pts.p;
    return v;
  }
};

struct GridPair { GridBuf g[2]; };

__device__ __forceinline__ bool grid_enabled(const GridBuf& g) { return g.enable == nullptr || *g.enable != 0; }

__device__ __forceinline__ int grid_cell_coord(float v, float o, float inv_h, int n) {
  int c = (int)floorf((v - o) * inv_h);
  return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

// bounding box (block reduction + flipped-float atomics)
__global__ void __launch_bounds__(256) grid_bbox(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  const int n = min(*g.n_ptr, g.cap);
  float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    float4 p = g.pts[i];
    mn[0] = fminf(mn[0], p.x); mx[0] = fmaxf(mx[0], p.x);
    mn[1] = fminf(mn[1], p.y); mx[1] = fmaxf(mx[1], p.y);
    mn[2] = fminf(mn[2], p.z); mx[2] = fmaxf(mx[2], p.z);
  }
  for (int a = 0; a < 3; a++) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  }
  // one set of atomics per CTA (per-warp atomics on six addresses serialise for ~10 us)
  __shared__ float s_mn[8][3], s_mx[8][3];
  const int warp = threadIdx.x >> 5;
  if (lane_id() == 0) for (int a = 0; a < 3; a++) { s_mn[warp][a] = mn[a]; s_mx[warp][a] = mx[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x;
    float lo = s_mn[0][a], hi = s_mx[0][a];
    for (int w = 1; w < 8; w++) { lo = fminf(lo, s_mn[w][a]); hi = fmaxf(hi, s_mx[w][a]); }
    if (lo <= hi) {
      atomicMin(&g.bbox[a], float_flip(lo));
      atomicMax(&g.bbox[3 + a], float_flip(hi));
    }
  }
}

__device__ __forceinline__ GridDesc grid_make_desc(const GridBuf& g, int n_given = -1) {
  GridDesc d;
  int n = min(n_given >= 0 ? n_given : *g.n_ptr, g.cap);
  d.n_points = n;
  d.valid = n > 0 ? 1 : 0;
  float mn[3], mx[3];
  for (int a = 0; a < 3; a++) { mn[a] = float_unflip(g.bbox[a]); mx[a] = float_unflip(g.bbox[3 + a]); }
  if (n <= 0) { mn[0] = mn[1] = mn[2] = 0.f; mx[0] = mx[1] = mx[2] = 0.f; }
  // an infinite coordinate (NaN is ignored by fminf / fmaxf) must not turn the cell count into garbage or keep the
  // loop below from ending: such a cloud gets one cell and no valid flag, every query then finds nothing
  bool finite = true;
  for (int a = 0; a < 3; a++) finite = finite && isfinite(mn[a]) && isfinite(mx[a]);
  if (!finite) { mn[0] = mn[1] = mn[2] = 0.f; mx[0] = mx[1] = mx[2] = 0.f; d.valid = 0; }
  float h = g.h0;
  int nx, ny, nz;
  while (true) {
    nx = (int)((mx[0] - mn[0]) / h) + 1; ny = (int)((mx[1] - mn[1]) / h) + 1; nz = (int)((mx[2] - mn[2]) / h) + 1;
    if ((long long)nx * ny * nz <= (long long)g.max_cells) break;
    h *= 2.f;
  }
  d.ox = mn[0]; d.oy = mn[1]; d.oz = mn[2]; d.h = h; d.inv_h = 1.0f / h;
  d.nx = nx; d.ny = ny; d.nz = nz; d.ncells = nx * ny * nz;
  return d;
}

// descriptor + zero the cell table (every block recomputes the descriptor from the bbox)
__global__ void __launch_bounds__(256) grid_setup(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  GridDesc d = grid_make_desc(g);
  for (int c = blockIdx.x * 256 + threadIdx.x; c < d.ncells; c += gridDim.x * 256) g.cell[c] = make_int2(0, 0);
  if (blockIdx.x == 0 && threadIdx.x == 0) { *g.desc = d; *g.cursor = 0; }
}

__global__ void __launch_bounds__(256) grid_count(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  const GridDesc d = *g.desc;
  if (blockIdx.x == 0 && threadIdx.x < 6) g.bbox[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;  // re-arm for next build
  for (int i = blockIdx.x * 256 + threadIdx.x; i < d.n_points; i += gridDim.x * 256) {
    float4 p = g.pts[i];
    int cx = grid_cell_coord(p.x, d.ox, d.inv_h, d.nx), cy = grid_cell_coord(p.y, d.oy, d.inv_h, d.ny), cz = grid_cell_coord(p.z, d.oz, d.inv_h, d.nz);
    int c = (cz * d.ny + cy) * d.nx + cx;
    g.pt_cell[i] = c;
    g.pt_rank[i] = atomicAdd(&g.cell[c].y, 1);
  }
}

// one warp per x-row (fixed y, z): the row's cells get consecutive slots, so a run of x-adjacent
// cells is ONE contiguous range of sorted_pts and a query needs one header load per row end.
__global__ void __launch_bounds__(256) grid_alloc(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  const GridDesc d = *g.desc;
  const int lane = lane_id();
  const int nrows = d.ny * d.nz;
  for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < nrows; row += gridDim.x * 8) {
    int2* cells = g.cell + (size_t)row * d.nx;
    int tot = 0;
    for (int x = lane; x < d.nx; x += 32) tot += cells[x].y;
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (tot == 0) continue;   // all starts stay 0, all counts 0: every range of this row is empty
    int base = 0;
    if (lane == 0) base = atomicAdd(g.cursor, tot);
    int run = __shfl_sync(0xffffffffu, base, 0);
    for (int x0 = 0; x0 < d.nx; x0 += 32) {
      int x = x0 + lane;
      int c = x < d.nx ? cells[x].y : 0;
      int incl = c;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (x < d.nx) cells[x].x = run + incl - c;
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
}

__global__ void __launch_bounds__(256) grid_scatter(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  const int n = g.desc->n_points;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    int dst = g.cell[g.pt_cell[i]].x + g.pt_rank[i];
    float4 p = g.pts[i];
    p.w = __int_as_float(i);
    g.sorted_pts[dst] = p;
  }
}

// ---- the same five phases as device functions, for kernels that build an index in their own tail
// (lo_optimize's epilogue, lm_front): `tid` / `nthr` = index and number of the threads taking part (all
// threads of all participating CTAs), the caller separates the phases with its own grid / cluster barrier.
// Phase 1: the caller folds the points' bounding box into g.bbox (flipped-float atomics, armed on entry).
__device__ __forceinline__ void bbox_block_commit(unsigned int* bbox, const float* mn, const float* mx, float (*s_mn)[3], float (*s_mx)[3], int nwarps) {
  float lo[3], hi[3];
  for (int a = 0; a < 3; a++) {
    lo[a] = mn[a]; hi[a] = mx[a];
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  }
  const int warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane_id() == 0) for (int a = 0; a < 3; a++) { s_mn[warp][a] = lo[a]; s_mx[warp][a] = hi[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x;
    float l = s_mn[0][a], h = s_mx[0][a];
    for (int w = 1; w < nwarps; w++) { l = fminf(l, s_mn[w][a]); h = fmaxf(h, s_mx[w][a]); }
    if (l <= h) {
      atomicMin(&bbox[a], float_flip(l));
      atomicMax(&bbox[3 + a], float_flip(h));
    }
  }
}
// Phase 2: descriptor from the bounding box (every thread computes the same one) + zeroed cell table
__device__ __forceinline__ GridDesc gridp_setup(const GridBuf& g, int n, int tid, int nthr) {
  GridBuf gl = g;
  unsigned int bb[6];
  for (int a = 0; a < 6; a++) bb[a] = __ldcg(&g.bbox[a]);
  gl.bbox = bb;
  GridDesc d = grid_make_desc(gl, n);
  int4* c4 = (int4*)g.cell;   // two cells per 16-byte store (the table is cudaMalloc-aligned)
  const int n4 = (d.ncells + 1) / 2;
  for (int c = tid; c < n4; c += nthr) c4[c] = make_int4(0, 0, 0, 0);
  if (tid == 0) { *g.desc = d; *g.cursor = 0; }
  return d;
}
// Phase 3: cell of every point and its rank inside the cell; re-arms the bounding box for the next build
__device__ __forceinline__ void gridp_count(const GridBuf& g, const GridDesc& d, int tid, int nthr) {
  if (tid < 6) g.bbox[tid] = tid < 3 ? 0xffffffffu : 0u;
  for (int i = tid; i < d.n_points; i += nthr) {
    float4 p = __ldcg(&g.pts[i]);
    int cx = grid_cell_coord(p.x, d.ox, d.inv_h, d.nx), cy = grid_cell_coord(p.y, d.oy, d.inv_h, d.ny), cz = grid_cell_coord(p.z, d.oz, d.inv_h, d.nz);
    int c = (cz * d.ny + cy) * d.nx + cx;
    g.pt_cell[i] = c;
    g.pt_rank[i] = atomicAdd(&g.cell[c].y, 1);
  }
}
// Phase 4: one warp per x-row, as grid_alloc
__device__ __forceinline__ void gridp_alloc(const GridBuf& g, const GridDesc& d, int wid, int nwarps) {
  const int lane = lane_id();
  const int nrows = d.ny * d.nz;
  for (int row = wid; row < nrows; row += nwarps) {
    int2* cells = g.cell + (size_t)row * d.nx;
    int tot = 0;
    for (int x = lane; x < d.nx; x += 32) tot += __ldcg(&cells[x].y);
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (tot == 0) continue;
    int base = 0;
    if (lane == 0) base = atomicAdd(g.cursor, tot);
    int run = __shfl_sync(0xffffffffu, base, 0);
    for (int x0 = 0; x0 < d.nx; x0 += 32) {
      int x = x0 + lane;
      int c = x < d.nx ? __ldcg(&cells[x].y) : 0;
      int incl = c;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (x < d.nx) cells[x].x = run + incl - c;
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
}
// Phase 5: counting-sort scatter
__device__ __forceinline__ void gridp_scatter(const GridBuf& g, const GridDesc& d, int tid, int nthr) {
  for (int i = tid; i < d.n_points; i += nthr) {
    int dst = __ldcg(&g.cell[g.pt_cell[i]].x) + g.pt_rank[i];
    float4 p = __ldcg(&g.pts[i]);
    p.w = __int_as_float(i);
    g.sorted_pts[dst] = p;
  }
}

// The five phases as ONE kernel for clouds of up to a few hundred thousand points: a 16-CTA thread-block cluster
// per index (blockIdx.y = which of the pair), phases separated by hardware cluster barriers instead of kernel
// boundaries.  A cluster is not gang-scheduled over the whole GPU like a cooperative grid: it needs 16 CTA slots
// of one GPC, so it runs beside the persistent LM kernels like any small kernel.
constexpr int GC_CLUSTER = 16, GC_THREADS = 512;
__global__ void __cluster_dims__(GC_CLUSTER, 1, 1) __launch_bounds__(GC_THREADS) grid_build_cluster(GridPair gp) {
  namespace cg = cooperative_groups;
  cg::cluster_group cl = cg::this_cluster();
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;     // uniform over the cluster
  __shared__ float s_mn[GC_THREADS / 32][3], s_mx[GC_THREADS / 32][3];
  const int tid = blockIdx.x * GC_THREADS + threadIdx.x, nthr = GC_CLUSTER * GC_THREADS;
  const int n = min(*g.n_ptr, g.cap);
  {
    float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int i = tid; i < n; i += nthr) {
      const float4 p = g.pts[i];
      mn[0] = fminf(mn[0], p.x); mx[0] = fmaxf(mx[0], p.x);
      mn[1] = fminf(mn[1], p.y); mx[1] = fmaxf(mx[1], p.y);
      mn[2] = fminf(mn[2], p.z); mx[2] = fmaxf(mx[2], p.z);
    }
    bbox_block_commit(g.bbox, mn, mx, s_mn, s_mx, GC_THREADS / 32);
  }
  __threadfence();
  cl.sync();
  const GridDesc d = gridp_setup(g, n, tid, nthr);
  __threadfence();
  cl.sync();
  gridp_count(g, d, tid, nthr);
  __threadfence();
  cl.sync();
  gridp_alloc(g, d, tid >> 5, nthr >> 5);
  __threadfence();
  cl.sync();
  gridp_scatter(g, d, tid, nthr);
}

// small = both clouds are known (host side) to hold at most a few hundred thousand points: one cluster kernel;
// otherwise the five grid-wide kernels, which scale to millions of points
inline int grid_build_pair(GridPair gp, int nblocks, cudaStream_t stream, bool small = false) {
  if (small) {
    static bool attr = false;
    if (!attr) { CUDA_OK(cudaFuncSetAttribute(grid_build_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1)); attr = true; }
    LAUNCH("grid_build_cluster", grid_build_cluster, dim3(GC_CLUSTER, 2), GC_THREADS, 0, stream, gp);
    CUDA_OK(cudaGetLastError());
    return LOAM_OK;
  }
  dim3 grid(nblocks, 2);
  LAUNCH("grid_bbox", grid_bbox, grid, 256, 0, stream, gp);
  LAUNCH("grid_setup", grid_setup, grid, 256, 0, stream, gp);
  LAUNCH("grid_count", grid_count, grid, 256, 0, stream, gp);
  LAUNCH("grid_alloc", grid_alloc, grid, 256, 0, stream, gp);
  LAUNCH("grid_scatter", grid_scatter, grid, 256, 0, stream, gp);
  CUDA_OK(cudaGetLastError());
  return LOAM_OK;
}

// ---- queries -----------------------------------------------------------------------------
__device__ __forceinline__ bool nn_less(float d, int i, float bd, int bi) { return d < bd || (d == bd && i < bi); }

// The 3x3x3 block of cells around (cx,cy,cz) as 9 x-rows = 9 contiguous ranges of sorted_pts.
// ONE round of header loads for the whole block: lane 2r loads the first cell of row r, lane
// 2r+1 the last one.  Afterwards every lane holds all nine (begin, prefix) pairs in registers.
struct GridRows9 {
  int beg[9];
  int pre[10];   // pre[r] = number of candidates in rows < r; pre[9] = total
};
// W = lanes cooperating on one query (32, 16 or 8); gl = lane index inside the group.  With W = 8
// lane 0 also loads the ninth row.
template <int W>
__device__ __forceinline__ void grid_rows9(const GridView& g, const GridDesc& d, int cx, int cy, int cz, GridRows9& R) {
  const int gl = lane_id() & (W - 1);
  const unsigned int gmask = (W == 32) ? 0xffffffffu : (((1u << W) - 1u) << (lane_id() & ~(W - 1)));
  const int x0 = max(cx - 1, 0), x1 = min(cx + 1, d.nx - 1);
  int vb = 0, ve = 0, vb8 = 0, ve8 = 0;
  if (gl < 9 && x0 <= x1) {
    const int y = cy + (gl % 3) - 1, z = cz + (gl / 3) - 1;
    if (y >= 0 && y < d.ny && z >= 0 && z < d.nz) {
      const int2* row = g.cell + (size_t)(z * d.ny + y) * d.nx;
      int2 c0 = __ldg(&row[x0]);
      int2 c1 = __ldg(&row[x1]);
      vb = c0.x; ve = c1.x + c1.y;
    }
  }
  if (W == 8 && gl == 0 && x0 <= x1) {
    const int y = cy + 1, z = cz + 1;   // row 8
    if (y >= 0 && y < d.ny && z >= 0 && z < d.nz) {
      const int2* row = g.cell + (size_t)(z * d.ny + y) * d.nx;
      int2 c0 = __ldg(&row[x0]);
      int2 c1 = __ldg(&row[x1]);
      vb8 = c0.x; ve8 = c1.x + c1.y;
    }
  }
  R.pre[0] = 0;
#pragma unroll
  for (int k = 0; k < 9; k++) {
    int b, en;
    if (W == 8 && k == 8) { b = __shfl_sync(gmask, vb8, 0, W); en = __shfl_sync(gmask, ve8, 0, W); }
    else { b = __shfl_sync(gmask, vb, k, W); en = __shfl_sync(gmask, ve, k, W); }
    R.beg[k] = b;
    R.pre[k + 1] = R.pre[k] + (en - b);
  }
}
// Same block of cells, restricted to the cells that can hold a point with d2 <= r2 of the query
// (qx,qy,qz).  r2 is an upper bound of the answer's largest distance that the caller KNOWS to be
// attained by real points of this cloud (e.g. the previous LM iteration's neighbours measured from
// the moved query), so dropping everything farther away cannot change an exact k-NN result.  Each
// row lane tests its own row (distance of the query to the row's y-z square) and trims the x range
// to the cells the ball reaches; a row is still ONE contiguous range, so nothing else changes.
// Cell membership was decided in fp32 (floor((p - o) * inv_h)): every per-axis distance is reduced
// by a margin far above that rounding before it is squared, which only makes the test keep more.
template <int W>
__device__ __forceinline__ void grid_rows9_bounded(const GridView& g, const GridDesc& d, int cx, int cy, int cz,
                                                   float qx, float qy, float qz, float r2, GridRows9& R) {
  static_assert(W >= 16, "one lane per row");
  const int gl = lane_id() & (W - 1);
  const unsigned int gmask = (W == 32) ? 0xffffffffu : (((1u << W) - 1u) << (lane_id() & ~(W - 1)));
  int vb = 0, ve = 0;
  if (gl < 9) {
    const int oy = (gl % 3) - 1, oz = (gl / 3) - 1;
    const int y = cy + oy, z = cz + oz;
    if (y >= 0 && y < d.ny && z >= 0 && z < d.nz) {
      const float margin = 4.0e-3f;
      const float fx = qx - (d.ox + (float)cx * d.h), fy = qy - (d.oy + (float)cy * d.h), fz = qz - (d.oz + (float)cz * d.h);
      const float dyv = fmaxf((oy < 0 ? fy : (oy > 0 ? d.h - fy : 0.f)) - margin, 0.f);
      const float dzv = fmaxf((oz < 0 ? fz : (oz > 0 ? d.h - fz : 0.f)) - margin, 0.f);
      const float base = dyv * dyv + dzv * dzv;
      if (base <= r2) {
        const float dxm = fmaxf(fx - margin, 0.f), dxp = fmaxf(d.h - fx - margin, 0.f);
        const int x0 = max((base + dxm * dxm <= r2) ? cx - 1 : cx, 0);
        const int x1 = min((base + dxp * dxp <= r2) ? cx + 1 : cx, d.nx - 1);
        if (x0 <= x1) {
          const int2* row = g.cell + (size_t)(z * d.ny + y) * d.nx;
          int2 c0 = __ldg(&row[x0]);
          int2 c1 = __ldg(&row[x1]);
          vb = c0.x; ve = c1.x + c1.y;
        }
      }
    }
  }
  R.pre[0] = 0;
#pragma unroll
  for (int k = 0; k < 9; k++) {
    const int b = __shfl_sync(gmask, vb, k, W), en = __shfl_sync(gmask, ve, k, W);
    R.beg[k] = b;
    R.pre[k + 1] = R.pre[k] + (en - b);
  }
}
// position in sorted_pts of flat candidate t (0 <= t < pre[9])
__device__ __forceinline__ int grid_rows9_pos(const GridRows9& R, int t) {
  int pos = R.beg[0] + t;
#pragma unroll
  for (int k = 1; k < 9; k++)
    if (t >= R.pre[k]) pos = R.beg[k] + (t - R.pre[k]);
  return pos;
}

// exact 1-NN, warp cooperative.  All lanes return the same (idx, d2); idx = -1 if the cloud is
// empty or nothing lies within sqrt(maxd2).  Tie-break: lowest original index.
// Stage 1 examines the 27 cells around the (clamped) query cell with one batch of header loads and
// independent point loads; almost every query ends there.  Stage 2 (rare) walks further
// Chebyshev shells cell by cell.
__device__ __forceinline__ void grid_nn1_warp(const GridView& g, float qx, float qy, float qz, float maxd2, int& out_idx, float& out_d2) {
  const GridDesc d = *g.desc;
  float bd = 3.0e38f;
  int bi = 0x7fffffff;
  if (d.valid) {
    int cx = (int)floorf((qx - d.ox) * d.inv_h), cy = (int)floorf((qy - d.oy) * d.inv_h), cz = (int)floorf((qz - d.oz) * d.inv_h);
    // the query may lie outside the grid: clamp its cell (the true distance to any unexamined
    // point is then even larger than the bound used below)
    int qcx = min(max(cx, 0), d.nx - 1), qcy = min(max(cy, 0), d.ny - 1), qcz = min(max(cz, 0), d.nz - 1);
    const float maxr = sqrtf(maxd2);
    const int lane = lane_id();
    {
      GridRows9 R;
      grid_rows9<32>(g, d, qcx, qcy, qcz, R);
      const int total = R.pre[9];
      for (int t0 = lane; t0 < total; t0 += 128) {
        float4 pc[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int t = t0 + u * 32;
          pc[u] = g.sorted_pts[grid_rows9_pos(R, t < total ? t : 0)];
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          if (t0 + u * 32 >= total) break;
          const float4 p = pc[u];
          float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
          float dd = 0.f;
          dd += dx * dx; dd += dy * dy; dd += dz * dz;
          int id = __float_as_int(p.w);
          if (nn_less(dd, id, bd, bi)) { bd = dd; bi = id; }
        }
      }
    }
    for (int k = 1;; k++) {
      int x0 = max(qcx - k, 0), x1 = min(qcx + k, d.nx - 1);
      int y0 = max(qcy - k, 0), y1 = min(qcy + k, d.ny - 1);
      int z0 = max(qcz - k, 0), z1 = min(qcz + k, d.nz - 1);
      if (k >= 2) {
        // shell k, one lane per (y, z) row of the (2k+1)^2 square: a face row is one contiguous
        // x-range, an interior row contributes only its two end cells
        const int side = 2 * k + 1;
        for (int r0 = 0; r0 < side * side; r0 += 32) {
          const int r = r0 + lane;
          int b0 = 0, n0 = 0, b1 = 0, n1 = 0;
          if (r < side * side) {
            const int dy = r % side - k, dz = r / side - k;
            const int y = qcy + dy, z = qcz + dz;
            if (y >= 0 && y < d.ny && z >= 0 && z < d.nz) {
              const int2* row = g.cell + (size_t)(z * d.ny + y) * d.nx;
              const bool face = (abs(dy) == k) || (abs(dz) == k);
              if (face) {
                int2 c0 = __ldg(&row[x0]);
                int2 c1 = __ldg(&row[x1]);
                b0 = c0.x; n0 = c1.x + c1.y - c0.x;
              } else {
                if (qcx - k >= 0) { int2 c0 = __ldg(&row[qcx - k]); b0 = c0.x; n0 = c0.y; }
                if (qcx + k < d.nx) { int2 c1 = __ldg(&row[qcx + k]); b1 = c1.x; n1 = c1.y; }
              }
            }
          }
          for (int t = 0; t < n0 + n1; t++) {
            float4 p = g.sorted_pts[t < n0 ? b0 + t : b1 + (t - n0)];
            float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
            float dd = 0.f;
            dd += dx * dx; dd += dy * dy; dd += dz * dz;
            int id = __float_as_int(p.w);
            if (nn_less(dd, id, bd, bi)) { bd = dd; bi = id; }
          }
        }
      }
      // warp-wide best so far
      float wb = bd;
      for (int o = 16; o > 0; o >>= 1) wb = fminf(wb, __shfl_xor_sync(0xffffffffu, wb, o));
      // every unexamined point is farther than k*h - slack along some axis from the CLAMPED cell
      float bound = (float)k * d.h - 0.01f;
      bool covered = (x0 == 0 && x1 == d.nx - 1 && y0 == 0 && y1 == d.ny - 1 && z0 == 0 && z1 == d.nz - 1);
      if (covered) break;
      if (bound > 0.f && (wb < bound * bound || bound > maxr)) break;
    }
    for (int o = 16; o > 0; o >>= 1) {
      float od = __shfl_xor_sync(0xffffffffu, bd, o);
      int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (nn_less(od, oi, bd, bi)) { bd = od; bi = oi; }
    }
  }
  if (bi == 0x7fffffff) { out_idx = -1; out_d2 = 3.0e38f; }
  else { out_idx = bi; out_d2 = bd; }
}

}  // namespace loamgpu
