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
  int* cell_count;       // max_cells
  int* cell_start;       // max_cells
  int* cursor;           // 1
  int* pt_cell;          // cap
  int* pt_rank;          // cap
  float4* sorted_pts;    // cap
  int* sorted_idx;       // cap
  const float4* pts;     // source cloud
  const int* n_ptr;      // device count of the source cloud
  const int* enable;     // optional device flag (nullptr = always build)
  int max_cells;
  int cap;
  float h0;
};

struct GridView {        // what queries need
  const GridDesc* desc;
  const int* cell_count;
  const int* cell_start;
  const float4* sorted_pts;
  const int* sorted_idx;
};

struct GridStorage {
  DevBuf<GridDesc> desc;
  DevBuf<unsigned int> bbox;
  DevBuf<int> cell_count, cell_start, cursor, pt_cell, pt_rank, sorted_idx;
  DevBuf<float4> sorted_pts;
  int max_cells = 0, cap = 0;
  int alloc(int cap_points, int max_cells_, cudaStream_t s) {
    cap = cap_points; max_cells = max_cells_;
    int r = LOAM_OK;
    r |= desc.alloc_zero(1, s); r |= bbox.alloc(6); r |= cell_count.alloc_zero(max_cells, s); r |= cell_start.alloc_zero(max_cells, s);
    r |= cursor.alloc_zero(1, s); r |= pt_cell.alloc(cap); r |= pt_rank.alloc(cap); r |= sorted_idx.alloc(cap); r |= sorted_pts.alloc(cap);
    if (r == LOAM_OK) {
      unsigned int init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
      if (cudaMemcpyAsync(bbox.p, init, sizeof(init), cudaMemcpyHostToDevice, s) != cudaSuccess) r = LOAM_ERR_CUDA;
      cudaStreamSynchronize(s);
    }
    return r;
  }
  GridBuf buf(const float4* pts, const int* n_ptr, const int* enable, float h0) {
    GridBuf b;
    b.desc = desc.p; b.bbox = bbox.p; b.cell_count = cell_count.p; b.cell_start = cell_start.p; b.cursor = cursor.p;
    b.pt_cell = pt_cell.p; b.pt_rank = pt_rank.p; b.sorted_pts = sorted_pts.p; b.sorted_idx = sorted_idx.p;
    b.pts = pts; b.n_ptr = n_ptr; b.enable = enable; b.max_cells = max_cells; b.cap = cap; b.h0 = h0;
    return b;
  }
  GridView view() const {
    GridView v;
    v.desc = desc.p; v.cell_count = cell_count.p; v.cell_start = cell_start.p; v.sorted_pts = sorted_pts.p; v.sorted_idx = sorted_idx.p;
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
  if (lane_id() == 0 && mn[0] <= mx[0]) {
    for (int a = 0; a < 3; a++) {
      atomicMin(&g.bbox[a], float_flip(mn[a]));
      atomicMax(&g.bbox[3 + a], float_flip(mx[a]));
    }
  }
}

__device__ __forceinline__ GridDesc grid_make_desc(const GridBuf& g) {
  GridDesc d;
  int n = min(*g.n_ptr, g.cap);
  d.n_points = n;
  d.valid = n > 0 ? 1 : 0;
  float mn[3], mx[3];
  for (int a = 0; a < 3; a++) { mn[a] = float_unflip(g.bbox[a]); mx[a] = float_unflip(g.bbox[3 + a]); }
  if (n <= 0) { mn[0] = mn[1] = mn[2] = 0.f; mx[0] = mx[1] = mx[2] = 0.f; }
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

// descriptor + zero the cell counters (every block recomputes the descriptor from the bbox)
__global__ void __launch_bounds__(256) grid_setup(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  GridDesc d = grid_make_desc(g);
  for (int c = blockIdx.x * 256 + threadIdx.x; c < d.ncells; c += gridDim.x * 256) g.cell_count[c] = 0;
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
    g.pt_rank[i] = atomicAdd(&g.cell_count[c], 1);
  }
}

__global__ void __launch_bounds__(256) grid_alloc(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  const int ncells = g.desc->ncells;
  for (int c = blockIdx.x * 256 + threadIdx.x; c < ncells; c += gridDim.x * 256) {
    int k = g.cell_count[c];
    g.cell_start[c] = k > 0 ? atomicAdd(g.cursor, k) : 0;
  }
}

__global__ void __launch_bounds__(256) grid_scatter(GridPair gp) {
  const GridBuf g = gp.g[blockIdx.y];
  if (!grid_enabled(g)) return;
  const int n = g.desc->n_points;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    int dst = g.cell_start[g.pt_cell[i]] + g.pt_rank[i];
    g.sorted_pts[dst] = g.pts[i];
    g.sorted_idx[dst] = i;
  }
}

inline int grid_build_pair(GridPair gp, int nblocks, cudaStream_t stream) {
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

// exact 1-NN, warp cooperative.  All lanes return the same (idx, d2); idx = -1 if the cloud is
// empty or nothing lies within sqrt(maxd2).  Tie-break: lowest original index.
__device__ __forceinline__ void grid_nn1_warp(const GridView& g, float qx, float qy, float qz, float maxd2, int& out_idx, float& out_d2) {
  const GridDesc d = *g.desc;
  float bd = 3.0e38f;
  int bi = 0x7fffffff;
  if (d.valid) {
    int cx = (int)floorf((qx - d.ox) * d.inv_h), cy = (int)floorf((qy - d.oy) * d.inv_h), cz = (int)floorf((qz - d.oz) * d.inv_h);
    // the query may lie outside the grid: clamp its cell and account for the gap in the bound
    int qcx = min(max(cx, 0), d.nx - 1), qcy = min(max(cy, 0), d.ny - 1), qcz = min(max(cz, 0), d.nz - 1);
    const float maxr = sqrtf(maxd2);
    const int lane = lane_id();
    for (int k = 0;; k++) {
      int x0 = max(qcx - k, 0), x1 = min(qcx + k, d.nx - 1);
      int y0 = max(qcy - k, 0), y1 = min(qcy + k, d.ny - 1);
      int z0 = max(qcz - k, 0), z1 = min(qcz + k, d.nz - 1);
      for (int z = z0; z <= z1; z++)
        for (int y = y0; y <= y1; y++)
          for (int x = x0; x <= x1; x++) {
            int cheb = max(max(abs(x - qcx), abs(y - qcy)), abs(z - qcz));
            if (cheb != k) continue;
            int c = (z * d.ny + y) * d.nx + x;
            int cnt = g.cell_count[c];
            if (cnt == 0) continue;
            int st = g.cell_start[c];
            for (int t = lane; t < cnt; t += 32) {
              float4 p = g.sorted_pts[st + t];
              float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
              float dd = 0.f;
              dd += dx * dx; dd += dy * dy; dd += dz * dz;
              int id = g.sorted_idx[st + t];
              if (nn_less(dd, id, bd, bi)) { bd = dd; bi = id; }
            }
          }
      // warp-wide best so far
      float wb = bd;
      for (int o = 16; o > 0; o >>= 1) wb = fminf(wb, __shfl_xor_sync(0xffffffffu, wb, o));
      // every unexamined point is farther than k*h - slack along some axis from the CLAMPED cell;
      // if the query is outside the grid the true distance is even larger.
      float bound = (float)k * d.h - 0.01f;
      bool covered = (x0 == 0 && x1 == d.nx - 1 && y0 == 0 && y1 == d.ny - 1 && z0 == 0 && z1 == d.nz - 1);
      if (covered) break;
      if (k >= 1 && bound > 0.f && (wb < bound * bound || bound > maxr)) break;
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
