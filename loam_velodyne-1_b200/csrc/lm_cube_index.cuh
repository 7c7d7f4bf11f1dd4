// lm_cube_index.cuh -- persistent per-cube spatial index of the cube map.
//
// What it replaces: laserMapping.cpp:707-708 builds both kd-trees over the WHOLE local map (the <= 125
// valid cubes, gathered at :674-679) on every mapped sweep, although a sweep only changes the handful of
// cubes it inserts points into (:980-1036).  The round-1 product did the same with its grid index
// (gather + bbox + count + alloc + scatter over every local-map point, every pass).
//
// Here every cube of the 5x5x5 window around the sensor keeps its OWN index between passes:
//   * one lattice for the whole map: cells of 50/48 m aligned with the cube boundaries, 48^3 cells per cube.
//     A query's 3x3x3 block of cells covers every point within 1 m (the only neighbours the reference
//     accepts, :719 / :826), whichever cubes the cells belong to;
//   * 125 slots per feature type, slot = cube world coordinates mod 5 (the window is 5 cubes wide, so two
//     cubes of one window never share a slot).  A slot holds {cell table, the cube's points in cell order}
//     and the signature {world coordinates, version, count} of the cube it was built from;
//   * every rewrite of a cube (lm_refilter_cube) stamps a fresh version into its descriptor; the window
//     shift moves descriptors (and versions) along.  lm_setup compares the signatures of the <= 250 cubes of
//     the local map with their slots and lists the stale ones; ci_build rebuilds exactly those -- the cubes
//     the previous sweep inserted into, plus the cubes that entered the window.  Everything else (typically
//     most of a large map) is reused as it is;
//   * the point regions live in one index arena per type; a slot keeps its region (12 % head room) while the
//     cube still fits, otherwise it takes a new one from the bump pointer; when the arena is exhausted the
//     whole table starts over (every slot stale once).
// Only the RESULT of nearestKSearch is part of the contract (grid_index.cuh): exact 5-NN in fp32, ascending
// (d2, index), index = position in the gathered local map.  A candidate carries (window position of its cube,
// position inside the cube); the window position orders like the cube's rank in laserCloudValidInd (both
// follow the (i, j, k) loop of :647-673), so ties resolve exactly as they do over the gathered cloud.
#pragma once
#include <cooperative_groups.h>
#include "grid_index.cuh"

namespace loamgpu {

constexpr int CUBE_W = 21, CUBE_H = 11, CUBE_D = 21;     // laserMapping.cpp:67-69
constexpr int CUBE_NUM = CUBE_W * CUBE_H * CUBE_D;       // :70

// ver: stamped by every kernel that rewrites the cube's points (lm_refilter.cuh); 0 = never written on the device
struct CubeDesc { int off; int cnt; int dirty; int ver; };   // dirty: not known to be one point per voxel in voxel order (lm_refilter.cuh)

constexpr int CI_N = 48;                          // cells per cube axis
constexpr int CI_TABLE = CI_N * CI_N * CI_N;      // cells per cube
constexpr int CI_WIN = 125;                       // cubes of the window = slots per type
constexpr int CI_SPAN = 5 * CI_N;                 // cells per axis of the window
constexpr float CI_H = 50.0f / 48.0f, CI_INV_H = 48.0f / 50.0f;
constexpr int CI_ADDR_BITS = 25;                  // index arena <= 2^25 points per type: (window position, address) share one word
constexpr unsigned int CI_ADDR_MASK = (1u << CI_ADDR_BITS) - 1u;

struct CISlot {
  int wi, wj, wk;      // world coordinates of the indexed cube (array index - laserCloudCen*)
  int ver, cnt;        // of its descriptor when the index was built
  int base, room;      // region of the index arena
  int fill;            // rows allocated so far (build scratch)
};

struct CIState {
  CISlot slot[2][CI_WIN];
  int win_slot[2][CI_WIN];    // window position q = (di * 5 + dj) * 5 + dk -> slot to search, -1 = no points there
  int win_cube[2][CI_WIN];    // -> cube array index
  int win_rank[CI_WIN];       // -> rank in laserCloudValidInd, -1 = not part of the local map
  float org[3];               // map coordinates of the window's minimum corner
  int nbuild;
  int build[2 * CI_WIN];      // T * 125 + q of the slots to rebuild this pass
  int cursor[2];              // bump pointers of the index arenas
  int cap;                    // points per index arena
  int stat[5];                // cumulative: slots built, points indexed, arena resets; last pass: slots built, points indexed
};

__device__ __forceinline__ int ci_pmod5(int v) { int m = v % 5; return m < 0 ? m + 5 : m; }

__device__ __forceinline__ int ci_local_cell(float v, float o) {
  int c = (int)floorf((v - o) * CI_INV_H);
  return c < 0 ? 0 : (c >= CI_N ? CI_N - 1 : c);   // a point stored a rounding error outside its cube stays in the edge cell
}

struct CIBuildArgs {
  CIState* ci;
  int* error;                 // LMState::error
  const CubeDesc* cubes[2];
  const float4* arena[2];
  int2* cell[2];              // [125][CI_TABLE] {end, count}: the cells of one x-row are contiguous in pts
  float4* pts[2];             // index arena: (x, y, z, position inside the cube as int bits)
};

// One 16-CTA cluster per stale slot (the list is walked with stride gridDim.y): zero the cell table, count,
// allocate row by row, scatter.  The phases are separated by hardware cluster barriers; clusters of different
// slots run side by side.
constexpr int CB_CLUSTER = 16, CB_THREADS = 512, CB_GRID_Y = 18;
constexpr int CB_ILP = 8;
constexpr int CB_ROWS = CI_N * CI_N / (CB_CLUSTER * CB_THREADS / 32);   // x-rows per warp
static_assert(CB_ROWS * (CB_CLUSTER * CB_THREADS / 32) == CI_N * CI_N, "rows divide evenly over the cluster's warps");
__global__ void __cluster_dims__(CB_CLUSTER, 1, 1) __launch_bounds__(CB_THREADS) ci_build(CIBuildArgs A) {
  namespace cg = cooperative_groups;
  cg::cluster_group cl = cg::this_cluster();
  CIState* ci = A.ci;
  const int nb = ci->nbuild;
  const int tid = blockIdx.x * CB_THREADS + threadIdx.x, nthr = CB_CLUSTER * CB_THREADS;
  const int lane = lane_id();
  for (int b = blockIdx.y; b < nb; b += gridDim.y) {
    const int e = ci->build[b];
    const int T = e / CI_WIN, q = e % CI_WIN;
    const int slot = ci->win_slot[T][q];
    if (slot < 0) continue;    // cluster-uniform
    const CISlot S = ci->slot[T][slot];
    const float4* __restrict__ src = A.arena[T] + A.cubes[T][ci->win_cube[T][q]].off;
    int2* cell = A.cell[T] + (size_t)slot * CI_TABLE;
    float4* __restrict__ out = A.pts[T];
    const float ox = 50.0f * (float)S.wi - 25.0f, oy = 50.0f * (float)S.wj - 25.0f, oz = 50.0f * (float)S.wk - 25.0f;
    const int n = S.cnt;
    // phase builds only: stamps of the cluster that builds the largest slot (first list entry), thread 0
    const bool stamp = PHASE_ON && b == 0 && tid == 0;
    if (stamp) { g_dbg_clk[960] = clock64(); g_dbg_clk[967] = n; g_dbg_clk[968] = nb; }
    {
      int4* c4 = (int4*)cell;
      for (int c = tid; c < CI_TABLE / 2; c += nthr) c4[c] = make_int4(0, 0, 0, 0);
    }
    __threadfence();
    cl.sync();
    if (stamp) g_dbg_clk[961] = clock64();
    for (int i0 = tid; i0 < n; i0 += CB_ILP * nthr) {   // independent loads in flight per thread
      float4 p[CB_ILP];
#pragma unroll
      for (int u = 0; u < CB_ILP; u++) p[u] = src[min(i0 + u * nthr, n - 1)];
#pragma unroll
      for (int u = 0; u < CB_ILP; u++) {
        if (i0 + u * nthr >= n) break;
        const int c = (ci_local_cell(p[u].z, oz) * CI_N + ci_local_cell(p[u].y, oy)) * CI_N + ci_local_cell(p[u].x, ox);
        atomicAdd(&cell[c].y, 1);
      }
    }
    __threadfence();
    cl.sync();
    if (stamp) g_dbg_clk[962] = clock64();
    {
      // one warp per x-row, CB_ROWS rows per warp: all counts are fetched first, the warp takes ONE range of the
      // slot's region for its rows (one atomic round trip instead of one per row)
      const int wid = tid >> 5;
      int c0[CB_ROWS], c1[CB_ROWS];
#pragma unroll
      for (int k = 0; k < CB_ROWS; k++) {
        const int2* cells = cell + (wid + k * (nthr >> 5)) * CI_N;
        c0[k] = __ldcg(&cells[lane].y);
        c1[k] = lane + 32 < CI_N ? __ldcg(&cells[lane + 32].y) : 0;
      }
      int mine = 0;
#pragma unroll
      for (int k = 0; k < CB_ROWS; k++) mine += c0[k] + c1[k];
      int incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      const int wtot = __shfl_sync(0xffffffffu, incl, 31);
      if (wtot > 0) {
        int base = 0;
        if (lane == 0) base = S.base + atomicAdd(&ci->slot[T][slot].fill, wtot);
        int run = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
        for (int k = 0; k < CB_ROWS; k++) {
          int i0 = c0[k], i1 = c1[k];
          for (int o = 1; o < 32; o <<= 1) {
            const int y0 = __shfl_up_sync(0xffffffffu, i0, o), y1 = __shfl_up_sync(0xffffffffu, i1, o);
            if (lane >= o) { i0 += y0; i1 += y1; }
          }
          const int t0 = __shfl_sync(0xffffffffu, i0, 31), t1 = __shfl_sync(0xffffffffu, i1, 31);
          if (t0 + t1 > 0) {   // an empty row keeps starts and counts 0: every range of it is empty
            int2* cells = cell + (wid + k * (nthr >> 5)) * CI_N;
            cells[lane].x = run + i0 - c0[k];
            if (lane + 32 < CI_N) cells[lane + 32].x = run + t0 + i1 - c1[k];
            run += t0 + t1;
          }
        }
      }
    }
    __threadfence();
    cl.sync();
    if (stamp) g_dbg_clk[963] = clock64();
    for (int i0 = tid; i0 < n; i0 += CB_ILP * nthr) {   // independent load -> atomic -> store chains per thread
      float4 p[CB_ILP];
      int dst[CB_ILP];
#pragma unroll
      for (int u = 0; u < CB_ILP; u++) p[u] = src[min(i0 + u * nthr, n - 1)];
#pragma unroll
      for (int u = 0; u < CB_ILP; u++) {
        dst[u] = -1;
        if (i0 + u * nthr < n) {
          const int c = (ci_local_cell(p[u].z, oz) * CI_N + ci_local_cell(p[u].y, oy)) * CI_N + ci_local_cell(p[u].x, ox);
          dst[u] = atomicAdd(&cell[c].x, 1);   // leaves {end, count} behind
        }
      }
#pragma unroll
      for (int u = 0; u < CB_ILP; u++) {
        if (dst[u] >= 0) { p[u].w = __int_as_float(i0 + u * nthr); out[dst[u]] = p[u]; }
      }
    }
    if (stamp) g_dbg_clk[964] = clock64();
  }
}

// ---- queries ----------------------------------------------------------------------------------
struct CIView {
  const int2* cell;
  const float4* pts;
};

struct CIRows9 {
  unsigned int beg[9];   // (window position << CI_ADDR_BITS) | first candidate in the index arena
  int pre[10];
};
__device__ __forceinline__ unsigned int ci_rows9_pos(const CIRows9& R, int t) {
  unsigned int pos = R.beg[0] + (unsigned int)t;
#pragma unroll
  for (int k = 1; k < 9; k++)
    if (t >= R.pre[k]) pos = R.beg[k] + (unsigned int)(t - R.pre[k]);
  return pos;
}

// The query's 3x3x3 block of lattice cells as 9 x-rows, the part of them inside cube column `di` (a row of
// three cells crosses at most one cube boundary in x: the caller makes a second pass with di + 1 when it does).
// bounded: only the cells a ball of squared radius r2 around the query reaches (grid_rows9_bounded).
__device__ __forceinline__ void ci_row_range(const CIView& g, const int* s_slot, const float* s_org, int cx, int cy, int cz, int di,
                                             bool bounded, float qx, float qy, float qz, float r2, int rowk, unsigned int& vb, unsigned int& ve) {
  vb = 0; ve = 0;
  const int oy = (rowk % 3) - 1, oz = (rowk / 3) - 1;
  const int y = cy + oy, z = cz + oz;
  if (y < 0 || y >= CI_SPAN || z < 0 || z >= CI_SPAN) return;
  int x0 = cx - 1, x1 = cx + 1;
  bool take = true;
  if (bounded) {
    const float margin = 4.0e-3f;
    const float fx = qx - (s_org[0] + (float)cx * CI_H), fy = qy - (s_org[1] + (float)cy * CI_H), fz = qz - (s_org[2] + (float)cz * CI_H);
    const float dyv = fmaxf((oy < 0 ? fy : (oy > 0 ? CI_H - fy : 0.f)) - margin, 0.f);
    const float dzv = fmaxf((oz < 0 ? fz : (oz > 0 ? CI_H - fz : 0.f)) - margin, 0.f);
    const float base = dyv * dyv + dzv * dzv;
    take = base <= r2;
    const float dxm = fmaxf(fx - margin, 0.f), dxp = fmaxf(CI_H - fx - margin, 0.f);
    if (!(base + dxm * dxm <= r2)) x0 = cx;
    if (!(base + dxp * dxp <= r2)) x1 = cx;
  }
  x0 = max(x0, di * CI_N);
  x1 = min(x1, di * CI_N + CI_N - 1);
  if (!take || x0 > x1) return;
  const int dj = y / CI_N, dk = z / CI_N;
  const int q = (di * 5 + dj) * 5 + dk;
  const int slot = s_slot[q];
  if (slot < 0) return;
  const int2* row = g.cell + (size_t)slot * CI_TABLE + ((z - dk * CI_N) * CI_N + (y - dj * CI_N)) * CI_N - di * CI_N;
  const int2 c0 = __ldg(&row[x0]);
  const int2 c1 = __ldg(&row[x1]);
  const unsigned int tag = (unsigned int)q << CI_ADDR_BITS;
  vb = tag | (unsigned int)(c0.x - c0.y);
  ve = tag | (unsigned int)c1.x;
}
// W = lanes cooperating on one query (32, 16 or 8): one lane per row; with W = 8 lane 0 also takes the ninth row
template <int W>
__device__ __forceinline__ void ci_rows9(const CIView& g, const int* s_slot, const float* s_org, int cx, int cy, int cz, int di,
                                         bool bounded, float qx, float qy, float qz, float r2, CIRows9& R) {
  static_assert(W == 8 || W == 16 || W == 32, "group size");
  const int gl = lane_id() & (W - 1);
  const unsigned int gmask = (W == 32) ? 0xffffffffu : (((1u << W) - 1u) << (lane_id() & ~(W - 1)));
  unsigned int vb = 0, ve = 0, vb8 = 0, ve8 = 0;
  if (gl < 9) ci_row_range(g, s_slot, s_org, cx, cy, cz, di, bounded, qx, qy, qz, r2, gl, vb, ve);
  if (W == 8 && gl == 0) ci_row_range(g, s_slot, s_org, cx, cy, cz, di, bounded, qx, qy, qz, r2, 8, vb8, ve8);
  R.pre[0] = 0;
#pragma unroll
  for (int k = 0; k < 9; k++) {
    unsigned int b, en;
    if (W == 8 && k == 8) { b = __shfl_sync(gmask, vb8, 0, W); en = __shfl_sync(gmask, ve8, 0, W); }
    else { b = __shfl_sync(gmask, vb, k, W); en = __shfl_sync(gmask, ve, k, W); }
    R.beg[k] = b;
    R.pre[k + 1] = R.pre[k] + (int)(en - b);
  }
}

// exact 5-NN inside radius 1 over the per-cube indexes: lm_knn5_group (lm_kernels.cuh) with the candidate
// source exchanged.  out[r].w = (window position << 24) | position inside the cube.
template <int W>
__device__ __forceinline__ bool ci_knn5_group(const CIView& g, const int* s_slot, const float* s_org, float qx, float qy, float qz,
                                              float4* out, bool have_prev = false) {
  const int lane = lane_id();
  const int gl = lane & (W - 1);
  const unsigned int gmask = (W == 32) ? 0xffffffffu : (((1u << W) - 1u) << (lane & ~(W - 1)));
  float ld[5]; int li[5]; float lx[5], ly[5], lz[5];
#pragma unroll
  for (int k = 0; k < 5; k++) { ld[k] = 3.0e38f; li[k] = 0x7fffffff; lx[k] = ly[k] = lz[k] = 0.f; }
  const int cx = (int)floorf((qx - s_org[0]) * CI_INV_H), cy = (int)floorf((qy - s_org[1]) * CI_INV_H), cz = (int)floorf((qz - s_org[2]) * CI_INV_H);
  float r2 = 0.f;
  if (have_prev) {   // uniform inside the group
#pragma unroll
    for (int k = 0; k < 5; k++) {
      const float4 p = out[k];
      float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
      float dd = 0.f;
      dd += dx * dx; dd += dy * dy; dd += dz * dz;
      r2 = fmaxf(r2, dd);
    }
    __syncwarp(gmask);   // every lane has read the old neighbours before any lane stores new ones
  }
  int total = 0;
  const int xa = max(cx - 1, 0), xb = min(cx + 1, CI_SPAN - 1);
  if (xa <= xb) {
    const int d0 = xa / CI_N, d1 = xb / CI_N;
    for (int di = d0; di <= d1; di++) {   // uniform inside the group
      CIRows9 R;
      ci_rows9<W>(g, s_slot, s_org, cx, cy, cz, di, have_prev, qx, qy, qz, r2, R);
      const int nc = R.pre[9];
      total += nc;
      // four independent candidate loads in flight per lane before the first one is consumed
      for (int t0 = gl; t0 < nc; t0 += 4 * W) {
        float4 pc[4];
        unsigned int ps[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int t = t0 + u * W;
          ps[u] = ci_rows9_pos(R, t < nc ? t : 0);
          pc[u] = g.pts[ps[u] & CI_ADDR_MASK];
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          if (t0 + u * W >= nc) break;
          const float4 p = pc[u];
          float dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;
          float dd = 0.f;
          dd += dx * dx; dd += dy * dy; dd += dz * dz;
          const int id = (int)((ps[u] >> CI_ADDR_BITS) << 24) | __float_as_int(p.w);
          if (nn_less(dd, id, ld[4], li[4])) {
            ld[4] = dd; li[4] = id; lx[4] = p.x; ly[4] = p.y; lz[4] = p.z;
#pragma unroll
            for (int k = 4; k > 0; k--) {
              if (nn_less(ld[k], li[k], ld[k - 1], li[k - 1])) {
                float td = ld[k]; ld[k] = ld[k - 1]; ld[k - 1] = td;
                int ti = li[k]; li[k] = li[k - 1]; li[k - 1] = ti;
                float tx = lx[k]; lx[k] = lx[k - 1]; lx[k - 1] = tx;
                float ty = ly[k]; ly[k] = ly[k - 1]; ly[k - 1] = ty;
                float tz = lz[k]; lz[k] = lz[k - 1]; lz[k - 1] = tz;
              }
            }
          }
        }
      }
    }
  }
  if (total < 5) return false;   // uniform inside the group
  // merge the W sorted lists: five rounds, the winner pops its head
  unsigned int bdu = 0; int bi = 0;
#pragma unroll
  for (int r = 0; r < 5; r++) {
    bdu = __reduce_min_sync(gmask, __float_as_uint(ld[0]));
    const bool tie = __float_as_uint(ld[0]) == bdu;
    bi = (int)__reduce_min_sync(gmask, tie ? (unsigned int)li[0] : 0x7fffffffu);
    if (tie && li[0] == bi && bi != 0x7fffffff) {   // ids are unique: exactly one owner
      out[r] = make_float4(lx[0], ly[0], lz[0], __int_as_float(bi));
#pragma unroll
      for (int k = 0; k < 4; k++) { ld[k] = ld[k + 1]; li[k] = li[k + 1]; lx[k] = lx[k + 1]; ly[k] = ly[k + 1]; lz[k] = lz[k + 1]; }
      ld[4] = 3.0e38f; li[4] = 0x7fffffff;
    }
  }
  return bi != 0x7fffffff && __uint_as_float(bdu) < 1.0;
}

}  // namespace loamgpu
