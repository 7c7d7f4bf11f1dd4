// lm_refilter.cuh -- laserMapping map update (LM-12, LM-13): insert the registered sweep into
// the cube map and re-run pcl::VoxelGrid on every valid cube (laserMapping.cpp:980-1036).
//
// The reference appends the new points to each cube and then filters the WHOLE cube again, every
// mapped sweep.  A cube that was filtered before holds exactly one point per occupied voxel, in
// ascending voxel order (PCL's idx = i + j*dx + k*dx*dy is the lexicographic order of (k, j, i)
// whatever the bounding box is, and voxel boundaries are absolute multiples of the leaf).  So the
// filter of (old ++ new) is a MERGE of an already sorted, duplicate-free list with the few new
// points of this sweep: only the new points are sorted (hundreds per cube, in shared memory), an
// old point whose voxel received nothing is copied bit for bit (0 + p = p, p / 1.0f = p), and a
// valid cube that received nothing is left alone (its filter is the identity).  One CTA per
// touched cube; no device-wide sort.
//
// Exactness guard: a centroid can round onto the far side of a voxel boundary, in which case the
// "one point per voxel, ascending" premise breaks for that cube.  Every CTA re-derives the voxel
// of each centroid it writes and marks the cube dirty when it differs; dirty cubes (and cubes
// whose keys turn out not to be strictly ascending, cubes that were only appended to while outside
// the valid set, and cubes with more new points than fit in shared memory) take the general path:
// all points of the cube sorted by (voxel, input order) with the single-CTA radix sort.
#pragma once
#include "lm_kernels.cuh"
#include "cta_radix.cuh"

namespace loamgpu {

constexpr int RF_THREADS = 512;
constexpr int RF_NEW_CAP = 6144;     // new points per cube per sweep held in shared memory (merge path)
constexpr size_t RF_SMEM_MERGE = (size_t)RF_NEW_CAP * (16 + 4 + 4 + 4 + 4 + 1);
constexpr size_t RF_SMEM = RF_SMEM_MERGE > sizeof(CtaRadixSmem) ? RF_SMEM_MERGE : sizeof(CtaRadixSmem);

// voxel coordinate of a map point relative to its cube, 8 bits per axis, PCL order (z, y, x)
__device__ __forceinline__ unsigned int lm_voxel_key(const float4 p, int cube, int cenW, int cenH, int cenD, int T, int* error) {
  const float inv = T == 0 ? (1.0f / 0.2f) : (1.0f / 0.4f);   // 5.0f / 2.5f exactly, as PCL computes 1/leaf
  const int cells = T == 0 ? 250 : 125;
  int ci = cube % CUBE_W, cj = (cube / CUBE_W) % CUBE_H, ck = cube / (CUBE_W * CUBE_H);
  int bi = (ci - cenW) * cells - (cells + 1) / 2 - 1;
  int bj = (cj - cenH) * cells - (cells + 1) / 2 - 1;
  int bk = (ck - cenD) * cells - (cells + 1) / 2 - 1;
  int gx = (int)floorf(p.x * inv) - bi, gy = (int)floorf(p.y * inv) - bj, gz = (int)floorf(p.z * inv) - bk;
  if ((gx | gy | gz) & ~255) {
    atomicOr(error, 4);
    gx = min(max(gx, 0), 255); gy = min(max(gy, 0), 255); gz = min(max(gz, 0), 255);
  }
  return ((unsigned int)gz << 16) | ((unsigned int)gy << 8) | (unsigned int)gx;
}

struct LMInsertArgs {
  LMState* st;
  const float4* stack[2];
  float4* ins_pts[2];
  int* ins_cube[2];
  unsigned int* ins_key[2];
  int* touch_cnt[2];
  int rec_slot;              // which LMState::rec the pass's record snapshot goes to
};
// K7a: stack points to the map with the refined pose, their cube and voxel (:980-1016)
__global__ void __launch_bounds__(256) lm_insert_ids(LMInsertArgs A) {
  LMState* st = A.st;
  __shared__ LMTrig s_g;
  __shared__ float s_T[6];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 6; i++) s_T[i] = st->transformTobeMapped[i];
    s_g = lm_trig(s_T);
  }
  __syncthreads();
  const LMTrig g = s_g;
  const int n0 = st->stackN[0], n1 = st->stackN[1];
  if (blockIdx.x == 0 && threadIdx.x == 0) {   // what the side-stream kernels of this pass read (LMState::insN / insT)
    st->insN[0] = n0; st->insN[1] = n1;
    for (int i = 0; i < 6; i++) st->insT[i] = s_T[i];
    LMState::Rec& R = st->rec[A.rec_slot & 1];
    for (int i = 0; i < 6; i++) { R.aft[i] = st->transformAftMapped[i]; R.bef[i] = st->transformBefMapped[i]; }
    R.iters = st->iters; R.rows = st->rows; R.deg = st->isDegenerate;
    R.mapN[0] = st->mapN[0]; R.mapN[1] = st->mapN[1]; R.stackN[0] = n0; R.stackN[1] = n1;
  }
  const int cenW = st->cenW, cenH = st->cenH, cenD = st->cenD;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n0 + n1; i += gridDim.x * 256) {
    int T = i < n0 ? 0 : 1;
    int k = T ? i - n0 : i;
    float4 pointSel = lm_to_map(A.stack[T][k], g, s_T);
    int cubeI = lm_cube_coord(pointSel.x, cenW), cubeJ = lm_cube_coord(pointSel.y, cenH), cubeK = lm_cube_coord(pointSel.z, cenD);
    int cube = -1;
    unsigned int key = 0;
    if (cubeI >= 0 && cubeI < CUBE_W && cubeJ >= 0 && cubeJ < CUBE_H && cubeK >= 0 && cubeK < CUBE_D) {
      cube = cubeI + CUBE_W * cubeJ + CUBE_W * CUBE_H * cubeK;
      atomicAdd(&A.touch_cnt[T][cube], 1);
      key = lm_voxel_key(pointSel, cube, cenW, cenH, cenD, T, &st->error);
    }
    A.ins_pts[T][k] = pointSel;
    A.ins_cube[T][k] = cube;
    A.ins_key[T][k] = key;
  }
}

struct LMPlanArgs {
  LMState* st;
  CubeDesc* cubes[2];
  int* touch_cnt[2];
  int* cube_rank[2];
};
// K7b: one CTA per feature type.  The cubes rewritten this sweep: every cube that received points
// plus every VALID cube that is dirty; procNewOff[] = where each cube's new points go in the bucket.
__global__ void __launch_bounds__(1024) lm_refilter_plan(LMPlanArgs A) {
  LMState* st = A.st;
  const int T = blockIdx.x;
  __shared__ int s_scan[32];
  __shared__ int s_carry;
  __shared__ unsigned char s_valid[CUBE_NUM];
  for (int c = threadIdx.x; c < CUBE_NUM; c += 1024) s_valid[c] = 0;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int nv = st->validNum;
  for (int r = threadIdx.x; r < nv; r += 1024) s_valid[st->validInd[r]] = 1;
  __syncthreads();
  for (int base = 0; base < CUBE_NUM; base += 1024) {
    int c = base + threadIdx.x;
    int flag = 0;
    if (c < CUBE_NUM) {
      int t = A.touch_cnt[T][c];
      flag = (t > 0 || (s_valid[c] && A.cubes[T][c].dirty && A.cubes[T][c].cnt > 0)) ? 1 : 0;
    }
    int x = flag;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane_id() >= o) x += y;
    }
    if (lane_id() == 31) s_scan[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = s_scan[threadIdx.x], s = w;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane_id() >= o) s += y;
      }
      s_scan[threadIdx.x] = s - w;
    }
    __syncthreads();
    int r = s_carry + s_scan[threadIdx.x >> 5] + x - flag;
    if (flag) {
      if (r < PROC_MAX) {
        A.cube_rank[T][c] = r;
        st->procCube[T][r] = c;
        st->procMode[T][r] = s_valid[c] ? 1 : 0;   // 1 = refilter (valid cube), 0 = append only
        st->procNewCnt[T][r] = A.touch_cnt[T][c];
      } else {
        atomicOr(&st->error, 2);
      }
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = r + flag;
    __syncthreads();
  }
  const int np = min(s_carry, PROC_MAX);
  if (threadIdx.x == 0) {
    st->nproc[T] = np;
    int off = 0;
    for (int r = 0; r < np; r++) { st->procNewOff[T][r] = off; off += st->procNewCnt[T][r]; }
    st->procNewOff[T][np] = off;
  }
  if (threadIdx.x == 0) st->scratchBump[T] = 0;
  if (threadIdx.x < 8) st->rfStat[T][threadIdx.x] = 0;
}

struct LMSortNewArgs {
  LMState* st;
  const int* ins_cube[2];
  const unsigned int* ins_key[2];
  const int* cube_rank[2];
  unsigned int* nk[2][2];   // ping-pong keys   (rank << 24 | voxel)
  unsigned int* nv[2][2];   // ping-pong values (stack index)
};
// K7c': the same list built per cube instead of by one global sort.  CTA (r, T) pulls the points
// of plan entry r out of the stack (stable compaction: stack order), sorts them by voxel in shared
// memory when the cube is going to be merged (three 8-bit passes over a few thousand pairs) and
// writes them at procNewOff[r].  Cubes are independent, so the sorts run side by side on different
// SMs; the one global 4-pass sort of all new points was 50 us of the mapping chain.  Append-only
// cubes and cubes with more new points than the merge path holds stay in stack order (the general
// path sorts everything again, stably).
constexpr int BN_THREADS = 512;
constexpr int BN_GRID = 16;
constexpr int BN_PER = 8, BN_TILE = BN_THREADS * BN_PER;
constexpr size_t BN_SMEM = (size_t)RF_NEW_CAP * 16 + sizeof(CtaRadixSmem);
__global__ void __launch_bounds__(BN_THREADS, 1) lm_bucket_new(LMSortNewArgs A) {
  extern __shared__ __align__(16) unsigned char bn_smem[];
  unsigned int* k0 = (unsigned int*)bn_smem;
  unsigned int* v0 = k0 + RF_NEW_CAP;
  unsigned int* k1 = v0 + RF_NEW_CAP;
  unsigned int* v1 = k1 + RF_NEW_CAP;
  CtaRadixSmem& S = *(CtaRadixSmem*)(v1 + RF_NEW_CAP);
  __shared__ int s_ws[BN_THREADS / 32];
  __shared__ int s_tile[BN_TILE];
  LMState* st = A.st;
  const int T = blockIdx.y;
  const int n = st->insN[T];
  const int np = st->nproc[T];
  if (blockIdx.x == 0 && threadIdx.x == 0) st->newSel[T] = 0;
  // The stack is walked in tiles of BN_TILE entries: a tile's cube ids are staged with coalesced loads, then every
  // thread owns BN_PER consecutive entries of the tile (so the compaction is stable: stack order) -- one L2 round
  // trip per tile instead of one per entry and thread.
  for (int r = blockIdx.x; r < np; r += gridDim.x) {
    const int cube = st->procCube[T][r];
    const int n_new = st->procNewCnt[T][r];
    const int mode = st->procMode[T][r];
    const bool sorted = mode != 0 && n_new <= RF_NEW_CAP;
    unsigned int* gk = A.nk[T][0] + st->procNewOff[T][r];
    unsigned int* gv = A.nv[T][0] + st->procNewOff[T][r];
    unsigned int* dk = sorted ? k0 : gk;
    unsigned int* dv = sorted ? v0 : gv;
    int done = 0;
    for (int t0 = 0; t0 < n; t0 += BN_TILE) {
#pragma unroll
      for (int u = 0; u < BN_PER; u++) {
        const int idx = t0 + u * BN_THREADS + (int)threadIdx.x;
        s_tile[u * BN_THREADS + threadIdx.x] = idx < n ? A.ins_cube[T][idx] : -1;
      }
      __syncthreads();
      unsigned int m = 0;
#pragma unroll
      for (int j = 0; j < BN_PER; j++) m |= (s_tile[threadIdx.x * BN_PER + j] == cube) ? (1u << j) : 0u;
      const int cnt = __popc(m);
      // exclusive prefix over the CTA
      int x = cnt;
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane_id() >= o) x += y;
      }
      if (lane_id() == 31) s_ws[threadIdx.x >> 5] = x;
      __syncthreads();
      int pos = done + x - cnt, tot = 0;
#pragma unroll
      for (int w = 0; w < BN_THREADS / 32; w++) { const int c = s_ws[w]; if (w < (int)(threadIdx.x >> 5)) pos += c; tot += c; }
      for (int j = 0; j < BN_PER; j++) {
        if ((m >> j) & 1u) {
          const int k = t0 + (int)threadIdx.x * BN_PER + j;
          dk[pos] = ((unsigned int)r << 24) | (mode ? A.ins_key[T][k] : 0u);
          dv[pos] = (unsigned int)k;
          pos++;
        }
      }
      done += tot;
      __syncthreads();
    }
    __syncthreads();
    if (sorted) {
      const int res = cta_radix_sort(k0, v0, k1, v1, n_new, 3, S);
      const unsigned int* sk = res ? k1 : k0;
      const unsigned int* sv = res ? v1 : v0;
      for (int j = threadIdx.x; j < n_new; j += BN_THREADS) { gk[j] = sk[j]; gv[j] = sv[j]; }
    }
    __syncthreads();
  }
}

// running fp32 centroid, PCL 1.7 VoxelGrid::applyFilter: centroid = 0; centroid += p ...; /= count
struct RFCentroid {
  float c0, c1, c2, c3;
  int n;
  __device__ __forceinline__ void init() { c0 = c1 = c2 = c3 = 0.f; n = 0; }
  __device__ __forceinline__ void add(const float4 p) { c0 += p.x; c1 += p.y; c2 += p.z; c3 += p.w; n++; }
  __device__ __forceinline__ float4 get() const {
    float f = (float)n;
    return make_float4(c0 / f, c1 / f, c2 / f, c3 / f);
  }
};

struct LMCubeArgs {
  LMState* st;
  CubeDesc* cubes[2];
  float4* arena[2];
  const float4* ins_pts[2];
  const unsigned int* nk[2][2];
  const unsigned int* nv[2][2];
  unsigned int* sk[2][2];     // general-path scratch (keys), ping-pong, scap elements each
  unsigned int* sv[2][2];
  int* touch_cnt[2];
  int* cube_rank[2];
  int scap;
};

// exclusive prefix of one int per thread over the CTA (RF_THREADS threads); *total = CTA sum
__device__ __forceinline__ int rf_block_scan(int v, int* s_ws, int* total) {
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  int x = v;
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();                 // s_ws may still be read from a previous call
  if (lane == 31) s_ws[warp] = x;
  __syncthreads();
  int off = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < RF_THREADS / 32; w++) {
    int c = s_ws[w];
    if (w < warp) off += c;
    tot += c;
  }
  *total = tot;
  return off + x - v;
}

// first position in a[0..n) with a[pos] >= key
__device__ __forceinline__ int rf_lower_bound(const unsigned int* a, int n, unsigned int key) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (a[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// the same for U keys at once: fixed trip count, the U probes of a step are independent loads, so
// their shared-memory latencies overlap (the merge path is a chain of such searches)
template <int U>
__device__ __forceinline__ void rf_lower_bound_multi(const unsigned int* a, int n, const unsigned int* key, int* pos) {
  int top = 1;
  while (top <= n) top <<= 1;
#pragma unroll
  for (int u = 0; u < U; u++) pos[u] = 0;
  for (int step = top >> 1; step > 0; step >>= 1) {
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int p = pos[u] + step;
      if (p <= n && a[p - 1] < key[u]) pos[u] = p;
    }
  }
}

// Lower bounds for the 4 x 32 keys a warp holds in one merge step.  Consecutive lanes hold
// consecutive old points, so key[u] ascends with the lane: eight lanes run the full binary
// searches that bracket the four ranges (first and last lane of every u; `lowkey[u]` <= the key of
// lane 0 replaces the lower end when the caller also wants that position), then every lane searches
// its own keys inside a bracket that is a handful of elements wide.  One 13-step probe sequence
// with 8 active lanes per warp instead of four with 32: the merge passes are bound by the shared
// memory pipe of the one SM that owns the cube.
__device__ __forceinline__ void rf_lower_bound_warp4(const unsigned int* a, int n, const unsigned int* key, const unsigned int* lowkey,
                                                     int* pos, int* lowpos) {
  const unsigned int full = 0xffffffffu;
  const int lane = lane_id();
  unsigned int mine = 0;
#pragma unroll
  for (int u = 0; u < 4; u++) {
    const unsigned int k_lo = lowkey ? __shfl_sync(full, lowkey[u], 0) : __shfl_sync(full, key[u], 0);
    const unsigned int k_hi = __shfl_sync(full, key[u], 31);
    if (lane == u) mine = k_lo;
    if (lane == 4 + u) mine = k_hi;
  }
  int p = 0;
  if (lane < 8) {
    int lo = 0, hi = n;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (a[mid] < mine) lo = mid + 1; else hi = mid;
    }
    p = lo;
  }
#pragma unroll
  for (int u = 0; u < 4; u++) {
    int lo = __shfl_sync(full, p, u);
    int hi = __shfl_sync(full, p, 4 + u);
    if (lowpos) lowpos[u] = lo;
    const unsigned int k = key[u];
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (a[mid] < k) lo = mid + 1; else hi = mid;
    }
    pos[u] = lo;
  }
}

// K7d: one CLUSTER of RF_SLICES CTAs per rewritten cube (r = rank in the plan, T = feature type).
// The merge path is sliced over the cluster: CTA s owns the old points [n_old*s/S, n_old*(s+1)/S)
// and the new points whose voxel lies between the first old voxel of its slice and of the next
// one (sorted lists: both are contiguous ranges).  The slices exchange three ints through
// distributed shared memory (premise violated, new-only voxel count, dirty) and CTA 0 allocates
// the output.  The work of a cube grows with the map (tens of thousands of old points in the
// cubes around the sensor) and is bound by shared-memory probes and dependent L2 loads of ONE SM;
// the cluster spreads it over eight.  Append-only cubes and the general path run on CTA 0 alone.
#ifndef LOAM_RF_SLICES
#define LOAM_RF_SLICES 8
#endif
#ifndef LOAM_RF_CLUSTERS
#define LOAM_RF_CLUSTERS 8
#endif
constexpr int RF_SLICES = LOAM_RF_SLICES;
__device__ __forceinline__ void rf_cube(const LMCubeArgs& A, const int T, const int r) {
  cg::cluster_group cluster = cg::this_cluster();
  const int srank = (int)cluster.block_rank();
  extern __shared__ __align__(16) unsigned char rf_smem[];
  float4* s_npt = (float4*)rf_smem;                               // RF_NEW_CAP  the new points, sorted by (voxel, stack order)
  unsigned int* s_nkey = (unsigned int*)(s_npt + RF_NEW_CAP);     // RF_NEW_CAP  their voxel keys
  unsigned int* s_nokey = s_nkey + RF_NEW_CAP;                    // RF_NEW_CAP  voxels that hold only new points
  int* s_nopos = (int*)(s_nokey + RF_NEW_CAP);                    // RF_NEW_CAP  first new point of each such voxel
  int* s_oldbefore = s_nopos + RF_NEW_CAP;                        // RF_NEW_CAP  old points in front of each such voxel
  unsigned char* s_found = (unsigned char*)(s_oldbefore + RF_NEW_CAP);   // RF_NEW_CAP
  CtaRadixSmem& S = *(CtaRadixSmem*)rf_smem;                      // general path (aliases the arrays above)
  __shared__ int s_ws[RF_THREADS / 32];
  __shared__ int s_base, s_dirty, s_sbase;
  __shared__ int s_x[4];          // read by the other CTAs of the cluster: bad, n_no, base, dirty
  __shared__ unsigned int s_kb[2];
  __shared__ int s_jb[2];
  LMState* st = A.st;
  const int cube = st->procCube[T][r];
  const int mode = st->procMode[T][r];
  const int n_new = st->procNewCnt[T][r];
  const int sel = st->newSel[T];
  const unsigned int* nkeys = A.nk[T][sel] + st->procNewOff[T][r];
  const unsigned int* nvals = A.nv[T][sel] + st->procNewOff[T][r];
  const CubeDesc desc = A.cubes[T][cube];
  const int n_old = desc.cnt;
  float4* arena = A.arena[T];
  const float4* old = arena + desc.off;
  const float4* ins_pts = A.ins_pts[T];
  const int cenW = st->cenW, cenH = st->cenH, cenD = st->cenD;
  if (threadIdx.x == 0) s_dirty = 0;
  __syncthreads();
  long long* clk = (threadIdx.x == 0 && srank == 0 && PHASE_ON) ? g_rf_clk + ((size_t)T * PROC_MAX + r) * 8 : nullptr;
#define RS(k) do { if (clk) clk[k] = clock64(); } while (0)
  if (clk) { for (int k = 0; k < 8; k++) clk[k] = 0; clk[7] = ((long long)n_old << 32) | (unsigned int)n_new; }
  RS(0);
  int n_out = 0;
  bool general = (mode == 1) && (desc.dirty != 0 || n_new > RF_NEW_CAP);   // cluster-uniform

  if (mode == 1 && !general) {
    // ---- merge path, slice `srank` ----
    const int i_a = (int)((long long)n_old * srank / RF_SLICES), i_b = (int)((long long)n_old * (srank + 1) / RF_SLICES);
    // all new keys (the slice bounds are found in them), flags cleared
    for (int j = threadIdx.x; j < n_new; j += RF_THREADS) { s_nkey[j] = nkeys[j] & 0x00ffffffu; s_found[j] = 0; }
    if (threadIdx.x < 2) {
      // voxel range of the slice: [first old voxel of this slice, first old voxel of the next one)
      const int i = threadIdx.x == 0 ? i_a : i_b;
      unsigned int k;
      if (threadIdx.x == 0 && srank == 0) k = 0u;
      else if (i >= n_old) k = 0xffffffffu;
      else k = lm_voxel_key(old[i], cube, cenW, cenH, cenD, T, &st->error);
      s_kb[threadIdx.x] = k;
    }
    __syncthreads();
    if (threadIdx.x < 2) s_jb[threadIdx.x] = rf_lower_bound(s_nkey, n_new, s_kb[threadIdx.x]);
    __syncthreads();
    const int j_a = s_jb[0], j_b = max(s_jb[0], s_jb[1]);
    // staging of the slice's new points: index loads, then point loads, four in flight per thread
    for (int j0 = j_a + threadIdx.x; j0 < j_b; j0 += 4 * RF_THREADS) {
      unsigned int vv[4];
      float4 pp[4];
#pragma unroll
      for (int u = 0; u < 4; u++) vv[u] = nvals[min(j0 + u * RF_THREADS, j_b - 1)];
#pragma unroll
      for (int u = 0; u < 4; u++) pp[u] = ins_pts[vv[u]];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int j = j0 + u * RF_THREADS;
        if (j < j_b) s_npt[j] = pp[u];
      }
    }
    __syncthreads();
    RS(5);
    // Pass A over the slice's old points: are they strictly ascending (premise, including the
    // step across the slice boundary), and which voxels of the new list already hold an old point
    int bad = 0;
    for (int ib = i_a; ib < i_b; ib += 4 * RF_THREADS) {   // warp-uniform trip count (shuffles inside)
      const int i0 = ib + threadIdx.x;
      unsigned int key[4];
      int jpos[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int i = i0 + u * RF_THREADS;
        key[u] = 0xffffffffu;
        if (i < i_b) {
          // both loads are independent and issued up front (a shuffle from lane - 1 would leave
          // lane 0 with a dependent global load per step)
          const float4 pc = old[i], pp = old[i > 0 ? i - 1 : 0];
          key[u] = lm_voxel_key(pc, cube, cenW, cenH, cenD, T, &st->error);
          if (i > 0 && key[u] <= lm_voxel_key(pp, cube, cenW, cenH, cenD, T, &st->error)) bad = 1;
        }
      }
      rf_lower_bound_warp4(s_nkey, n_new, key, nullptr, jpos, nullptr);
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (i0 + u * RF_THREADS < i_b && jpos[u] < n_new && s_nkey[jpos[u]] == key[u]) s_found[jpos[u]] = 1;   // first new point of that voxel
    }
    bad = __syncthreads_or(bad);
    RS(1);
    // voxels that hold only new points: segment heads of the slice's new range without an old point
    int n_no = 0;
    for (int base = j_a; base < j_b; base += RF_THREADS) {
      const int j = base + threadIdx.x;
      int flag = 0;
      if (j < j_b) flag = ((j == 0 || s_nkey[j - 1] != s_nkey[j]) && !s_found[j]) ? 1 : 0;
      int tot;
      const int rank = n_no + rf_block_scan(flag, s_ws, &tot);
      if (flag) { s_nokey[rank] = s_nkey[j]; s_nopos[rank] = j; }
      n_no += tot;
    }
    RS(6);
    for (int q = threadIdx.x; q < n_no; q += RF_THREADS) s_oldbefore[q] = i_b;
    if (threadIdx.x == 0) { s_x[0] = bad; s_x[1] = n_no; s_x[3] = 0; }
    cluster.sync();
    int pre_no = 0, tot_no = 0, any_bad = 0;
#pragma unroll
    for (int t = 0; t < RF_SLICES; t++) {
      const int* x = cluster.map_shared_rank(s_x, t);
      const int b = x[0], c = x[1];
      any_bad |= b;
      if (t < srank) pre_no += c;
      tot_no += c;
    }
    if (any_bad) {
      general = true;      // cluster-uniform; CTA 0 sorts the whole cube below
      cluster.sync();      // everyone has read s_x before a CTA that returns rewrites it for its next cube
    } else {
      n_out = n_old + tot_no;
      if (srank == 0 && threadIdx.x == 0) {
        int b = atomicAdd(&st->bump[T], n_out);
        if (b + n_out > st->arenaCap) { atomicOr(&st->error, 1); b = -1; }
        s_x[2] = b;
      }
      cluster.sync();
      RS(2);
      const int base = *cluster.map_shared_rank(&s_x[2], 0);
      int dirty = 0;
      if (base >= 0) {
        float4* out = arena + base + pre_no;
        // Pass B: every old point moves up by the number of new-only voxels in front of it and
        // absorbs the new points of its own voxel (old first, then stack order: the reference's
        // input order)
        for (int ib = i_a; ib < i_b; ib += 4 * RF_THREADS) {
          const int i0 = ib + threadIdx.x;
          float4 p[4];
          unsigned int key[8];   // [0..3] this point, [4..7] its predecessor (inside the slice)
          int c[8], jpos[4];
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int i = i0 + u * RF_THREADS;
            key[u] = 0xffffffffu; key[4 + u] = 0u;
            if (i < i_b) {
              p[u] = old[i];
              const float4 pp = old[i > 0 ? i - 1 : 0];
              key[u] = lm_voxel_key(p[u], cube, cenW, cenH, cenD, T, &st->error);
              key[4 + u] = i > i_a ? lm_voxel_key(pp, cube, cenW, cenH, cenD, T, &st->error) : 0u;
            }
          }
          rf_lower_bound_warp4(s_nokey, n_no, key, key + 4, c, c + 4);
          // the predecessor of lane > 0 is the previous lane's point; lane 0 got its own from the bracket
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int up = __shfl_up_sync(0xffffffffu, c[u], 1);
            if (lane_id() > 0) c[4 + u] = up;
          }
          rf_lower_bound_warp4(s_nkey, n_new, key, nullptr, jpos, nullptr);
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int i = i0 + u * RF_THREADS;
            if (i >= i_b) break;
            // new-only voxels between the predecessor and this point sit right in front of it
            const int cprev = i > i_a ? c[4 + u] : 0;
            for (int q = cprev; q < c[u]; q++) s_oldbefore[q] = i;
            int j = jpos[u];
            float4 po = p[u];
            if (j < n_new && s_nkey[j] == key[u]) {
              RFCentroid cen;
              cen.init();
              cen.add(po);
              for (; j < n_new && s_nkey[j] == key[u]; j++) cen.add(s_npt[j]);
              po = cen.get();
              if (lm_voxel_key(po, cube, cenW, cenH, cenD, T, &st->error) != key[u]) dirty = 1;
            }
            out[i + c[u]] = po;
          }
        }
        __syncthreads();
        RS(3);
        for (int q = threadIdx.x; q < n_no; q += RF_THREADS) {
          const unsigned int key = s_nokey[q];
          RFCentroid cen;
          cen.init();
          for (int j = s_nopos[q]; j < n_new && s_nkey[j] == key; j++) cen.add(s_npt[j]);
          const float4 p = cen.get();
          if (lm_voxel_key(p, cube, cenW, cenH, cenD, T, &st->error) != key) dirty = 1;
          out[s_oldbefore[q] + q] = p;
        }
      }
      dirty = __syncthreads_or(dirty);
      if (threadIdx.x == 0) s_x[3] = dirty;
      cluster.sync();
      if (srank == 0 && threadIdx.x == 0) {
        int d_any = 0;
        for (int t = 0; t < RF_SLICES; t++) d_any |= cluster.map_shared_rank(s_x, t)[3];
        s_dirty = d_any;
        if (base >= 0) { CubeDesc d; d.off = base; d.cnt = n_out; d.dirty = d_any; d.ver = atomicAdd(&st->verNext, 1) + 1; A.cubes[T][cube] = d; }
      }
      cluster.sync();      // s_x stays valid until CTA 0 has read it
      if (srank != 0) return;
      goto done;
    }
  }
  // append-only cubes and the general path: CTA 0 of the cluster alone
  if (srank != 0) return;
  if (mode == 0) {
    // ---- cube outside the valid set: the reference only appends (:1000-1016), in stack order.
    // Marked dirty so the first refilter after it becomes valid sorts everything. ----
    n_out = n_old + n_new;
    if (threadIdx.x == 0) {
      int b = atomicAdd(&st->bump[T], n_out);
      if (b + n_out > st->arenaCap) { atomicOr(&st->error, 1); b = -1; }
      s_base = b;
    }
    __syncthreads();
    if (s_base < 0) goto done;
    for (int i = threadIdx.x; i < n_old; i += RF_THREADS) arena[s_base + i] = old[i];
    for (int j = threadIdx.x; j < n_new; j += RF_THREADS) arena[s_base + n_old + j] = ins_pts[nvals[j]];
    if (threadIdx.x == 0) { CubeDesc d; d.off = s_base; d.cnt = n_out; d.dirty = 1; d.ver = atomicAdd(&st->verNext, 1) + 1; A.cubes[T][cube] = d; }
    goto done;
  }

  {
    // ---- general path: every point of the cube sorted by (voxel, input order) ----
    const int n_all = n_old + n_new;
    if (threadIdx.x == 0) {
      int b = atomicAdd(&st->scratchBump[T], n_all);
      if (b + n_all > A.scap) { atomicOr(&st->error, 16); b = -1; }
      s_sbase = b;
    }
    __syncthreads();
    if (s_sbase < 0) goto done;
    unsigned int* k0 = A.sk[T][0] + s_sbase; unsigned int* v0 = A.sv[T][0] + s_sbase;
    unsigned int* k1 = A.sk[T][1] + s_sbase; unsigned int* v1 = A.sv[T][1] + s_sbase;
    for (int e = threadIdx.x; e < n_all; e += RF_THREADS) {
      // payload = input order: old points first, then the new ones in stack order (the new list
      // is sorted by (voxel, stack index), so its position order is stack order inside a voxel)
      k0[e] = e < n_old ? lm_voxel_key(old[e], cube, cenW, cenH, cenD, T, &st->error) : (nkeys[e - n_old] & 0x00ffffffu);
      v0[e] = (unsigned int)e;
    }
    __threadfence_block();
    __syncthreads();
    const int res = cta_radix_sort(k0, v0, k1, v1, n_all, 3, S);
    const unsigned int* sk = res ? k1 : k0;
    const unsigned int* sv = res ? v1 : v0;
    int heads = 0;
    for (int e = threadIdx.x; e < n_all; e += RF_THREADS) heads += (e == 0 || sk[e - 1] != sk[e]) ? 1 : 0;
    int tot;
    rf_block_scan(heads, s_ws, &tot);
    n_out = tot;
    if (threadIdx.x == 0) {
      int b = atomicAdd(&st->bump[T], n_out);
      if (b + n_out > st->arenaCap) { atomicOr(&st->error, 1); b = -1; }
      s_base = b;
    }
    __syncthreads();
    if (s_base < 0) goto done;
    float4* out = arena + s_base;
    int dirty = 0, carry = 0;
    for (int base = 0; base < n_all; base += RF_THREADS) {
      const int e = base + threadIdx.x;
      int head = 0;
      unsigned int key = 0;
      if (e < n_all) {
        key = sk[e];
        head = (e == 0 || sk[e - 1] != key) ? 1 : 0;
      }
      int t2;
      const int slot = carry + rf_block_scan(head, s_ws, &t2);
      carry += t2;
      if (head) {
        RFCentroid cen;
        cen.init();
        for (int q = e; q < n_all && sk[q] == key; q++) {
          const unsigned int ord = sv[q];
          cen.add(ord < (unsigned int)n_old ? old[ord] : ins_pts[nvals[ord - n_old]]);
        }
        const float4 p = cen.get();
        if (lm_voxel_key(p, cube, cenW, cenH, cenD, T, &st->error) != key) dirty = 1;
        out[slot] = p;
      }
    }
    if (dirty) s_dirty = 1;
    __syncthreads();
    if (threadIdx.x == 0) { CubeDesc d; d.off = s_base; d.cnt = n_out; d.dirty = s_dirty; d.ver = atomicAdd(&st->verNext, 1) + 1; A.cubes[T][cube] = d; }
  }
done:
  RS(4);
#undef RS
  if (threadIdx.x == 0) {
    A.touch_cnt[T][cube] = 0; A.cube_rank[T][cube] = -1;
    atomicAdd(&st->rfStat[T][mode == 0 ? 2 : (general ? 1 : 0)], 1);
    atomicMax(&st->rfStat[T][3], n_old); atomicMax(&st->rfStat[T][4], n_new);
    atomicAdd(&st->rfStat[T][5], n_out);
    if (s_dirty) atomicAdd(&st->rfStat[T][7], 1);
  }
}
// one wave of clusters; a cluster walks the plan with stride (number of clusters)
constexpr int RF_GRID = LOAM_RF_CLUSTERS * RF_SLICES;
__global__ void __cluster_dims__(RF_SLICES, 1, 1) __launch_bounds__(RF_THREADS, 1) lm_refilter_cube(LMCubeArgs A) {
  const int T = blockIdx.y;
  const int np = A.st->nproc[T];
  for (int r = blockIdx.x / RF_SLICES; r < np; r += gridDim.x / RF_SLICES) {
    rf_cube(A, T, r);
    __syncthreads();
  }
}

}  // namespace loamgpu
