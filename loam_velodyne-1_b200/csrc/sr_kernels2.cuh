// sr_kernels2.cuh -- scanRegistration, second half: curvature / rejection / sector sort / greedy
// selection / less-flat voxel filter / compaction (scanRegistration.cpp:359-582).
#pragma once
#include "sr_kernels.cuh"
#include "cta_radix.cuh"

namespace loamgpu {

constexpr int CR_TILE = 256;
constexpr int CR_HALO = 8;  // >= 6 (rejection reaches i+6), 5 for the stencil

// K-E: curvature (:359-381), flag reset, rejection (:395-452) in gather form, and the
// "consecutive gap^2 > 0.05" flags used by the suppression loops (:495-520).
// picked[j] = C(j) | OR_{i=j..j+5} A(i) | OR_{i=j-6..j-1} B(i), i restricted to [5, n-7], where
//   A(i): gap(i,i+1)>0.1, depth1>depth2, ratio test  -> marks i-5..i
//   B(i): gap(i,i+1)>0.1, depth1<=depth2, ratio test -> marks i+1..i+6
//   C(i): parallel-beam test                          -> marks i
__device__ __forceinline__ void sr_curv_reject_body(const float4* __restrict__ full, const SRCounts* __restrict__ cnt,
                                                          float* __restrict__ curvature, int* __restrict__ sort_ind,
                                                          unsigned char* __restrict__ picked, signed char* __restrict__ label,
                                                          unsigned char* __restrict__ gapflag) {
  __shared__ float sx[CR_TILE + 2 * CR_HALO], sy[CR_TILE + 2 * CR_HALO], sz[CR_TILE + 2 * CR_HALO];
  __shared__ unsigned char sA[CR_TILE + 2 * CR_HALO], sB[CR_TILE + 2 * CR_HALO];
  const int n = cnt->n_full;
  for (int tile0 = blockIdx.x * CR_TILE; tile0 < n; tile0 += gridDim.x * CR_TILE) {
    __syncthreads();
    for (int k = threadIdx.x; k < CR_TILE + 2 * CR_HALO; k += CR_TILE) {
      int g = tile0 - CR_HALO + k;
      float4 p = (g >= 0 && g < n) ? full[g] : make_float4(0.f, 0.f, 0.f, 0.f);
      sx[k] = p.x; sy[k] = p.y; sz[k] = p.z;
    }
    __syncthreads();
    // occlusion conditions for every i in tile +- halo (needs i and i+1)
    for (int k = threadIdx.x; k < CR_TILE + 2 * CR_HALO - 1; k += CR_TILE) {
      int i = tile0 - CR_HALO + k;
      unsigned char a = 0, b = 0;
      if (i >= 5 && i < n - 6) {
        float diffX = sx[k + 1] - sx[k];
        float diffY = sy[k + 1] - sy[k];
        float diffZ = sz[k + 1] - sz[k];
        float diff = diffX * diffX + diffY * diffY + diffZ * diffZ;
        if (diff > 0.1) {
          float depth1 = dm::sqrt(sx[k] * sx[k] + sy[k] * sy[k] + sz[k] * sz[k]);
          float depth2 = dm::sqrt(sx[k + 1] * sx[k + 1] + sy[k + 1] * sy[k + 1] + sz[k + 1] * sz[k + 1]);
          if (depth1 > depth2) {
            diffX = sx[k + 1] - sx[k] * depth2 / depth1;
            diffY = sy[k + 1] - sy[k] * depth2 / depth1;
            diffZ = sz[k + 1] - sz[k] * depth2 / depth1;
            if (dm::sqrt(diffX * diffX + diffY * diffY + diffZ * diffZ) / depth2 < 0.1) a = 1;
          } else {
            diffX = sx[k + 1] * depth1 / depth2 - sx[k];
            diffY = sy[k + 1] * depth1 / depth2 - sy[k];
            diffZ = sz[k + 1] * depth1 / depth2 - sz[k];
            if (dm::sqrt(diffX * diffX + diffY * diffY + diffZ * diffZ) / depth1 < 0.1) b = 1;
          }
        }
      }
      sA[k] = a; sB[k] = b;
    }
    __syncthreads();
    int j = tile0 + threadIdx.x;
    int k = threadIdx.x + CR_HALO;
    if (j < n) {
      unsigned char pk = 0;
      if (j >= 5 && j < n - 6) {  // C(j), :440-451
        float dX = sx[k + 1] - sx[k], dY = sy[k + 1] - sy[k], dZ = sz[k + 1] - sz[k];
        float diff = dX * dX + dY * dY + dZ * dZ;
        float diffX2 = sx[k] - sx[k - 1];
        float diffY2 = sy[k] - sy[k - 1];
        float diffZ2 = sz[k] - sz[k - 1];
        float diff2 = diffX2 * diffX2 + diffY2 * diffY2 + diffZ2 * diffZ2;
        float dis = sx[k] * sx[k] + sy[k] * sy[k] + sz[k] * sz[k];
        if (diff > 0.0002 * dis && diff2 > 0.0002 * dis) pk = 1;
      }
#pragma unroll
      for (int d = 0; d <= 5; d++) pk |= sA[k + d];
#pragma unroll
      for (int d = 1; d <= 6; d++) pk |= sB[k - d];
      picked[j] = pk;
      // gap flag between j-1 and j, for the +-5 suppression loops
      unsigned char gf = 1;
      if (j >= 1) {
        float diffX = sx[k] - sx[k - 1];
        float diffY = sy[k] - sy[k - 1];
        float diffZ = sz[k] - sz[k - 1];
        gf = (diffX * diffX + diffY * diffY + diffZ * diffZ > 0.05) ? 1 : 0;
      }
      gapflag[j] = gf;
      label[j] = 0;
      sort_ind[j] = j;
      float c = 0.f;
      if (j >= 5 && j < n - 5) {
        float diffX = sx[k - 5] + sx[k - 4] + sx[k - 3] + sx[k - 2] + sx[k - 1] - 10 * sx[k]
                    + sx[k + 1] + sx[k + 2] + sx[k + 3] + sx[k + 4] + sx[k + 5];
        float diffY = sy[k - 5] + sy[k - 4] + sy[k - 3] + sy[k - 2] + sy[k - 1] - 10 * sy[k]
                    + sy[k + 1] + sy[k + 2] + sy[k + 3] + sy[k + 4] + sy[k + 5];
        float diffZ = sz[k - 5] + sz[k - 4] + sz[k - 3] + sz[k - 2] + sz[k - 1] - 10 * sz[k]
                    + sz[k + 1] + sz[k + 2] + sz[k + 3] + sz[k + 4] + sz[k + 5];
        c = diffX * diffX + diffY * diffY + diffZ * diffZ;
      }
      curvature[j] = c;
    }
  }
}

__device__ __forceinline__ void sector_bounds(const int* scan_start, const int* scan_end, int ring, int j, int& sp, int& ep) {
  int s = scan_start[ring], e = scan_end[ring];
  sp = (s * (6 - j) + e * j) / 6;             // :463
  ep = (s * (5 - j) + e * (j + 1)) / 6 - 1;   // :464
}

// K-F: grid (6, R).  The reference's bubble sort (:466-474) swaps on strict '<' only, i.e. it is
// a stable ascending sort by curvature == ascending sort of the key (curvature bits, index).
__device__ __forceinline__ void sr_sort_sectors_body(const float* __restrict__ curvature, int* __restrict__ sort_ind,
                                                       const int* __restrict__ scan_start, const int* __restrict__ scan_end,
                                                       SRCounts* cnt) {
  __shared__ unsigned long long keys[SR_SECTOR_CAP];
  int ring = blockIdx.y, j = blockIdx.x;
  int sp, ep;
  sector_bounds(scan_start, scan_end, ring, j, sp, ep);
  int m = ep - sp + 1;
  int nfull = cnt->n_full;
  if (m <= 1 || sp < 0 || ep >= nfull) return;
  if (m > SR_SECTOR_CAP) { if (threadIdx.x == 0) atomicOr(&cnt->error, 1); return; }
  int np2 = 2;
  while (np2 < m) np2 <<= 1;
  for (int k = threadIdx.x; k < np2; k += 512) {
    unsigned long long key = ~0ull;
    if (k < m) {
      int idx = sp + k;
      key = ((unsigned long long)__float_as_uint(curvature[idx]) << 32) | (unsigned int)idx;
    }
    keys[k] = key;
  }
  __syncthreads();
  for (int size = 2; size <= np2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < (np2 >> 1); t += 512) {
        int lo = 2 * t - (t & (stride - 1));
        int hi = lo + stride;
        bool up = ((lo & size) == 0);
        unsigned long long a = keys[lo], b = keys[hi];
        if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (int k = threadIdx.x; k < m; k += 512) sort_ind[sp + k] = (int)(keys[k] & 0xffffffffu);
}

// K-G: one CTA per ring.  Greedy selection :476-566 with the suppression flags shared by the six
// sectors of the ring (they spill up to 5 positions across sector boundaries, App. A.1 §14).
// The picks of a sector are strictly sequential, so the whole working set of the ring is staged in
// shared memory first (sorted (curvature, index) pairs, picked flags, a running count of the
// "gap" flags) and the greedy scans run without touching global memory:
//   position j is suppressed by a pick p  <=>  |j - p| <= 5 and no gap flag lies between them
//                                         <=>  |j - p| <= 5 and seg[j] == seg[p]   (seg = prefix count of gap flags)
// which every lane evaluates locally for its own candidate after the pick is broadcast.
//
// The six sectors of a ring depend on each other only through the flags a sector raises in the first
// five positions of the NEXT one (its "forward spill"; the backward spill lands in a sector that is
// already finished and only shows in the final flags).  So six warps run the six sectors at once, each
// as if nothing spilled into it, keeping their out-of-sector marks aside; warp 0 then walks the
// sectors in order: a sector's speculative result is exactly the sequential one unless one of ITS
// picks sits on a position the previous sector's final spill newly flags (a flagged position is only
// ever skipped, and a position examined while unflagged is either picked or ends the scan), in which
// case that sector alone is redone with the right flags.  Rings with a sector shorter than 11 points
// (a spill could reach the sector after next) take the plain sequential path.
constexpr int SEL_THREADS = 256;
constexpr int SEL_CAP = SR_RING_CAP + 32;   // ring window: [scanStart - 5, scanEnd + 5)
struct SelShared {
  unsigned char picked[SEL_CAP];
  unsigned char picked0[SEL_CAP];   // flags as staged (before any pick of this ring)
  unsigned short seg[SEL_CAP];
  int ind[SEL_CAP];      // sorted order, window-relative
  float curv[SEL_CAP];   // curvature in sorted order
  short cpick[6][20], fpick[6][4];   // picks of every sector in emission order (window-relative)
  int nc[6], nf[6];
  unsigned int fwd[6], bwd[6];       // out-of-sector marks: bit b of fwd = position ep+1+b, of bwd = position sp-1-b
};

// one warp: the two greedy scans of sector [sp, ep] (window-relative positions = index - w0).
// direct: marks go straight into the flags wherever they land (sequential path); otherwise marks
// outside [sp, ep] are collected in fwd / bwd.
__device__ __forceinline__ void sel_sector(SelShared& S, int j, int sp, int ep, int w0, int wn, bool direct) {
  const int lane = threadIdx.x & 31;
  int nc = 0, nf = 0;
  if (lane == 0) { S.fwd[j] = 0u; S.bwd[j] = 0u; }
  __syncwarp();
  // ---- corners: k = ep .. sp ----
  int largest = 0;
  bool done = false;
  for (int base = ep; base >= sp && !done; base -= 32) {
    const int k = base - lane;
    const bool valid = k >= sp;
    const int ind = valid ? S.ind[k - w0] : 0;
    const float c = valid ? S.curv[k - w0] : 0.f;
    const int seg = S.seg[ind];
    bool ok = valid && (c > 0.1) && S.picked[ind] == 0;
    const unsigned int stop = __ballot_sync(0xffffffffu, valid && !(c > 0.1));
    while (true) {
      const unsigned int mk = __ballot_sync(0xffffffffu, ok);
      if (mk == 0) break;
      const int L = __ffs(mk) - 1;
      const int pind = __shfl_sync(0xffffffffu, ind, L);
      const int pseg = __shfl_sync(0xffffffffu, seg, L);
      largest++;
      if (largest <= 20) {          // :484-493: the first two are sharp (label 2), the rest less sharp (label 1)
        if (lane == 0) S.cpick[j][nc] = (short)pind;
        nc++;
      } else {
        done = true;
        break;
      }
      if (lane < 11) {   // :495-520, positions pind-5 .. pind+5
        const int q = pind - 5 + lane;
        if (q >= 0 && q < wn && S.seg[q] == pseg) {
          const int qa = q + w0;
          if (direct || (qa >= sp && qa <= ep)) S.picked[q] = 1;
          else if (qa > ep) atomicOr(&S.fwd[j], 1u << (qa - ep - 1));
          else atomicOr(&S.bwd[j], 1u << (sp - 1 - qa));
        }
      }
      if (abs(ind - pind) <= 5 && seg == pseg) ok = false;
      if (lane <= L) ok = false;
    }
    __syncwarp();
    if (stop) done = true;  // ascending sort: everything further down has c <= 0.1
  }
  // ---- flat: k = sp .. ep ----
  int smallest = 0;
  done = false;
  for (int base = sp; base <= ep && !done; base += 32) {
    const int k = base + lane;
    const bool valid = k <= ep;
    const int ind = valid ? S.ind[k - w0] : 0;
    const float c = valid ? S.curv[k - w0] : 0.f;
    const int seg = S.seg[ind];
    bool ok = valid && (c < 0.1) && S.picked[ind] == 0;
    const unsigned int stop = __ballot_sync(0xffffffffu, valid && !(c < 0.1));
    while (true) {
      const unsigned int mk = __ballot_sync(0xffffffffu, ok);
      if (mk == 0) break;
      const int L = __ffs(mk) - 1;
      const int pind = __shfl_sync(0xffffffffu, ind, L);
      const int pseg = __shfl_sync(0xffffffffu, seg, L);
      if (lane == 0) S.fpick[j][nf] = (short)pind;
      nf++;
      smallest++;
      if (smallest >= 4) { done = true; break; }  // :534-536 breaks BEFORE marking
      if (lane < 11) {
        const int q = pind - 5 + lane;
        if (q >= 0 && q < wn && S.seg[q] == pseg) {
          const int qa = q + w0;
          if (direct || (qa >= sp && qa <= ep)) S.picked[q] = 1;
          else if (qa > ep) atomicOr(&S.fwd[j], 1u << (qa - ep - 1));
          else atomicOr(&S.bwd[j], 1u << (sp - 1 - qa));
        }
      }
      if (abs(ind - pind) <= 5 && seg == pseg) ok = false;
      if (lane <= L) ok = false;
    }
    __syncwarp();
    if (stop) done = true;
  }
  if (lane == 0) { S.nc[j] = nc; S.nf[j] = nf; }
  __syncwarp();
}

__device__ __forceinline__ void sr_select_body(const float* __restrict__ curvature, const int* __restrict__ sort_ind,
                                                         const int* __restrict__ scan_start, const int* __restrict__ scan_end,
                                                         const int* __restrict__ ring_start, const unsigned char* __restrict__ gapflag,
                                                         unsigned char* __restrict__ picked, signed char* __restrict__ label,
                                                         int* __restrict__ sharp_idx, int* __restrict__ less_sharp_idx,
                                                         int* __restrict__ flat_idx, int* __restrict__ ring_counts,
                                                         SRCounts* __restrict__ cnt) {
  extern __shared__ unsigned char sel_smem[];
  SelShared& S = *reinterpret_cast<SelShared*>(sel_smem);
  __shared__ int s_wsum[SEL_THREADS / 32];
  __shared__ int s_parallel;
  const int ring = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nfull = cnt->n_full;
  const int Sx = scan_start[ring], E = scan_end[ring];
  int* my_sharp = sharp_idx + ring * SR_SHARP_PER_RING;
  int* my_less = less_sharp_idx + ring * SR_LESS_SHARP_PER_RING;
  int* my_flat = flat_idx + ring * SR_FLAT_PER_RING;
  const int w0 = max(Sx - 5, 0), w1 = min(E + 5, nfull);   // window of positions this ring can touch
  const int wn = w1 - w0;
  if (E <= Sx || wn <= 0 || Sx < 0) {   // no sector holds a point (:463-464 give ep < sp)
    if (threadIdx.x == 0) { ring_counts[ring * 4 + 0] = 0; ring_counts[ring * 4 + 1] = 0; ring_counts[ring * 4 + 2] = 0; }
    return;
  }
  if (wn > SEL_CAP) {
    if (threadIdx.x == 0) { atomicOr(&cnt->error, 2); ring_counts[ring * 4 + 0] = 0; ring_counts[ring * 4 + 1] = 0; ring_counts[ring * 4 + 2] = 0; }
    return;
  }
  DBG_STAMP(25);
  // ---- stage: picked flags, prefix count of gap flags, sorted (index, curvature) ----
  // every thread owns a contiguous chunk of the window: serial count of its gap flags, ONE block-wide
  // scan of the 256 chunk totals, then the running count inside the chunk
  {
    const int C = (wn + SEL_THREADS - 1) / SEL_THREADS;   // <= 17
    const int k0 = threadIdx.x * C, k1 = min(k0 + C, wn);
    int tot = 0;
    for (int k = k0; k < k1; k++) {
      const unsigned char f = picked[w0 + k];
      S.picked[k] = f;
      S.picked0[k] = f;
      const int gf = (k > 0 && gapflag[w0 + k]) ? 1 : 0;   // gap between w0+k-1 and w0+k
      tot += gf;
      S.seg[k] = (unsigned short)gf;
    }
    int x = tot;
    for (int o = 1; o < 32; o <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_wsum[warp] = x;
    __syncthreads();
    int run = x - tot;
    for (int w = 0; w < warp; w++) run += s_wsum[w];
    for (int k = k0; k < k1; k++) { run += S.seg[k]; S.seg[k] = (unsigned short)run; }
  }
  for (int k = Sx + threadIdx.x; k < min(E, nfull); k += SEL_THREADS) {   // sectors tile [S, E-1]
    const int ind = sort_ind[k];
    S.ind[k - w0] = ind - w0;
    S.curv[k - w0] = curvature[ind];
  }
  if (threadIdx.x == 0) {
    int par = 1;
    for (int j = 0; j < 6; j++) {
      int sp, ep;
      sector_bounds(scan_start, scan_end, ring, j, sp, ep);
      if (sp < 0 || ep >= nfull || ep - sp + 1 < 11) par = 0;
      S.nc[j] = 0; S.nf[j] = 0; S.fwd[j] = 0u; S.bwd[j] = 0u;
    }
    s_parallel = par;
  }
  __syncthreads();
  const bool parallel = s_parallel != 0;
  DBG_STAMP(26);
  if (parallel) {
    // ---- six sectors at once, each blind to its neighbours' marks ----
    if (warp < 6) {
      int sp, ep;
      sector_bounds(scan_start, scan_end, ring, warp, sp, ep);
      sel_sector(S, warp, sp, ep, w0, wn, false);
    }
    __syncthreads();
    DBG_STAMP(27);
    // ---- warp 0: accept or redo the sectors in order ----
    if (warp == 0) {
      for (int j = 1; j < 6; j++) {
        int sp, ep;
        sector_bounds(scan_start, scan_end, ring, j, sp, ep);
        const unsigned int in = S.fwd[j - 1];          // final forward spill of the previous sector
        if (in == 0u) continue;
        // positions it newly flags, and whether this sector picked one of them
        unsigned int fresh = 0u;
        for (int b = 0; b < 5; b++)
          if ((in >> b & 1u) && S.picked0[sp + b - w0] == 0) fresh |= 1u << b;
        bool hit = false;
        if (fresh) {
          const int nc = S.nc[j], nf = S.nf[j];
          int pos = -1;
          if (lane < nc) pos = S.cpick[j][lane];
          else if (lane - 20 >= 0 && lane - 20 < nf) pos = S.fpick[j][lane - 20];
          const int b = pos < 0 ? -1 : pos + w0 - sp;
          hit = __any_sync(0xffffffffu, b >= 0 && b < 5 && (fresh >> b & 1u));
        }
        if (!hit) {
          if (lane < 5 && (in >> lane & 1u)) S.picked[sp + lane - w0] = 1;
          __syncwarp();
          continue;
        }
        for (int q = sp - w0 + lane; q <= ep - w0; q += 32) S.picked[q] = S.picked0[q];
        __syncwarp();
        if (lane < 5 && (in >> lane & 1u)) S.picked[sp + lane - w0] = 1;
        __syncwarp();
        sel_sector(S, j, sp, ep, w0, wn, false);
      }
      // out-of-sector marks that no later scan needed: backward spills, and the last sector's forward one
      if (lane < 30) {
        const int j = lane / 5, b = lane % 5;
        int sp, ep;
        sector_bounds(scan_start, scan_end, ring, j, sp, ep);
        if (S.bwd[j] >> b & 1u) { const int q = sp - 1 - b - w0; if (q >= 0 && q < wn) S.picked[q] = 1; }
        if (j == 5 && (S.fwd[5] >> b & 1u)) { const int q = ep + 1 + b - w0; if (q >= 0 && q < wn) S.picked[q] = 1; }
      }
    }
  } else if (warp == 0) {
    for (int j = 0; j < 6; j++) {
      int sp, ep;
      sector_bounds(scan_start, scan_end, ring, j, sp, ep);
      if (sp < 0 || ep >= nfull || ep < sp) continue;
      sel_sector(S, j, sp, ep, w0, wn, true);
    }
  }
  __syncthreads();
  DBG_STAMP(28);
  // ---- emission in the reference's order: sector by sector, corners then flats ----
  if (warp == 0) {
    int n_sharp = 0, n_less = 0, n_flat = 0;
    for (int j = 0; j < 6; j++) {
      const int nc = S.nc[j], nf = S.nf[j];
      if (lane < nc) {
        const int g = w0 + S.cpick[j][lane];
        label[g] = lane < 2 ? 2 : 1;
        if (lane < 2) my_sharp[n_sharp + lane] = g;
        my_less[n_less + lane] = g;
      }
      if (lane < nf) {
        const int g = w0 + S.fpick[j][lane];
        label[g] = -1;
        my_flat[n_flat + lane] = g;
      }
      n_sharp += min(nc, 2); n_less += nc; n_flat += nf;
    }
    if (lane == 0) {
      ring_counts[ring * 4 + 0] = n_sharp;
      ring_counts[ring * 4 + 1] = n_less;
      ring_counts[ring * 4 + 2] = n_flat;
    }
  }
  // cloudNeighborPicked after selection (only ever raised; windows of different rings are disjoint)
  for (int k = threadIdx.x; k < wn; k += SEL_THREADS) if (S.picked[k]) picked[w0 + k] = 1;
  DBG_STAMP(29);
}

// K-H: one block per ring.  surfPointsLessFlatScan = positions k in [scanStart, scanEnd-1] with
// cloudLabel[k] <= 0 in index order (:568-572), then pcl::VoxelGrid(0.2) (:575-581, App. B.1):
// cell = (int)(floor(p*5) - (float)min_b), idx = i + j*dx + k*dx*dy, output ascending idx,
// centroid of all four fields = running fp32 sum in input order / (float)count.
struct LessFlatSmem {
  float4 pts[SR_RING_CAP];
  unsigned int k[2][SR_RING_CAP], v[2][SR_RING_CAP];
  CtaRadixSmem R;
};
__device__ __forceinline__ void sr_lessflat_body(const float4* __restrict__ full, const signed char* __restrict__ label,
                                                    const int* __restrict__ scan_start, const int* __restrict__ scan_end,
                                                    float4* __restrict__ out_tmp, int* __restrict__ ring_counts,
                                                    SRCounts* cnt) {
  extern __shared__ unsigned char smem_raw[];
  LessFlatSmem& LS = *reinterpret_cast<LessFlatSmem*>(smem_raw);
  float4* pts = LS.pts;
  __shared__ int s_scan[32];
  __shared__ int s_carry, s_m, s_out;
  __shared__ unsigned int s_mn[3], s_mx[3];
  const int ring = blockIdx.x;
  const int nfull = cnt->n_full;
  int lo = scan_start[ring], hi = scan_end[ring] - 1;  // sectors tile [start, end-1]
  if (threadIdx.x == 0) { s_carry = 0; s_out = 0; }
  if (threadIdx.x < 3) { s_mn[threadIdx.x] = 0xffffffffu; s_mx[threadIdx.x] = 0u; }
  __syncthreads();
  if (lo < 0) lo = 0;
  if (hi >= nfull) hi = nfull - 1;
  DBG_STAMP(20);
  // ordered compaction of candidates into pts[]
  for (int base = lo; base <= hi; base += 1024) {
    int k = base + threadIdx.x;
    int flag = (k <= hi && label[k] <= 0) ? 1 : 0;
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
    int pos = s_carry + s_scan[threadIdx.x >> 5] + x - flag;
    if (flag) {
      if (pos < SR_RING_CAP) {
        float4 p = full[k];
        pts[pos] = p;
        if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
          atomicMin(&s_mn[0], float_flip(p.x)); atomicMax(&s_mx[0], float_flip(p.x));
          atomicMin(&s_mn[1], float_flip(p.y)); atomicMax(&s_mx[1], float_flip(p.y));
          atomicMin(&s_mn[2], float_flip(p.z)); atomicMax(&s_mx[2], float_flip(p.z));
        }
      }
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = pos + flag;
    __syncthreads();
  }
  int m = s_carry;
  DBG_STAMP(21);
  if (m > SR_RING_CAP) { if (threadIdx.x == 0) atomicOr(&cnt->error, 2); m = SR_RING_CAP; }
  if (m == 0) { if (threadIdx.x == 0) ring_counts[ring * 4 + 3] = 0; return; }
  const float inv = 1.0f / 0.2f;  // inverse_leaf_size = 1/leaf in fp32
  int minb[3], divb[3];
  for (int a = 0; a < 3; a++) {
    float mn = float_unflip(s_mn[a]), mx = float_unflip(s_mx[a]);
    minb[a] = (int)floorf(mn * inv);
    int maxb = (int)floorf(mx * inv);
    divb[a] = maxb - minb[a] + 1;
  }
  for (int k = threadIdx.x; k < m; k += 1024) {
    float4 p = pts[k];
    int i0 = (int)(floorf(p.x * inv) - (float)minb[0]);
    int i1 = (int)(floorf(p.y * inv) - (float)minb[1]);
    int i2 = (int)(floorf(p.z * inv) - (float)minb[2]);
    LS.k[0][k] = (unsigned int)(i0 + i1 * divb[0] + i2 * divb[0] * divb[1]);
    LS.v[0][k] = (unsigned int)k;
  }
  __syncthreads();
  DBG_STAMP(22);
  // ascending (voxel index, position): stable LSD radix sort over the bytes the largest index needs
  // (a 66-step bitonic network on 2048 keys took 17 of this kernel's 25 us)
  const unsigned int max_idx = (unsigned int)divb[0] * (unsigned int)divb[1] * (unsigned int)divb[2] - 1u;
  const int passes = max_idx < (1u << 8) ? 1 : (max_idx < (1u << 16) ? 2 : (max_idx < (1u << 24) ? 3 : 4));
  int cur = 0;
  for (int ps = 0; ps < passes; ps++) {
    cta_radix_pass_reg<SR_RING_CAP / 1024>(LS.k[cur], LS.v[cur], LS.k[cur ^ 1], LS.v[cur ^ 1], m, 8 * ps, LS.R);
    cur ^= 1;
  }
  const unsigned int* kk = LS.k[cur];
  const unsigned int* vv = LS.v[cur];
  DBG_STAMP(23);
  // segment heads -> ordered output slots
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < m; base += 1024) {
    int k = base + threadIdx.x;
    int head = 0;
    if (k < m) head = (k == 0) || (kk[k] != kk[k - 1]);
    int x = head;
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
    int slot = s_carry + s_scan[threadIdx.x >> 5] + x - head;
    if (head) {
      const unsigned int hk = kk[k];
      float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
      int q = k;
      while (q < m && kk[q] == hk) {
        float4 p = pts[(int)vv[q]];
        c0 += p.x; c1 += p.y; c2 += p.z; c3 += p.w;
        q++;
      }
      float cntf = (float)(q - k);
      c0 /= cntf; c1 /= cntf; c2 /= cntf; c3 /= cntf;
      out_tmp[(size_t)ring * SR_RING_CAP + slot] = make_float4(c0, c1, c2, c3);
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = slot + head;
    __syncthreads();
  }
  if (threadIdx.x == 0) ring_counts[ring * 4 + 3] = s_carry;
  DBG_STAMP(24);
}

// K-I: one block per ring: offsets = exclusive scan over rings of the four per-ring counts, then
// copy.  Emission order = ring -> sector -> pick order (:485-489,:531), less-flat ring -> voxel.
__device__ __forceinline__ void sr_compact_body(const float4* __restrict__ full, const int* __restrict__ ring_counts,
                                                  int n_scans, const int* __restrict__ sharp_idx,
                                                  const int* __restrict__ less_sharp_idx, const int* __restrict__ flat_idx,
                                                  const float4* __restrict__ lessflat_tmp, float4* __restrict__ sharp,
                                                  float4* __restrict__ less_sharp, float4* __restrict__ flat,
                                                  float4* __restrict__ less_flat, int* __restrict__ sharp_out_idx,
                                                  int* __restrict__ less_sharp_out_idx, int* __restrict__ flat_out_idx,
                                                  SRCounts* cnt, int max_points) {
  __shared__ int off[4], tot[4];
  const int ring = blockIdx.x;
  if (threadIdx.x < 4) {
    int o = 0, t = 0;
    for (int r = 0; r < n_scans; r++) {
      int c = ring_counts[r * 4 + threadIdx.x];
      if (r < ring) o += c;
      t += c;
    }
    off[threadIdx.x] = o;
    tot[threadIdx.x] = t;
  }
  __syncthreads();
  if (ring == 0 && threadIdx.x == 0) {
    cnt->n_sharp = tot[0]; cnt->n_less_sharp = tot[1]; cnt->n_flat = tot[2]; cnt->n_less_flat = tot[3];
  }
  int c0 = ring_counts[ring * 4 + 0], c1 = ring_counts[ring * 4 + 1], c2 = ring_counts[ring * 4 + 2], c3 = ring_counts[ring * 4 + 3];
  for (int k = threadIdx.x; k < c0; k += 128) {
    int idx = sharp_idx[ring * SR_SHARP_PER_RING + k];
    sharp[off[0] + k] = full[idx]; sharp_out_idx[off[0] + k] = idx;
  }
  for (int k = threadIdx.x; k < c1; k += 128) {
    int idx = less_sharp_idx[ring * SR_LESS_SHARP_PER_RING + k];
    less_sharp[off[1] + k] = full[idx]; less_sharp_out_idx[off[1] + k] = idx;
  }
  for (int k = threadIdx.x; k < c2; k += 128) {
    int idx = flat_idx[ring * SR_FLAT_PER_RING + k];
    flat[off[2] + k] = full[idx]; flat_out_idx[off[2] + k] = idx;
  }
  for (int k = threadIdx.x; k < c3; k += 128) {
    if (off[3] + k < max_points) less_flat[off[3] + k] = lessflat_tmp[(size_t)ring * SR_RING_CAP + k];
  }
}

}  // namespace loamgpu
