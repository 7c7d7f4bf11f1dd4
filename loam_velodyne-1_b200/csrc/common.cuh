// common.cuh -- shared host/device utilities for the sm_100a LOAM kernels.
// Device TUs are compiled with -fmad=false: every float expression that feeds an index-selecting
// comparison must round exactly like the reference's x86-64 SSE2 code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <map>
#include <atomic>
#include <vector>
#include "loam_b200.h"
#include "loam_b200_debug.h"
#include "loam/detmath.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace loamgpu {
namespace dm = loam::dm;

void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;
extern std::atomic<long long> g_host_wait_ns;   // host time spent blocked on back-pressure events (staging / parameter slots)
long long now_ns();
extern std::atomic<long long> g_host_node_ns[3];  // host time the pipeline spent enqueueing scanRegistration / laserOdometry / laserMapping
struct HostNode { int i; long long t0; HostNode(int i_) : i(i_), t0(now_ns()) {} ~HostNode() { g_host_node_ns[i] += now_ns() - t0; } };
// the capacity flags the kernels raise (SRCounts::error, LMState::error), spelled out for the caller
inline void describe_sr_error(int flags) {
  set_error("scanRegistration capacity exceeded:%s%s%s%s (flags %d; limits: include/loam_b200.h, \"Hard limits\")",
            (flags & 1) ? " more than 1024 points in one sixth of a ring (LOAM_LIMIT_SECTOR_POINTS);" : "",
            (flags & 2) ? " more than 4096 less-flat candidates on one ring (LOAM_LIMIT_RING_POINTS);" : "",
            (flags & 4) ? " more points than loam_params.max_points;" : "",
            (flags & ~7) ? " other;" : "", flags);
}
inline void describe_lm_error(int flags) {
  set_error("laserMapping capacity exceeded:%s%s%s%s%s%s (flags %d; limits: include/loam_b200.h, \"Hard limits\")",
            (flags & 1) ? " cube-map arena full (raise loam_params.map_arena_points);" : "",
            (flags & 2) ? " more than 250 cubes rewritten by one sweep (LOAM_LIMIT_CUBES_PER_SWEEP);" : "",
            (flags & 4) ? " a map point lies outside its cube's voxel range;" : "",
            (flags & 8) ? " local map larger than map_arena_points / 4;" : "",
            (flags & 16) ? " refilter scratch full (map_arena_points / 4);" : "",
            (flags & 32) ? " per-cube index arena full or a cube above 16 Mi points;" : "", flags);
}
struct HostWait { long long t0; HostWait() : t0(now_ns()) {} ~HostWait() { g_host_wait_ns += now_ns() - t0; } };
extern int g_profile;   // 0 off, 1 per-kernel CUDA events (serialising), 2 timeline stamps around every launch (non-serialising), 3 stamps at node boundaries only

struct ProfEntry { double ms = 0; long long calls = 0; };
std::map<std::string, ProfEntry>& prof_table();
void prof_begin(cudaStream_t s);
void prof_end(const char* name, cudaStream_t s);
void prof_node(const char* name, cudaStream_t s, bool begin);   // mode 3: stamps at node boundaries only

#define CUDA_OK(expr)                                                                     \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      loamgpu::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return LOAM_ERR_CUDA;                                                               \
    }                                                                                     \
  } while (0)

// every kernel launch of the library goes through this macro: counts launches (bench.py's
// gpu_launches) and, in profile mode, brackets the launch with CUDA events on its own stream
#define LAUNCH(name, kern, grid, block, smem, stream, ...)                 \
  do {                                                                     \
    if (loamgpu::g_profile && loamgpu::g_profile != 3) loamgpu::prof_begin(stream); \
    kern<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);              \
    loamgpu::g_launches++;                                                 \
    if (loamgpu::g_profile && loamgpu::g_profile != 3) loamgpu::prof_end(name, stream); \
  } while (0)

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  int alloc(size_t count) {
    free();
    n = count;
    if (count == 0) return LOAM_OK;
    cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
    if (e != cudaSuccess) { set_error("cudaMalloc(%zu bytes) failed: %s", count * sizeof(T), cudaGetErrorString(e)); p = nullptr; return LOAM_ERR_CUDA; }
    return LOAM_OK;
  }
  int alloc_zero(size_t count, cudaStream_t s = 0) {
    int r = alloc(count);
    if (r != LOAM_OK) return r;
    if (count && cudaMemsetAsync(p, 0, count * sizeof(T), s) != cudaSuccess) return LOAM_ERR_CUDA;
    return LOAM_OK;
  }
  void free() { if (p) cudaFree(p); p = nullptr; n = 0; }
  ~DevBuf() { free(); }
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
};

template <typename T>
struct PinnedBuf {
  T* p = nullptr;
  size_t n = 0;
  int alloc(size_t count) {
    free();
    n = count;
    if (cudaMallocHost((void**)&p, count * sizeof(T)) != cudaSuccess) { set_error("cudaMallocHost failed"); p = nullptr; return LOAM_ERR_CUDA; }
    return LOAM_OK;
  }
  void free() { if (p) cudaFreeHost(p); p = nullptr; n = 0; }
  ~PinnedBuf() { free(); }
};

// word copy executed by one CTA; src may be pinned host memory (zero-copy read over PCIe/C2C)
__global__ void __launch_bounds__(128) pinned_fetch(unsigned int* __restrict__ dst, const unsigned int* __restrict__ src, int nwords) {
  for (int i = threadIdx.x; i < nwords; i += 128) dst[i] = src[i];
}

// small ring of pinned staging slots for host->device parameter uploads that may still be in
// flight when the host prepares the next one (asynchronous pipeline mode)
template <typename T, int N>
struct PinnedRing {
  T* p[N];
  cudaEvent_t ev[N];
  int next = 0;
  bool ok = false;
  int alloc() {
    for (int i = 0; i < N; i++) {
      if (cudaMallocHost((void**)&p[i], sizeof(T)) != cudaSuccess) { set_error("cudaMallocHost failed"); return LOAM_ERR_CUDA; }
      if (cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) != cudaSuccess) return LOAM_ERR_CUDA;
    }
    ok = true;
    return LOAM_OK;
  }
  // copies `v` to the device through the next free slot
  int upload(T* dst, const T& v, cudaStream_t s) {
    int i = next;
    next = (next + 1) % N;
    { HostWait hw; if (cudaEventSynchronize(ev[i]) != cudaSuccess) return LOAM_ERR_CUDA; }
    *p[i] = v;
    // a kernel reads the pinned slot instead of a copy-engine transfer: the copy engines execute
    // their queue in order, so a small H2D that has to wait for its stream (e.g. the mapping
    // stream, which runs a sweep or two behind) would hold back the next sweeps' point transfers
    static_assert(sizeof(T) % 4 == 0, "word-sized parameter blocks only");
    pinned_fetch<<<1, 128, 0, s>>>((unsigned int*)dst, (const unsigned int*)p[i], (int)(sizeof(T) / 4));
    g_launches++;
    if (cudaEventRecord(ev[i], s) != cudaSuccess) return LOAM_ERR_CUDA;
    return LOAM_OK;
  }
  ~PinnedRing() {
    if (!ok) return;
    for (int i = 0; i < N; i++) { cudaFreeHost(p[i]); cudaEventDestroy(ev[i]); }
  }
};

// order-preserving float <-> uint mapping for atomicMin/Max on floats
__device__ __forceinline__ unsigned int float_flip(float f) {
  unsigned int u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float float_unflip(unsigned int u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// lanes of the warp holding the same 8-bit digit (0 for idle lanes): eight ballots.  match.any
// serialises over the distinct values of the warp (~0.3 us with random digits on sm_100), this
// does not.
__device__ __forceinline__ unsigned int warp_match8(int d, bool in) {
  unsigned int m = __ballot_sync(0xffffffffu, in);
#pragma unroll
  for (int b = 0; b < 8; b++) {
    const bool bit = (d >> b) & 1;
    const unsigned int bal = __ballot_sync(0xffffffffu, bit);
    m &= bit ? bal : ~bal;
  }
  return in ? m : 0u;
}

// phase clocks of the two persistent kernels (test hook, loam_debug_phase_clocks): when enabled,
// thread 0 of CTA 0 stamps clock64() at the phase boundaries of every LM iteration
// Compiled in only with -DLOAM_PHASE_CLOCKS (LOAM_PHASE_CLOCKS=1 python loam_velodyne-1_b200/build.py --force,
// scripts/dev_phase_clocks.py): in the default build PHASE_ON is the constant 0 and every stamp, the flag load
// in the per-feature loops included, is removed by the compiler.
constexpr int PHASE_CLK_N = 512;
#ifdef LOAM_PHASE_CLOCKS
__device__ int g_phase_on = 0;
#define PHASE_ON (g_phase_on)
#else
#define PHASE_ON 0
#endif
__device__ long long g_phase_clk[2][PHASE_CLK_N];
__device__ long long g_rf_clk[2 * 250 * 8];   // lm_refilter_cube: 8 stamps per (type, rank)
__device__ long long g_dbg_clk[1024];           // ad-hoc stamps of single-CTA kernels (dev only)
#define DBG_STAMP(slot) do { if (PHASE_ON && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_dbg_clk[slot] = clock64(); } while (0)
__device__ __forceinline__ void phase_stamp(int which, int& slot) {
  if (slot >= 0 && slot < PHASE_CLK_N) g_phase_clk[which][slot++] = clock64();
}

}  // namespace loamgpu
