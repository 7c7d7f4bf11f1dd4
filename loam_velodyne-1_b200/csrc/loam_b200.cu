// loam_b200.cu -- the one translation unit of libloam_b200.so: globals, profiling hooks and the
// extern "C" boundary declared in include/loam_b200.h.  Compiled for sm_100a only, -fmad=false.
#include <stdarg.h>
#include <mutex>
#include "common.cuh"
#include "loam_sr.cuh"
#include "loam_lo.cuh"
#include "loam_lm.cuh"
#include "lm_persist.cuh"
#include "loam_pipeline.cuh"
#include "debug_hooks.cuh"

namespace loamgpu {

std::atomic<long long> g_launches(0);
std::atomic<long long> g_host_wait_ns(0);
std::atomic<long long> g_host_node_ns[3];
long long now_ns() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (long long)ts.tv_sec * 1000000000ll + ts.tv_nsec; }
int g_profile = 0;
static thread_local char g_err[512] = "";
static thread_local cudaEvent_t g_ev0 = nullptr, g_ev1 = nullptr;   // per launching thread (the batch's workers launch too)
static std::mutex g_prof_mu;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

std::map<std::string, ProfEntry>& prof_table() {
  static std::map<std::string, ProfEntry> t;
  return t;
}
// mode 2: a one-thread kernel writes %globaltimer before and after every launch on the launch's own
// stream; nothing is synchronised, so the stamps show how the streams really overlap
__global__ void trace_stamp(unsigned long long* out) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  *out = t;
}
struct TraceRec { const char* name; void* stream; int i0, i1; };
static unsigned long long* g_trace_buf = nullptr;
static const int TRACE_CAP = 1 << 17;
static int g_trace_n = 0, g_trace_i0 = -1;
static std::vector<TraceRec> g_trace_recs;

void prof_node(const char* name, cudaStream_t s, bool begin) {
  if (g_profile != 3 || !g_trace_buf || g_trace_n + 1 > TRACE_CAP) return;
  static std::map<std::string, int> open_idx;
  const int i = g_trace_n++;
  trace_stamp<<<1, 1, 0, s>>>(g_trace_buf + i);
  if (begin) { open_idx[name] = i; return; }
  TraceRec r;
  r.name = name; r.stream = (void*)s; r.i0 = open_idx.count(name) ? open_idx[name] : i; r.i1 = i;
  g_trace_recs.push_back(r);
}
void prof_begin(cudaStream_t s) {
  if (g_profile == 3) return;
  if (g_profile == 2) {
    g_trace_i0 = -1;
    if (g_trace_buf && g_trace_n + 2 <= TRACE_CAP) {
      g_trace_i0 = g_trace_n++;
      trace_stamp<<<1, 1, 0, s>>>(g_trace_buf + g_trace_i0);
    }
    return;
  }
  if (!g_ev0) { cudaEventCreate(&g_ev0); cudaEventCreate(&g_ev1); }
  cudaEventRecord(g_ev0, s);
}
void prof_end(const char* name, cudaStream_t s) {
  if (g_profile == 3) return;
  if (g_profile == 2) {
    if (g_trace_i0 >= 0) {
      TraceRec r;
      r.name = name; r.stream = (void*)s; r.i0 = g_trace_i0; r.i1 = g_trace_n++;
      trace_stamp<<<1, 1, 0, s>>>(g_trace_buf + r.i1);
      g_trace_recs.push_back(r);
    }
    return;
  }
  cudaEventRecord(g_ev1, s);
  cudaEventSynchronize(g_ev1);
  float ms = 0;
  cudaEventElapsedTime(&ms, g_ev0, g_ev1);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfEntry& e = prof_table()[name];
  e.ms += ms;
  e.calls++;
}

template <typename T>
static int download_array(const T* dev, int count, T* out, int cap, int* n, cudaStream_t s) {
  if (n) *n = count;
  int c = count < cap ? count : cap;
  if (c > 0 && out) {
    CUDA_OK(cudaMemcpyAsync(out, dev, sizeof(T) * (size_t)c, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
  }
  return LOAM_OK;
}

}  // namespace loamgpu

using namespace loamgpu;

// test hook: phase clocks of lo_optimize (which = 0) / lm_optimize (1); see common.cuh
static_assert(loamgpu::SR_MAX_RINGS == LOAM_LIMIT_RINGS && loamgpu::SR_RING_CAP == LOAM_LIMIT_RING_POINTS &&
              loamgpu::SR_SECTOR_CAP == LOAM_LIMIT_SECTOR_POINTS && loamgpu::PROC_MAX == LOAM_LIMIT_CUBES_PER_SWEEP,
              "the limits documented in include/loam_b200.h are the kernels' constants");

extern "C" int loam_debug_phase_enable(int on) {
#ifdef LOAM_PHASE_CLOCKS
  CUDA_OK(cudaMemcpyToSymbol(g_phase_on, &on, sizeof(int)));
  return LOAM_OK;
#else
  (void)on;
  set_error("phase clocks are compiled out of this build (rebuild with -DLOAM_PHASE_CLOCKS)");
  return LOAM_ERR_ARG;
#endif
}
extern "C" int loam_debug_phase_clocks(int which, long long* out, int n) {
  if (which == 3) {
    if (n > 1024) n = 1024;
    CUDA_OK(cudaDeviceSynchronize());
    CUDA_OK(cudaMemcpyFromSymbol(out, g_dbg_clk, sizeof(long long) * n, 0));
    return LOAM_OK;
  }
  if (which < 0 || which > 2 || !out || n < 0) return LOAM_ERR_ARG;
  CUDA_OK(cudaDeviceSynchronize());
  if (which == 2) {
    if (n > 2 * 250 * 8) n = 2 * 250 * 8;
    CUDA_OK(cudaMemcpyFromSymbol(out, g_rf_clk, sizeof(long long) * n, 0));
    return LOAM_OK;
  }
  if (n > PHASE_CLK_N) n = PHASE_CLK_N;
  CUDA_OK(cudaMemcpyFromSymbol(out, g_phase_clk, sizeof(long long) * n, sizeof(long long) * PHASE_CLK_N * which));
  return LOAM_OK;
}

struct loam_sr { SRHandle h; };
struct loam_lo { LOHandle h; };
struct loam_lm { LMHandle h; };
struct loam_tm { TMHost h; };
struct loam_pipeline { PipelineHandle h; };
struct loam_pipeline_batch { PipelineBatch b; };

extern "C" {

int loam_params_preset(loam_params* p, int preset) {
  if (!p) return LOAM_ERR_ARG;
  memset(p, 0, sizeof(*p));
  p->system_delay = 20;
  p->lo_max_iter = 25;
  p->lm_max_iter = 10;
  p->map_arena_points = 8 << 20;
  switch (preset) {
    case 1: p->n_scans = 16; p->ring_model = 0; p->ring_lower_deg = -15.f; p->ring_upper_deg = 15.f; p->max_points = 40000; break;
    case 2: p->n_scans = 64; p->ring_model = 1; p->ring_lower_deg = -24.8f; p->ring_upper_deg = 2.0f; p->max_points = 160000; break;
    case 4: p->n_scans = 128; p->ring_model = 1; p->ring_lower_deg = -25.f; p->ring_upper_deg = 15.f; p->max_points = 300000; break;
    default: return LOAM_ERR_ARG;
  }
  return LOAM_OK;
}

const char* loam_last_error(void) { return g_err; }
long long loam_kernel_launch_count(void) { return g_launches.load(); }
long long loam_debug_host_wait_ns(void) { return g_host_wait_ns.load(); }
void loam_debug_host_node_ns(long long out[3]) { for (int i = 0; i < 3; i++) out[i] = g_host_node_ns[i].load(); }
int loam_profile_enable(int on) {
  g_profile = on != 0 ? 1 : 0;
  prof_table().clear();
  return LOAM_OK;
}
int loam_debug_trace_enable(int on) {
  if (on && !g_trace_buf) CUDA_OK(cudaMalloc((void**)&g_trace_buf, sizeof(unsigned long long) * TRACE_CAP));
  CUDA_OK(cudaDeviceSynchronize());
  g_trace_n = 0;
  g_trace_recs.clear();
  g_profile = on == 3 ? 3 : (on ? 2 : 0);
  return LOAM_OK;
}
int loam_debug_trace_dump(const char* path) {
  if (!path || !g_trace_buf) { set_error("trace is not enabled"); return LOAM_ERR_ARG; }
  CUDA_OK(cudaDeviceSynchronize());
  std::vector<unsigned long long> t((size_t)g_trace_n + 1);
  if (g_trace_n > 0) CUDA_OK(cudaMemcpy(t.data(), g_trace_buf, sizeof(unsigned long long) * g_trace_n, cudaMemcpyDeviceToHost));
  FILE* f = fopen(path, "w");
  if (!f) { set_error("cannot open %s", path); return LOAM_ERR_ARG; }
  fprintf(f, "kernel,stream,start_ns,end_ns\n");
  for (const TraceRec& r : g_trace_recs) fprintf(f, "%s,%p,%llu,%llu\n", r.name, r.stream, t[r.i0], t[r.i1]);
  fclose(f);
  return (int)g_trace_recs.size();
}
int loam_profile_report(char* buf, int cap) {
  std::string s;
  char line[256];
  for (auto& kv : prof_table()) {
    snprintf(line, sizeof(line), "%s %lld %.6f\n", kv.first.c_str(), kv.second.calls, kv.second.ms);
    s += line;
  }
  if (buf && cap > 0) {
    strncpy(buf, s.c_str(), cap - 1);
    buf[cap - 1] = 0;
  }
  return (int)s.size();
}

// ---------------- scanRegistration ----------------
int loam_sr_create(const loam_params* p, int device, loam_sr** out) {
  if (!p || !out) return LOAM_ERR_ARG;
  loam_sr* h = new loam_sr();
  int r = h->h.init(p, device, nullptr);
  if (r != LOAM_OK) { delete h; return r; }
  *out = h;
  return LOAM_OK;
}
void loam_sr_destroy(loam_sr* h) { delete h; }
int loam_sr_imu(loam_sr* h, double stamp, const double quat[4], const float acc[3]) {
  if (!h) return LOAM_ERR_ARG;
  return h->h.on_imu(stamp, quat, acc);
}
int loam_sr_process(loam_sr* h, const float* xyz, int n, int stride_bytes, double stamp) {
  if (!h) return LOAM_ERR_ARG;
  int r = h->h.process(xyz, n, stride_bytes, stamp);
  if (r != LOAM_OK) return r;
  return h->h.sync_counts();
}
int loam_sr_process_packets(loam_sr* h, const unsigned char* packets, int n_packets, double stamp) {
  if (!h) return LOAM_ERR_ARG;
  int r = h->h.process_packets(packets, n_packets, stamp);
  if (r != LOAM_OK) return r;
  return h->h.sync_counts();
}
int loam_sr_download(loam_sr* hh, int which, float* xyzi, int cap, int* n) {
  if (!hh) return LOAM_ERR_ARG;
  SRHandle& h = hh->h;
  const SRCounts& c = h.last_counts;
  const float4* src; int count;
  switch (which) {
    case LOAM_CLOUD_FULL: src = h.d_full.p; count = c.n_full; break;
    case LOAM_CLOUD_SHARP: src = h.d_sharp.p; count = c.n_sharp; break;
    case LOAM_CLOUD_LESS_SHARP: src = h.d_less_sharp.p; count = c.n_less_sharp; break;
    case LOAM_CLOUD_FLAT: src = h.d_flat.p; count = c.n_flat; break;
    case LOAM_CLOUD_LESS_FLAT: src = h.d_less_flat.p; count = c.n_less_flat; break;
    default: return LOAM_ERR_ARG;
  }
  return download_array((const float4*)src, count, (float4*)xyzi, cap, n, h.stream);
}
int loam_sr_download_ints(loam_sr* hh, int which, int* out, int cap, int* n) {
  if (!hh) return LOAM_ERR_ARG;
  SRHandle& h = hh->h;
  const SRCounts& c = h.last_counts;
  switch (which) {
    case LOAM_INTS_SORT_IND: return download_array(h.d_sort_ind.p, c.n_full, out, cap, n, h.stream);
    case LOAM_INTS_SHARP_IDX: return download_array(h.d_sharp_idx.p, c.n_sharp, out, cap, n, h.stream);
    case LOAM_INTS_LESS_SHARP_IDX: return download_array(h.d_less_sharp_idx.p, c.n_less_sharp, out, cap, n, h.stream);
    case LOAM_INTS_FLAT_IDX: return download_array(h.d_flat_idx.p, c.n_flat, out, cap, n, h.stream);
    case LOAM_INTS_SCAN_START: return download_array(h.d_scan_start.p, h.P.n_scans, out, cap, n, h.stream);
    case LOAM_INTS_SCAN_END: return download_array(h.d_scan_end.p, h.P.n_scans, out, cap, n, h.stream);
    case LOAM_INTS_LABEL:
    case LOAM_INTS_PICKED: {
      int cnt = c.n_full;
      if (n) *n = cnt;
      int m = cnt < cap ? cnt : cap;
      std::vector<signed char> tmp(m > 0 ? m : 1);
      const void* src = which == LOAM_INTS_LABEL ? (const void*)h.d_label.p : (const void*)h.d_picked.p;
      if (m > 0) {
        CUDA_OK(cudaMemcpyAsync(tmp.data(), src, m, cudaMemcpyDeviceToHost, h.stream));
        CUDA_OK(cudaStreamSynchronize(h.stream));
        for (int i = 0; i < m; i++) out[i] = tmp[i];
      }
      return LOAM_OK;
    }
  }
  return LOAM_ERR_ARG;
}
int loam_sr_download_floats(loam_sr* hh, int which, float* out, int cap, int* n) {
  if (!hh) return LOAM_ERR_ARG;
  if (which != LOAM_FLOATS_CURVATURE) return LOAM_ERR_ARG;
  return download_array(hh->h.d_curv.p, hh->h.last_counts.n_full, out, cap, n, hh->h.stream);
}
int loam_sr_imu_trans(loam_sr* h, float out12[12]) {
  if (!h) return LOAM_ERR_ARG;
  h->h.imu_trans(out12);
  return LOAM_OK;
}

#include "loam_capi_rest.inc"

}  // extern "C"

// ---------------- allocation-free voxel filter (no counterpart in the reference) ----------------
// pcl::VoxelGrid(leaf) of one device-resident cloud through the same one-cluster kernel as
// loam_debug_voxel_filter_coop (vx_coop.cuh), for callers that cannot call cudaMalloc (CUDA graph
// capture, serving loops, harnesses that forbid it).  All scratch the kernel needs -- ping-pong
// keys / payloads, per-tile digit histograms, segment heads, the sorted copy of the points, the
// bounding box and three scalars -- lives in ONE file-scope __device__ array of fixed size; every
// GPU that loads the library holds it until the process exits.  In host code `g_na_scratch` is only
// a host-side stand-in for the symbol, never a device address: the address is fetched once per
// device with cudaGetSymbolAddress and that pointer is what the kernels receive.
// Every call shares the array and it keeps the previous call's contents: na_setup re-arms what must
// start from a known value (point count, bounding box, overflow flag, output count); everything
// else is fully rewritten by the kernel before it is read.  Calls that overlap on the GPU (two
// streams, two host threads) race on the array -- serialise them.
#define NA_MAX_POINTS LOAM_NOALLOC_MAX_POINTS
constexpr size_t NA_TILE_CAP = (NA_MAX_POINTS + loamgpu::VC_TILE - 1) / loamgpu::VC_TILE + 1;
constexpr size_t na_align(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t NA_OFF_K0 = 0;
constexpr size_t NA_OFF_K1 = NA_OFF_K0 + na_align(sizeof(unsigned int) * NA_MAX_POINTS);
constexpr size_t NA_OFF_V0 = NA_OFF_K1 + na_align(sizeof(unsigned int) * NA_MAX_POINTS);
constexpr size_t NA_OFF_V1 = NA_OFF_V0 + na_align(sizeof(unsigned int) * NA_MAX_POINTS);
constexpr size_t NA_OFF_SORTED = NA_OFF_V1 + na_align(sizeof(unsigned int) * NA_MAX_POINTS);
constexpr size_t NA_OFF_HIST = NA_OFF_SORTED + na_align(sizeof(float4) * NA_MAX_POINTS);
constexpr size_t NA_OFF_HEADS = NA_OFF_HIST + na_align(sizeof(int) * 256 * NA_TILE_CAP);
constexpr size_t NA_OFF_BBOX = NA_OFF_HEADS + na_align(sizeof(int) * NA_TILE_CAP);
constexpr size_t NA_OFF_SCAL = NA_OFF_BBOX + 256;   // int n, overflow, out_count
constexpr size_t NA_SCRATCH_BYTES = NA_OFF_SCAL + 256;
static __device__ __align__(256) unsigned char g_na_scratch[NA_SCRATCH_BYTES];

// device code may name the array directly; only host code needs the fetched address
__global__ void na_setup(int n, int* out_count) {
  unsigned int* bbox = (unsigned int*)(g_na_scratch + NA_OFF_BBOX);
  int* scal = (int*)(g_na_scratch + NA_OFF_SCAL);
  if (threadIdx.x < 6) bbox[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;
  if (threadIdx.x == 6) { scal[0] = n; scal[1] = 0; scal[2] = 0; *out_count = 0; }
}

constexpr int NA_MAX_DEVICES = 64;
static unsigned char* g_na_ptr[NA_MAX_DEVICES];   // device address of g_na_scratch, per device
static std::mutex g_na_mu;

extern "C" int loam_voxel_filter_noalloc(const float* d_xyzi, int n, float leaf, float* d_out, int cap_out, int* d_n_out, void* stream) {
  if (!d_xyzi || !d_out || !d_n_out || n < 0 || !(leaf > 0.f)) { set_error("loam_voxel_filter_noalloc: bad argument"); return LOAM_ERR_ARG; }
  if (n > NA_MAX_POINTS) {
    set_error("loam_voxel_filter_noalloc: %d points exceed the static scratch limit LOAM_NOALLOC_MAX_POINTS = %d", n, (int)NA_MAX_POINTS);
    return LOAM_ERR_CAPACITY;
  }
  if (cap_out < n) {   // every point can be its own voxel; the kernel does not clip its output
    set_error("loam_voxel_filter_noalloc: cap_out (%d) must be at least n (%d)", cap_out, n);
    return LOAM_ERR_ARG;
  }
  int dev = -1;
  CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NA_MAX_DEVICES) { set_error("loam_voxel_filter_noalloc: device ordinal %d out of range", dev); return LOAM_ERR_ARG; }
  unsigned char* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[dev]) {   // set-up, once per device, with that device current
      cudaDeviceProp prop;
      CUDA_OK(cudaGetDeviceProperties(&prop, dev));
      if (prop.major < 10) { set_error("loam_voxel_filter_noalloc: device %d is sm_%d%d, sm_100 needed (no CPU fallback)", dev, prop.major, prop.minor); return LOAM_ERR_CUDA; }
      unsigned char* p = nullptr;
      CUDA_OK(cudaGetSymbolAddress((void**)&p, g_na_scratch));
      CUDA_OK(cudaFuncSetAttribute(vx_filter_coop, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      g_na_ptr[dev] = p;
    }
    base = g_na_ptr[dev];
  }
  cudaStream_t s = (cudaStream_t)stream;
  int* scal = (int*)(base + NA_OFF_SCAL);
  VxCoopArgs VA;
  VxCoopJob& j = VA.j[0];
  j.pts = (const float4*)d_xyzi; j.n_ptr = scal; j.bbox = (unsigned int*)(base + NA_OFF_BBOX);
  j.keys[0] = (unsigned int*)(base + NA_OFF_K0); j.keys[1] = (unsigned int*)(base + NA_OFF_K1);
  j.vals[0] = (unsigned int*)(base + NA_OFF_V0); j.vals[1] = (unsigned int*)(base + NA_OFF_V1);
  j.hist = (int*)(base + NA_OFF_HIST); j.heads = (int*)(base + NA_OFF_HEADS); j.sorted = (float4*)(base + NA_OFF_SORTED);
  j.out = (float4*)d_out; j.out_count = d_n_out; j.overflow = scal + 1; j.leaf = leaf; j.tile_cap = (int)NA_TILE_CAP;
  VA.j[1] = j;
  na_setup<<<1, 32, 0, s>>>(n, d_n_out);
  g_launches++;
  vx_filter_coop<<<dim3(VC_CLUSTER, 1), VC_THREADS, 0, s>>>(VA);
  g_launches++;
  CUDA_OK(cudaGetLastError());
  return LOAM_OK;
}
