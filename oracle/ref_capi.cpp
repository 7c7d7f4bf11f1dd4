// ORACLE (test infrastructure).  Checker API (checker_api.h) implemented by the reference's OWN
// node code: loads oracle/_ref/libref_{sr_N,lo,lm,tm}.so (built by build_ref.sh from
// /root/reference/src/*.cpp) and routes messages between them exactly as the topic graph of
// SURVEY.md §3.0 does.  Each handle loads private copies of the node libraries (the nodes keep
// their state in file-scope globals), so several handles can coexist in one process.
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <string>
#include <vector>
#include <atomic>
#include "checker_api.h"

namespace {

struct Node {
  void* dl = nullptr;
  std::string tmp;
  void (*start)() = nullptr;
  void (*stop)() = nullptr;
  void (*post_cloud)(const char*, const float*, int, int, double) = nullptr;
  void (*post_imu)(const char*, double, const double*, const double*) = nullptr;
  void (*post_odom)(const char*, const double*) = nullptr;
  void (*cycle)() = nullptr;
  int (*out_count)(const char*) = nullptr;
  void (*out_clear)() = nullptr;
  int (*get_cloud)(const char*, float*, int) = nullptr;
  int (*get_odom)(const char*, double*) = nullptr;
  double (*busy)() = nullptr;
  int (*get_ints)(int, int*, int) = nullptr;
  int (*get_floats)(int, float*, int) = nullptr;
  void* sym(const char* n) { return dlsym(dl, n); }
};

std::atomic<int> g_serial(0);

std::vector<std::string>& doomed() { static std::vector<std::string> v; return v; }
void remove_doomed() { for (const std::string& p : doomed()) unlink(p.c_str()); }
void remove_at_exit(const std::string& path) {
  static bool armed = false;
  if (!armed) { atexit(remove_doomed); armed = true; }
  doomed().push_back(path);
}

std::string self_dir() {
  Dl_info info;
  if (dladdr((void*)&self_dir, &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    size_t k = p.rfind('/');
    return k == std::string::npos ? "." : p.substr(0, k);
  }
  return ".";
}

bool load_node(Node& n, const std::string& name) {
#ifdef REF_LIBM
  std::string src = self_dir() + "/libref_" + name + "_libm.so";
#else
  std::string src = self_dir() + "/libref_" + name + ".so";
#endif
  // private copy => private globals
  char tmpl[256];
  // unique per process, per driver library (two builds of this driver can live in one process: the
  // reference's nodes and tests/adapter's) and per load
  snprintf(tmpl, sizeof(tmpl), "/tmp/loamref_%d_%lx_%d_%s.so", (int)getpid(), (unsigned long)(uintptr_t)&g_serial, g_serial.fetch_add(1), name.c_str());
  std::string cmd = "cp '" + src + "' '" + tmpl + "'";
  if (system(cmd.c_str()) != 0) { fprintf(stderr, "ref_capi: cannot copy %s\n", src.c_str()); return false; }
  n.tmp = tmpl;
  n.dl = dlopen(tmpl, RTLD_NOW | RTLD_LOCAL);
  if (!n.dl) { fprintf(stderr, "ref_capi: dlopen failed: %s\n", dlerror()); unlink(tmpl); return false; }
  // the copy stays on disk until the node is unloaded: glibc identifies loaded objects by (device,
  // inode), and an unlinked copy's inode can be handed to the next copy, which dlopen would then
  // take for the library that is already loaded
  n.start = (void (*)())n.sym("node_start");
  n.stop = (void (*)())n.sym("node_stop");
  n.post_cloud = (void (*)(const char*, const float*, int, int, double))n.sym("node_post_cloud");
  n.post_imu = (void (*)(const char*, double, const double*, const double*))n.sym("node_post_imu");
  n.post_odom = (void (*)(const char*, const double*))n.sym("node_post_odom");
  n.cycle = (void (*)())n.sym("node_cycle");
  n.out_count = (int (*)(const char*))n.sym("node_out_count");
  n.out_clear = (void (*)())n.sym("node_out_clear");
  n.get_cloud = (int (*)(const char*, float*, int))n.sym("node_get_cloud");
  n.get_odom = (int (*)(const char*, double*))n.sym("node_get_odom");
  n.busy = (double (*)())n.sym("node_busy");
  n.get_ints = (int (*)(int, int*, int))n.sym("node_get_ints");
  n.get_floats = (int (*)(int, float*, int))n.sym("node_get_floats");
  return n.start && n.cycle;
}

struct Handle {
  loam_params P;
  Node sr, lo, lm, tm;
  loam_sweep_result res;
  std::vector<float> buf, full;
  std::vector<float> c_sharp, c_less_sharp, c_flat, c_less_flat, c_corner_last, c_surf_last, c_full_end, c_registered, c_surround;
  int n_full = 0;
  bool ok = false;
};

int fetch(Node& n, const char* topic, std::vector<float>& dst, std::vector<float>& scratch) {
  int cnt = n.get_cloud(topic, scratch.data(), (int)(scratch.size() / 4));
  if (cnt * 4 > (int)scratch.size()) {
    scratch.resize((size_t)cnt * 4);
    cnt = n.get_cloud(topic, scratch.data(), cnt);
  }
  dst.assign(scratch.begin(), scratch.begin() + (size_t)cnt * 4);
  return cnt;
}

// nav_msgs/Odometry orientation -> (rx, ry, rz) as laserMapping.cpp:304-318 decodes it is done
// inside the receiving NODE; here poses are only read back for reporting from node globals.

}  // namespace

extern "C" {

const char* chk_kind(void) {
#ifdef REF_LIBM
  return "reference-libm";
#else
  return "reference";
#endif
}

void* chk_create(const loam_params* p) {
  Handle* h = new Handle();
  h->P = *p;
  memset(&h->res, 0, sizeof(h->res));
  std::string srname;
  if (p->n_scans == 16 && p->ring_model == 0) srname = "sr_1";
  else if (p->n_scans == 64 && p->ring_model == 1) srname = "sr_2";
  else if (p->n_scans == 128 && p->ring_model == 1) srname = "sr_4";
  else { fprintf(stderr, "ref_capi: no reference build for n_scans=%d ring_model=%d\n", p->n_scans, p->ring_model); delete h; return nullptr; }
  if (!load_node(h->sr, srname) || !load_node(h->lo, "lo") || !load_node(h->lm, "lm") || !load_node(h->tm, "tm")) { delete h; return nullptr; }
  void (*setdelay)(int) = (void (*)(int))h->sr.sym("node_set_system_delay");
  if (setdelay) setdelay(p->system_delay);
  h->sr.start(); h->lo.start(); h->lm.start(); h->tm.start();
  h->buf.resize((size_t)4 * 400000);
  h->ok = true;
  return h;
}

void chk_destroy(void* hv) {
  Handle* h = (Handle*)hv;
  if (!h) return;
  Node* ns[4] = {&h->sr, &h->lo, &h->lm, &h->tm};
  for (Node* n : ns) {
    if (n->dl) {
      if (n->stop) n->stop();
      dlclose(n->dl);
      // dlclose may leave the object mapped (libstdc++ unique symbols make it NODELETE); its file is
      // therefore only removed when the process exits, so that its inode cannot be handed to a later copy
      if (!n->tmp.empty()) remove_at_exit(n->tmp);
    }
  }
  delete h;
}

int chk_imu(void* hv, double stamp, const double quat[4], const float acc[3]) {
  Handle* h = (Handle*)hv;
  double a[3] = {acc[0], acc[1], acc[2]};
  h->sr.post_imu("/imu/data", stamp, quat, a);
  h->sr.cycle();
  h->lm.post_imu("/imu/data", stamp, quat, a);
  h->lm.cycle();
  return 0;
}

// the six scanRegistration topics held in h->full / c_sharp / ... / res.imu_trans -> laserOdometry ->
// transformMaintenance -> laserMapping, routed as the topic graph does
static void run_from_features(Handle* h, double stamp) {
  loam_sweep_result& r = h->res;
  Node &lo = h->lo, &lm = h->lm, &tm = h->tm;
  {
    r.sr_status = LOAM_OK;
    // -> laserOdometry (six topics, same stamp)
    lo.out_clear();
    lo.post_cloud("/laser_cloud_sharp", h->c_sharp.data(), r.n_sharp, 4, stamp);
    lo.post_cloud("/laser_cloud_less_sharp", h->c_less_sharp.data(), r.n_less_sharp, 4, stamp);
    lo.post_cloud("/laser_cloud_flat", h->c_flat.data(), r.n_flat, 4, stamp);
    lo.post_cloud("/laser_cloud_less_flat", h->c_less_flat.data(), r.n_less_flat, 4, stamp);
    lo.post_cloud("/velodyne_cloud_2", h->full.data(), r.n_full, 4, stamp);
    {
      std::vector<float> t3(12);
      for (int i = 0; i < 4; i++) { t3[3 * i] = r.imu_trans[3 * i]; t3[3 * i + 1] = r.imu_trans[3 * i + 1]; t3[3 * i + 2] = r.imu_trans[3 * i + 2]; }
      lo.post_cloud("/imu_trans", t3.data(), 4, 3, stamp);
    }
    lo.cycle();
    lo.get_floats(0, r.transform, 6);
    lo.get_floats(1, r.transform_sum, 6);
    bool lo_clouds = lo.out_count("/laser_cloud_corner_last") > 0;
    lm.out_clear();
    if (lo.out_count("/laser_odom_to_init") > 0) {
      r.lo_status = LOAM_OK;
      double od[14];
      lo.get_odom("/laser_odom_to_init", od);
      tm.out_clear();
      tm.post_odom("/laser_odom_to_init", od);
      tm.cycle();
      double im[14];
      if (tm.get_odom("/integrated_to_init", im)) {
        // decode as the consumers do (laserMapping.cpp:304-318): read TM's globals instead
        if (tm.get_floats) tm.get_floats(0, r.transform_mapped, 6);
      }
      lm.post_odom("/laser_odom_to_init", od);
    }
    if (lo_clouds) {
      fetch(lo, "/laser_cloud_corner_last", h->c_corner_last, h->buf);
      fetch(lo, "/laser_cloud_surf_last", h->c_surf_last, h->buf);
    }
    bool full3 = lo.out_count("/velodyne_cloud_3") > 0;
    if (full3) fetch(lo, "/velodyne_cloud_3", h->c_full_end, h->buf);
    // -> laserMapping (corner_last / surf_last are published on the init sweep too)
    if (lo_clouds) {
      lm.post_cloud("/laser_cloud_corner_last", h->c_corner_last.data(), (int)h->c_corner_last.size() / 4, 4, stamp);
      lm.post_cloud("/laser_cloud_surf_last", h->c_surf_last.data(), (int)h->c_surf_last.size() / 4, 4, stamp);
    }
    if (full3) lm.post_cloud("/velodyne_cloud_3", h->c_full_end.data(), (int)h->c_full_end.size() / 4, 4, stamp);
    lm.cycle();
    if (lm.out_count("/aft_mapped_to_init") > 0) {
      r.lm_status = LOAM_OK;
      lm.get_floats(0, r.transform_aft_mapped, 6);
      lm.get_floats(1, r.transform_bef_mapped, 6);
      fetch(lm, "/velodyne_cloud_registered", h->c_registered, h->buf);
      if (lm.out_count("/laser_cloud_surround") > 0) fetch(lm, "/laser_cloud_surround", h->c_surround, h->buf);
      double am[14];
      lm.get_odom("/aft_mapped_to_init", am);
      tm.post_odom("/aft_mapped_to_init", am);
      tm.cycle();
      int (*gci)(int, float*, int) = (int (*)(int, float*, int))lm.sym("node_get_cloud_internal");
      if (gci) {
        r.n_map_corner = gci(0, h->buf.data(), 0); r.n_map_surf = gci(1, h->buf.data(), 0);
        r.n_stack_corner = gci(2, h->buf.data(), 0); r.n_stack_surf = gci(3, h->buf.data(), 0);
      }
    }
  }
}

int chk_push_sweep(void* hv, const float* xyz, int n, double stamp, loam_sweep_result* out) {
  Handle* h = (Handle*)hv;
  loam_sweep_result& r = h->res;
  r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
  Node& sr = h->sr;
  sr.out_clear();
  sr.post_cloud("/velodyne_points", xyz, n, 3, stamp);
  sr.cycle();
  if (sr.out_count("/velodyne_cloud_2") > 0) {
    r.n_full = h->n_full = fetch(sr, "/velodyne_cloud_2", h->full, h->buf);
    r.n_sharp = fetch(sr, "/laser_cloud_sharp", h->c_sharp, h->buf);
    r.n_less_sharp = fetch(sr, "/laser_cloud_less_sharp", h->c_less_sharp, h->buf);
    r.n_flat = fetch(sr, "/laser_cloud_flat", h->c_flat, h->buf);
    r.n_less_flat = fetch(sr, "/laser_cloud_less_flat", h->c_less_flat, h->buf);
    std::vector<float> it;
    int ni = fetch(sr, "/imu_trans", it, h->buf);
    for (int i = 0; i < 4 && i < ni; i++) { r.imu_trans[3 * i] = it[4 * i]; r.imu_trans[3 * i + 1] = it[4 * i + 1]; r.imu_trans[3 * i + 2] = it[4 * i + 2]; }
    run_from_features(h, stamp);
  }
  if (out) *out = r;
  return r.sr_status;
}

// the same chain fed with HOST copies of scanRegistration's six topics (checker_api.h)
int chk_lo_push(void* hv, const float* full, int n_full, const float* sharp, int n_sharp, const float* less_sharp, int n_less_sharp,
                const float* flat, int n_flat, const float* less_flat, int n_less_flat, const float imu_trans[12], double stamp,
                loam_sweep_result* out) {
  Handle* h = (Handle*)hv;
  loam_sweep_result& r = h->res;
  r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
  h->full.assign(full, full + (size_t)4 * n_full);
  h->c_sharp.assign(sharp, sharp + (size_t)4 * n_sharp);
  h->c_less_sharp.assign(less_sharp, less_sharp + (size_t)4 * n_less_sharp);
  h->c_flat.assign(flat, flat + (size_t)4 * n_flat);
  h->c_less_flat.assign(less_flat, less_flat + (size_t)4 * n_less_flat);
  r.n_full = h->n_full = n_full; r.n_sharp = n_sharp; r.n_less_sharp = n_less_sharp; r.n_flat = n_flat; r.n_less_flat = n_less_flat;
  for (int i = 0; i < 12; i++) r.imu_trans[i] = imu_trans ? imu_trans[i] : 0.f;
  run_from_features(h, stamp);
  if (out) *out = r;
  return r.sr_status;
}


int chk_get_cloud(void* hv, int which, float* out, int cap) {
  Handle* h = (Handle*)hv;
  const std::vector<float>* v = nullptr;
  switch (which) {
    case LOAM_CLOUD_FULL: v = &h->full; break;
    case LOAM_CLOUD_SHARP: v = &h->c_sharp; break;
    case LOAM_CLOUD_LESS_SHARP: v = &h->c_less_sharp; break;
    case LOAM_CLOUD_FLAT: v = &h->c_flat; break;
    case LOAM_CLOUD_LESS_FLAT: v = &h->c_less_flat; break;
    case LOAM_CLOUD_CORNER_LAST: v = &h->c_corner_last; break;
    case LOAM_CLOUD_SURF_LAST: v = &h->c_surf_last; break;
    case LOAM_CLOUD_FULL_END: v = &h->c_full_end; break;
    case LOAM_CLOUD_REGISTERED: v = &h->c_registered; break;
    case LOAM_CLOUD_SURROUND: v = &h->c_surround; break;
    case LOAM_CLOUD_MAP_CORNER: case LOAM_CLOUD_MAP_SURF: case LOAM_CLOUD_STACK_CORNER: case LOAM_CLOUD_STACK_SURF: {
      int (*gci)(int, float*, int) = (int (*)(int, float*, int))h->lm.sym("node_get_cloud_internal");
      return gci ? gci(which - LOAM_CLOUD_MAP_CORNER, out, cap) : -1;
    }
    default: return -1;
  }
  int n = (int)v->size() / 4;
  memcpy(out, v->data(), sizeof(float) * 4 * (size_t)(n < cap ? n : cap));
  return n;
}

int chk_get_ints(void* hv, int which, int* out, int cap) {
  Handle* h = (Handle*)hv;
  if (which <= LOAM_INTS_SORT_IND) {
    int n = h->n_full < cap ? h->n_full : cap;
    h->sr.get_ints(which, out, n);
    return h->n_full;
  }
  if (which >= LOAM_INTS_CORNER_IND1 && which <= LOAM_INTS_SURF_IND3) {
    int nq = which <= LOAM_INTS_CORNER_IND2 ? h->res.n_sharp : h->res.n_flat;
    h->lo.get_ints(which, out, nq < cap ? nq : cap);
    return nq;
  }
  return -1;  // emission-order index lists are not observable from outside the reference
}

int chk_get_floats(void* hv, int which, float* out, int cap) {
  Handle* h = (Handle*)hv;
  if (which == LOAM_FLOATS_CURVATURE) {
    int n = h->n_full < cap ? h->n_full : cap;
    h->sr.get_floats(0, out, n);
    return h->n_full;
  }
  return -1;
}

void chk_busy_seconds(void* hv, double out[4]) {
  Handle* h = (Handle*)hv;
  out[0] = h->sr.busy(); out[1] = h->lo.busy(); out[2] = h->lm.busy(); out[3] = h->tm.busy();
}

void chk_lo_set_state(void* hv, const float transform[6], const float transformSum[6]) {
  Handle* h = (Handle*)hv;
  void (*f)(const float*, const float*) = (void (*)(const float*, const float*))h->lo.sym("node_lo_set_state");
  if (f) f(transform, transformSum);
}

}  // extern "C"

// ---- the reference as it is deployed: one thread per node, topic queues of depth 2 between them ----------
// (scanRegistration / laserOdometry / laserMapping are separate processes in the reference, "the program
// consumes two cores", README.md:46; transformMaintenance's two callbacks run under one lock)
#include <condition_variable>
#include <chrono>
#include <deque>
#include <mutex>
#include <thread>
namespace {
template <class T> struct Mailbox {
  std::deque<T> q; std::mutex m; std::condition_variable cv; size_t cap = 2; bool closed = false;
  void push(T&& v) { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return q.size() < cap; }); q.push_back(std::move(v)); cv.notify_all(); }
  bool pop(T& v) { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return !q.empty() || closed; }); if (q.empty()) return false; v = std::move(q.front()); q.pop_front(); cv.notify_all(); return true; }
  void close() { std::unique_lock<std::mutex> l(m); closed = true; cv.notify_all(); }
};
struct SixTopics { int k; double stamp; std::vector<float> full, sharp, less_sharp, flat, less_flat, imu_trans; };
struct OdomTopics { int k; double stamp; double odom[14]; bool have_odom, clouds, full3; std::vector<float> corner_last, surf_last, full_end; };
}  // namespace

extern "C" double chk_run_pipelined(void* hv, const float* xyz, const int* counts, const double* stamps, int n_sweeps, float* poses) {
  Handle* h = (Handle*)hv;
  Node &sr = h->sr, &lo = h->lo, &lm = h->lm, &tm = h->tm;
  Mailbox<SixTopics> q1;
  Mailbox<OdomTopics> q2;
  std::mutex tm_mu;
  auto t0 = std::chrono::steady_clock::now();
  std::thread t_sr([&] {
    std::vector<float> scratch((size_t)4 * 400000);
    size_t off = 0;
    for (int k = 0; k < n_sweeps; k++) {
      sr.out_clear();
      sr.post_cloud("/velodyne_points", xyz + off * 3, counts[k], 3, stamps[k]);
      off += (size_t)counts[k];
      sr.cycle();
      if (sr.out_count("/velodyne_cloud_2") <= 0) continue;
      SixTopics m; m.k = k; m.stamp = stamps[k];
      fetch(sr, "/velodyne_cloud_2", m.full, scratch); fetch(sr, "/laser_cloud_sharp", m.sharp, scratch);
      fetch(sr, "/laser_cloud_less_sharp", m.less_sharp, scratch); fetch(sr, "/laser_cloud_flat", m.flat, scratch);
      fetch(sr, "/laser_cloud_less_flat", m.less_flat, scratch);
      std::vector<float> it;
      int ni = fetch(sr, "/imu_trans", it, scratch);
      m.imu_trans.assign(12, 0.f);
      for (int i = 0; i < 4 && i < ni; i++) { m.imu_trans[3 * i] = it[4 * i]; m.imu_trans[3 * i + 1] = it[4 * i + 1]; m.imu_trans[3 * i + 2] = it[4 * i + 2]; }
      q1.push(std::move(m));
    }
    q1.close();
  });
  std::thread t_lo([&] {
    std::vector<float> scratch((size_t)4 * 400000);
    SixTopics m;
    while (q1.pop(m)) {
      lo.out_clear();
      lo.post_cloud("/laser_cloud_sharp", m.sharp.data(), (int)m.sharp.size() / 4, 4, m.stamp);
      lo.post_cloud("/laser_cloud_less_sharp", m.less_sharp.data(), (int)m.less_sharp.size() / 4, 4, m.stamp);
      lo.post_cloud("/laser_cloud_flat", m.flat.data(), (int)m.flat.size() / 4, 4, m.stamp);
      lo.post_cloud("/laser_cloud_less_flat", m.less_flat.data(), (int)m.less_flat.size() / 4, 4, m.stamp);
      lo.post_cloud("/velodyne_cloud_2", m.full.data(), (int)m.full.size() / 4, 4, m.stamp);
      lo.post_cloud("/imu_trans", m.imu_trans.data(), 4, 3, m.stamp);
      lo.cycle();
      OdomTopics o; o.k = m.k; o.stamp = m.stamp;
      o.have_odom = lo.out_count("/laser_odom_to_init") > 0;
      o.clouds = lo.out_count("/laser_cloud_corner_last") > 0;
      o.full3 = lo.out_count("/velodyne_cloud_3") > 0;
      if (!o.have_odom && !o.clouds) continue;
      if (o.have_odom) {
        lo.get_odom("/laser_odom_to_init", o.odom);
        std::lock_guard<std::mutex> l(tm_mu);
        tm.out_clear();
        tm.post_odom("/laser_odom_to_init", o.odom);
        tm.cycle();
        if (tm.get_floats) tm.get_floats(0, poses + 6 * (size_t)m.k, 6);
      }
      if (o.clouds) { fetch(lo, "/laser_cloud_corner_last", o.corner_last, scratch); fetch(lo, "/laser_cloud_surf_last", o.surf_last, scratch); }
      if (o.full3) fetch(lo, "/velodyne_cloud_3", o.full_end, scratch);
      // an odometry-only sweep (clouds not published) still reaches laserMapping, as on the topic graph
      q2.push(std::move(o));
    }
    q2.close();
  });
  std::thread t_lm([&] {
    OdomTopics o;
    while (q2.pop(o)) {
      lm.out_clear();
      if (o.have_odom) lm.post_odom("/laser_odom_to_init", o.odom);
      if (o.clouds) {
        lm.post_cloud("/laser_cloud_corner_last", o.corner_last.data(), (int)o.corner_last.size() / 4, 4, o.stamp);
        lm.post_cloud("/laser_cloud_surf_last", o.surf_last.data(), (int)o.surf_last.size() / 4, 4, o.stamp);
      }
      if (o.full3) lm.post_cloud("/velodyne_cloud_3", o.full_end.data(), (int)o.full_end.size() / 4, 4, o.stamp);
      lm.cycle();
      if (lm.out_count("/aft_mapped_to_init") > 0) {
        double am[14];
        lm.get_odom("/aft_mapped_to_init", am);
        std::lock_guard<std::mutex> l(tm_mu);
        tm.post_odom("/aft_mapped_to_init", am);
        tm.cycle();
      }
    }
  });
  t_sr.join(); t_lo.join(); t_lm.join();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
