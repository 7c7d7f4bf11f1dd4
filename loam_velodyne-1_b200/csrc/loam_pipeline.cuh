// loam_pipeline.cuh -- the four nodes wired as the reference's topic graph on
// one device: sr -> lo -> tm.onOdometry -> (lm -> tm.onMapping).  Like the reference's separate
// ROS nodes, the stages of consecutive sweeps overlap: scanRegistration of sweep k+1 runs on its
// own stream while laserOdometry / laserMapping of sweep k run on the pipeline stream.
//
// All per-sweep state lives in device memory and every kernel takes its sizes from device
// counters, so a sweep is a fixed launch sequence with NO host round trip: push_async() only
// enqueues, a one-thread kernel snapshots the poses/counters of the sweep into a device ring, and
// collect() reads the ring back (one small D2H) and runs transformMaintenance on the host.
#pragma once
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include "loam_lm.cuh"

namespace loamgpu {

// odometry part of a sweep's record (PipeRecord, pipe_record_body: lo_kernels2.cuh) as a launch of its own -- only when
// the odometry pass could not carry it (first sweep, in-kernel tail) ...
__global__ void pipe_record(PipeRecord* out, const SRCounts* cnt, const ImuState* imu, const LOState* lo) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  pipe_record_body(out, cnt, imu, lo);
}
// ... and its mapping part, appended on the mapping stream when the sweep was mapped
// (second mapping side stream, behind the pass's map update: poses and counters come from the snapshot lm_insert_ids
// took, because the next pass is already rewriting the live fields; error flags and arena pointers are read as they are)
__global__ void pipe_record_lm(PipeRecord* out, const LMState* lm, int rec_slot) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const LMState::Rec& R = lm->rec[rec_slot & 1];
  for (int i = 0; i < 6; i++) { out->aft[i] = R.aft[i]; out->bef[i] = R.bef[i]; }
  out->lm_iters = R.iters; out->lm_rows = R.rows; out->lm_deg = R.deg; out->lm_error = lm->error;
  for (int T = 0; T < 2; T++) { out->mapN[T] = R.mapN[T]; out->stackN[T] = R.stackN[T]; out->bump[T] = lm->bump[T]; }
}
// ... and the size of the surround map, on the stream that filters it
__global__ void pipe_record_surround(PipeRecord* out, const LMState* lm) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  out->surroundN = lm->surroundN;
}

struct PipelineHandle {
  static const int RING = 1024;
  SRHandle sr;
  LOHandle lo;
  LMHandle lm;
  TMHost tm;
  cudaStream_t stream = nullptr;      // laserOdometry, pose record
  cudaStream_t sr_stream = nullptr;   // scanRegistration of the NEXT sweep runs beside them (banked outputs, loam_sr.cuh)
  cudaStream_t lm_stream = nullptr;   // laserMapping of sweep k runs beside the odometry of sweeps k+1, k+2 (input banks, loam_lm.cuh)
  cudaEvent_t ev_join[2] = {nullptr, nullptr};
  // result ring in pinned host memory, written by pipe_record / pipe_record_lm directly (no copy
  // engine: a D2H queued behind the mapping stream would stall the records of the next sweeps)
  PinnedBuf<PipeRecord> h_ring;
  struct Pending { int slot; int sr_status, lo_status, lm_status; };
  std::deque<Pending> pending;
  int next_slot = 0;
  loam_sweep_result res;

  bool sr_stream_shared = false;
  // shared_sr_stream: the registration stream of a batch of pipelines (loam_pipeline_create_batch); the handle
  // then does not own it
  int init(const loam_params* p, int dev, cudaStream_t shared_sr_stream = nullptr) {
    CUDA_OK(cudaSetDevice(dev));
    CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    if (shared_sr_stream) { sr_stream = shared_sr_stream; sr_stream_shared = true; }
    else CUDA_OK(cudaStreamCreateWithFlags(&sr_stream, cudaStreamNonBlocking));
    int r = sr.init(p, dev, sr_stream);
    if (r != LOAM_OK) return r;
    r = lo.init(p, dev, stream);
    if (r != LOAM_OK) return r;
    {
      // the mapping chain is the longest of the three: its kernels are scheduled first
      int lo_pri = 0, hi_pri = 0;
      CUDA_OK(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
      CUDA_OK(cudaStreamCreateWithPriority(&lm_stream, cudaStreamNonBlocking, hi_pri));
    }
    for (int i = 0; i < 2; i++) CUDA_OK(cudaEventCreateWithFlags(&ev_join[i], cudaEventDisableTiming));
    r = lm.init(p, dev, lm_stream);
    if (r != LOAM_OK) return r;
    if (h_ring.alloc(RING) != LOAM_OK) return LOAM_ERR_CUDA;
    memset(&res, 0, sizeof(res));
    return LOAM_OK;
  }
  ~PipelineHandle() {
    if (sr.copy_stream) cudaStreamSynchronize(sr.copy_stream);
    if (sr_stream) { cudaStreamSynchronize(sr_stream); if (!sr_stream_shared) cudaStreamDestroy(sr_stream); }
    if (lm.aux) cudaStreamSynchronize(lm.aux);
    if (lm.aux2) cudaStreamSynchronize(lm.aux2);
    if (lm_stream) { cudaStreamSynchronize(lm_stream); cudaStreamDestroy(lm_stream); }
    if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); }
    for (int i = 0; i < 2; i++) if (ev_join[i]) cudaEventDestroy(ev_join[i]);
  }

  // make the pipeline stream wait for everything enqueued so far on the other streams (no host sync)
  int join() {
    CUDA_OK(cudaEventRecord(ev_join[0], sr_stream));
    CUDA_OK(cudaEventRecord(ev_join[1], lm_stream));
    CUDA_OK(cudaStreamWaitEvent(stream, ev_join[0], 0));
    CUDA_OK(cudaStreamWaitEvent(stream, ev_join[1], 0));
    return lm.join_side(stream, ev_join[1]);                 // the map update / surround map of the last mapped sweeps
  }

  int imu(double stamp, const double quat[4], const float acc[3]) {
    sr.on_imu(stamp, quat, acc);
    lm.on_imu(stamp, quat);
    return LOAM_OK;
  }

  // enqueue one sweep.  src_device: xyz is packed device memory and aux = {first xyz, last xyz}
  int push_async(const float* xyz, int n, int stride_bytes, double stamp, bool src_device, const float* aux, int first_idx) {
    if ((int)pending.size() >= RING) { set_error("result ring full: call collect"); return LOAM_ERR_CAPACITY; }
    int rc;
    { HostNode hn(0); rc = src_device ? sr.process_resident(xyz, n, aux, aux + 3, first_idx, stamp) : sr.process(xyz, n, stride_bytes, stamp); }
    return after_registration(rc, stamp);
  }
  // same for a sweep that arrives as raw Velodyne VLP-16 data packets (decoded on the device, pk_decode.cuh)
  int push_packets_async(const unsigned char* packets, int n_packets, double stamp) {
    if ((int)pending.size() >= RING) { set_error("result ring full: call collect"); return LOAM_ERR_CAPACITY; }
    return after_registration(sr.process_packets(packets, n_packets, stamp), stamp);
  }
  int after_registration(int rc, double stamp) {
    Pending pd;
    pd.slot = next_slot;
    pd.sr_status = pd.lo_status = pd.lm_status = LOAM_NOT_READY;
    if (rc < 0) return rc;
    if (g_profile == 3) prof_node("scanRegistration", sr_stream, false);
    if (rc == LOAM_OK) {
      pd.sr_status = LOAM_OK;
      bool pub = false;
      long long t_lo = now_ns();
      lo.record_out = h_ring.p + pd.slot;
      lo.record_written = false;
      int lrc = lo.process(sr, &pub);
      lo.record_out = nullptr;
      if (lrc < 0) return lrc;
      if (g_profile == 3) prof_node("laserOdometry", stream, false);
      pd.lo_status = lrc;
      // record + release of the scanRegistration bank BEFORE the mapping is enqueued: the next
      // sweeps' registration / odometry do not wait for it
      if (!lo.record_written) LAUNCH("pipe_record", pipe_record, 1, 32, 0, stream, h_ring.p + pd.slot, sr.d_cnt.p, sr.imu_state_dev(), lo.d_state.p);
      int rr = sr.release_bank(stream);
      if (rr < 0) return rr;
      g_host_node_ns[1] += now_ns() - t_lo;
      HostNode hn(2);
      if (lrc == LOAM_OK && pub && !(lo.P.dev_flags & 1024)) {
        int mrc = lm.process(lo, stamp);
        if (mrc < 0) return mrc;
        if (g_profile == 3) prof_node("laserMapping", lm_stream, false);
        pd.lm_status = mrc;
      }
      if (pd.lm_status == LOAM_OK) {
        // behind the pass's map update, but on the second side stream: the next pass's window shift + index refresh (first
        // side stream) do not queue behind a one-thread kernel
        CUDA_OK(cudaEventRecord(lm.ev_upd, lm.aux));
        CUDA_OK(cudaStreamWaitEvent(lm.aux2, lm.ev_upd, 0));
        LAUNCH("pipe_record_lm", pipe_record_lm, 1, 32, 0, lm.aux2, h_ring.p + pd.slot, lm.d_state.p, lm.last_rec_slot);
        CUDA_OK(cudaEventRecord(lm.ev_rec[lm.last_rec_slot], lm.aux2));
        if (lm.last_surround) LAUNCH("pipe_record_surround", pipe_record_surround, 1, 32, 0, lm.aux2, h_ring.p + pd.slot, lm.d_state.p);
      }
    }
    next_slot = (next_slot + 1) % RING;
    pending.push_back(pd);
    return rc;
  }

  // wait for everything enqueued so far; fills out[0..*n) in push order (transformMaintenance runs here)
  int collect(loam_sweep_result* out, int cap, int* n) {
    CUDA_OK(cudaStreamSynchronize(stream));
    CUDA_OK(cudaStreamSynchronize(lm_stream));
    { int sr_ = lm.sync_side(); if (sr_ != LOAM_OK) return sr_; }
    int k = 0;
    while (!pending.empty()) {
      Pending pd = pending.front();
      pending.pop_front();
      loam_sweep_result& r = res;
      r.sr_status = pd.sr_status; r.lo_status = pd.lo_status; r.lm_status = pd.lm_status;
      if (pd.sr_status == LOAM_OK) {
        const PipeRecord& q = h_ring.p[pd.slot];
        if (q.cnt.error) { describe_sr_error(q.cnt.error); return LOAM_ERR_CAPACITY; }
        sr.last_counts = q.cnt;
        lo.last_n_sharp = q.cnt.n_sharp; lo.last_n_flat = q.cnt.n_flat;
        r.n_full = q.cnt.n_full; r.n_sharp = q.cnt.n_sharp; r.n_less_sharp = q.cnt.n_less_sharp;
        r.n_flat = q.cnt.n_flat; r.n_less_flat = q.cnt.n_less_flat;
        memcpy(r.imu_trans, q.imu_trans, sizeof(r.imu_trans));
        r.lo_iters = q.lo_iters; r.lo_rows = q.lo_rows; r.lo_degenerate = q.lo_deg;
        memcpy(r.transform, q.transform, sizeof(float) * 6);
        memcpy(r.transform_sum, q.transformSum, sizeof(float) * 6);
        memcpy(lo.last_state.transform, q.transform, sizeof(float) * 6);
        memcpy(lo.last_state.transformSum, q.transformSum, sizeof(float) * 6);
        lo.last_state.iters = q.lo_iters; lo.last_state.rows = q.lo_rows; lo.last_state.isDegenerate = q.lo_deg;
        lo.last_state.n_corner_last = q.cnt.n_less_sharp; lo.last_state.n_surf_last = q.cnt.n_less_flat;
        if (pd.lm_status == LOAM_OK) lo.last_state.n_full_end = q.cnt.n_full;
        if (pd.lo_status == LOAM_OK) {
          tm.on_odometry(q.transformSum, r.transform_mapped);
          if (pd.lm_status == LOAM_OK) {
            if (q.lm_error) { describe_lm_error(q.lm_error); return LOAM_ERR_CAPACITY; }
            r.lm_iters = q.lm_iters; r.lm_rows = q.lm_rows; r.lm_degenerate = q.lm_deg;
            r.n_map_corner = q.mapN[0]; r.n_map_surf = q.mapN[1]; r.n_stack_corner = q.stackN[0]; r.n_stack_surf = q.stackN[1];
            memcpy(r.transform_aft_mapped, q.aft, sizeof(float) * 6);
            memcpy(r.transform_bef_mapped, q.bef, sizeof(float) * 6);
            LMState& m = lm.last_state;
            memcpy(m.transformAftMapped, q.aft, sizeof(float) * 6);
            memcpy(m.transformBefMapped, q.bef, sizeof(float) * 6);
            m.iters = q.lm_iters; m.rows = q.lm_rows; m.isDegenerate = q.lm_deg;
            m.mapN[0] = q.mapN[0]; m.mapN[1] = q.mapN[1]; m.stackN[0] = q.stackN[0]; m.stackN[1] = q.stackN[1];
            m.bump[0] = q.bump[0]; m.bump[1] = q.bump[1]; m.registeredN = q.cnt.n_full;
            if (q.surroundN >= 0) m.surroundN = q.surroundN;
            tm.on_mapping(q.aft, q.bef);
          }
        }
      }
      if (out && k < cap) out[k] = r;
      k++;
    }
    if (n) *n = k;
    return LOAM_OK;
  }

  int push_packets(const unsigned char* packets, int n_packets, double stamp, loam_sweep_result* out) {
    int rc = push_packets_async(packets, n_packets, stamp);
    if (rc < 0) return rc;
    int got = 0;
    int c = collect(nullptr, 0, &got);
    if (c < 0) return c;
    if (out) *out = res;
    return res.sr_status;
  }
  int push(const float* xyz, int n, int stride_bytes, double stamp, loam_sweep_result* out) {
    int rc = push_async(xyz, n, stride_bytes, stamp, false, nullptr, 0);
    if (rc < 0) return rc;
    int got = 0;
    int c = collect(nullptr, 0, &got);
    if (c < 0) return c;
    if (out) *out = res;
    return res.sr_status;
  }
};

// A batch of pipelines of one GPU whose scanRegistration passes are issued as ONE launch chain per step
// (BASELINE.json configs[4]); laserOdometry / laserMapping stay per sequence on their own streams.  The
// per-sequence host work (staging, the odometry / mapping launches) is spread over a few worker threads of
// the library: the handles are independent, only the registration chain is common.
struct BatchWorkers {
  std::vector<std::thread> th;
  std::mutex mu;
  std::condition_variable cv, cv_done;
  std::function<void(int)> job;
  int gen = 0, pending = 0, T = 0, device = 0;
  bool stop = false;
  void start(int threads, int dev) {
    T = threads; device = dev;
    for (int t = 0; t < T; t++) th.emplace_back([this, t] { loop(t); });
  }
  void loop(int t) {
    cudaSetDevice(device);
    int seen = 0;
    for (;;) {
      std::function<void(int)> f;
      {
        std::unique_lock<std::mutex> l(mu);
        cv.wait(l, [&] { return stop || gen != seen; });
        if (stop) return;
        seen = gen;
        f = job;
      }
      f(t);
      {
        std::unique_lock<std::mutex> l(mu);
        if (--pending == 0) cv_done.notify_all();
      }
    }
  }
  void run(const std::function<void(int)>& f) {
    std::unique_lock<std::mutex> l(mu);
    job = f; gen++; pending = T;
    cv.notify_all();
    cv_done.wait(l, [&] { return pending == 0; });
  }
  ~BatchWorkers() {
    { std::unique_lock<std::mutex> l(mu); stop = true; }
    cv.notify_all();
    for (auto& t : th) t.join();
  }
};

struct PipelineBatch {
  std::vector<PipelineHandle*> pipes;
  SRBatch srb;
  BatchWorkers workers;
  cudaStream_t sr_stream = nullptr;
  int init(const loam_params* p, int dev, int count) {
    CUDA_OK(cudaSetDevice(dev));
    CUDA_OK(cudaStreamCreateWithFlags(&sr_stream, cudaStreamNonBlocking));
    for (int i = 0; i < count; i++) {
      PipelineHandle* h = new PipelineHandle();
      pipes.push_back(h);
      int r = h->init(p, dev, sr_stream);
      if (r != LOAM_OK) return r;
    }
    int r = srb.init(count);
    if (r != LOAM_OK) return r;
    workers.start(count < 8 ? count : 8, dev);
    return SRBatch::set_attributes();
  }
  ~PipelineBatch() {
    for (PipelineHandle* h : pipes) delete h;
    if (sr_stream) { cudaStreamSynchronize(sr_stream); cudaStreamDestroy(sr_stream); }
  }
  // one sweep of every sequence; enqueue only
  // first_last / first_index != nullptr: the sweeps are packed xyz in DEVICE memory (loam_pipeline_push_device's form)
  int push_async(const float* const* xyz, const int* n, int stride_bytes, const double* stamps,
                 const float* const* first_last = nullptr, const int* first_index = nullptr) {
    const int B = (int)pipes.size(), T = workers.T;
    std::vector<SRHandle*> hs(B);
    std::vector<int> rcs(B, LOAM_NOT_READY), rcs2(B, LOAM_OK);
    std::vector<std::string> errs(T);
    for (int i = 0; i < B; i++) {
      if ((int)pipes[i]->pending.size() >= PipelineHandle::RING) { set_error("result ring full: call collect"); return LOAM_ERR_CAPACITY; }
      hs[i] = &pipes[i]->sr;
    }
    long long t0 = now_ns();
    int r = srb.begin();
    if (r < 0) return r;
    workers.run([&](int t) {
      for (int i = t; i < B; i += T) {
        rcs[i] = first_last ? srb.prepare_device(hs[i], i, xyz[i], n[i], first_last[i], first_index[i], stamps[i])
                            : srb.prepare(hs[i], i, xyz[i], n[i], stride_bytes, stamps[i]);
        if (rcs[i] < 0) errs[t] = loam_last_error();
      }
    });
    r = srb.launch(hs.data(), n, rcs.data());
    g_host_node_ns[0] += now_ns() - t0;
    if (r < 0) { for (auto& e : errs) if (!e.empty()) set_error("%s", e.c_str()); return r; }
    workers.run([&](int t) {
      for (int i = t; i < B; i += T) {
        rcs2[i] = pipes[i]->after_registration(rcs[i], stamps[i]);
        if (rcs2[i] < 0) errs[t] = loam_last_error();
      }
    });
    for (int i = 0; i < B; i++)
      if (rcs2[i] < 0) { for (auto& e : errs) if (!e.empty()) set_error("%s", e.c_str()); return rcs2[i]; }
    return LOAM_OK;
  }
};

}  // namespace loamgpu
