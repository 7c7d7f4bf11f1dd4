// loam_sr.cuh -- host side of the scanRegistration node: owns the device buffers, keeps the IMU
// ring buffer (sequential, host; scanRegistration.cpp:638-660,173-209) and enqueues the kernels.
#pragma once
#include <limits.h>
#include <vector>
#include "sr_batch.cuh"
#include "pose_math.cuh"
#include "pk_decode.cuh"

namespace loamgpu {

struct SRHandle {
  loam_params P;
  SRParamsDev Pd;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int cap = 0;       // max_points
  int nblk_cap = 0;
  // system delay (:213-219)
  int systemInitCount = 0;
  bool systemInited = false;
  // device buffers
  // incoming sweeps: RAW_DEPTH device/pinned buffer pairs filled by a separate copy stream, so the
  // H2D transfer of sweep k+1 overlaps the kernels of sweep k (ev_copied: copy -> compute,
  // ev_free: compute -> copy, both also guard the reuse of the pinned staging buffer)
  static const int RAW_DEPTH = 4;
  DevBuf<float> d_raw2[RAW_DEPTH];
  DevBuf<float> d_decoded;             // xyz of a sweep that arrived as raw packets (pk_decode.cuh)
  float* d_raw_cur = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_copied[RAW_DEPTH] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_free[RAW_DEPTH] = {nullptr, nullptr, nullptr, nullptr};
  int raw_sel = 0;
  DevBuf<short> d_ring;
  DevBuf<float> d_ori, d_reltime, d_curv;
  DevBuf<int> d_hist, d_ring_start, d_scan_start, d_scan_end, d_pos, d_sort_ind, d_ring_tot;
  DevBuf<unsigned int> d_scan_ticket;
  DevBuf<unsigned char> d_picked, d_gap, d_steps;
  DevBuf<signed char> d_label;
  DevBuf<int> d_sharp_ring, d_less_sharp_ring, d_flat_ring, d_blockmax;
  DevBuf<int> d_sharp_idx, d_less_sharp_idx, d_flat_idx;
  DevBuf<float4> d_lessflat_tmp;
  // What the node publishes (the five clouds :584-612, their counters, per-ring counts) lives in
  // two banks, so that sweep k+1 can be registered on this node's stream while the consumers of
  // sweep k (laserOdometry, the pose record) still read bank k&1 on another stream.  d_full ...
  // d_cnt are views of the bank of the most recent sweep.
  struct Bank {
    DevBuf<float4> full, sharp, less_sharp, flat, less_flat;
    DevBuf<SRCounts> cnt;
    DevBuf<int> ring_counts;
    cudaEvent_t released = nullptr;   // recorded by the consumer when it is done with the bank
  } bank[2];
  int cur_bank = 0;
  template <typename T> struct View { T* p = nullptr; };
  View<float4> d_full, d_sharp, d_less_sharp, d_flat, d_less_flat;
  View<SRCounts> d_cnt;
  View<int> d_ring_counts;
  void select_bank(int b) {
    cur_bank = b;
    Bank& k = bank[b];
    d_full.p = k.full.p; d_sharp.p = k.sharp.p; d_less_sharp.p = k.less_sharp.p; d_flat.p = k.flat.p; d_less_flat.p = k.less_flat.p;
    d_cnt.p = k.cnt.p; d_ring_counts.p = k.ring_counts.p;
  }
  // called by the consumer of the current bank after its last read has been enqueued on `s`
  int release_bank(cudaStream_t s) {
    if (s != stream) CUDA_OK(cudaEventRecord(bank[cur_bank].released, s));
    return LOAM_OK;
  }
  DevBuf<ImuSnapshot> d_imu;
  DevBuf<ImuState> d_imu_state;  // [2]
  PinnedBuf<float> h_raw2[RAW_DEPTH];
  PinnedBuf<SRCounts> h_cnt;
  PinnedRing<ImuSnapshot, 4> h_imu;
  PinnedRing<SRCounts, 4> h_cnt_init;
  PinnedBuf<ImuState> h_imu_state;
  int imu_parity = 0;
  // host IMU ring buffer
  ImuSnapshot imu;
  float imuAcc[3][IMU_Q];
  bool has_imu = false;
  int last_n = 0;
  double last_stamp = 0;
  SRCounts last_counts;
  bool processed = false;

  // carrier_only: the handle will only ever hold features handed in from the host (set_features: laserOdometry fed from
  // host copies of the six topics).  It then needs the two output banks, the counters and the IMU state -- not the raw
  // staging ring (4 x 12 B x max_points pinned + device) nor the per-point scratch of the registration kernels.
  bool carrier_only = false;
  int init(const loam_params* p, int dev, cudaStream_t s, bool carrier = false) {
    P = *p;
    device = dev;
    carrier_only = carrier;
    if (P.n_scans < 1 || P.n_scans > SR_MAX_RINGS || P.max_points < 64) {
      set_error("n_scans = %d / max_points = %d: n_scans must be 1 .. %d (LOAM_LIMIT_RINGS), max_points at least 64", P.n_scans, P.max_points, SR_MAX_RINGS);
      return LOAM_ERR_ARG;
    }
    CUDA_OK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) { set_error("device %d is sm_%d%d; this library is built for sm_100a only", dev, prop.major, prop.minor); return LOAM_ERR_CUDA; }
    if (s) stream = s; else { CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)); own_stream = true; }
    Pd.n_scans = P.n_scans; Pd.ring_model = P.ring_model; Pd.ring_lower = P.ring_lower_deg; Pd.ring_upper = P.ring_upper_deg;
    Pd.max_points = P.max_points;
    cap = P.max_points;
    nblk_cap = (cap + SR_BLK - 1) / SR_BLK;
    int R = P.n_scans;
    int r = LOAM_OK;
    const size_t pc = carrier ? 16 : (size_t)cap;          // per-point scratch of the registration kernels
    const size_t pb = carrier ? 1 : (size_t)nblk_cap;
    for (int b = 0; b < RAW_DEPTH; b++) { r |= d_raw2[b].alloc(pc * 3); r |= h_raw2[b].alloc(pc * 3); }
    r |= d_ring.alloc(pc); r |= d_ori.alloc(pc); r |= d_reltime.alloc(pc);
    r |= d_curv.alloc_zero(pc, stream); r |= d_hist.alloc((size_t)R * pb); r |= d_ring_start.alloc(R + 1);
    r |= d_scan_start.alloc(R); r |= d_scan_end.alloc(R); r |= d_ring_tot.alloc_zero(SR_MAX_RINGS, stream); r |= d_scan_ticket.alloc_zero(1, stream); r |= d_pos.alloc(pc); r |= d_sort_ind.alloc_zero(pc, stream);
    r |= d_picked.alloc_zero(pc + 16, stream); r |= d_gap.alloc_zero(pc + 16, stream); r |= d_steps.alloc(pc);
    r |= d_label.alloc_zero(pc, stream);
    r |= d_sharp_ring.alloc((size_t)R * SR_SHARP_PER_RING); r |= d_less_sharp_ring.alloc((size_t)R * SR_LESS_SHARP_PER_RING);
    r |= d_flat_ring.alloc((size_t)R * SR_FLAT_PER_RING);
    r |= d_blockmax.alloc(pb);
    r |= d_sharp_idx.alloc((size_t)R * SR_SHARP_PER_RING); r |= d_less_sharp_idx.alloc((size_t)R * SR_LESS_SHARP_PER_RING);
    r |= d_flat_idx.alloc((size_t)R * SR_FLAT_PER_RING);
    for (int b = 0; b < 2; b++) {
      Bank& k = bank[b];
      r |= k.full.alloc(cap); r |= k.sharp.alloc((size_t)R * SR_SHARP_PER_RING); r |= k.less_sharp.alloc((size_t)R * SR_LESS_SHARP_PER_RING);
      r |= k.flat.alloc((size_t)R * SR_FLAT_PER_RING); r |= k.less_flat.alloc(cap);
      r |= k.cnt.alloc_zero(1, stream); r |= k.ring_counts.alloc_zero((size_t)R * 4, stream);
      if (cudaEventCreateWithFlags(&k.released, cudaEventDisableTiming) != cudaSuccess) r = LOAM_ERR_CUDA;
    }
    select_bank(0);
    r |= d_lessflat_tmp.alloc(carrier ? 16 : (size_t)R * SR_RING_CAP);
    r |= d_imu.alloc(1); r |= d_imu_state.alloc_zero(2, stream);
    r |= h_cnt.alloc(1); r |= h_imu.alloc(); r |= h_cnt_init.alloc(); r |= h_imu_state.alloc(1);
    if (r != LOAM_OK) return LOAM_ERR_CUDA;
    CUDA_OK(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
    for (int b = 0; b < RAW_DEPTH; b++) {
      CUDA_OK(cudaEventCreateWithFlags(&ev_copied[b], cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_free[b], cudaEventDisableTiming));
    }
    memset(&imu, 0, sizeof(imu));
    memset(imuAcc, 0, sizeof(imuAcc));
    imu.pointer_last = -1;
    memset(&last_counts, 0, sizeof(last_counts));
    memset(h_imu_state.p, 0, sizeof(ImuState));
    CUDA_OK(cudaFuncSetAttribute(sr_select, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SelShared)));
    CUDA_OK(cudaFuncSetAttribute(sr_lessflat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LessFlatSmem)));
    CUDA_OK(cudaStreamSynchronize(stream));
    return LOAM_OK;
  }
  ~SRHandle() {
    if (copy_stream) { cudaStreamSynchronize(copy_stream); cudaStreamDestroy(copy_stream); }
    for (int b = 0; b < 2; b++) if (bank[b].released) cudaEventDestroy(bank[b].released);
    for (int b = 0; b < RAW_DEPTH; b++) {
      if (ev_copied[b]) cudaEventDestroy(ev_copied[b]);
      if (ev_free[b]) cudaEventDestroy(ev_free[b]);
    }
    if (own_stream && stream) cudaStreamDestroy(stream);
  }

  // imuHandler :638-660 + AccumulateIMUShift :173-209 (host, sequential)
  int on_imu(double stamp, const double quat[4], const float acc[3]) {
    double roll, pitch, yaw;
    pm::rpy_from_quat(quat[0], quat[1], quat[2], quat[3], roll, pitch, yaw);
    double lax = acc[0], lay = acc[1], laz = acc[2];
    float accX = lay - dm::sin(roll) * dm::cos(pitch) * 9.81;
    float accY = laz - dm::cos(roll) * dm::cos(pitch) * 9.81;
    float accZ = lax + dm::sin(pitch) * 9.81;
    int last = (imu.pointer_last + 1) % IMU_Q;
    imu.pointer_last = last;
    imu.time[last] = stamp;
    imu.roll[last] = roll; imu.pitch[last] = pitch; imu.yaw[last] = yaw;
    imuAcc[0][last] = accX; imuAcc[1][last] = accY; imuAcc[2][last] = accZ;
    {
      float r_ = imu.roll[last], p_ = imu.pitch[last], y_ = imu.yaw[last];
      float aX = imuAcc[0][last], aY = imuAcc[1][last], aZ = imuAcc[2][last];
      float x1 = dm::cos(r_) * aX - dm::sin(r_) * aY;
      float y1 = dm::sin(r_) * aX + dm::cos(r_) * aY;
      float z1 = aZ;
      float x2 = x1;
      float y2 = dm::cos(p_) * y1 - dm::sin(p_) * z1;
      float z2 = dm::sin(p_) * y1 + dm::cos(p_) * z1;
      aX = dm::cos(y_) * x2 + dm::sin(y_) * z2;
      aY = y2;
      aZ = -dm::sin(y_) * x2 + dm::cos(y_) * z2;
      int back = (last + IMU_Q - 1) % IMU_Q;
      double timeDiff = imu.time[last] - imu.time[back];
      const double scanPeriod = 0.1;
      if (timeDiff < scanPeriod) {
        imu.shift[0][last] = imu.shift[0][back] + imu.velo[0][back] * timeDiff + aX * timeDiff * timeDiff / 2;
        imu.shift[1][last] = imu.shift[1][back] + imu.velo[1][back] * timeDiff + aY * timeDiff * timeDiff / 2;
        imu.shift[2][last] = imu.shift[2][back] + imu.velo[2][back] * timeDiff + aZ * timeDiff * timeDiff / 2;
        imu.velo[0][last] = imu.velo[0][back] + aX * timeDiff;
        imu.velo[1][last] = imu.velo[1][back] + aY * timeDiff;
        imu.velo[2][last] = imu.velo[2][back] + aZ * timeDiff;
      }
    }
    has_imu = true;
    return LOAM_OK;
  }

  // laserCloudHandler :211-636.  Enqueues everything on `stream`; no host synchronisation.
  // batch_item != nullptr: do everything up to (and excluding) the kernel launches and describe the pass in
  // *batch_item; the caller launches the batched kernels for all sequences at once (SRBatch below)
  int process(const float* xyz, int n, int stride_bytes, double stamp, SRDev* batch_item = nullptr) {
    if (carrier_only) { set_error("this handle only carries host-side features (no registration scratch)"); return LOAM_ERR_ARG; }
    processed = false;
    if (!systemInited) {  // :213-219
      systemInitCount++;
      if (systemInitCount >= P.system_delay) systemInited = true;
      return LOAM_NOT_READY;
    }
    if (n <= 0 || !xyz || stride_bytes < 12 || (stride_bytes & 3)) { set_error("bad sweep"); return LOAM_ERR_ARG; }
    if (n > cap) { set_error("sweep of %d points exceeds max_points=%d", n, cap); return LOAM_ERR_CAPACITY; }
    CUDA_OK(cudaSetDevice(device));
    // A CUDA-pinned, packed (stride 12) caller buffer is read by the DMA engine in place (the caller
    // keeps it unchanged until collect() / the synchronous call returns); anything else is staged
    // through one of RAW_DEPTH pinned buffers.
    const int b = raw_sel;
    raw_sel = (raw_sel + 1) % RAW_DEPTH;
    const int sf = stride_bytes / 4;
    const float* hr = nullptr;
    if (sf == 3) {
      cudaPointerAttributes at;
      if (cudaPointerGetAttributes(&at, xyz) == cudaSuccess && at.type == cudaMemoryTypeHost) hr = xyz;
      else cudaGetLastError();
    }
    if (!hr) {
      { HostWait hw; CUDA_OK(cudaEventSynchronize(ev_copied[b])); }   // previous transfer out of this staging buffer is done
      float* st = h_raw2[b].p;
      if (sf == 3) memcpy(st, xyz, sizeof(float) * 3 * (size_t)n);
      else
        for (int i = 0; i < n; i++) {
          const float* p = xyz + (size_t)i * sf;
          st[3 * i] = p[0]; st[3 * i + 1] = p[1]; st[3 * i + 2] = p[2];
        }
      hr = st;
    }
    int first = -1, lastf = -1;
    for (int i = 0; i < n; i++) {
      const float* p = hr + 3 * i;
      if (isfinite(p[0]) && isfinite(p[1]) && isfinite(p[2])) { first = i; break; }
    }
    for (int i = n - 1; i >= 0; i--) {
      const float* p = hr + 3 * i;
      if (isfinite(p[0]) && isfinite(p[1]) && isfinite(p[2])) { lastf = i; break; }
    }
    if (first < 0) return LOAM_NOT_READY;
    float startOri, endOri;
    sweep_orientation(hr + 3 * first, hr + 3 * lastf, startOri, endOri);
    d_raw_cur = d_raw2[b].p;
    CUDA_OK(cudaStreamWaitEvent(copy_stream, ev_free[b], 0));   // kernels that read d_raw2[b] RAW_DEPTH sweeps ago
    CUDA_OK(cudaMemcpyAsync(d_raw_cur, hr, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice, copy_stream));
    CUDA_OK(cudaEventRecord(ev_copied[b], copy_stream));
    CUDA_OK(cudaStreamWaitEvent(stream, ev_copied[b], 0));
    if (batch_item) {
      if (has_imu && imu.pointer_last >= 0) { set_error("batched scanRegistration does not deskew with IMU data (use the per-sequence entry points)"); return LOAM_ERR_ARG; }
      batch_ev_free = ev_free[b];
      return prepare_batch_item(n, startOri, endOri, stamp, batch_item);
    }
    int rc = process_device(n, first, startOri, endOri, stamp);
    CUDA_OK(cudaEventRecord(ev_free[b], stream));
    return rc;
  }
  cudaEvent_t batch_ev_free = nullptr;
  int prepare_batch_item(int n, float startOri, float endOri, double stamp, SRDev* d) {
    select_bank(cur_bank ^ 1);
    CUDA_OK(cudaStreamWaitEvent(stream, bank[cur_bank].released, 0));   // consumers of the sweep before last
    d->raw = d_raw_cur; d->n = n; d->nblk = (n + SR_BLK - 1) / SR_BLK; d->startOri = startOri; d->endOri = endOri; d->meta = 0;
    d->ring = d_ring.p; d->ori = d_ori.p; d->hist = d_hist.p; d->ring_start = d_ring_start.p; d->scan_start = d_scan_start.p; d->scan_end = d_scan_end.p;
    d->cnt = d_cnt.p; d->ring_tot = d_ring_tot.p; d->ticket = d_scan_ticket.p;
    d->full = d_full.p; d->pos = d_pos.p; d->reltime = d_reltime.p; d->curv = d_curv.p; d->sort_ind = d_sort_ind.p;
    d->picked = d_picked.p; d->label = d_label.p; d->gap = d_gap.p;
    d->sharp_ring = d_sharp_ring.p; d->less_sharp_ring = d_less_sharp_ring.p; d->flat_ring = d_flat_ring.p; d->ring_counts = d_ring_counts.p;
    d->lessflat_tmp = d_lessflat_tmp.p; d->sharp = d_sharp.p; d->less_sharp = d_less_sharp.p; d->flat = d_flat.p; d->less_flat = d_less_flat.p;
    d->sharp_idx = d_sharp_idx.p; d->less_sharp_idx = d_less_sharp_idx.p; d->flat_idx = d_flat_idx.p;
    last_n = n;
    last_stamp = stamp;
    processed = true;
    return LOAM_OK;
  }

  // raw-packet front end: `n_packets` Velodyne data packets (1206 B each, HOST memory) holding one revolution;
  // VLP-16 format for a 16-ring handle, HDL-64E format for a 64-ring one.  The bytes go through the same
  // staging ring / copy stream as an xyz sweep.  packet_metadata = 0: pk_decode writes the xyz cloud and the
  // unchanged node follows; 1: pk_ring_meta also delivers ring ids and relTime from the packet fields, and
  // sr_ring_ori (atan / atan2 per point, halfPassed) is not run at all.
  bool packet_sensor(pk::SensorDesc* S) const {
    S->n_scans = P.n_scans; S->lower_deg = P.ring_lower_deg; S->upper_deg = P.ring_upper_deg;
    if (P.n_scans == 16 && P.ring_model == 0) { S->kind = pk::VLP16; return true; }
    if (P.n_scans == 64 && P.ring_model == 1) { S->kind = pk::HDL64E; return true; }
    return false;
  }
  int process_packets(const unsigned char* packets, int n_packets, double stamp) {
    if (carrier_only) { set_error("this handle only carries host-side features (no registration scratch)"); return LOAM_ERR_ARG; }
    processed = false;
    pk::SensorDesc S;
    if (!packet_sensor(&S)) { set_error("no packet format for n_scans=%d ring_model=%d (VLP-16: 16 / 0, HDL-64E: 64 / 1)", P.n_scans, P.ring_model); return LOAM_ERR_ARG; }
    if (!systemInited) {  // :213-219
      systemInitCount++;
      if (systemInitCount >= P.system_delay) systemInited = true;
      return LOAM_NOT_READY;
    }
    if (n_packets <= 0 || !packets) { set_error("bad packet buffer"); return LOAM_ERR_ARG; }
    const long long n_ll = (long long)n_packets * pk::POINTS_PER_PACKET;
    if (n_ll > cap) { set_error("%d packets = %lld points exceed max_points=%d", n_packets, n_ll, cap); return LOAM_ERR_CAPACITY; }
    const int n = (int)n_ll;
    const size_t bytes = (size_t)n_packets * pk::PACKET_BYTES;
    CUDA_OK(cudaSetDevice(device));
    if (!d_decoded.p && d_decoded.alloc((size_t)cap * 3) != LOAM_OK) return LOAM_ERR_CUDA;
    const int b = raw_sel;
    raw_sel = (raw_sel + 1) % RAW_DEPTH;
    CUDA_OK(cudaEventSynchronize(ev_copied[b]));
    unsigned char* st = (unsigned char*)h_raw2[b].p;
    memcpy(st, packets, bytes);
    // first / last return of the revolution (scanRegistration.cpp:230-232 reads them on the host)
    int first = -1, lastf = -1;
    pk::Slot s0, sl;
    for (int i = 0; i < n && first < 0; i++) {
      s0 = pk::decode_slot(st + (size_t)(i / pk::POINTS_PER_PACKET) * pk::PACKET_BYTES, i % pk::POINTS_PER_PACKET, S);
      if (s0.valid) first = i;
    }
    for (int i = n - 1; i >= 0 && lastf < 0; i--) {
      sl = pk::decode_slot(st + (size_t)(i / pk::POINTS_PER_PACKET) * pk::PACKET_BYTES, i % pk::POINTS_PER_PACKET, S);
      if (sl.valid) lastf = i;
    }
    if (first < 0) return LOAM_NOT_READY;
    float startOri, endOri;
    sweep_orientation(s0.xyz, sl.xyz, startOri, endOri);
    unsigned char* d_pk = (unsigned char*)d_raw2[b].p;
    CUDA_OK(cudaStreamWaitEvent(copy_stream, ev_free[b], 0));
    CUDA_OK(cudaMemcpyAsync(d_pk, st, bytes, cudaMemcpyHostToDevice, copy_stream));
    CUDA_OK(cudaEventRecord(ev_copied[b], copy_stream));
    CUDA_OK(cudaStreamWaitEvent(stream, ev_copied[b], 0));
    d_raw_cur = d_decoded.p;
    if (P.packet_metadata) {
      pk::MetaSpan M;
      M.first = first; M.last = lastf; M.az0 = s0.az; M.span = 0;
      M.span = pk::meta_daz(lastf, sl.az, M);
      meta_packets = d_pk; meta_n_packets = n_packets; meta_sensor = S; meta_span = M;
      int rc = process_device(n, first, startOri, endOri, stamp, true);
      CUDA_OK(cudaEventRecord(ev_free[b], stream));   // the packet bytes are consumed by pk_ring_meta
      return rc;
    }
    LAUNCH("pk_decode", pk_decode, (n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184, 256, 0, stream, d_pk, n_packets, S, d_decoded.p);
    CUDA_OK(cudaEventRecord(ev_free[b], stream));   // the packet bytes are consumed; d_decoded is ordered on `stream`
    return process_device(n, first, startOri, endOri, stamp);
  }
  const unsigned char* meta_packets = nullptr;
  int meta_n_packets = 0;
  pk::SensorDesc meta_sensor;
  pk::MetaSpan meta_span;

  // :230-238
  static void sweep_orientation(const float* p0, const float* pl, float& startOri, float& endOri) {
    startOri = -dm::atan2(p0[1], p0[0]);
    endOri = -dm::atan2(pl[1], pl[0]) + 2 * M_PI;
    if (endOri - startOri > 3 * M_PI) {
      endOri -= 2 * M_PI;
    } else if (endOri - startOri < M_PI) {
      endOri += 2 * M_PI;
    }
  }

  // same, for a sweep that is already resident in device memory (packed xyz).  first_xyz/last_xyz:
  // host copies of the first / last finite point, first = index of the first finite point.
  int process_resident(const float* d_xyz, int n, const float* first_xyz, const float* last_xyz, int first, double stamp, SRDev* batch_item = nullptr) {
    if (carrier_only) { set_error("this handle only carries host-side features (no registration scratch)"); return LOAM_ERR_ARG; }
    processed = false;
    if (!systemInited) {
      systemInitCount++;
      if (systemInitCount >= P.system_delay) systemInited = true;
      return LOAM_NOT_READY;
    }
    if (n <= 0 || !d_xyz) { set_error("bad sweep"); return LOAM_ERR_ARG; }
    if (n > cap) { set_error("sweep of %d points exceeds max_points=%d", n, cap); return LOAM_ERR_CAPACITY; }
    CUDA_OK(cudaSetDevice(device));
    float startOri, endOri;
    sweep_orientation(first_xyz, last_xyz, startOri, endOri);
    d_raw_cur = const_cast<float*>(d_xyz);
    if (batch_item) {
      if (has_imu && imu.pointer_last >= 0) { set_error("batched scanRegistration does not deskew with IMU data (use the per-sequence entry points)"); return LOAM_ERR_ARG; }
      batch_ev_free = nullptr;
      return prepare_batch_item(n, startOri, endOri, stamp, batch_item);
    }
    return process_device(n, first, startOri, endOri, stamp);
  }

  int process_device(int n, int first, float startOri, float endOri, double stamp, bool from_packet_metadata = false) {
    const int R = P.n_scans;
    const int nblk = (n + SR_BLK - 1) / SR_BLK;
    select_bank(cur_bank ^ 1);
    CUDA_OK(cudaStreamWaitEvent(stream, bank[cur_bank].released, 0));   // consumers of the sweep before last
    if (g_profile == 3) prof_node("scanRegistration", stream, true);
    SRCounts init;
    memset(&init, 0, sizeof(init));
    init.h_min = INT_MAX;
    init.last_valid = -1;
    if (h_cnt_init.upload(d_cnt.p, init, stream) != LOAM_OK) { set_error("counter upload failed"); return LOAM_ERR_CUDA; }
    if (from_packet_metadata)
      LAUNCH("pk_ring_meta", pk_ring_meta, nblk, SR_BLK, 0, stream, meta_packets, meta_n_packets, meta_sensor, meta_span, d_decoded.p, d_ring.p, d_ori.p,
             d_hist.p, nblk, d_cnt.p);
    else
      LAUNCH("sr_ring_ori", sr_ring_ori, nblk, SR_BLK, 0, stream, d_raw_cur, n, 3, Pd, startOri, d_ring.p, d_ori.p, d_hist.p, nblk, d_cnt.p);
    LAUNCH("sr_ring_scan", sr_ring_scan, R, RSC_THREADS, 0, stream, d_hist.p, nblk, R, d_ring_start.p, d_scan_start.p, d_scan_end.p, d_cnt.p, cap,
           d_ring_tot.p, d_scan_ticket.p);
    LAUNCH("sr_scatter", sr_scatter, nblk, SR_BLK, 0, stream, d_raw_cur, n, 3, Pd, startOri, endOri, d_ring.p, d_ori.p, d_hist.p, d_ring_start.p, nblk,
           d_cnt.p, d_full.p, d_pos.p, d_reltime.p, from_packet_metadata ? 1 : 0);
    if (has_imu && imu.pointer_last >= 0) {  // :286
      if (h_imu.upload(d_imu.p, imu, stream) != LOAM_OK) { set_error("imu upload failed"); return LOAM_ERR_CUDA; }
      ImuState* st_in = d_imu_state.p + imu_parity;
      ImuState* st_out = d_imu_state.p + (imu_parity ^ 1);
      CUDA_OK(cudaMemcpyAsync(st_out, st_in, sizeof(ImuState), cudaMemcpyDeviceToDevice, stream));
      LAUNCH("sr_imu_steps", sr_imu_steps_kernel, nblk, SR_BLK, 0, stream, n, d_ring.p, d_reltime.p, stamp, d_imu.p, st_in, d_steps.p, d_blockmax.p);
      LAUNCH("sr_imu_deskew", sr_imu_deskew, nblk, SR_BLK, 0, stream, n, first, d_ring.p, d_reltime.p, d_pos.p, stamp, d_imu.p, st_in, st_out,
             d_steps.p, d_blockmax.p, nblk, d_full.p, cap, d_cnt.p);
      imu_parity ^= 1;
    }
    int crblocks = (n + CR_TILE - 1) / CR_TILE;
    LAUNCH("sr_curv_reject", sr_curv_reject, crblocks, CR_TILE, 0, stream, d_full.p, d_cnt.p, d_curv.p, d_sort_ind.p, d_picked.p, d_label.p, d_gap.p);
    LAUNCH("sr_sort_sectors", sr_sort_sectors, dim3(6, R), 512, 0, stream, d_curv.p, d_sort_ind.p, d_scan_start.p, d_scan_end.p, d_cnt.p);
    LAUNCH("sr_select", sr_select, R, SEL_THREADS, sizeof(SelShared), stream, d_curv.p, d_sort_ind.p, d_scan_start.p, d_scan_end.p, d_ring_start.p, d_gap.p,
           d_picked.p, d_label.p, d_sharp_ring.p, d_less_sharp_ring.p, d_flat_ring.p, d_ring_counts.p, d_cnt.p);
    size_t smem = sizeof(LessFlatSmem);
    LAUNCH("sr_lessflat", sr_lessflat, R, 1024, smem, stream, d_full.p, d_label.p, d_scan_start.p, d_scan_end.p, d_lessflat_tmp.p, d_ring_counts.p, d_cnt.p);
    LAUNCH("sr_compact", sr_compact, R, 128, 0, stream, d_full.p, d_ring_counts.p, R, d_sharp_ring.p, d_less_sharp_ring.p, d_flat_ring.p,
           d_lessflat_tmp.p, d_sharp.p, d_less_sharp.p, d_flat.p, d_less_flat.p, d_sharp_idx.p, d_less_sharp_idx.p, d_flat_idx.p, d_cnt.p, cap);
    CUDA_OK(cudaGetLastError());
    last_n = n;
    last_stamp = stamp;
    processed = true;
    return LOAM_OK;
  }

  const ImuState* imu_state_dev() const { return d_imu_state.p + imu_parity; }

  // Feature carrier: fills the next output bank from HOST copies of the node's six topics (the five
  // clouds, x,y,z,intensity per point, and the 12 floats of /imu_trans, :614-629), as if this node had
  // produced them -- the form laserOdometry is fed in when the features were computed / recorded
  // elsewhere.  The clouds must be ring-major as scanRegistration emits them (the reference's window
  // scans, laserOdometry.cpp:484-525, rely on the same order).
  int set_features(const float* const clouds[5], const int counts[5], const float imu_trans[12]) {
    processed = false;
    const int R = P.n_scans;
    for (int c = 0; c < 5; c++) {
      if (counts[c] < 0 || (counts[c] > 0 && !clouds[c])) { set_error("bad feature cloud %d", c); return LOAM_ERR_ARG; }
      if (counts[c] > cap) { set_error("feature cloud %d of %d points exceeds max_points=%d", c, counts[c], cap); return LOAM_ERR_CAPACITY; }
    }
    {  // the sharp / less-sharp / flat banks hold what scanRegistration can emit: 12 / 120 / 24 points per ring
      const int lim[5] = {cap, R * SR_SHARP_PER_RING, R * SR_LESS_SHARP_PER_RING, R * SR_FLAT_PER_RING, cap};
      const char* nm[5] = {"full", "sharp", "less-sharp", "flat", "less-flat"};
      for (int c = 1; c < 4; c++)
        if (counts[c] > lim[c]) {
          set_error("%s feature cloud of %d points exceeds the bank of %d points (%d rings x %d per ring)", nm[c], counts[c], lim[c], R, lim[c] / R);
          return LOAM_ERR_CAPACITY;
        }
    }
    std::vector<int> rc((size_t)R * 4, 0);
    const int col_of[5] = {-1, 0, 1, 2, 3};   // full, sharp, less sharp, flat, less flat
    for (int c = 1; c < 5; c++) {
      int prev = 0;
      for (int i = 0; i < counts[c]; i++) {
        const int ring = (int)clouds[c][4 * (size_t)i + 3];
        if (ring < prev || ring >= R) { set_error("feature cloud %d is not ring-major (point %d, ring %d)", c, i, ring); return LOAM_ERR_ARG; }
        prev = ring;
        rc[(size_t)ring * 4 + col_of[c]]++;
      }
    }
    CUDA_OK(cudaSetDevice(device));
    select_bank(cur_bank ^ 1);
    CUDA_OK(cudaStreamWaitEvent(stream, bank[cur_bank].released, 0));
    float4* dst[5] = {d_full.p, d_sharp.p, d_less_sharp.p, d_flat.p, d_less_flat.p};
    for (int c = 0; c < 5; c++)
      if (counts[c]) CUDA_OK(cudaMemcpyAsync(dst[c], clouds[c], sizeof(float4) * (size_t)counts[c], cudaMemcpyHostToDevice, stream));
    SRCounts cn;
    memset(&cn, 0, sizeof(cn));
    cn.n_full = counts[0]; cn.n_sharp = counts[1]; cn.n_less_sharp = counts[2]; cn.n_flat = counts[3]; cn.n_less_flat = counts[4];
    cn.h_min = INT_MAX; cn.last_valid = counts[0] - 1;
    CUDA_OK(cudaMemcpyAsync(d_cnt.p, &cn, sizeof(cn), cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaMemcpyAsync(d_ring_counts.p, rc.data(), sizeof(int) * rc.size(), cudaMemcpyHostToDevice, stream));
    ImuState is;
    memset(&is, 0, sizeof(is));
    is.rpy_start[1] = imu_trans[0]; is.rpy_start[2] = imu_trans[1]; is.rpy_start[0] = imu_trans[2];
    is.rpy_cur[1] = imu_trans[3]; is.rpy_cur[2] = imu_trans[4]; is.rpy_cur[0] = imu_trans[5];
    for (int a = 0; a < 3; a++) { is.shift_from_start[a] = imu_trans[6 + a]; is.velo_from_start[a] = imu_trans[9 + a]; }
    imu_parity ^= 1;
    CUDA_OK(cudaMemcpyAsync(d_imu_state.p + imu_parity, &is, sizeof(is), cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaStreamSynchronize(stream));   // the sources are pageable host memory of the caller
    last_counts = cn;
    *h_imu_state.p = is;
    processed = true;
    return LOAM_OK;
  }

  // blocking: fetch the counters of the last sweep
  int sync_counts() {
    CUDA_OK(cudaMemcpyAsync(h_cnt.p, d_cnt.p, sizeof(SRCounts), cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaMemcpyAsync(h_imu_state.p, imu_state_dev(), sizeof(ImuState), cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    last_counts = *h_cnt.p;
    if (last_counts.error) { describe_sr_error(last_counts.error); return LOAM_ERR_CAPACITY; }
    return LOAM_OK;
  }
  void imu_trans(float out[12]) const {  // :614-629
    const ImuState& s = *h_imu_state.p;
    out[0] = s.rpy_start[1]; out[1] = s.rpy_start[2]; out[2] = s.rpy_start[0];
    out[3] = s.rpy_cur[1]; out[4] = s.rpy_cur[2]; out[5] = s.rpy_cur[0];
    for (int a = 0; a < 3; a++) { out[6 + a] = s.shift_from_start[a]; out[9 + a] = s.velo_from_start[a]; }
  }
};

// One launch chain for the scanRegistration passes of many sequences (sr_batch.cuh).  The handles must have
// been created on one device with the same parameters and a SHARED stream (loam_pipeline_create_batch).
struct SRBatch {
  static const int SLOTS = 4;
  int B = 0;
  DevBuf<SRDev> d_items;
  SRDev* h_items[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
  int next = 0;
  int init(int count) {
    B = count;
    if (d_items.alloc((size_t)count) != LOAM_OK) return LOAM_ERR_CUDA;
    for (int i = 0; i < SLOTS; i++) {
      CUDA_OK(cudaMallocHost((void**)&h_items[i], sizeof(SRDev) * (size_t)count));
      CUDA_OK(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
    }
    return LOAM_OK;
  }
  ~SRBatch() {
    for (int i = 0; i < SLOTS; i++) { if (h_items[i]) cudaFreeHost(h_items[i]); if (ev[i]) cudaEventDestroy(ev[i]); }
  }
  // rcs[i] = what the per-sequence call would have returned (LOAM_NOT_READY: swallowed by systemDelay / empty).
  // begin() picks the staging slot, prepare(i) does sequence i's host side (any thread; one call per sequence),
  // launch() issues the chain for the sequences that are ready.
  SRDev* cur_items = nullptr;
  int cur_slot = 0;
  std::vector<SRDev> tmp_items;
  int begin() {
    cur_slot = next;
    next = (next + 1) % SLOTS;
    { HostWait hw; CUDA_OK(cudaEventSynchronize(ev[cur_slot])); }
    cur_items = h_items[cur_slot];
    tmp_items.resize(B);
    return LOAM_OK;
  }
  int prepare(SRHandle* h, int i, const float* xyz, int n, int stride_bytes, double stamp) {
    return h->process(xyz, n, stride_bytes, stamp, &tmp_items[i]);
  }
  int prepare_device(SRHandle* h, int i, const float* d_xyz, int n, const float* first_last_xyz, int first_index, double stamp) {
    return h->process_resident(d_xyz, n, first_last_xyz, first_last_xyz + 3, first_index, stamp, &tmp_items[i]);
  }
  int launch(SRHandle* const* hs, const int* n, const int* rcs) {
    SRDev* items = cur_items;
    int m = 0, max_nblk = 1, max_cr = 1;
    std::vector<int> who;
    for (int i = 0; i < B; i++) {
      if (rcs[i] < 0) return rcs[i];
      if (rcs[i] == LOAM_OK) {
        items[m] = tmp_items[i];
        max_nblk = max_nblk > items[m].nblk ? max_nblk : items[m].nblk;
        const int cr = (n[i] + CR_TILE - 1) / CR_TILE;
        max_cr = max_cr > cr ? max_cr : cr;
        who.push_back(i);
        m++;
      }
    }
    if (m == 0) return LOAM_OK;
    SRHandle& h0 = *hs[who[0]];
    cudaStream_t s = h0.stream;
    const int R = h0.P.n_scans;
    CUDA_OK(cudaMemcpyAsync(d_items.p, items, sizeof(SRDev) * (size_t)m, cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaEventRecord(ev[cur_slot], s));
    if (g_profile == 3) prof_node("scanRegistration", s, true);
    LAUNCH("sr_batch_init_b", sr_batch_init_b, m, 32, 0, s, d_items.p);
    LAUNCH("sr_ring_ori_b", sr_ring_ori_b, dim3(max_nblk, 1, m), SR_BLK, 0, s, d_items.p, h0.Pd);
    LAUNCH("sr_ring_scan_b", sr_ring_scan_b, dim3(R, 1, m), RSC_THREADS, 0, s, d_items.p, R, h0.cap);
    LAUNCH("sr_scatter_b", sr_scatter_b, dim3(max_nblk, 1, m), SR_BLK, 0, s, d_items.p, h0.Pd);
    LAUNCH("sr_curv_reject_b", sr_curv_reject_b, dim3(max_cr, 1, m), CR_TILE, 0, s, d_items.p);
    LAUNCH("sr_sort_sectors_b", sr_sort_sectors_b, dim3(6, R, m), 512, 0, s, d_items.p);
    LAUNCH("sr_select_b", sr_select_b, dim3(R, 1, m), SEL_THREADS, sizeof(SelShared), s, d_items.p);
    LAUNCH("sr_lessflat_b", sr_lessflat_b, dim3(R, 1, m), 1024, sizeof(LessFlatSmem), s, d_items.p);
    LAUNCH("sr_compact_b", sr_compact_b, dim3(R, 1, m), 128, 0, s, d_items.p, R, h0.cap);
    CUDA_OK(cudaGetLastError());
    for (int i : who)
      if (hs[i]->batch_ev_free) CUDA_OK(cudaEventRecord(hs[i]->batch_ev_free, s));   // the raw staging buffers are consumed
    return LOAM_OK;
  }
  static int set_attributes() {
    CUDA_OK(cudaFuncSetAttribute(sr_select_b, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SelShared)));
    CUDA_OK(cudaFuncSetAttribute(sr_lessflat_b, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LessFlatSmem)));
    return LOAM_OK;
  }
};

}  // namespace loamgpu
