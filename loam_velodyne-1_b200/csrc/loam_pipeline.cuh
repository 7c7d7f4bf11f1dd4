// loam_pipeline.cuh -- the four nodes wired as the reference's topic graph on
// one device and one CUDA stream: sr -> lo -> tm.onOdometry -> (lm -> tm.onMapping).
#pragma once
#include "loam_lm.cuh"

namespace loamgpu {

struct PipelineHandle {
  SRHandle sr;
  LOHandle lo;
  LMHandle lm;
  TMHost tm;
  cudaStream_t stream = nullptr;
  loam_sweep_result res;

  int init(const loam_params* p, int dev) {
    CUDA_OK(cudaSetDevice(dev));
    CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    int r = sr.init(p, dev, stream);
    if (r != LOAM_OK) return r;
    r = lo.init(p, dev, stream);
    if (r != LOAM_OK) return r;
    r = lm.init(p, dev, stream);
    if (r != LOAM_OK) return r;
    memset(&res, 0, sizeof(res));
    return LOAM_OK;
  }
  ~PipelineHandle() { if (stream) { cudaStreamSynchronize(stream); } }

  int imu(double stamp, const double quat[4], const float acc[3]) {
    sr.on_imu(stamp, quat, acc);
    lm.on_imu(stamp, quat);
    return LOAM_OK;
  }

  int push(const float* xyz, int n, int stride_bytes, double stamp, loam_sweep_result* out) {
    loam_sweep_result& r = res;
    r.sr_status = r.lo_status = r.lm_status = LOAM_NOT_READY;
    int rc = sr.process(xyz, n, stride_bytes, stamp);
    if (rc < 0) return rc;
    if (rc == LOAM_OK) {
      r.sr_status = LOAM_OK;
      bool pub = false;
      int lrc = lo.process(sr, &pub);
      if (lrc < 0) return lrc;
      int mrc = LOAM_NOT_READY;
      if (lrc == LOAM_OK && pub) {
        mrc = lm.process(lo, stamp);
        if (mrc < 0) return mrc;
      }
      // one synchronisation per sweep: counters + poses
      CUDA_OK(cudaMemcpyAsync(sr.h_cnt.p, sr.d_cnt.p, sizeof(SRCounts), cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaMemcpyAsync(sr.h_imu_state.p, sr.imu_state_dev(), sizeof(ImuState), cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaMemcpyAsync(lo.h_state.p, lo.d_state.p, sizeof(LOState), cudaMemcpyDeviceToHost, stream));
      if (mrc == LOAM_OK) CUDA_OK(cudaMemcpyAsync(lm.h_state.p, lm.d_state.p, sizeof(LMState), cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaStreamSynchronize(stream));
      sr.last_counts = *sr.h_cnt.p;
      if (sr.last_counts.error) { set_error("scanRegistration capacity exceeded (flags %d)", sr.last_counts.error); return LOAM_ERR_CAPACITY; }
      lo.last_state = *lo.h_state.p;
      lo.last_n_sharp = sr.last_counts.n_sharp; lo.last_n_flat = sr.last_counts.n_flat;
      r.n_full = sr.last_counts.n_full; r.n_sharp = sr.last_counts.n_sharp; r.n_less_sharp = sr.last_counts.n_less_sharp;
      r.n_flat = sr.last_counts.n_flat; r.n_less_flat = sr.last_counts.n_less_flat;
      sr.imu_trans(r.imu_trans);
      r.lo_iters = lo.last_state.iters; r.lo_rows = lo.last_state.rows; r.lo_degenerate = lo.last_state.isDegenerate;
      for (int i = 0; i < 6; i++) { r.transform[i] = lo.last_state.transform[i]; r.transform_sum[i] = lo.last_state.transformSum[i]; }
      if (lrc == LOAM_OK) {
        r.lo_status = LOAM_OK;
        tm.on_odometry(lo.last_state.transformSum, r.transform_mapped);
        if (mrc == LOAM_OK) {
          lm.last_state = *lm.h_state.p;
          const LMState& m = lm.last_state;
          if (m.error) { set_error("laserMapping capacity exceeded (flags %d)", m.error); return LOAM_ERR_CAPACITY; }
          r.lm_status = LOAM_OK;
          r.lm_iters = m.iters; r.lm_rows = m.rows; r.lm_degenerate = m.isDegenerate;
          r.n_map_corner = m.mapN[0]; r.n_map_surf = m.mapN[1]; r.n_stack_corner = m.stackN[0]; r.n_stack_surf = m.stackN[1];
          for (int i = 0; i < 6; i++) { r.transform_aft_mapped[i] = m.transformAftMapped[i]; r.transform_bef_mapped[i] = m.transformBefMapped[i]; }
          tm.on_mapping(m.transformAftMapped, m.transformBefMapped);
          if (m.bump[0] > lm.arena_cap / 2 || m.bump[1] > lm.arena_cap / 2) {
            int cr = lm.compact();
            if (cr != LOAM_OK) return cr;
          }
        }
      }
    }
    if (out) *out = r;
    return r.sr_status;
  }
};

}  // namespace loamgpu
