// loam_lo.cuh -- host side of the laserOdometry node (laserOdometry.cpp main loop :410-931).
#pragma once
#include "lo_kernels2.cuh"
#include "loam_sr.cuh"

namespace loamgpu {

// cell sizes of the two odometry indexes (any size gives the same exact 1-NN; it only trades candidates per cell
// against how often a query has to walk further shells)
#ifndef LOAM_LO_CELL_CORNER
#define LOAM_LO_CELL_CORNER 1.0f
#endif
#ifndef LOAM_LO_CELL_SURF
#define LOAM_LO_CELL_SURF 1.0f
#endif
constexpr float LO_CELL_CORNER = LOAM_LO_CELL_CORNER, LO_CELL_SURF = LOAM_LO_CELL_SURF;

struct LOHandle {
  loam_params P;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  cudaEvent_t ev_dep = nullptr;
  int cap = 0, fcap = 0, coop_blocks = 0;
  bool systemInited = false;      // :52
  int frameCount = 1;             // :407 (= skipFrameNum)
  static const int skipFrameNum = 1;
  DevBuf<LOState> d_state;
  DevBuf<float4> d_corner_last, d_surf_last, d_full_end, d_coeff;
  DevBuf<unsigned char> d_row_valid;
  DevBuf<int> d_c1, d_c2, d_s1, d_s2, d_s3, d_corner_roff, d_surf_roff;
  DevBuf<double> d_partials;
  GridStorage g_corner, g_surf;
  PinnedBuf<LOState> h_state;
  LOState last_state;
  bool last_published = false;
  int last_n_sharp = 0, last_n_flat = 0;

  int init(const loam_params* p, int dev, cudaStream_t s) {
    P = *p;
    device = dev;
    // one lane per accumulated row slab (lo_optimize), one claim counter per search: at most 32 iterations
    if (P.lo_max_iter < 1 || P.lo_max_iter > 32) { set_error("lo_max_iter = %d is outside 1 .. 32 (the reference uses 25)", P.lo_max_iter); return LOAM_ERR_ARG; }
    if (P.n_scans < 1 || P.n_scans > SR_MAX_RINGS || P.max_points < 64) {
      set_error("n_scans = %d / max_points = %d: n_scans must be 1 .. %d (LOAM_LIMIT_RINGS), max_points at least 64", P.n_scans, P.max_points, SR_MAX_RINGS);
      return LOAM_ERR_ARG;
    }
    CUDA_OK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) { set_error("device %d is not sm_100", dev); return LOAM_ERR_CUDA; }
    if (s) stream = s; else { CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)); own_stream = true; }
    CUDA_OK(cudaEventCreateWithFlags(&ev_dep, cudaEventDisableTiming));
    cap = P.max_points;
    fcap = (36 * P.n_scans + 127) / 128 * 128;   // 2 + 4 features per sector and ring (scanRegistration.cpp:470-560), rounded to whole L1 lines
    int per_sm = 0;
    CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lo_optimize, LO_THREADS, 0));
    if (per_sm < 1) { set_error("lo_optimize cannot be made resident"); return LOAM_ERR_CUDA; }
    // An SM cannot change its L1 / shared-memory split while a CTA is resident.  The persistent kernel itself needs
    // 15 KB, but it sits on every SM for hundreds of microseconds: with the split sized for it alone, the other nodes'
    // kernels that need 50 .. 200 KB (sr_select, sr_lessflat, lm_bucket_new, lm_refilter_cube) could not start
    // anywhere until it had finished.  (dev_flags bit 9 keeps the driver's choice, for comparison.)
    if ((P.dev_flags & 512) == 0)
      CUDA_OK(cudaFuncSetAttribute(lo_optimize, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    coop_blocks = prop.multiProcessorCount;   // one CTA per SM unless the caller asks for a smaller grid
    if (P.coop_ctas > 0 && P.coop_ctas < coop_blocks) coop_blocks = P.coop_ctas;
    int r = LOAM_OK;
    r |= d_state.alloc_zero(1, stream);
    r |= d_corner_last.alloc(cap); r |= d_surf_last.alloc(cap); r |= d_full_end.alloc(cap);
    r |= d_coeff.alloc((size_t)P.lo_max_iter * fcap); r |= d_row_valid.alloc_zero((size_t)P.lo_max_iter * fcap, stream);
    r |= d_c1.alloc_zero(fcap, stream); r |= d_c2.alloc_zero(fcap, stream);
    r |= d_s1.alloc_zero(fcap, stream); r |= d_s2.alloc_zero(fcap, stream); r |= d_s3.alloc_zero(fcap, stream);
    r |= d_partials.alloc_zero((size_t)2 * coop_blocks * LO_NRED, stream);
    r |= d_corner_roff.alloc_zero(SR_MAX_RINGS + 1, stream); r |= d_surf_roff.alloc_zero(SR_MAX_RINGS + 1, stream);
    r |= g_corner.alloc(cap, 1 << 21, stream); r |= g_surf.alloc(cap, 1 << 21, stream);
    r |= h_state.alloc(1);
    if (r != LOAM_OK) return LOAM_ERR_CUDA;
    memset(&last_state, 0, sizeof(last_state));
    CUDA_OK(cudaStreamSynchronize(stream));
    return LOAM_OK;
  }
  ~LOHandle() {
    if (ev_dep) cudaEventDestroy(ev_dep);
    if (carrier) { cudaStreamSynchronize(stream); delete carrier; }
    if (own_stream && stream) cudaStreamDestroy(stream);
  }

  // one loop-body pass for the six topics scanRegistration just produced; enqueue only
  int process(SRHandle& sr, bool* published) {
    *published = false;
    if (!sr.processed) return LOAM_NOT_READY;
    CUDA_OK(cudaSetDevice(device));
    if (sr.stream != stream) {
      CUDA_OK(cudaEventRecord(ev_dep, sr.stream));
      CUDA_OK(cudaStreamWaitEvent(stream, ev_dep, 0));
    }
    if (g_profile == 3) prof_node("laserOdometry", stream, true);
    const ImuState* imu = sr.imu_state_dev();
    GridPair gp;
    gp.g[0] = g_corner.buf(d_corner_last.p, &d_state.p->n_corner_last, &d_state.p->index_enable, LO_CELL_CORNER);
    gp.g[1] = g_surf.buf(d_surf_last.p, &d_state.p->n_surf_last, &d_state.p->index_enable, LO_CELL_SURF);
    const int nb = 148;
    if (!systemInited) {  // :427-456
      LAUNCH("lo_adopt_first", lo_adopt_first, nb, 256, 0, stream, sr.d_cnt.p, sr.d_less_sharp.p, sr.d_less_flat.p, d_corner_last.p, d_surf_last.p);
      LAUNCH("lo_finish", lo_finish, 1, 32, 0, stream, d_state.p, sr.d_cnt.p, imu, 1, 0, sr.d_ring_counts.p, P.n_scans, d_corner_roff.p, d_surf_roff.p);
      int r = grid_build_pair(gp, nb, stream, (P.dev_flags & 16) == 0);
      if (r != LOAM_OK) return r;
      systemInited = true;
      last_published = false;
      return LOAM_NOT_READY;
    }
    LOArgs A;
    A.st = d_state.p; A.cnt = sr.d_cnt.p; A.sharp = sr.d_sharp.p; A.flat = sr.d_flat.p;
    A.corner_last = d_corner_last.p; A.surf_last = d_surf_last.p;
    A.corner_grid = g_corner.view(); A.surf_grid = g_surf.view();
    A.corner_ind1 = d_c1.p; A.corner_ind2 = d_c2.p; A.surf_ind1 = d_s1.p; A.surf_ind2 = d_s2.p; A.surf_ind3 = d_s3.p;
    A.coeff = d_coeff.p; A.row_valid = d_row_valid.p; A.partials = d_partials.p; A.imu = imu;
    A.corner_roff = d_corner_roff.p; A.surf_roff = d_surf_roff.p; A.n_scans = P.n_scans;
    A.fcap = fcap; A.max_iter = P.lo_max_iter; A.contiguous = (P.dev_flags & 1) ? 0 : ((P.dev_flags & 2) ? 2 : 1); A.dynsearch = (P.dev_flags & 256) ? 0 : 1;   // dev flag bit 0: strided assignment (measured slower: neighbouring features share cells)
    frameCount++;  // :885
    const int do_full = frameCount >= skipFrameNum + 1 ? 1 : 0;
    A.less_sharp = sr.d_less_sharp.p; A.less_flat = sr.d_less_flat.p; A.full = sr.d_full.p;
    A.corner_last_out = d_corner_last.p; A.surf_last_out = d_surf_last.p; A.full_end = d_full_end.p; A.do_full = do_full;
    // The tail (TransformToEnd, swap, bookkeeping, index rebuild) either runs inside lo_optimize -- four more
    // grid barriers, seven launches fewer, lower latency: the choice when many sequences share a GPU and the host
    // launch rate is the limit (coop_ctas > 0) -- or as separate small kernels that trickle into whatever the two
    // gang-scheduled persistent kernels leave free, which is +5 % throughput for ONE stream on a whole GPU
    // (measured: 4130-4190 vs 3930 sweeps/s, HDL-64).  Dev flags: bit 2 forces the separate kernels, bit 3 the epilogue.
    A.epilogue = (P.dev_flags & 4) ? 0 : ((P.dev_flags & 8) ? 1 : (P.coop_ctas > 0 ? 1 : 0));
    A.gb[0] = gp.g[0]; A.gb[1] = gp.g[1];
    A.ring_counts = sr.d_ring_counts.p; A.corner_roff_out = d_corner_roff.p; A.surf_roff_out = d_surf_roff.p;
    {
      void* args[] = {&A};
      if (g_profile) prof_begin(stream);
      CUDA_OK(cudaLaunchCooperativeKernel((void*)lo_optimize, dim3(coop_blocks), dim3(LO_THREADS), args, 0, stream));
      g_launches++;
      if (g_profile) prof_end("lo_optimize", stream);
    }
    // TransformToEnd, the swap, the bookkeeping and the rebuild of both indices run in lo_optimize's epilogue
    if (!A.epilogue) {   // A/B switch (dev flag bit 2): the same tail as separate launches
      ToEndArgs T;
      T.st = d_state.p; T.cnt = sr.d_cnt.p; T.imu = imu;
      T.less_sharp = sr.d_less_sharp.p; T.less_flat = sr.d_less_flat.p; T.full = sr.d_full.p;
      T.corner_last = d_corner_last.p; T.surf_last = d_surf_last.p; T.full_end = d_full_end.p; T.do_full = do_full;
      T.ring_counts = sr.d_ring_counts.p; T.n_scans = P.n_scans; T.corner_roff = d_corner_roff.p; T.surf_roff = d_surf_roff.p;
      T.record = record_out;     // bookkeeping (lo_finish) and the pipeline's pose record ride on this launch
      record_written = record_out != nullptr;
      LAUNCH("lo_transform_to_end", lo_transform_to_end_kernel, 2 * nb, 256, 0, stream, T);
      int r = grid_build_pair(gp, nb, stream, (P.dev_flags & 16) == 0);   // the last clouds are feature clouds: always small
      if (r != LOAM_OK) return r;
    }
    if (do_full) { frameCount = 0; *published = true; }  // :910-911
    last_published = *published;
    return LOAM_OK;
  }

  // the same pass fed from HOST copies of the six topics: a private scanRegistration handle on this
  // node's stream serves as the feature carrier (SRHandle::set_features)
  SRHandle* carrier = nullptr;
  // set by a pipeline before process(): where the sweep's pose record goes; record_written tells it afterwards whether
  // the pass wrote it (it does not on the first sweep and with the in-kernel tail)
  PipeRecord* record_out = nullptr;
  bool record_written = false;
  int process_host(const float* const clouds[5], const int counts[5], const float imu_trans[12], bool* published) {
    *published = false;
    if (!carrier) {
      carrier = new SRHandle();
      int r = carrier->init(&P, device, stream, true);
      if (r != LOAM_OK) { delete carrier; carrier = nullptr; return r; }
    }
    int r = carrier->set_features(clouds, counts, imu_trans);
    if (r != LOAM_OK) return r;
    r = process(*carrier, published);
    if (r < 0) return r;
    int rr = carrier->release_bank(stream);
    return rr < 0 ? rr : r;
  }

  int sync_state() {
    CUDA_OK(cudaMemcpyAsync(h_state.p, d_state.p, sizeof(LOState), cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    last_state = *h_state.p;
    return LOAM_OK;
  }
};

}  // namespace loamgpu
