// loam_lm.cuh -- host side of the laserMapping node (laserMapping.cpp main loop :408-1096) and
// the host-only transformMaintenance node (transformMaintenance.cpp).
#pragma once
#include "lm_kernels2.cuh"
#include "vx_coop.cuh"
#include "loam_lo.cuh"

namespace loamgpu {

struct LMHandle {
  loam_params P;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  cudaEvent_t ev_dep = nullptr;
  // side stream for the branches of a pass that do not depend on each other: the local-map gather +
  // index build (K4, K5) beside the incoming-sweep filters (K2, K3), and the registered cloud (K8)
  // beside the map update (K7)
  cudaStream_t aux = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_aux = nullptr;
  // second side stream (low priority) for the products of a pass that nothing on the device waits for:
  // the registered cloud and the voxel filter of the surround map.  They would otherwise sit on `aux`
  // between the map update of pass k and the window shift of pass k+1, i.e. on the cycle
  // LM loop -> map update -> window shift -> gather -> index -> LM loop that sets the mapping rate.
  cudaStream_t aux2 = nullptr;
  cudaEvent_t ev_reg = nullptr, ev_sur_in = nullptr, ev_sur_done = nullptr;
  // pose record of a pass (loam_pipeline.cuh): snapshot slot of the pass, "map update enqueued" and "slot read" events
  cudaEvent_t ev_upd = nullptr, ev_rec[2] = {nullptr, nullptr};
  int rec_slot = 0, last_rec_slot = 0;
  int cap = 0, ecap = 0, arena_cap = 0, coop_blocks = 0;
  int cur_arena = 0;
  int sweeps_since_compact = 0;
  static const int stackFrameNum = 1, mapFrameNum = 5;  // :51-52
  int frameCount = stackFrameNum - 1, mapFrameCount = mapFrameNum - 1;  // :404-405
  DevBuf<LMState> d_state;
  DevBuf<LMImu> d_imu;
  DevBuf<CubeDesc> d_cubes[2];
  DevBuf<int> d_touch[2], d_rank[2], d_newoff[2];
  DevBuf<float4> d_arena[2][2];        // [type][A/B]
  DevBuf<float4> d_stack2[2], d_stack[2], d_E[2], d_ins_pts[2], d_registered, d_surround;
  DevBuf<int> d_ins_cube[2], d_nn[2];
  DevBuf<unsigned int> d_ins_key[2], d_nk[2][2], d_nv[2][2], d_sk[2][2], d_sv[2][2];
  DevBuf<unsigned int> d_skeys[2][2], d_svals[2][2], d_bbox[2];
  DevBuf<int> d_shist[2], d_sdig[2], d_sheads[2], d_uhist, d_udig, d_uheads;
  DevBuf<unsigned int> d_ukeys[2], d_uvals[2], d_ubbox;
  DevBuf<float4> d_S, d_ssorted[2];
  int vc_tile_cap = 0;
  DevBuf<double> d_partials;
  GridStorage g_map[2];
  // per-cube persistent index (lm_cube_index.cuh); dev_flags bit 5 (32) selects the round-1 scheme
  // instead: gather the local map and index all of it, every pass
  bool use_ci = false;
  int ci_cap = 0;
  bool full_grid = true, concurrent_nodes = false;   // block size of the LM kernel: see lm_kernels2.cuh
  DevBuf<CIState> d_ci;
  DevBuf<int2> d_ci_cell[2];
  DevBuf<float4> d_ci_pts[2];
  // the gathered local map (laserCloudCornerFromMap / SurfFromMap) is only materialised when somebody asks for it
  bool snap_pending = false;
  int snap_arena = 0;
  PinnedBuf<LMState> h_state;
  PinnedRing<LMImu, 4> h_imu;
  LMImu imu;
  bool has_imu = false;
  LMState last_state;
  long long map_hint = 0;   // points of a map loaded from a file (the host's only knowledge of the map size before the first sync)
  bool last_surround = false;
  int last_n_registered = 0;
  // What the node receives from laserOdometry (/laser_cloud_corner_last, _surf_last,
  // /velodyne_cloud_3, /laser_odom_to_init).  When the odometry runs on another stream the
  // messages are copied into one of two input banks (lm_handoff, enqueued on the odometry's
  // stream), so the odometry of the next sweeps runs beside this sweep's mapping -- the topic
  // queue between the reference's two nodes.
  struct Input { const LOState* lo; const float4* corner_last; const float4* surf_last; const float4* full_end; };
  DevBuf<LOState> d_in_state[2];
  DevBuf<float4> d_in_corner[2], d_in_surf[2], d_in_full[2];
  cudaEvent_t ev_in_ready[2] = {nullptr, nullptr}, ev_in_done[2] = {nullptr, nullptr};
  int in_bank = 0;

  int init(const loam_params* p, int dev, cudaStream_t s) {
    P = *p;
    device = dev;
    if (P.lm_max_iter < 1 || P.lm_max_iter > 1000) { set_error("lm_max_iter = %d is outside 1 .. 1000 (the reference uses 10)", P.lm_max_iter); return LOAM_ERR_ARG; }
    if (P.map_arena_points < 0 || (P.map_arena_points > 0 && P.map_arena_points < 4096)) {
      set_error("map_arena_points = %d: 0 (default, 8 Mi points) or at least 4096", P.map_arena_points); return LOAM_ERR_ARG;
    }
    CUDA_OK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) { set_error("device %d is not sm_100", dev); return LOAM_ERR_CUDA; }
    if (s) stream = s; else { CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)); own_stream = true; }
    CUDA_OK(cudaEventCreateWithFlags(&ev_dep, cudaEventDisableTiming));
    {
      int pri = 0;
      CUDA_OK(cudaStreamGetPriority(stream, &pri));
      CUDA_OK(cudaStreamCreateWithPriority(&aux, cudaStreamNonBlocking, pri));
      CUDA_OK(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_aux, cudaEventDisableTiming));
      int lo_pri = 0, hi_pri = 0;
      CUDA_OK(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
      CUDA_OK(cudaStreamCreateWithPriority(&aux2, cudaStreamNonBlocking, lo_pri));
      CUDA_OK(cudaEventCreateWithFlags(&ev_reg, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_sur_in, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_sur_done, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_upd, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_rec[0], cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&ev_rec[1], cudaEventDisableTiming));
    }
    cap = P.max_points;
    arena_cap = P.map_arena_points > 0 ? P.map_arena_points : (8 << 20);
    ecap = arena_cap / 4;
    if (ecap < 4 * cap) ecap = 4 * cap;
    coop_blocks = prop.multiProcessorCount;   // one CTA per SM unless the caller asks for a smaller grid
    if (P.coop_ctas > 0 && P.coop_ctas < coop_blocks) coop_blocks = P.coop_ctas;
    {
      long long want = 3LL * ecap + 65536;
      if (P.dev_flags & 64) want = (long long)ecap + ecap / 8 + 256 * CI_WIN + 1024;   // test hook: the smallest legal arena (frequent resets)
      if (want > (1LL << CI_ADDR_BITS)) want = 1LL << CI_ADDR_BITS;
      ci_cap = (int)want;
      // a local map of `ecap` points must fit the index arena with its head room, twice over
      use_ci = (P.dev_flags & 32) == 0 && (long long)ecap + ecap / 8 + 256 * CI_WIN <= ci_cap;
    }
    int per_sm = 0;
    if (use_ci) CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lm_optimize_ci, LM_THREADS_ALONE, 0));
    else CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lm_optimize, LM_THREADS_ALONE, 0));
    if (per_sm < 1) { set_error("lm_optimize cannot be made resident"); return LOAM_ERR_CUDA; }
    full_grid = coop_blocks == prop.multiProcessorCount;
    if ((P.dev_flags & 512) == 0) {   // see loam_lo.cuh: the persistent kernel must not pin the SMs to a small shared-memory split
      CUDA_OK(cudaFuncSetAttribute(lm_optimize, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
      CUDA_OK(cudaFuncSetAttribute(lm_optimize_ci, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    }
    if (use_ci) CUDA_OK(cudaFuncSetAttribute(ci_build, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CUDA_OK(cudaFuncSetAttribute(lm_refilter_cube, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RF_SMEM));
    if (RF_SLICES > 8) CUDA_OK(cudaFuncSetAttribute(lm_refilter_cube, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CUDA_OK(cudaFuncSetAttribute(lm_bucket_new, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BN_SMEM));
    CUDA_OK(cudaFuncSetAttribute(vx_filter_coop, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    int r = LOAM_OK;
    r |= d_state.alloc_zero(1, stream); r |= d_imu.alloc_zero(1, stream);
    int snblk = (cap + RS_TILE - 1) / RS_TILE + 1, enblk = (ecap + RS_TILE - 1) / RS_TILE + 1;
    int svblk = (cap + VX_TILE - 1) / VX_TILE + 1, evblk = (ecap + VX_TILE - 1) / VX_TILE + 1;
    for (int T = 0; T < 2; T++) {
      r |= d_cubes[T].alloc_zero(CUBE_NUM, stream); r |= d_touch[T].alloc_zero(CUBE_NUM, stream);
      r |= d_rank[T].alloc(CUBE_NUM); r |= d_newoff[T].alloc(CUBE_NUM);
      r |= d_arena[T][0].alloc(arena_cap); r |= d_arena[T][1].alloc(arena_cap);
      r |= d_stack2[T].alloc(cap); r |= d_stack[T].alloc(cap); r |= d_E[T].alloc(ecap); r |= d_ins_pts[T].alloc(cap);
      r |= d_ins_cube[T].alloc(cap); r |= d_nn[T].alloc((size_t)cap * 5);
      r |= d_ins_key[T].alloc(cap);
      for (int k = 0; k < 2; k++) { r |= d_nk[T][k].alloc(cap); r |= d_nv[T][k].alloc(cap); r |= d_sk[T][k].alloc(ecap); r |= d_sv[T][k].alloc(ecap); }
      for (int k = 0; k < 2; k++) {
        r |= d_skeys[T][k].alloc(cap); r |= d_svals[T][k].alloc(cap);
      }
      r |= d_bbox[T].alloc(6);
      vc_tile_cap = (cap + VC_TILE - 1) / VC_TILE + 1;
      r |= d_shist[T].alloc((size_t)256 * vc_tile_cap); r |= d_ssorted[T].alloc(cap);
      r |= d_sdig[T].alloc(256); r |= d_sheads[T].alloc(svblk);
      if (use_ci) { r |= d_ci_cell[T].alloc((size_t)CI_WIN * CI_TABLE); r |= d_ci_pts[T].alloc(ci_cap); }
      else r |= g_map[T].alloc(ecap, 1 << 22, stream);
      if (r == LOAM_OK) {
        if (cudaMemsetAsync(d_rank[T].p, 0xff, sizeof(int) * CUBE_NUM, stream) != cudaSuccess) r = LOAM_ERR_CUDA;
        unsigned int init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
        if (cudaMemcpyAsync(d_bbox[T].p, init, sizeof(init), cudaMemcpyHostToDevice, stream) != cudaSuccess) r = LOAM_ERR_CUDA;
        cudaStreamSynchronize(stream);
      }
    }
    if (use_ci) r |= d_ci.alloc_zero(1, stream);
    r |= d_S.alloc(ecap); r |= d_surround.alloc(ecap); r |= d_ubbox.alloc(6);
    for (int k = 0; k < 2; k++) { r |= d_ukeys[k].alloc(ecap); r |= d_uvals[k].alloc(ecap); }
    r |= d_uhist.alloc((size_t)256 * enblk); r |= d_udig.alloc(256); r |= d_uheads.alloc(evblk);
    if (r == LOAM_OK) {
      unsigned int binit[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
      if (cudaMemcpyAsync(d_ubbox.p, binit, sizeof(binit), cudaMemcpyHostToDevice, stream) != cudaSuccess) r = LOAM_ERR_CUDA;
      cudaStreamSynchronize(stream);
    }
    r |= d_registered.alloc(cap); r |= d_partials.alloc_zero((size_t)2 * coop_blocks * LO_NRED, stream);
    r |= h_state.alloc(1); r |= h_imu.alloc();
    for (int b = 0; b < 2; b++) {
      r |= d_in_state[b].alloc_zero(1, stream); r |= d_in_corner[b].alloc(cap); r |= d_in_surf[b].alloc(cap); r |= d_in_full[b].alloc(cap);
      if (cudaEventCreateWithFlags(&ev_in_ready[b], cudaEventDisableTiming) != cudaSuccess) r = LOAM_ERR_CUDA;
      if (cudaEventCreateWithFlags(&ev_in_done[b], cudaEventDisableTiming) != cudaSuccess) r = LOAM_ERR_CUDA;
    }
    if (r != LOAM_OK) return LOAM_ERR_CUDA;
    LMState init;
    memset(&init, 0, sizeof(init));
    init.cenW = 10; init.cenH = 5; init.cenD = 10;  // :64-66
    init.arenaCap = arena_cap;
    *h_state.p = init;
    CUDA_OK(cudaMemcpyAsync(d_state.p, h_state.p, sizeof(LMState), cudaMemcpyHostToDevice, stream));
    memset(&imu, 0, sizeof(imu));
    imu.pointer_last = -1;
    last_state = init;
    CUDA_OK(cudaStreamSynchronize(stream));
    return reset_index();
  }
  // every slot of the per-cube index forgets what it holds (start-up, map load)
  int reset_index() {
    snap_pending = false;
    if (!use_ci) return LOAM_OK;
    std::vector<CIState> h(1);
    memset(h.data(), 0, sizeof(CIState));
    for (int T = 0; T < 2; T++)
      for (int s = 0; s < CI_WIN; s++) { h[0].slot[T][s].ver = -1; h[0].win_slot[T][s] = -1; }
    h[0].cap = ci_cap;
    CUDA_OK(cudaMemcpy(d_ci.p, h.data(), sizeof(CIState), cudaMemcpyHostToDevice));
    return LOAM_OK;
  }
  // laserCloudCornerFromMap / SurfFromMap of the last pass, gathered when a download asks for them (the pass
  // itself searches the per-cube indexes and never needs the concatenated cloud)
  int materialize_local_map() {
    if (!snap_pending) return LOAM_OK;
    LMGatherArgs GA;
    GA.st = d_state.p; GA.ecap = ecap;
    for (int T = 0; T < 2; T++) { GA.cubes[T] = d_cubes[T].p; GA.arena[T] = d_arena[T][snap_arena].p; GA.E[T] = d_E[T].p; }
    LAUNCH("lm_gather_map", lm_gather_map, dim3(4 * 148, 2), 256, 0, aux, GA);
    CUDA_OK(cudaGetLastError());
    snap_pending = false;
    return LOAM_OK;
  }
  int ensure_local_map() {
    int r = materialize_local_map();
    if (r != LOAM_OK) return r;
    CUDA_OK(cudaStreamSynchronize(aux));
    return LOAM_OK;
  }
  ~LMHandle() {
    if (ev_dep) cudaEventDestroy(ev_dep);
    if (aux) { cudaStreamSynchronize(aux); cudaStreamDestroy(aux); }
    if (aux2) { cudaStreamSynchronize(aux2); cudaStreamDestroy(aux2); }
    if (ev_fork) cudaEventDestroy(ev_fork);
    if (ev_aux) cudaEventDestroy(ev_aux);
    if (ev_reg) cudaEventDestroy(ev_reg);
    if (ev_sur_in) cudaEventDestroy(ev_sur_in);
    if (ev_sur_done) cudaEventDestroy(ev_sur_done);
    if (ev_upd) cudaEventDestroy(ev_upd);
    for (int b = 0; b < 2; b++) if (ev_rec[b]) cudaEventDestroy(ev_rec[b]);
    for (int b = 0; b < 2; b++) {
      if (ev_in_ready[b]) cudaEventDestroy(ev_in_ready[b]);
      if (ev_in_done[b]) cudaEventDestroy(ev_in_done[b]);
    }
    if (own_stream && stream) cudaStreamDestroy(stream);
  }

  // imuHandler :323-335
  int on_imu(double stamp, const double quat[4]) {
    double roll, pitch, yaw;
    pm::rpy_from_quat(quat[0], quat[1], quat[2], quat[3], roll, pitch, yaw);
    int last = (imu.pointer_last + 1) % IMU_Q;
    imu.pointer_last = last;
    imu.time[last] = stamp;
    imu.roll[last] = roll;
    imu.pitch[last] = pitch;
    has_imu = true;
    return LOAM_OK;
  }

  SortPair sort_pair() {
    SortPair sp;
    sp.njobs = 2;
    for (int T = 0; T < 2; T++) {
      SortJob& j = sp.j[T];
      j.keys[0] = d_skeys[T][0].p; j.keys[1] = d_skeys[T][1].p; j.vals[0] = d_svals[T][0].p; j.vals[1] = d_svals[T][1].p;
      j.hist = d_shist[T].p; j.digit_total = d_sdig[T].p; j.n_ptr = &d_state.p->stackN2[T]; j.nblk_cap = (cap + RS_TILE - 1) / RS_TILE + 1;
    }
    return sp;
  }

  // one loop-body pass; consumes corner_last / surf_last / full_end / transformSum of `lo`
  int process(LOHandle& lo, double stamp) {
    if (!lo.last_published) return LOAM_NOT_READY;
    frameCount++;
    if (frameCount < stackFrameNum) return LOAM_NOT_READY;
    frameCount = 0;
    CUDA_OK(cudaSetDevice(device));
    Input in;
    in.lo = lo.d_state.p; in.corner_last = lo.d_corner_last.p; in.surf_last = lo.d_surf_last.p; in.full_end = lo.d_full_end.p;
    int bank = -1;
    concurrent_nodes = lo.stream != stream;
    if (lo.stream != stream) {
      bank = in_bank;
      in_bank ^= 1;
      CUDA_OK(cudaStreamWaitEvent(lo.stream, ev_in_done[bank], 0));   // the mapping pass that last read this bank
      LMHandoffArgs HA;
      HA.lo = in.lo; HA.src[0] = in.corner_last; HA.src[1] = in.surf_last; HA.src[2] = in.full_end;
      HA.dst[0] = d_in_corner[bank].p; HA.dst[1] = d_in_surf[bank].p; HA.dst[2] = d_in_full[bank].p; HA.lo_out = d_in_state[bank].p; HA.cap = cap;
      LAUNCH("lm_handoff", lm_handoff, 148, 256, 0, lo.stream, HA);
      CUDA_OK(cudaEventRecord(ev_in_ready[bank], lo.stream));
      CUDA_OK(cudaStreamWaitEvent(stream, ev_in_ready[bank], 0));
      in.lo = d_in_state[bank].p; in.corner_last = d_in_corner[bank].p; in.surf_last = d_in_surf[bank].p; in.full_end = d_in_full[bank].p;
    }
    if (g_profile == 3) prof_node("laserMapping", stream, true);
    int rc = run(in, stamp);
    if (bank >= 0) CUDA_OK(cudaEventRecord(ev_in_done[bank], aux));   // the side stream holds the pass's last reader of the bank
    return rc;
  }

  int run(const Input& in, double stamp) {
    LMState* st = d_state.p;
    const int nb = 148;
    float4* arena[2] = {d_arena[0][cur_arena].p, d_arena[1][cur_arena].p};
    // Stream plan of one pass (`stream` = main, `aux` = side; the side stream still holds the map update
    // of the PREVIOUS pass, which therefore runs beside this pass's pose prediction and incoming filters):
    //   main: K1a pose prediction -> K2 stack transform -> K3 incoming filters -> [wait side] K6 LM loop -> K7a insert ids
    //   side: [map update of pass k-1] -> [wait K1a] K1b window shift + cube selection -> K4 gather -> K5 index
    //         ... [wait K7a] K8 registered cloud, K7b-d map update, compaction, surround, pose record
    // K1a
    LMSetupArgs SA;
    SA.st = st; SA.lo = in.lo; SA.cubes[0] = d_cubes[0].p; SA.cubes[1] = d_cubes[1].p; SA.phase = 1;
    SA.ci = use_ci ? d_ci.p : nullptr;
    LAUNCH("lm_setup", lm_setup, 1, 32, 0, stream, SA);
    CUDA_OK(cudaEventRecord(ev_fork, stream));
    CUDA_OK(cudaStreamWaitEvent(aux, ev_fork, 0));
    // K1b on the side stream, behind the previous pass's map update (both work on the cube descriptors)
    SA.phase = 2;
    LAUNCH("lm_setup_cubes", lm_setup_cubes, 1, 256, 0, aux, SA);
    // K2
    LMStackArgs KA;
    KA.st = st; KA.in[0] = in.corner_last; KA.in[1] = in.surf_last; KA.out[0] = d_stack2[0].p; KA.out[1] = d_stack2[1].p;
    LAUNCH("lm_stack_transform", lm_stack_transform, nb, 256, 0, stream, KA);
    // K3: downSizeFilterCorner (0.2) / downSizeFilterSurf (0.4) on the incoming sweep: one cooperative kernel
    {
      VxCoopArgs VA;
      for (int T = 0; T < 2; T++) {
        VxCoopJob& j = VA.j[T];
        j.pts = d_stack2[T].p; j.n_ptr = &st->stackN2[T]; j.bbox = d_bbox[T].p;
        j.keys[0] = d_skeys[T][0].p; j.keys[1] = d_skeys[T][1].p; j.vals[0] = d_svals[T][0].p; j.vals[1] = d_svals[T][1].p;
        j.hist = d_shist[T].p; j.heads = d_sheads[T].p; j.sorted = d_ssorted[T].p; j.out = d_stack[T].p; j.out_count = &st->stackN[T];
        j.overflow = &st->vxOverflow[T]; j.leaf = T == 0 ? 0.2f : 0.4f; j.tile_cap = vc_tile_cap;
      }
      LAUNCH("vx_filter_coop", vx_filter_coop, dim3(VC_CLUSTER, 2), VC_THREADS, 0, stream, VA);
    }
    if (use_ci) {
      // K4 + K5: the slots K1b found stale (cubes the last sweep was inserted into, cubes that entered the window)
      // are re-indexed; nothing is gathered
      CIBuildArgs BA;
      BA.ci = d_ci.p; BA.error = &st->error;
      for (int T = 0; T < 2; T++) { BA.cubes[T] = d_cubes[T].p; BA.arena[T] = arena[T]; BA.cell[T] = d_ci_cell[T].p; BA.pts[T] = d_ci_pts[T].p; }
      LAUNCH("ci_build", ci_build, dim3(CB_CLUSTER, CB_GRID_Y), CB_THREADS, 0, aux, BA);
      CUDA_OK(cudaEventRecord(ev_aux, aux));
      CUDA_OK(cudaStreamWaitEvent(stream, ev_aux, 0));
      snap_pending = true;
      snap_arena = cur_arena;
    } else {
    // K4
    LMGatherArgs GA;
    GA.st = st; GA.ecap = ecap;
    for (int T = 0; T < 2; T++) { GA.cubes[T] = d_cubes[T].p; GA.arena[T] = arena[T]; GA.E[T] = d_E[T].p; }
    LAUNCH("lm_gather_map", lm_gather_map, dim3(4 * nb, 2), 256, 0, aux, GA);
    // K5: kdtreeCornerFromMap / kdtreeSurfFromMap -> grid index, cell 1.02 m
    {
      GridPair gp;
      for (int T = 0; T < 2; T++) gp.g[T] = g_map[T].buf(d_E[T].p, &st->mapN[T], &st->optimize_enable, 1.02f);
      // the local map grows slowly: the size the host saw last (sync / collect) decides between the one-cluster
      // build and the five grid-wide kernels (mapping-only stress: > 1 M points)
      const bool small = (P.dev_flags & 16) == 0 && last_state.mapN[0] + last_state.mapN[1] < 400000 && map_hint < 400000;
      int r = grid_build_pair(gp, 2 * nb, aux, small);
      if (r != LOAM_OK) return r;
      CUDA_OK(cudaEventRecord(ev_aux, aux));
      CUDA_OK(cudaStreamWaitEvent(stream, ev_aux, 0));
    }
    }
    // K6
    if (has_imu) {
      if (h_imu.upload(d_imu.p, imu, stream) != LOAM_OK) { set_error("imu upload failed"); return LOAM_ERR_CUDA; }
    }
    {
      LMOptArgs OA;
      OA.st = st;
      for (int T = 0; T < 2; T++) {
        OA.stack[T] = d_stack[T].p; OA.map[T] = d_E[T].p; OA.grid[T] = g_map[T].view(); OA.nn_dbg[T] = d_nn[T].p;
        OA.ci[T].cell = d_ci_cell[T].p; OA.ci[T].pts = d_ci_pts[T].p;
      }
      OA.cis = d_ci.p;
      OA.partials = d_partials.p; OA.imu = d_imu.p; OA.has_imu = has_imu ? 1 : 0; OA.timeLaserOdometry = stamp; OA.max_iter = P.lm_max_iter;
      void* args[] = {&OA};
      if (g_profile) prof_begin(stream);
      CUDA_OK(cudaLaunchCooperativeKernel(use_ci ? (void*)lm_optimize_ci : (void*)lm_optimize, dim3(coop_blocks),
                                          dim3((full_grid && !concurrent_nodes && (P.dev_flags & 128) == 0) ? LM_THREADS_ALONE : LM_THREADS), args, 0, stream));
      g_launches++;
      if (g_profile) prof_end("lm_optimize", stream);
    }
    // K7: insert + per-cube refilter (lm_refilter.cuh).  K7a stays on the main stream (it reads the stack
    // and the refined pose, and snapshots both for the side stream); everything after it is the side stream's.
    {
      LMInsertArgs IA;
      IA.st = st;
      last_rec_slot = rec_slot;
      rec_slot ^= 1;
      IA.rec_slot = last_rec_slot;
      CUDA_OK(cudaStreamWaitEvent(stream, ev_rec[last_rec_slot], 0));   // the record of the pass before last has read this slot
      for (int T = 0; T < 2; T++) { IA.stack[T] = d_stack[T].p; IA.ins_pts[T] = d_ins_pts[T].p; IA.ins_cube[T] = d_ins_cube[T].p; IA.ins_key[T] = d_ins_key[T].p; IA.touch_cnt[T] = d_touch[T].p; }
      LAUNCH("lm_insert_ids", lm_insert_ids, nb, 256, 0, stream, IA);
      CUDA_OK(cudaEventRecord(ev_fork, stream));
      CUDA_OK(cudaStreamWaitEvent(aux, ev_fork, 0));
      CUDA_OK(cudaStreamWaitEvent(aux2, ev_fork, 0));
      // K8: registered cloud (second side stream: beside the map update)
      LAUNCH("lm_registered", lm_registered, 2 * nb, 256, 0, aux2, st, in.full_end, &in.lo->n_full_end, d_registered.p);
      CUDA_OK(cudaEventRecord(ev_reg, aux2));
      LMPlanArgs PA;
      PA.st = st;
      for (int T = 0; T < 2; T++) { PA.cubes[T] = d_cubes[T].p; PA.touch_cnt[T] = d_touch[T].p; PA.cube_rank[T] = d_rank[T].p; }
      LAUNCH("lm_refilter_plan", lm_refilter_plan, 2, 1024, 0, aux, PA);
      LMSortNewArgs NA;
      NA.st = st;
      for (int T = 0; T < 2; T++) {
        NA.ins_cube[T] = d_ins_cube[T].p; NA.ins_key[T] = d_ins_key[T].p; NA.cube_rank[T] = d_rank[T].p;
        for (int k = 0; k < 2; k++) { NA.nk[T][k] = d_nk[T][k].p; NA.nv[T][k] = d_nv[T][k].p; }
      }
      LAUNCH("lm_bucket_new", lm_bucket_new, dim3(BN_GRID, 2), BN_THREADS, BN_SMEM, aux, NA);
      LMCubeArgs CA;
      CA.st = st; CA.scap = ecap;
      for (int T = 0; T < 2; T++) {
        CA.cubes[T] = d_cubes[T].p; CA.arena[T] = arena[T]; CA.ins_pts[T] = d_ins_pts[T].p;
        CA.touch_cnt[T] = d_touch[T].p; CA.cube_rank[T] = d_rank[T].p;
        for (int k = 0; k < 2; k++) { CA.nk[T][k] = d_nk[T][k].p; CA.nv[T][k] = d_nv[T][k].p; CA.sk[T][k] = d_sk[T][k].p; CA.sv[T][k] = d_sv[T][k].p; }
      }
      LAUNCH("lm_refilter_cube", lm_refilter_cube, dim3(RF_GRID, 2), RF_THREADS, RF_SMEM, aux, CA);
    }
    if (++sweeps_since_compact >= 8) {  // bounded garbage in the log-structured arena
      int cr = compact();
      if (cr != LOAM_OK) return cr;
    }
    mapFrameCount++;  // :1038-1040
    last_surround = false;
    if (mapFrameCount >= mapFrameNum) {
      mapFrameCount = 0;
      last_surround = true;
      int sr_ = surround();
      if (sr_ != LOAM_OK) return sr_;
    }
    // the registered cloud was the pass's last reader of the input bank: `aux` (on which the caller
    // releases the bank) picks it up here, long after the 5 us kernel has finished
    CUDA_OK(cudaStreamWaitEvent(aux, ev_reg, 0));
    // no join here: the side stream's work is ordered before the next pass's K1b, and everything that
    // needs the pass to be complete (sync_state, the pipeline's collect / join, the destructor) waits for
    // `aux` and `aux2` (sync_side / join_side)
    CUDA_OK(cudaGetLastError());
    return LOAM_OK;
  }

  // LM-14 (:1042-1057): gather the surround cubes, whole-cloud VoxelGrid(0.2)
  int surround() {
    LMState* st = d_state.p;
    LMSurroundArgs SA;
    SA.st = st; SA.S = d_S.p; SA.scap = ecap;
    for (int T = 0; T < 2; T++) { SA.cubes[T] = d_cubes[T].p; SA.arena[T] = d_arena[T][cur_arena].p; }
    // the gather copies the cubes out of the arena on `aux` (ordered with the map updates); the voxel
    // filter of the copy runs on `aux2`, beside the next passes
    CUDA_OK(cudaStreamWaitEvent(aux, ev_sur_done, 0));   // the previous surround filter still reads d_S / surroundIn
    LAUNCH("lm_surround_plan", lm_surround_plan, 1, 256, 0, aux, SA);
    LAUNCH("lm_surround_gather", lm_surround_gather, 4 * 148, 256, 0, aux, SA);
    CUDA_OK(cudaEventRecord(ev_sur_in, aux));
    CUDA_OK(cudaStreamWaitEvent(aux2, ev_sur_in, 0));
    cudaStream_t aux = this->aux2;   // everything below: second side stream
    CloudKeyPair cp;
    cp.njobs = 1;
    CloudKeyJob& cj = cp.j[0];
    cj.pts = d_S.p; cj.n_ptr = &st->surroundIn; cj.bbox = d_ubbox.p; cj.keys = d_ukeys[0].p; cj.vals = d_uvals[0].p;
    cj.overflow = &st->vxOverflow[2]; cj.leaf = 0.2f;
    cp.j[1] = cj;
    LAUNCH("vx_bbox", vx_bbox, dim3(296, 1), 256, 0, aux, cp);
    LAUNCH("vx_cloud_keys", vx_cloud_keys, dim3(296, 1), 256, 0, aux, cp);
    LAUNCH("vx_rearm_bbox", vx_rearm_bbox, 1, 32, 0, aux, cp);
    SortPair sp;
    sp.njobs = 1;
    SortJob& sj = sp.j[0];
    sj.keys[0] = d_ukeys[0].p; sj.keys[1] = d_ukeys[1].p; sj.vals[0] = d_uvals[0].p; sj.vals[1] = d_uvals[1].p;
    sj.hist = d_uhist.p; sj.digit_total = d_udig.p; sj.n_ptr = &st->surroundIn; sj.nblk_cap = (ecap + RS_TILE - 1) / RS_TILE + 1;
    sp.j[1] = sj;
    int r = radix_sort_pairs(sp, ecap, aux);
    if (r != LOAM_OK) return r;
    VoxelPair vp;
    vp.njobs = 1;
    VoxelJob& vj = vp.j[0];
    vj.keys = d_ukeys[0].p; vj.vals = d_uvals[0].p; vj.pts = d_S.p; vj.n_ptr = &st->surroundIn;
    vj.block_heads = d_uheads.p; vj.out = d_surround.p; vj.out_base_ptr = nullptr; vj.out_count = &st->surroundN;
    vj.seg_count = nullptr; vj.seg_shift = 0; vj.nblk_cap = (ecap + VX_TILE - 1) / VX_TILE + 1;
    vp.j[1] = vj;
    r = voxel_reduce_sorted(vp, ecap, aux);
    if (r != LOAM_OK) return r;
    CUDA_OK(cudaEventRecord(ev_sur_done, aux));
    return LOAM_OK;
  }

  // host waits for both side streams
  int sync_side() {
    CUDA_OK(cudaStreamSynchronize(aux));
    CUDA_OK(cudaStreamSynchronize(aux2));
    return LOAM_OK;
  }
  // stream `s` waits for both side streams (no host sync); `ev` is a scratch event of the caller
  int join_side(cudaStream_t s, cudaEvent_t ev) {
    CUDA_OK(cudaEventRecord(ev, aux));
    CUDA_OK(cudaStreamWaitEvent(s, ev, 0));
    CUDA_OK(cudaEventRecord(ev, aux2));
    CUDA_OK(cudaStreamWaitEvent(s, ev, 0));
    return LOAM_OK;
  }

  int compact() {
    // the copy below overwrites the other arena: if the local-map snapshot of the last pass still points into it
    // (two compactions since that pass), gather it first
    if (snap_pending && snap_arena == (cur_arena ^ 1)) { int r = materialize_local_map(); if (r != LOAM_OK) return r; }
    LMCompactArgs CA;
    CA.st = d_state.p;
    for (int T = 0; T < 2; T++) {
      CA.cubes[T] = d_cubes[T].p; CA.src[T] = d_arena[T][cur_arena].p; CA.dst[T] = d_arena[T][cur_arena ^ 1].p; CA.new_off[T] = d_newoff[T].p;
    }
    LAUNCH("lm_compact_plan", lm_compact_plan, 2, 1024, 0, aux, CA);
    LAUNCH("lm_compact_copy", lm_compact_copy, dim3(296, 2), 256, 0, aux, CA);
    LAUNCH("lm_compact_commit", lm_compact_commit, dim3(19, 2), 256, 0, aux, CA);
    CUDA_OK(cudaGetLastError());
    cur_arena ^= 1;
    sweeps_since_compact = 0;
    return LOAM_OK;
  }

  int sync_state() {
    { int sr_ = sync_side(); if (sr_ != LOAM_OK) return sr_; }   // map update, registered cloud, surround map of the last pass
    CUDA_OK(cudaMemcpyAsync(h_state.p, d_state.p, sizeof(LMState), cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    last_state = *h_state.p;
    if (last_state.error) { describe_lm_error(last_state.error); return LOAM_ERR_CAPACITY; }
    if (last_state.bump[0] > arena_cap / 2 || last_state.bump[1] > arena_cap / 2) {
      int r = compact();
      if (r != LOAM_OK) return r;
    }
    return LOAM_OK;
  }
};

// transformMaintenance.cpp, host only (north_star: "transformMaintenance stays on the host")
struct TMHost {
  float transformSum[6], transformIncre[6], transformMapped[6], transformBefMapped[6], transformAftMapped[6];  // :49-53
  TMHost() {
    for (int i = 0; i < 6; i++) transformSum[i] = transformIncre[i] = transformMapped[i] = transformBefMapped[i] = transformAftMapped[i] = 0;
  }
  // odomAftMappedHandler :182-203: pose arrives as a quaternion (decoded here), bef as twist floats
  // decoded = the caller (a ROS node) already took getRPY of the message's quaternion
  void on_mapping(const float aft[6], const float bef[6], bool decoded = false) {
    float a[6];
    if (decoded) { for (int i = 0; i < 6; i++) a[i] = aft[i]; }
    else pm::odometry_msg_roundtrip(aft, a);
    for (int i = 0; i < 6; i++) { transformAftMapped[i] = a[i]; transformBefMapped[i] = bef[i]; }
  }
  // laserOdometryHandler :147-180
  void on_odometry(const float sum[6], float integrated[6], bool decoded = false) {
    if (decoded) { for (int i = 0; i < 6; i++) transformSum[i] = sum[i]; }
    else pm::odometry_msg_roundtrip(sum, transformSum);
    pm::transformAssociateToMap(transformSum, transformBefMapped, transformAftMapped, transformIncre, transformMapped);
    for (int i = 0; i < 6; i++) integrated[i] = transformMapped[i];
  }
};

}  // namespace loamgpu
