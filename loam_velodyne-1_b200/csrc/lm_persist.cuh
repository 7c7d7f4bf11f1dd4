// lm_persist.cuh -- host-side entry points of the laserMapping node that the reference does not
// have but its users ask for:
//   * lm_process_host: one loop-body pass (laserMapping.cpp:408-1096) fed from HOST copies of the
//     four topics (/laser_cloud_corner_last, /laser_cloud_surf_last, /velodyne_cloud_3,
//     /laser_odom_to_init) -- teacher forcing for the mapping-only configuration;
//   * lm_save_map / lm_load_map: checkpoint / resume of the cube map (laserCloudCornerArray /
//     laserCloudSurfArray :74-81), the window centre (:64-66), the five transforms (:96-100), the
//     frame counters (:404-405) and the IMU ring (:102-108);
//   * lm_export_pcd: a cloud of the node as a binary PCD v0.7 file (x y z intensity).
// The file format is defined here and restated independently by the oracle (oracle/port_capi.cpp).
#pragma once
#include <cstdio>
#include <vector>
#include "loam_lm.cuh"

namespace loamgpu {

// "LOAMMAP1" file: header, IMU ring, then per type (corner, surf): counts[4851] (int32),
// dirty[4851] (uint8: 1 = cube not known to be one point per voxel in voxel order), points
// (x,y,z,intensity float32) of all cubes back to back in cube-index order i + 21*j + 21*11*k.
struct MapFileHeader {
  char magic[8];
  int version;
  int dims[3];
  int cen[3];
  int frameCount, mapFrameCount;
  int isDegenerate;
  float transformSum[6], transformIncre[6], transformTobeMapped[6], transformBefMapped[6], transformAftMapped[6];
  int imuPointerFront, imuPointerLast, hasImu, imuQueLength;
  long long total[2];
};

// on_device = false: the clouds are HOST arrays, copied in stream order into one of the node's two
// input banks; true: DEVICE arrays that the caller keeps unchanged until the pass has run (nothing
// is copied but the 100-byte odometry state).  Enqueue only: the caller synchronises through
// LMHandle::sync_state().
inline int lm_process_clouds(LMHandle& h, const float* corner, int nc, const float* surf, int ns, const float* full, int nf,
                             const float transform_sum[6], double stamp, bool on_device, bool sum_decoded = false) {
  if (nc < 0 || ns < 0 || nf < 0 || (nc && !corner) || (ns && !surf) || (nf && !full) || !transform_sum) return LOAM_ERR_ARG;
  if (nc > h.cap || ns > h.cap || nf > h.cap) { set_error("laserMapping input cloud exceeds max_points (%d)", h.cap); return LOAM_ERR_CAPACITY; }
  h.frameCount++;
  if (h.frameCount < LMHandle::stackFrameNum) return LOAM_NOT_READY;
  h.frameCount = 0;
  CUDA_OK(cudaSetDevice(h.device));
  int bank = h.in_bank;
  h.in_bank ^= 1;
  LOState ls;
  memset(&ls, 0, sizeof(ls));
  for (int i = 0; i < 6; i++) ls.transformSum[i] = transform_sum[i];
  ls.n_corner_last = nc; ls.n_surf_last = ns; ls.n_full_end = nf;
  ls.cornerLastNum = nc; ls.surfLastNum = ns;
  ls.sum_decoded = sum_decoded ? 1 : 0;
  LMHandle::Input in;
  in.lo = h.d_in_state[bank].p;
  if (on_device) {
    in.corner_last = (const float4*)corner; in.surf_last = (const float4*)surf; in.full_end = (const float4*)full;
  } else {
    // pageable source: cudaMemcpyAsync stages it before returning; the copy itself is ordered behind the previous pass
    if (nc) CUDA_OK(cudaMemcpyAsync(h.d_in_corner[bank].p, corner, sizeof(float4) * (size_t)nc, cudaMemcpyHostToDevice, h.stream));
    if (ns) CUDA_OK(cudaMemcpyAsync(h.d_in_surf[bank].p, surf, sizeof(float4) * (size_t)ns, cudaMemcpyHostToDevice, h.stream));
    if (nf) CUDA_OK(cudaMemcpyAsync(h.d_in_full[bank].p, full, sizeof(float4) * (size_t)nf, cudaMemcpyHostToDevice, h.stream));
    in.corner_last = h.d_in_corner[bank].p; in.surf_last = h.d_in_surf[bank].p; in.full_end = h.d_in_full[bank].p;
  }
  CUDA_OK(cudaMemcpyAsync(h.d_in_state[bank].p, &ls, sizeof(ls), cudaMemcpyHostToDevice, h.stream));
  if (g_profile == 3) prof_node("laserMapping", h.stream, true);
  int rc = h.run(in, stamp);
  if (g_profile == 3) prof_node("laserMapping", h.stream, false);
  return rc;
}

inline int lm_save_map(LMHandle& h, const char* path) {
  if (!path) return LOAM_ERR_ARG;
  CUDA_OK(cudaSetDevice(h.device));
  int arena_before = h.cur_arena;
  int r = h.sync_state();
  if (r == LOAM_OK && h.cur_arena != arena_before) r = h.sync_state();   // a compaction was due: read the state it left
  if (r != LOAM_OK) return r;
  CUDA_OK(cudaStreamSynchronize(h.aux));
  CUDA_OK(cudaStreamSynchronize(h.aux2));
  const LMState& s = h.last_state;
  MapFileHeader H;
  memset(&H, 0, sizeof(H));
  memcpy(H.magic, "LOAMMAP1", 8);
  H.version = 1;
  H.dims[0] = CUBE_W; H.dims[1] = CUBE_H; H.dims[2] = CUBE_D;
  H.cen[0] = s.cenW; H.cen[1] = s.cenH; H.cen[2] = s.cenD;
  H.frameCount = h.frameCount; H.mapFrameCount = h.mapFrameCount;
  H.isDegenerate = s.isDegenerate;
  for (int i = 0; i < 6; i++) {
    H.transformSum[i] = s.transformSum[i]; H.transformIncre[i] = s.transformIncre[i]; H.transformTobeMapped[i] = s.transformTobeMapped[i];
    H.transformBefMapped[i] = s.transformBefMapped[i]; H.transformAftMapped[i] = s.transformAftMapped[i];
  }
  H.imuPointerFront = s.imuPointerFront; H.imuPointerLast = h.imu.pointer_last; H.hasImu = h.has_imu ? 1 : 0; H.imuQueLength = IMU_Q;
  std::vector<CubeDesc> cubes[2];
  std::vector<float4> pts[2];
  for (int T = 0; T < 2; T++) {
    cubes[T].resize(CUBE_NUM);
    CUDA_OK(cudaMemcpy(cubes[T].data(), h.d_cubes[T].p, sizeof(CubeDesc) * CUBE_NUM, cudaMemcpyDeviceToHost));
    int used = s.bump[T];
    pts[T].resize(used > 0 ? used : 1);
    if (used > 0) CUDA_OK(cudaMemcpy(pts[T].data(), h.d_arena[T][h.cur_arena].p, sizeof(float4) * (size_t)used, cudaMemcpyDeviceToHost));
    long long tot = 0;
    for (int c = 0; c < CUBE_NUM; c++) {
      const CubeDesc& d = cubes[T][c];
      if (d.cnt < 0 || (d.cnt > 0 && (d.off < 0 || d.off + d.cnt > used))) { set_error("cube descriptor %d out of the arena", c); return LOAM_ERR_CUDA; }
      tot += d.cnt;
    }
    H.total[T] = tot;
  }
  FILE* f = fopen(path, "wb");
  if (!f) { set_error("cannot open %s for writing", path); return LOAM_ERR_ARG; }
  bool ok = fwrite(&H, sizeof(H), 1, f) == 1;
  ok = ok && fwrite(h.imu.time, sizeof(double), IMU_Q, f) == (size_t)IMU_Q;
  ok = ok && fwrite(h.imu.roll, sizeof(float), IMU_Q, f) == (size_t)IMU_Q;
  ok = ok && fwrite(h.imu.pitch, sizeof(float), IMU_Q, f) == (size_t)IMU_Q;
  for (int T = 0; T < 2 && ok; T++) {
    std::vector<int> cnt(CUBE_NUM);
    std::vector<unsigned char> dirty(CUBE_NUM);
    for (int c = 0; c < CUBE_NUM; c++) { cnt[c] = cubes[T][c].cnt; dirty[c] = cubes[T][c].cnt > 0 && cubes[T][c].dirty ? 1 : 0; }
    ok = ok && fwrite(cnt.data(), sizeof(int), CUBE_NUM, f) == (size_t)CUBE_NUM;
    ok = ok && fwrite(dirty.data(), 1, CUBE_NUM, f) == (size_t)CUBE_NUM;
    for (int c = 0; c < CUBE_NUM && ok; c++) {
      const CubeDesc& d = cubes[T][c];
      if (d.cnt > 0) ok = fwrite(&pts[T][d.off], sizeof(float4), d.cnt, f) == (size_t)d.cnt;
    }
  }
  ok = (fclose(f) == 0) && ok;
  if (!ok) { set_error("short write to %s", path); return LOAM_ERR_ARG; }
  return LOAM_OK;
}

inline int lm_load_map(LMHandle& h, const char* path) {
  if (!path) return LOAM_ERR_ARG;
  CUDA_OK(cudaSetDevice(h.device));
  FILE* f = fopen(path, "rb");
  if (!f) { set_error("cannot open %s", path); return LOAM_ERR_ARG; }
  MapFileHeader H;
  LMImu imu;
  memset(&imu, 0, sizeof(imu));
  bool ok = fread(&H, sizeof(H), 1, f) == 1;
  if (!ok || memcmp(H.magic, "LOAMMAP1", 8) != 0 || H.version != 1 || H.dims[0] != CUBE_W || H.dims[1] != CUBE_H || H.dims[2] != CUBE_D ||
      H.imuQueLength != IMU_Q) {
    fclose(f);
    set_error("%s is not a LOAMMAP1 file of a %dx%dx%d cube map", path, CUBE_W, CUBE_H, CUBE_D);
    return LOAM_ERR_ARG;
  }
  // header fields that index arrays later on: refuse a file that would send them out of range
  if (H.cen[0] < 0 || H.cen[0] >= CUBE_W || H.cen[1] < 0 || H.cen[1] >= CUBE_H || H.cen[2] < 0 || H.cen[2] >= CUBE_D ||
      H.imuPointerLast < -1 || H.imuPointerLast >= IMU_Q || H.imuPointerFront < 0 || H.imuPointerFront >= IMU_Q ||
      H.frameCount < 0 || H.mapFrameCount < 0) {
    fclose(f);
    set_error("%s: header fields out of range (window centre %d %d %d, imu pointers %d %d)", path, H.cen[0], H.cen[1], H.cen[2],
              H.imuPointerLast, H.imuPointerFront);
    return LOAM_ERR_ARG;
  }
  ok = ok && fread(imu.time, sizeof(double), IMU_Q, f) == (size_t)IMU_Q;
  ok = ok && fread(imu.roll, sizeof(float), IMU_Q, f) == (size_t)IMU_Q;
  ok = ok && fread(imu.pitch, sizeof(float), IMU_Q, f) == (size_t)IMU_Q;
  imu.pointer_last = H.imuPointerLast;
  std::vector<CubeDesc> cubes[2];
  std::vector<float4> pts[2];
  for (int T = 0; T < 2 && ok; T++) {
    if (H.total[T] < 0 || H.total[T] > h.arena_cap / 2) {
      fclose(f);
      set_error("map of %lld points does not fit the arena (map_arena_points = %d)", H.total[T], h.arena_cap);
      return LOAM_ERR_CAPACITY;
    }
    std::vector<int> cnt(CUBE_NUM);
    std::vector<unsigned char> dirty(CUBE_NUM);
    ok = ok && fread(cnt.data(), sizeof(int), CUBE_NUM, f) == (size_t)CUBE_NUM;
    ok = ok && fread(dirty.data(), 1, CUBE_NUM, f) == (size_t)CUBE_NUM;
    cubes[T].resize(CUBE_NUM);
    long long off = 0;
    for (int c = 0; c < CUBE_NUM && ok; c++) {
      if (cnt[c] < 0) { ok = false; break; }
      CubeDesc d;
      d.off = (int)off; d.cnt = cnt[c]; d.dirty = cnt[c] > 0 && dirty[c] ? 1 : 0; d.ver = 0;
      cubes[T][c] = d;
      off += cnt[c];
    }
    ok = ok && off == H.total[T];
    pts[T].resize(off > 0 ? off : 1);
    if (ok && off > 0) ok = fread(pts[T].data(), sizeof(float4), (size_t)off, f) == (size_t)off;
  }
  fclose(f);
  if (!ok) { set_error("%s is truncated or inconsistent", path); return LOAM_ERR_ARG; }
  CUDA_OK(cudaStreamSynchronize(h.stream));
  CUDA_OK(cudaStreamSynchronize(h.aux));
  CUDA_OK(cudaStreamSynchronize(h.aux2));
  h.cur_arena = 0;
  h.sweeps_since_compact = 0;
  LMState s;
  memset(&s, 0, sizeof(s));
  s.cenW = H.cen[0]; s.cenH = H.cen[1]; s.cenD = H.cen[2];
  s.isDegenerate = H.isDegenerate;
  for (int i = 0; i < 6; i++) {
    s.transformSum[i] = H.transformSum[i]; s.transformIncre[i] = H.transformIncre[i]; s.transformTobeMapped[i] = H.transformTobeMapped[i];
    s.transformBefMapped[i] = H.transformBefMapped[i]; s.transformAftMapped[i] = H.transformAftMapped[i];
  }
  s.imuPointerFront = H.imuPointerFront;
  s.arenaCap = h.arena_cap;
  for (int T = 0; T < 2; T++) {
    s.bump[T] = (int)H.total[T];
    if (T == 0) h.map_hint = 0;
    h.map_hint += H.total[T];
    CUDA_OK(cudaMemcpy(h.d_cubes[T].p, cubes[T].data(), sizeof(CubeDesc) * CUBE_NUM, cudaMemcpyHostToDevice));
    if (H.total[T] > 0) CUDA_OK(cudaMemcpy(h.d_arena[T][0].p, pts[T].data(), sizeof(float4) * (size_t)H.total[T], cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemset(h.d_touch[T].p, 0, sizeof(int) * CUBE_NUM));
    CUDA_OK(cudaMemset(h.d_rank[T].p, 0xff, sizeof(int) * CUBE_NUM));
  }
  CUDA_OK(cudaMemcpy(h.d_state.p, &s, sizeof(s), cudaMemcpyHostToDevice));
  h.last_state = s;
  h.frameCount = H.frameCount;
  h.mapFrameCount = H.mapFrameCount;
  h.imu = imu;
  h.has_imu = H.hasImu != 0;
  return h.reset_index();
}

// binary PCD v0.7, FIELDS x y z intensity (what pcl::io::savePCDFileBinary writes for PointXYZI
// without the padding words).  which: a loam_cloud selector of the node, or 20 / 21 = every cube
// of the corner / surf map back to back in cube-index order.
inline int lm_export_pcd(LMHandle& h, int which, const char* path) {
  if (!path) return LOAM_ERR_ARG;
  CUDA_OK(cudaSetDevice(h.device));
  int arena_before = h.cur_arena;
  int r = h.sync_state();
  if (r == LOAM_OK && h.cur_arena != arena_before) r = h.sync_state();
  if (r != LOAM_OK) return r;
  CUDA_OK(cudaStreamSynchronize(h.aux));
  CUDA_OK(cudaStreamSynchronize(h.aux2));
  const LMState& s = h.last_state;
  if (which == LOAM_CLOUD_MAP_CORNER || which == LOAM_CLOUD_MAP_SURF) { int er = h.ensure_local_map(); if (er != LOAM_OK) return er; }
  std::vector<float4> pts;
  const float4* src = nullptr;
  int n = 0;
  switch (which) {
    case LOAM_CLOUD_MAP_CORNER: src = h.d_E[0].p; n = s.mapN[0]; break;
    case LOAM_CLOUD_MAP_SURF: src = h.d_E[1].p; n = s.mapN[1]; break;
    case LOAM_CLOUD_STACK_CORNER: src = h.d_stack[0].p; n = s.stackN[0]; break;
    case LOAM_CLOUD_STACK_SURF: src = h.d_stack[1].p; n = s.stackN[1]; break;
    case LOAM_CLOUD_REGISTERED: src = h.d_registered.p; n = s.registeredN; break;
    case LOAM_CLOUD_SURROUND: src = h.d_surround.p; n = s.surroundN; break;
    case 20: case 21: {
      int T = which - 20;
      std::vector<CubeDesc> cubes(CUBE_NUM);
      CUDA_OK(cudaMemcpy(cubes.data(), h.d_cubes[T].p, sizeof(CubeDesc) * CUBE_NUM, cudaMemcpyDeviceToHost));
      int used = s.bump[T];
      std::vector<float4> arena(used > 0 ? used : 1);
      if (used > 0) CUDA_OK(cudaMemcpy(arena.data(), h.d_arena[T][h.cur_arena].p, sizeof(float4) * (size_t)used, cudaMemcpyDeviceToHost));
      for (int c = 0; c < CUBE_NUM; c++)
        if (cubes[c].cnt > 0 && cubes[c].off >= 0 && cubes[c].off + cubes[c].cnt <= used)
          pts.insert(pts.end(), arena.begin() + cubes[c].off, arena.begin() + cubes[c].off + cubes[c].cnt);
      n = (int)pts.size();
      break;
    }
    default: return LOAM_ERR_ARG;
  }
  if (src) {
    pts.resize(n > 0 ? n : 1);
    if (n > 0) CUDA_OK(cudaMemcpy(pts.data(), src, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost));
  }
  FILE* f = fopen(path, "wb");
  if (!f) { set_error("cannot open %s for writing", path); return LOAM_ERR_ARG; }
  fprintf(f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z intensity\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
             "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA binary\n", n, n);
  bool ok = n == 0 || fwrite(pts.data(), sizeof(float4), n, f) == (size_t)n;
  ok = (fclose(f) == 0) && ok;
  if (!ok) { set_error("short write to %s", path); return LOAM_ERR_ARG; }
  return LOAM_OK;
}

}  // namespace loamgpu
