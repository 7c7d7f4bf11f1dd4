// Drives the reference's topic graph through the C++ facade (include/loam/loam.hpp), node by
// node, on sweeps read from a binary file: for every sweep  sr.process -> lo.process ->
// tm.onOdometry -> (lm.process -> tm.onMapping).  Prints one line of hex floats per sweep; the
// Python test compares them bit for bit with the ctypes Pipeline.
// With a third argument (a writable file name) a second LaserMapping is driven beside the first
// from HOST copies of laserOdometry's clouds, check-pointed to that file half way and resumed in
// a fresh object; its poses are printed as "lm2" and must equal "lm".
//   facade_demo <preset> <sweeps.bin> [map file]     sweeps.bin: int32 n_sweeps, then per sweep: int32 n, float64 stamp, n * 3 float32
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "loam/loam.hpp"

static void hex6(const float* v) {
  for (int i = 0; i < 6; i++) {
    unsigned int u;
    memcpy(&u, &v[i], 4);
    printf(" %08x", u);
  }
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: facade_demo <preset> <sweeps.bin>\n"); return 2; }
  FILE* f = fopen(argv[2], "rb");
  if (!f) { perror("open"); return 2; }
  int nsw = 0;
  if (fread(&nsw, 4, 1, f) != 1) return 2;
  try {
    loam::Params P = loam::Params::preset(atoi(argv[1]), /*system_delay=*/0);
    loam::ScanRegistration sr(P);
    loam::LaserOdometry lo(P);
    loam::LaserMapping lm(P);
    loam::TransformMaintenance tm;
    const char* map_path = argc > 3 ? argv[3] : NULL;
    loam::LaserMapping* lm2 = map_path ? new loam::LaserMapping(P) : NULL;
    for (int k = 0; k < nsw; k++) {
      int n = 0;
      double stamp = 0;
      if (fread(&n, 4, 1, f) != 1 || fread(&stamp, 8, 1, f) != 1) return 2;
      std::vector<float> xyz((size_t)n * 3);
      if (fread(xyz.data(), 4, xyz.size(), f) != xyz.size()) return 2;
      pcl::PointCloud<pcl::PointXYZI> sweep;   // what a /velodyne_points subscriber holds
      sweep.resize((size_t)n);
      for (int i = 0; i < n; i++) { sweep[i].x = xyz[3 * i]; sweep[i].y = xyz[3 * i + 1]; sweep[i].z = xyz[3 * i + 2]; }
      loam::Features feat;
      int src = sr.process(sweep, stamp, &feat);
      printf("%d sr=%d", k, src);
      if (src != LOAM_OK) { printf("\n"); continue; }
      printf(" n=%zu/%zu/%zu/%zu/%zu", feat.full.size(), feat.sharp.size(), feat.less_sharp.size(), feat.flat.size(), feat.less_flat.size());
      loam::OdomResult od;
      int lrc = lo.process(sr, &od);
      printf(" lo=%d", lrc);
      if (lrc == LOAM_OK) {
        hex6(od.transform_sum);
        loam::Pose6 integrated = tm.onOdometry(od.transform_sum);
        if (od.has_mapping_input) {
          loam::MapResult mr;
          int mrc = lm.process(lo, stamp, &mr);
          printf(" lm=%d", mrc);
          if (mrc == LOAM_OK) {
            hex6(mr.transform_aft_mapped);
            tm.onMapping(mr.transform_aft_mapped, mr.transform_bef_mapped);
          }
          if (lm2) {
            loam::Cloud corner, surf, full;
            lo.cloud(LOAM_CLOUD_CORNER_LAST, &corner); lo.cloud(LOAM_CLOUD_SURF_LAST, &surf); lo.cloud(LOAM_CLOUD_FULL_END, &full);
            loam::MapResult m2;
            int rc2 = lm2->process(corner, surf, full, od.transform_sum, stamp, &m2);
            printf(" lm2=%d", rc2);
            if (rc2 == LOAM_OK) hex6(m2.transform_aft_mapped);
            if (k == nsw / 2 || k == nsw / 2 + 1) {   // checkpoint, throw the node away, resume in a new one
              lm2->saveMap(map_path);
              delete lm2;
              lm2 = new loam::LaserMapping(P);
              lm2->loadMap(map_path);
            }
          }
        }
        printf(" tm");
        hex6(integrated.v);
      }
      printf("\n");
    }
    delete lm2;
  } catch (const loam::Error& e) {
    fprintf(stderr, "error %d: %s\n", e.code, e.what());
    return 1;
  }
  fclose(f);
  return 0;
}
