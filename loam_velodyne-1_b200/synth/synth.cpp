// synth.cpp -- seeded synthetic lidar sweep generator (host C++, no libm dependence in outputs).
//
// BASELINE.json configs have no bags and no network, so both the CPU oracle and the CUDA path
// consume sweeps ray-cast here.  A sweep is emitted the way /velodyne_points
// delivers it to scanRegistration.cpp:211-228: XYZ floats in the sensor frame (x fwd, y left,
// z up), firing order azimuth-major, no-return rays omitted, every ray cast from the sensor pose
// at ITS firing time (so motion distortion is physically present).  All trigonometry goes
// through loam/detmath.h and the noise is an Irwin-Hall sum of hashed uniforms, so the bytes are
// identical on every machine.
#include <stdint.h>
#include <string.h>
#include <vector>
#include <thread>
#include <algorithm>
#include "loam/detmath.h"

namespace {
using namespace loam;

struct Box { double lo[3], hi[3]; };
struct Cyl { double cx, cy, r, z0, z1; };

struct Scene {
  bool has_room = false;
  Box room;                 // sensor is inside: rays hit the inner faces
  bool has_ground = false;
  double ground_z = 0;
  std::vector<Box> boxes;
  std::vector<Cyl> cyls;
  // 1-D broad phase along x
  double bin0 = 0, binw = 8.0;
  std::vector<std::vector<int> > box_bins, cyl_bins;
  double max_range = 120.0;
  void build_bins() {
    if (boxes.empty() && cyls.empty()) return;
    double xmin = 1e30, xmax = -1e30;
    for (auto& b : boxes) { xmin = std::min(xmin, b.lo[0]); xmax = std::max(xmax, b.hi[0]); }
    for (auto& c : cyls) { xmin = std::min(xmin, c.cx - c.r); xmax = std::max(xmax, c.cx + c.r); }
    bin0 = xmin;
    int nb = (int)((xmax - xmin) / binw) + 1;
    box_bins.assign(nb, {});
    cyl_bins.assign(nb, {});
    for (int i = 0; i < (int)boxes.size(); i++) {
      int a = (int)((boxes[i].lo[0] - bin0) / binw), b = (int)((boxes[i].hi[0] - bin0) / binw);
      for (int k = std::max(a, 0); k <= std::min(b, nb - 1); k++) box_bins[k].push_back(i);
    }
    for (int i = 0; i < (int)cyls.size(); i++) {
      int a = (int)((cyls[i].cx - cyls[i].r - bin0) / binw), b = (int)((cyls[i].cx + cyls[i].r - bin0) / binw);
      for (int k = std::max(a, 0); k <= std::min(b, nb - 1); k++) cyl_bins[k].push_back(i);
    }
  }
};

struct Sensor {
  int rings = 16, az = 1800;
  std::vector<double> elev_deg;  // firing order within one azimuth step; ring id == position
  double noise_sigma = 0.01;
};

struct Config {
  int preset;   // sensor + trajectory: 1 VLP-16, 2 HDL-64E, 4 128 rings + IMU
  int scene_id = 0;  // 0 = the BASELINE.json scene of the preset; 1.. = branch-coverage scenes (VLP-16 only), see build()
  uint64_t seed;
  Sensor sensor;
  Scene scene;
  bool imu = false;
};

inline uint64_t splitmix(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

inline double gauss(uint64_t seed, uint64_t sweep, uint64_t idx) {
  uint64_t s = seed * 0xD1342543DE82EF95ull + sweep * 0x2545F4914F6CDD1Dull + idx * 0x9E3779B97F4A7C15ull + 1;
  double acc = 0;
  for (int k = 0; k < 6; k++) {
    uint64_t r = splitmix(s);
    acc += (double)(r >> 40) * (1.0 / 16777216.0);
    acc += (double)((r >> 8) & 0xFFFFFF) * (1.0 / 16777216.0);
  }
  return acc - 6.0;  // Irwin-Hall(12): mean 0, variance 1
}

const double kPi = 3.14159265358979323846;

// ground-truth sensor pose at time t: position + roll/pitch/yaw (world = velodyne axes)
void gt_pose(int preset, double t, double p[3], double rpy[3]) {
  if (preset == 1) {
    p[0] = 1.0 * t; p[1] = 0.15 * dm::dsin(0.3 * t); p[2] = 0.0;
    rpy[0] = 0; rpy[1] = 0; rpy[2] = 0.1 * dm::dsin(0.5 * t);
  } else if (preset == 4) {
    p[0] = 8.0 * t; p[1] = 1.0 * dm::dsin(0.2 * t); p[2] = 0.05 * dm::dsin(1.1 * t);
    rpy[0] = 0.02 * dm::dsin(1.3 * t); rpy[1] = 0.015 * dm::dsin(0.9 * t + 0.5); rpy[2] = 0.08 * dm::dsin(0.25 * t);
  } else {
    p[0] = 10.0 * t; p[1] = 1.0 * dm::dsin(0.2 * t); p[2] = 0.0;
    rpy[0] = 0; rpy[1] = 0; rpy[2] = 0.05 * dm::dsin(0.2 * t);
  }
}

void rot_from_rpy(const double rpy[3], double R[9]) {
  double cr = dm::dcos(rpy[0]), sr = dm::dsin(rpy[0]);
  double cp = dm::dcos(rpy[1]), sp = dm::dsin(rpy[1]);
  double cy = dm::dcos(rpy[2]), sy = dm::dsin(rpy[2]);
  R[0] = cy * cp; R[1] = cy * sp * sr - sy * cr; R[2] = cy * sp * cr + sy * sr;
  R[3] = sy * cp; R[4] = sy * sp * sr + cy * cr; R[5] = sy * sp * cr - cy * sr;
  R[6] = -sp;     R[7] = cp * sr;                R[8] = cp * cr;
}

inline bool ray_box(const double o[3], const double d[3], const Box& b, double& tnear, double& tfar) {
  double t0 = -1e300, t1 = 1e300;
  for (int a = 0; a < 3; a++) {
    if (d[a] == 0.0) {
      if (o[a] < b.lo[a] || o[a] > b.hi[a]) return false;
    } else {
      double inv = 1.0 / d[a];
      double ta = (b.lo[a] - o[a]) * inv, tb = (b.hi[a] - o[a]) * inv;
      if (ta > tb) { double s = ta; ta = tb; tb = s; }
      if (ta > t0) t0 = ta;
      if (tb < t1) t1 = tb;
      if (t0 > t1) return false;
    }
  }
  tnear = t0; tfar = t1;
  return true;
}

double cast(const Scene& sc, const double o[3], const double d[3]) {
  double best = sc.max_range;
  double tn, tf;
  if (sc.has_room && ray_box(o, d, sc.room, tn, tf) && tf > 1e-6 && tf < best) best = tf;
  if (sc.has_ground && d[2] < 0) {
    double t = (sc.ground_z - o[2]) / d[2];
    if (t > 1e-6 && t < best) best = t;
  }
  if (!sc.box_bins.empty()) {
    double xa = o[0], xb = o[0] + d[0] * best;
    if (xa > xb) { double s = xa; xa = xb; xb = s; }
    int nb = (int)sc.box_bins.size();
    int ba = (int)((xa - sc.bin0) / sc.binw), bb = (int)((xb - sc.bin0) / sc.binw);
    if (xa < sc.bin0) ba = 0;
    ba = std::max(ba, 0); bb = std::min(bb, nb - 1);
    for (int k = ba; k <= bb; k++) {
      for (int i : sc.box_bins[k])
        if (ray_box(o, d, sc.boxes[i], tn, tf) && tn > 1e-6 && tn < best) best = tn;
      for (int i : sc.cyl_bins[k]) {
        const Cyl& c = sc.cyls[i];
        double ox = o[0] - c.cx, oy = o[1] - c.cy;
        double A = d[0] * d[0] + d[1] * d[1];
        if (A < 1e-12) continue;
        double B = ox * d[0] + oy * d[1];
        double C = ox * ox + oy * oy - c.r * c.r;
        double disc = B * B - A * C;
        if (disc < 0) continue;
        double t = (-B - dm::dsqrt(disc)) / A;
        if (t > 1e-6 && t < best) {
          double z = o[2] + d[2] * t;
          if (z >= c.z0 && z <= c.z1) best = t;
        }
      }
    }
  }
  return best;
}

void add_box(Scene& s, double x0, double x1, double y0, double y1, double z0, double z1) {
  Box b;
  b.lo[0] = x0; b.hi[0] = x1; b.lo[1] = y0; b.hi[1] = y1; b.lo[2] = z0; b.hi[2] = z1;
  s.boxes.push_back(b);
}

void build(Config& c) {
  Sensor& s = c.sensor;
  Scene& sc = c.scene;
  if (c.preset == 1 && c.scene_id != 0) {
    // Scenes that drive the reference into its guard branches (laserOdometry.cpp:697-700, :770-811,
    // laserMapping.cpp:886-889, :927-954), which the BASELINE scenes never take:
    //  1 "tunnel": smooth open-ended tube, nothing along the axis within range -> translation along x
    //     is unobservable -> isDegenerate in laserOdometry and laserMapping
    //  2 "ribbed tunnel": the tube with longitudinal ribs -> large corner clouds, x still unobservable
    //  3 "sparse": ground plane seen by the three lowest rings only + groups of thin poles -> a handful of
    //     features: fewer than 10 / 50 selected rows, corner clouds of <= 10 points
    s.rings = 16; s.az = 1800; s.noise_sigma = 0.01;
    const double e[16] = {-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15};
    s.elev_deg.assign(e, e + 16);
    if (c.scene_id == 1) {
      sc.has_room = true;
      sc.room.lo[0] = -5000; sc.room.hi[0] = 5000; sc.room.lo[1] = -2; sc.room.hi[1] = 2;
      sc.room.lo[2] = -1.2; sc.room.hi[2] = 1.8;
      sc.max_range = 60.0;
    } else if (c.scene_id == 2) {
      // the same tube with longitudinal ribs (hand rail, cable tray, kerb, ceiling rail): plenty of true
      // edges, every one parallel to the axis, so x stays unobservable while the corner clouds are large
      sc.has_room = true;
      sc.room.lo[0] = -5000; sc.room.hi[0] = 5000; sc.room.lo[1] = -2; sc.room.hi[1] = 2;
      sc.room.lo[2] = -1.2; sc.room.hi[2] = 1.8;
      sc.max_range = 60.0;
      sc.binw = 500.0;
      add_box(sc, -5000, 5000, 1.8, 2.0, 0.2, 0.4);
      add_box(sc, -5000, 5000, -2.0, -1.75, -0.6, -0.35);
      add_box(sc, -5000, 5000, -0.3, 0.3, 1.6, 1.8);
      add_box(sc, -5000, 5000, 1.2, 1.5, -1.2, -1.0);
      s.noise_sigma = 0.005;
    } else {
      sc.has_ground = true; sc.ground_z = -1.2; sc.max_range = 7.5;   // -15 deg ring at 4.6 m, -13 deg at 5.3 m, -11 deg at 6.3 m
      // groups of seven thin poles: each crossing ring picks up 4+ corner points inside one 60-degree
      // sector, of which only two become "sharp" -- so the less-sharp cloud stays well above the next
      // sweep's sharp count (the reference reads laserCloudCornerLast[j] for j < cornerPointsSharpNum,
      // laserOdometry.cpp:486: past the end of the cloud otherwise, which no test may rely on)
      for (int k = 0; k < 40; k++) {
        const double cx = 3.0 + 6.0 * k, cy = (k % 2) ? -2.5 : 2.0;
        for (int q = 0; q < 7; q++) {
          Cyl p; p.cx = cx + 0.35 * q; p.cy = cy; p.r = 0.06; p.z0 = -1.2; p.z1 = 3.0; sc.cyls.push_back(p);
        }
      }
    }
  } else if (c.preset == 1) {  // VLP-16 in a closed corridor (BASELINE config 1)
    s.rings = 16; s.az = 1800; s.noise_sigma = 0.01;
    const double e[16] = {-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15};
    s.elev_deg.assign(e, e + 16);
    sc.has_room = true;
    sc.room.lo[0] = -30; sc.room.hi[0] = 50; sc.room.lo[1] = -2; sc.room.hi[1] = 2;
    sc.room.lo[2] = -1.2; sc.room.hi[2] = 1.8;
    sc.max_range = 100.0;
    for (int k = 0; k < 17; k++) {  // pilasters every 5 m on both walls
      double x = -30 + 2.5 + 5.0 * k;
      add_box(sc, x, x + 0.3, 1.7, 2.0, -1.2, 1.8);
      add_box(sc, x + 1.1, x + 1.4, -2.0, -1.7, -1.2, 1.8);
    }
    for (int k = 0; k < 8; k++) {  // cabinets / door frames
      double x = -28 + 10.0 * k;
      add_box(sc, x, x + 0.9, 1.45, 2.0, -1.2, 0.9);
      add_box(sc, x + 4.0, x + 4.6, -2.0, -1.5, -1.2, 0.3);
    }
    add_box(sc, -30.0, 50.0, -0.3, 0.3, 1.5, 1.8);  // ceiling beam along the corridor
  } else {  // urban canyon (BASELINE configs 2 and 4)
    if (c.preset == 4) {
      s.rings = 128; s.az = 2048; s.noise_sigma = 0.02;
      for (int r = 0; r < 128; r++) s.elev_deg.push_back(-25.0 + 40.0 * r / 127.0);
      c.imu = true;
    } else {
      s.rings = 64; s.az = 2000; s.noise_sigma = 0.02;
      for (int r = 0; r < 64; r++) s.elev_deg.push_back(-24.8 + 26.8 * r / 63.0);
    }
    sc.has_ground = true; sc.ground_z = -1.73; sc.max_range = 120.0;
    double x0 = -160, x1 = 700;
    // facades at y = +-8 m: solid slabs with 1 m deep recesses every 4 m (recess = gap between slabs)
    for (double x = x0; x < x1; x += 4.0) {
      add_box(sc, x, x + 2.6, 8.0, 12.0, -1.73, 14.0);
      add_box(sc, x + 2.6, x + 4.0, 9.0, 12.0, -1.73, 14.0);
      add_box(sc, x + 1.0, x + 3.4, -12.0, -8.0, -1.73, 11.0);
      add_box(sc, x + 3.4, x + 5.0, -12.0, -9.0, -1.73, 11.0);
    }
    for (double x = x0; x < x1; x += 10.0) {  // poles
      Cyl p; p.cx = x + 3.0; p.cy = 6.0; p.r = 0.2; p.z0 = -1.73; p.z1 = 6.0; sc.cyls.push_back(p);
      Cyl q; q.cx = x + 8.0; q.cy = -6.2; q.r = 0.2; q.z0 = -1.73; q.z1 = 6.0; sc.cyls.push_back(q);
    }
    for (double x = x0; x < x1; x += 14.0) {  // parked "cars" and kerb-side boxes
      add_box(sc, x, x + 4.2, 4.0, 5.8, -1.73, -0.25);
      add_box(sc, x + 7.0, x + 11.0, -5.9, -4.1, -1.73, -0.15);
      add_box(sc, x + 5.0, x + 5.8, 6.6, 7.4, -1.73, 0.4);
    }
    for (double x = x0 + 60; x < x1; x += 90.0)  // gantries crossing the street
      add_box(sc, x, x + 1.0, -8.0, 8.0, 5.0, 6.0);
  }
  sc.build_bins();
}

struct Handle { Config cfg; };

}  // namespace

extern "C" {

void* synth_create(int preset, uint64_t seed) {
  Handle* h = new Handle();
  h->cfg.preset = preset % 16;     // preset + 16 * scene
  h->cfg.scene_id = preset / 16;
  h->cfg.seed = seed;
  build(h->cfg);
  return h;
}
void synth_destroy(void* hv) { delete (Handle*)hv; }
int synth_rings(void* hv) { return ((Handle*)hv)->cfg.sensor.rings; }
int synth_azimuth_steps(void* hv) { return ((Handle*)hv)->cfg.sensor.az; }
int synth_has_imu(void* hv) { return ((Handle*)hv)->cfg.imu ? 1 : 0; }

// ground truth in the world frame (velodyne axes): pos[3], rpy[3]
void synth_gt_pose(void* hv, double t, double* pos, double* rpy) { gt_pose(((Handle*)hv)->cfg.preset, t, pos, rpy); }

// One sweep (index k, stamp = 0.1*k).  xyz: 3 floats per point.  Returns the number of points.
static int synth_sweep_impl(void* hv, int k, float* xyz, int cap, double* stamp, int nthreads, bool grid);
int synth_sweep(void* hv, int k, float* xyz, int cap, double* stamp, int nthreads) { return synth_sweep_impl(hv, k, xyz, cap, stamp, nthreads, false); }
// the same sweep as the full firing grid: azimuth-major, rings x azimuth steps slots, NaN where a ray gave no
// return (what a packetiser needs: a data packet has a slot for every firing)
int synth_sweep_grid(void* hv, int k, float* xyz, int cap, double* stamp, int nthreads) { return synth_sweep_impl(hv, k, xyz, cap, stamp, nthreads, true); }
static int synth_sweep_impl(void* hv, int k, float* xyz, int cap, double* stamp, int nthreads, bool grid) {
  const Config& c = ((Handle*)hv)->cfg;
  const Sensor& s = c.sensor;
  const int R = s.rings, A = s.az;
  const double t0 = 0.1 * k;
  if (stamp) *stamp = t0;
  std::vector<float> tmp((size_t)R * A * 3);
  std::vector<unsigned char> ok((size_t)R * A, 0);
  std::vector<double> ce(R), se(R);
  for (int r = 0; r < R; r++) {
    double e = s.elev_deg[r] * kPi / 180.0;
    ce[r] = dm::dcos(e); se[r] = dm::dsin(e);
  }
  if (nthreads < 1) nthreads = 1;
  auto work = [&](int a0, int a1) {
    for (int a = a0; a < a1; a++) {
      double phi = kPi - (kPi / A) - 2.0 * kPi * a / A;  // clockwise spin, seam behind the sensor
      double cph = dm::dcos(phi), sph = dm::dsin(phi);
      for (int r = 0; r < R; r++) {
        size_t i = (size_t)a * R + r;
        double t = t0 + 0.1 * ((double)i / ((double)A * R));
        double p[3], rpy[3], Rm[9];
        gt_pose(c.preset, t, p, rpy);
        rot_from_rpy(rpy, Rm);
        double ds[3] = {ce[r] * cph, ce[r] * sph, se[r]};
        double dw[3] = {Rm[0] * ds[0] + Rm[1] * ds[1] + Rm[2] * ds[2],
                        Rm[3] * ds[0] + Rm[4] * ds[1] + Rm[5] * ds[2],
                        Rm[6] * ds[0] + Rm[7] * ds[1] + Rm[8] * ds[2]};
        double rho = cast(c.scene, p, dw);
        if (rho >= c.scene.max_range) continue;
        rho += s.noise_sigma * gauss(c.seed, (uint64_t)k, (uint64_t)i);
        if (rho < 0.3) continue;
        tmp[3 * i] = (float)(rho * ds[0]);
        tmp[3 * i + 1] = (float)(rho * ds[1]);
        tmp[3 * i + 2] = (float)(rho * ds[2]);
        ok[i] = 1;
      }
    }
  };
  if (nthreads == 1) work(0, A);
  else {
    std::vector<std::thread> th;
    for (int w = 0; w < nthreads; w++) th.emplace_back(work, (int)((long)A * w / nthreads), (int)((long)A * (w + 1) / nthreads));
    for (auto& x : th) x.join();
  }
  int n = 0;
  if (grid) {
    const float nanv = __builtin_nanf("");
    for (size_t i = 0; i < (size_t)R * A; i++) {
      if ((int)i >= cap) break;
      if (ok[i]) { xyz[3 * i] = tmp[3 * i]; xyz[3 * i + 1] = tmp[3 * i + 1]; xyz[3 * i + 2] = tmp[3 * i + 2]; }
      else { xyz[3 * i] = nanv; xyz[3 * i + 1] = nanv; xyz[3 * i + 2] = nanv; }
    }
    return R * A;
  }
  for (size_t i = 0; i < (size_t)R * A; i++) {
    if (!ok[i]) continue;
    if (n < cap) { xyz[3 * n] = tmp[3 * i]; xyz[3 * n + 1] = tmp[3 * i + 1]; xyz[3 * n + 2] = tmp[3 * i + 2]; }
    n++;
  }
  return n;
}

// IMU sample j at 200 Hz (t = 0.005*j): orientation quaternion (x,y,z,w) of the sensor in the
// world and the specific force in the sensor frame, i.e. what sensor_msgs/Imu carries into
// scanRegistration.cpp:638-660.
void synth_imu(void* hv, int j, double* t_out, double* quat, float* acc) {
  const Config& c = ((Handle*)hv)->cfg;
  double t = 0.005 * j;
  *t_out = t;
  double p[3], rpy[3];
  gt_pose(c.preset, t, p, rpy);
  double hr = rpy[0] * 0.5, hp = rpy[1] * 0.5, hy = rpy[2] * 0.5;
  double cr = dm::dcos(hr), sr = dm::dsin(hr), cp = dm::dcos(hp), sp = dm::dsin(hp), cy = dm::dcos(hy), sy = dm::dsin(hy);
  quat[0] = sr * cp * cy - cr * sp * sy;
  quat[1] = cr * sp * cy + sr * cp * sy;
  quat[2] = cr * cp * sy - sr * sp * cy;
  quat[3] = cr * cp * cy + sr * sp * sy;
  const double h = 1e-3;
  double pm[3], p0[3], pp[3], d[3];
  gt_pose(c.preset, t - h, pm, d);
  gt_pose(c.preset, t, p0, d);
  gt_pose(c.preset, t + h, pp, d);
  double aw[3];
  for (int a = 0; a < 3; a++) aw[a] = (pp[a] - 2 * p0[a] + pm[a]) / (h * h);
  aw[2] += 9.81;  // specific force = a - g, g = (0,0,-9.81)
  double Rm[9];
  rot_from_rpy(rpy, Rm);
  for (int a = 0; a < 3; a++) acc[a] = (float)(Rm[a] * aw[0] + Rm[3 + a] * aw[1] + Rm[6 + a] * aw[2]);
}

}  // extern "C"
