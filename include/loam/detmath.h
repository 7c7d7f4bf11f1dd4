// detmath.h -- deterministic transcendental functions shared by host C++ and sm_100a device code.
//
// Why this exists: the reference calls libm sin/cos/atan/atan2/asin on
// floats (e.g. scanRegistration.cpp:248,262; laserOdometry.cpp:112-122,259-272;
// laserMapping.cpp:112-196,236-250).  glibc's sinf/cosf and CUDA's are both within 1 ulp but not
// bit-identical, and glibc even selects FMA / non-FMA variants per CPU.  "Feature and
// correspondence indices bit-exact" therefore needs ONE definition of these functions used by
// the CPU oracle and by the kernels.  This header is that definition: every function is built
// from IEEE-754 +,-,*,/ and sqrt on doubles only (no FMA contraction: host TUs are compiled
// with -ffp-contract=off, device TUs with -fmad=false), so host and device produce the same
// bits.  The float entry points evaluate in double and round once, which makes them the
// correctly rounded float result except with probability ~1e-8 per call; tests/test_detmath.py
// measures the distance to glibc (<= 1 ulp, equal in the overwhelming majority of cases).
//
// Polynomials are the classic fdlibm (Sun Microsystems, 1993, freely distributable) kernels
// __kernel_sin / __kernel_cos / atan, restated.
#pragma once
#include <math.h>

#if defined(__FAST_MATH__)
#error "detmath.h must not be compiled with fast-math"
#endif

#if defined(__CUDACC__)
#define LOAM_HD __host__ __device__ __forceinline__
#else
#define LOAM_HD inline
#endif

namespace loam {
namespace dm {

// round-to-nearest-even integer of v (exact operation on both sides: roundsd / cvt.rni.f64)
LOAM_HD double round_even(double v) {
#if defined(__CUDA_ARCH__)
  return ::rint(v);
#else
  return __builtin_rint(v);
#endif
}

LOAM_HD double kernel_sin(double x) {
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
               S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
               S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  double z = x * x;
  double v = z * x;
  double r = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
  return x + v * (S1 + z * r);
}

LOAM_HD double kernel_cos(double x) {
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
               C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
               C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double z = x * x;
  double r = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
  return 1.0 - (0.5 * z - z * r);
}

// Cody-Waite reduction to [-pi/4, pi/4]; returns quadrant in *q (0..3)
LOAM_HD double reduce_pio2(double x, int* q) {
  const double INVPIO2 = 6.36619772367581382433e-01;
  const double PIO2_1 = 1.57079632673412561417e+00;   // first 33 bits of pi/2
  const double PIO2_1T = 6.07710050650619224932e-11;  // pi/2 - PIO2_1
  double k = round_even(x * INVPIO2);
  double r = (x - k * PIO2_1) - k * PIO2_1T;
  long long ki = (long long)k;
  *q = (int)(ki & 3);
  return r;
}

LOAM_HD double dsin(double x) {
  int q;
  double r = reduce_pio2(x, &q);
  switch (q) {
    case 0: return kernel_sin(r);
    case 1: return kernel_cos(r);
    case 2: return -kernel_sin(r);
    default: return -kernel_cos(r);
  }
}

LOAM_HD double dcos(double x) {
  int q;
  double r = reduce_pio2(x, &q);
  switch (q) {
    case 0: return kernel_cos(r);
    case 1: return -kernel_sin(r);
    case 2: return -kernel_cos(r);
    default: return kernel_sin(r);
  }
}

LOAM_HD double dabs(double x) { return x < 0 ? -x : x; }

LOAM_HD double datan(double x) {
  const double hi[4] = {4.63647609000806093515e-01, 7.85398163397448278999e-01,
                        9.82793723247329054082e-01, 1.57079632679489655800e+00};
  const double lo[4] = {2.26987774529616870924e-17, 3.06161699786838301793e-17,
                        1.39033110312309984516e-17, 6.12323399573676603587e-17};
  const double aT[11] = {3.33333333333329318027e-01,  -1.99999999998764832476e-01,
                         1.42857142725034663711e-01,  -1.11111104054623557880e-01,
                         9.09088713343650656196e-02,  -7.69187620504482999495e-02,
                         6.66107313738753120669e-02,  -5.83357013379057348645e-02,
                         4.97687799461593236017e-02,  -3.65315727442169155270e-02,
                         1.62858201153657823623e-02};
  if (x != x) return x;
  bool neg = x < 0;
  double ax = dabs(x);
  int id;
  if (ax >= 7.3786976294838206464e19) {  // 2^66
    double z = hi[3] + lo[3];
    return neg ? -z : z;
  }
  if (ax < 0.4375) {
    if (ax < 1.862645149230957e-09) return x;  // 2^-29
    id = -1;
  } else if (ax < 1.1875) {
    if (ax < 0.6875) { id = 0; ax = (2.0 * ax - 1.0) / (2.0 + ax); }
    else             { id = 1; ax = (ax - 1.0) / (ax + 1.0); }
  } else {
    if (ax < 2.4375) { id = 2; ax = (ax - 1.5) / (1.0 + 1.5 * ax); }
    else             { id = 3; ax = -1.0 / ax; }
  }
  double z = ax * ax;
  double w = z * z;
  double s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
  double s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
  if (id < 0) {
    double r = ax - ax * (s1 + s2);
    return neg ? -r : r;
  }
  double r = hi[id] - ((ax * (s1 + s2) - lo[id]) - ax);
  return neg ? -r : r;
}

LOAM_HD double datan2(double y, double x) {
  const double PI = 3.1415926535897931160E+00, PI_LO = 1.2246467991473531772E-16;
  if (x != x || y != y) return x + y;
  if (y == 0.0) {
    if (x > 0.0 || (x == 0.0 && 1.0 / x > 0.0)) return y;  // +-0 for x >= +0
    return (1.0 / y < 0.0) ? -PI : PI;                      // x < 0 or x == -0
  }
  if (x == 0.0) return y < 0.0 ? -0.5 * PI : 0.5 * PI;
  double z = datan(dabs(y / x));
  if (x > 0.0) return y < 0.0 ? -z : z;
  return y < 0.0 ? (z - PI_LO) - PI : PI - (z - PI_LO);
}

LOAM_HD double dsqrt(double v) {
#if defined(__CUDA_ARCH__)
  return __dsqrt_rn(v);
#else
  return __builtin_sqrt(v);
#endif
}

LOAM_HD double dasin(double x) {
  if (x != x) return x;
  if (x > 1.0 || x < -1.0) return (x - x) / (x - x);  // NaN, as libm
  return datan2(x, dsqrt((1.0 - x) * (1.0 + x)));
}

// ---- float entry points: what the reference's float overloads of <math.h> are defined as ----
LOAM_HD float sin(float x) { return (float)dsin((double)x); }
LOAM_HD float cos(float x) { return (float)dcos((double)x); }
LOAM_HD float atan(float x) { return (float)datan((double)x); }
LOAM_HD float atan2(float y, float x) { return (float)datan2((double)y, (double)x); }
LOAM_HD float asin(float x) { return (float)dasin((double)x); }
LOAM_HD float sqrt(float v) {
#if defined(__CUDA_ARCH__)
  return __fsqrt_rn(v);
#else
  return __builtin_sqrtf(v);
#endif
}
LOAM_HD float fabs(float v) { return ::fabsf(v); }

// ---- double overloads (IMU handler, tf quaternion/RPY hand-off run in double) ----
LOAM_HD double sin(double x) { return dsin(x); }
LOAM_HD double cos(double x) { return dcos(x); }
LOAM_HD double atan(double x) { return datan(x); }
LOAM_HD double atan2(double y, double x) { return datan2(y, x); }
LOAM_HD double asin(double x) { return dasin(x); }
LOAM_HD double sqrt(double v) { return dsqrt(v); }
LOAM_HD double fabs(double v) { return dabs(v); }

}  // namespace dm
}  // namespace loam
