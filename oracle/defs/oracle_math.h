// ORACLE (test infrastructure).  Which libm the oracle uses.
// Default: loam/detmath.h (deterministic, identical bits on host and device -- the project's
// adopted definition of the reference's <math.h> calls).
// -DORACLE_LIBM: glibc, i.e. exactly what the reference binary would call on this machine; used
// by tests/test_oracle_libm.py to bound the effect of the substitution (poses within the
// north_star tolerance, feature sets compared).
#pragma once
#include <math.h>
#include <cmath>
#include "loam/detmath.h"

namespace omath {
#ifdef ORACLE_LIBM
inline float sin(float x) { return ::sinf(x); }
inline float cos(float x) { return ::cosf(x); }
inline float atan(float x) { return ::atanf(x); }
inline float atan2(float y, float x) { return ::atan2f(y, x); }
inline float asin(float x) { return ::asinf(x); }
inline double sin(double x) { return ::sin(x); }
inline double cos(double x) { return ::cos(x); }
inline double atan(double x) { return ::atan(x); }
inline double atan2(double y, double x) { return ::atan2(y, x); }
inline double asin(double x) { return ::asin(x); }
#else
using loam::dm::sin;
using loam::dm::cos;
using loam::dm::atan;
using loam::dm::atan2;
using loam::dm::asin;
#endif
inline float sqrt(float x) { return ::sqrtf(x); }
inline double sqrt(double x) { return ::sqrt(x); }
inline float fabs(float x) { return ::fabsf(x); }
inline double fabs(double x) { return ::fabs(x); }
// integer arguments promote to double as in <cmath>
inline double sqrt(int x) { return ::sqrt((double)x); }
// pow(x, 2) as used by the convergence tests (laserOdometry.cpp:815-822): exact square in double
inline double pow(double x, int e) {
  double r = 1.0;
  for (int i = 0; i < e; i++) r *= x;
  return r;
}
}  // namespace omath
