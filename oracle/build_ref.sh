#!/bin/bash
# ORACLE (test infrastructure).  Compiles the reference's OWN node sources, read where they lie
# under /root/reference, into oracle/_ref/*.so.  The sources are never copied into the repo: a
# scripted patch (sed) is applied on the fly and the result is piped straight into g++.
# Patch, limited to what BASELINE.json's configs need to vary:
#   scanRegistration.cpp:61      const int N_SCANS = 16      -> REF_N_SCANS
#   scanRegistration.cpp:57      const int systemDelay = 20  -> runtime int (benchmarks use 0)
#   *.cpp                        [40000] array capacity      -> [REF_MAX_POINTS]
#   scanRegistration.cpp:250-256 ring-id block               -> #if REF_RING_MODEL==0 <unchanged>
#                                #else <the HDL-64E block of the bk variant, bounds as macros>
#   main                         -> ref_node_main (-Dmain=...)
# Everything else -- every numeric statement cited in SURVEY.md §8a -- is the reference's text.
set -euo pipefail
REF=${1:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
mkdir -p "$OUT"
CXX=${CXX:-g++}
FLAGS="-O2 -ffp-contract=off -fPIC -shared -std=c++14 -w -fvisibility=hidden -fno-gnu-unique -pthread -I$HERE/shims -I$HERE -I$HERE/../include -I$REF/include -Dmain=ref_node_main -include $HERE/shims/ref_shim.h"

# sanity: the lines the patch addresses must still read as expected
sed -n 250p "$REF/src/scanRegistration.cpp" | grep -q 'int roundedAngle = int(angle + (angle<0.0?-0.5:+0.5));'
sed -n 256p "$REF/src/scanRegistration.cpp" | grep -q '^    }$'
grep -q 'const int N_SCANS = 16;' "$REF/src/scanRegistration.cpp"

RING_ELSE='#else\n    { const float angleLowerBoundDeg = REF_RING_LOWER; const float angleUpperBoundDeg = REF_RING_UPPER;\n      const float angleSpan = angleUpperBoundDeg - angleLowerBoundDeg; const float angleStep = angleSpan / (N_SCANS - 1);\n      float angleID = (angle - angleLowerBoundDeg) / angleStep; scanID = int(angleID + 0.5f); }\n#endif'

build_sr () {  # name n_scans ring_model lower upper
  sed -e 's/const int N_SCANS = 16;/const int N_SCANS = REF_N_SCANS;/' \
      -e 's/const int systemDelay = 20;/int systemDelay = 20;/' \
      -e 's/\[40000\]/[REF_MAX_POINTS]/g' \
      -e '250i #if REF_RING_MODEL == 0' \
      -e "256a $RING_ELSE" \
      "$REF/src/scanRegistration.cpp" | cat - "$HERE/ref_glue.inc" | \
  $CXX $FLAGS $EXTRA -DREF_NODE_SR -DREF_N_SCANS=$2 -DREF_RING_MODEL=$3 -DREF_RING_LOWER=$4 -DREF_RING_UPPER=$5 -DREF_MAX_POINTS=300000 -x c++ - -o "$OUT/libref_sr_$1$SUFFIX.so"
}
build_node () {  # file define name
  sed -e 's/\[40000\]/[REF_MAX_POINTS]/g' "$REF/src/$1.cpp" | cat - "$HERE/ref_glue.inc" | \
  $CXX $FLAGS $EXTRA -D$2 -DREF_MAX_POINTS=300000 -x c++ - -o "$OUT/libref_$3$SUFFIX.so"
}
for variant in det libm; do
  if [ $variant = libm ]; then EXTRA="-DORACLE_LIBM"; SUFFIX="_libm"; else EXTRA=""; SUFFIX=""; fi
  build_sr 1 16 0 0.0f 0.0f &
  build_sr 2 64 1 -24.8f 2.0f &
  build_sr 4 128 1 -25.0f 15.0f &
  build_node laserOdometry REF_NODE_LO lo &
  build_node laserMapping REF_NODE_LM lm &
  build_node transformMaintenance REF_NODE_TM tm &
  wait
done
$CXX -O2 -fPIC -shared -std=c++14 -Wall -I$HERE -I$HERE/../include "$HERE/ref_capi.cpp" -o "$OUT/libloam_ref.so" -ldl -pthread
$CXX -O2 -fPIC -shared -std=c++14 -Wall -DREF_LIBM -I$HERE -I$HERE/../include "$HERE/ref_capi.cpp" -o "$OUT/libloam_ref_libm.so" -ldl -pthread
ls -la "$OUT"
