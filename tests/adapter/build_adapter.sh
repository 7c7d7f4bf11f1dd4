#!/bin/bash
# TEST harness.  Compiles the four ROS nodes of adapters/ros/ against the in-process ROS stand-ins
# of oracle/shims/ (ROS is not installed in this image) into node libraries with the interface
# oracle/ref_capi.cpp drives, plus a copy of that driver: tests/adapter/_build/libloam_ref.so then
# behaves like oracle/_ref/libloam_ref.so with this repository's nodes in place of the reference's.
set -euo pipefail
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
OUT=$HERE/_build
LIBDIR=$ROOT/loam_velodyne-1_b200
mkdir -p "$OUT"
CXX=${CXX:-g++}
FLAGS="-O1 -ffp-contract=off -fPIC -shared -std=c++14 -w -fvisibility=hidden -fno-gnu-unique -pthread -I$ROOT/oracle/shims -I$ROOT/oracle -I$ROOT/include -I$ROOT/adapters/ros -Dmain=ref_node_main -include $ROOT/oracle/shims/ref_shim.h"
LINK="-L$LIBDIR -lloam_b200 -Wl,-rpath,$LIBDIR"
node () {  # source define output [extra]
  cat "$ROOT/adapters/ros/$1.cpp" "$ROOT/oracle/ref_glue.inc" "$HERE/adapter_glue.inc" | $CXX $FLAGS -D$2 ${4:-} -x c++ - -o "$OUT/libref_$3.so" $LINK
}
node scan_registration_node ADAPTER_NODE_SR sr_1 -DLOAM_ADAPTER_PRESET=1 &
node scan_registration_node ADAPTER_NODE_SR sr_2 -DLOAM_ADAPTER_PRESET=2 &
node scan_registration_node ADAPTER_NODE_SR sr_4 -DLOAM_ADAPTER_PRESET=4 &
node laser_odometry_node ADAPTER_NODE_LO lo &
node laser_mapping_node ADAPTER_NODE_LM lm &
node transform_maintenance_node ADAPTER_NODE_TM tm &
wait
$CXX -O2 -fPIC -shared -std=c++14 -Wall -I"$ROOT/oracle" -I"$ROOT/include" "$ROOT/oracle/ref_capi.cpp" -o "$OUT/libloam_ref.so" -ldl -pthread
ls "$OUT"
