"""The C++ class facade (include/loam/loam.hpp) over the C ABI: compiles and links everywhere;
on a GPU it must reproduce the ctypes Pipeline bit for bit when driven node by node."""
import importlib
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "loam_velodyne-1_b200"
SRC = os.path.join(ROOT, "tests", "cpp", "facade_demo.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "facade_demo")


def _build(pkg):
    if not os.path.exists(pkg.LIB_PATH):
        importlib.import_module(PKG + ".build").build_cuda()
    libdir = os.path.dirname(pkg.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-DLOAM_FORCE_PCL_MIN", "-I", os.path.join(ROOT, "include"), SRC, "-o", EXE,
           "-L", libdir, "-lloam_b200", "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return EXE


def test_facade_compiles_and_links(pkg):
    exe = _build(pkg)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr   # loads libloam_b200.so, no compute call


def test_pcl_min_layout():
    """the in-repo PCL stand-in must keep PCL's ABI: PointXYZ 16 B, PointXYZI 32 B, 16-aligned"""
    src = '#include "loam/pcl_min.h"\nstatic_assert(sizeof(pcl::PointXYZ)==16 && alignof(pcl::PointXYZ)==16,"");\n' \
          'static_assert(sizeof(pcl::PointXYZI)==32 && alignof(pcl::PointXYZI)==16 && offsetof(pcl::PointXYZI,intensity)==16,"");\nint main(){return 0;}\n'
    r = subprocess.run(["g++", "-std=c++17", "-DLOAM_FORCE_PCL_MIN", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), "-x", "c++", "-"],
                       input=src, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.gpu
def test_facade_node_by_node_equals_pipeline(pkg, synth, tmp_path):
    exe = _build(pkg)
    preset, nsw = 1, 12
    S = synth.Synth(preset)
    sweeps = [S.sweep(k) for k in range(nsw)]
    path = os.path.join(tmp_path, "sweeps.bin")
    with open(path, "wb") as f:
        f.write(struct.pack("<i", nsw))
        for xyz, st in sweeps:
            f.write(struct.pack("<id", xyz.shape[0], st))
            f.write(np.ascontiguousarray(xyz, np.float32).tobytes())
    r = subprocess.run([exe, str(preset), path, os.path.join(tmp_path, "facade.loammap")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    assert len(lines) == nsw
    G = pkg.Pipeline(pkg.make_params(preset, system_delay=0))
    hexes = lambda v: " ".join("%08x" % x for x in np.array(v[:], np.float32).view(np.uint32))
    mapped = 0
    for k, (xyz, st) in enumerate(sweeps):
        g = G.push(xyz, st)
        line = lines[k]
        assert line.startswith(f"{k} sr={g.sr_status}")
        if g.sr_status != 0:
            continue
        assert f" n={g.n_full}/{g.n_sharp}/{g.n_less_sharp}/{g.n_flat}/{g.n_less_flat}" in line, (k, line)
        assert f" lo={g.lo_status}" in line
        if g.lo_status == 0:
            assert f" lo=0 {hexes(g.transform_sum)}" in line, (k, line)
            assert f" tm {hexes(g.transform_mapped)}" in line, (k, line)
            if g.lm_status == 0:
                mapped += 1
                assert f" lm=0 {hexes(g.transform_aft_mapped)}" in line, (k, line)
                # the second LaserMapping: host clouds in, check-pointed to a file and resumed half way
                assert f" lm2=0 {hexes(g.transform_aft_mapped)}" in line, (k, line)
    assert mapped >= 4
    assert os.path.getsize(os.path.join(tmp_path, "facade.loammap")) > 4851 * 10


def test_ros_adapter_nodes_compile_against_the_ros_stand_ins(pkg):
    """adapters/ros/*.cpp build (and link against libloam_b200.so) with the oracle's ROS / tf / pcl_conversions
    stand-ins; running them needs a GPU (tests/test_gpu_ros_adapter.py)"""
    if not os.path.exists(pkg.LIB_PATH):
        importlib.import_module(PKG + ".build").build_cuda()
    r = subprocess.run(["bash", os.path.join(ROOT, "tests", "adapter", "build_adapter.sh")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    for name in ("libref_sr_1.so", "libref_lo.so", "libref_lm.so", "libref_tm.so", "libloam_ref.so"):
        assert os.path.exists(os.path.join(ROOT, "tests", "adapter", "_build", name))
