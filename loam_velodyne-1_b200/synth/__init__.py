"""Seeded synthetic lidar sweeps (ctypes wrapper over synth.cpp).

BASELINE.json configs: preset 1 = VLP-16 corridor (seed 1234), 2 = HDL-64E urban canyon (seed
2345), 4 = 128-ring + IMU (seed 3456).  Output = what /velodyne_points carries into
scanRegistration.cpp:211: (n, 3) float32 XYZ in the sensor frame, firing order.
"""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libloam_synth.so")
DEFAULT_SEED = {1: 1234, 2: 2345, 4: 3456}


def build():
    inc = os.path.join(_HERE, "..", "..", "include")
    cmd = ["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c++14", "-I", inc,
           os.path.join(_HERE, "synth.cpp"), "-o", _LIB, "-lpthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("synth build failed:\n" + r.stderr)


_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = ctypes.CDLL(_LIB)
        L.synth_create.restype = ctypes.c_void_p
        L.synth_create.argtypes = [ctypes.c_int, ctypes.c_uint64]
        L.synth_destroy.argtypes = [ctypes.c_void_p]
        L.synth_rings.argtypes = [ctypes.c_void_p]
        L.synth_azimuth_steps.argtypes = [ctypes.c_void_p]
        L.synth_has_imu.argtypes = [ctypes.c_void_p]
        L.synth_gt_pose.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p]
        L.synth_sweep.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                  ctypes.c_void_p, ctypes.c_int]
        L.synth_imu.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


class Synth:
    def __init__(self, preset, seed=None, threads=None):
        self.L = _load()
        self.preset = preset
        self.seed = DEFAULT_SEED[preset] if seed is None else seed
        self.h = self.L.synth_create(preset, self.seed)
        self.rings = self.L.synth_rings(self.h)
        self.az = self.L.synth_azimuth_steps(self.h)
        self.has_imu = bool(self.L.synth_has_imu(self.h))
        self.threads = threads or min(os.cpu_count() or 1, 16)
        self._buf = np.empty((self.rings * self.az, 3), dtype=np.float32)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.synth_destroy(self.h)
            self.h = None

    def sweep(self, k):
        """returns (xyz (n,3) float32, stamp)"""
        stamp = ctypes.c_double()
        n = self.L.synth_sweep(self.h, k, self._buf.ctypes.data, self._buf.shape[0], ctypes.byref(stamp), self.threads)
        return self._buf[:n].copy(), stamp.value

    def imu(self, j):
        """IMU sample j at 200 Hz -> (t, quat xyzw float64[4], acc float32[3])"""
        t = ctypes.c_double()
        q = np.zeros(4, dtype=np.float64)
        a = np.zeros(3, dtype=np.float32)
        self.L.synth_imu(self.h, j, ctypes.byref(t), q.ctypes.data, a.ctypes.data)
        return t.value, q, a

    def gt_pose(self, t):
        p = np.zeros(3)
        r = np.zeros(3)
        self.L.synth_gt_pose(self.h, float(t), p.ctypes.data, r.ctypes.data)
        return p, r
