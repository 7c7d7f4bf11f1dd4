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
        L.synth_sweep_grid.argtypes = L.synth_sweep.argtypes
        L.synth_imu.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


class Synth:
    def __init__(self, preset, seed=None, threads=None, scene=0):
        """scene 0 = the BASELINE.json scene of the preset; preset 1 also has the branch-coverage scenes
        1 "tunnel", 2 "plain", 3 "sparse" (synth.cpp build())"""
        self.L = _load()
        self.preset = preset
        self.scene = scene
        self.seed = DEFAULT_SEED[preset] if seed is None else seed
        self.h = self.L.synth_create(preset + 16 * scene, self.seed)
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

    def sweep_grid(self, k):
        """the same sweep as the full firing grid: (azimuth steps, rings, 3) float32, NaN = no return"""
        stamp = ctypes.c_double()
        buf = np.empty((self.az, self.rings, 3), dtype=np.float32)
        self.L.synth_sweep_grid(self.h, k, buf.ctypes.data, self.az * self.rings, ctypes.byref(stamp), self.threads)
        return buf, stamp.value

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


# ---------------------------------------------------------------------------------------------
# Velodyne VLP-16 data packets from a synthetic sweep (raw-packet front end)
# ---------------------------------------------------------------------------------------------
VLP16_PACKET_BYTES = 1206
VLP16_LASER_DEG = np.array([-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15], np.float64)


def vlp16_packets(xyz, stamp=0.0, drop=None):
    """Encodes a preset-1 sweep (16 lasers x an even number of azimuth steps, every ray returned,
    firing order as the generator emits it) as the sensor's own UDP payloads: (n_packets, 1206) uint8.
    One generator azimuth step = one firing sequence; two per data block; 12 blocks per packet
    (the last packet is padded with no-return blocks that continue the azimuth ramp).
    drop: optional boolean mask (n,) of returns to encode as "no return" (distance 0)."""
    xyz = np.asarray(xyz, np.float64)
    n = xyz.shape[0]
    assert n % 32 == 0, "needs an even number of 16-laser firing sequences"
    steps = n // 16
    d = np.sqrt((xyz ** 2).sum(axis=1))
    raw = np.clip(np.rint(d / 0.002), 1, 65535).astype(np.uint16)
    if drop is not None:
        raw = np.where(np.asarray(drop, bool), np.uint16(0), raw)
    first = xyz.reshape(steps, 16, 3)[:, 0, :]
    az = np.degrees(np.arctan2(-first[:, 1], first[:, 0])) % 360.0          # clockwise from +x (forward)
    az_blocks = np.rint(az[0::2] * 100.0).astype(np.int64) % 36000
    nblk = steps // 2
    npk = (nblk + 11) // 12
    step_az = int((az_blocks[1] - az_blocks[0]) % 36000) if nblk > 1 else 40
    pk = np.zeros((npk, VLP16_PACKET_BYTES), np.uint8)
    raw_b = raw.reshape(nblk, 32)
    for b in range(npk * 12):
        p, k = divmod(b, 12)
        o = 100 * k
        pk[p, o], pk[p, o + 1] = 0xFF, 0xEE
        a = int(az_blocks[b]) if b < nblk else int((az_blocks[-1] + step_az * (b - nblk + 1)) % 36000)
        pk[p, o + 2], pk[p, o + 3] = a & 0xFF, a >> 8
        if b < nblk:
            body = np.zeros((32, 3), np.uint8)
            body[:, 0] = raw_b[b] & 0xFF
            body[:, 1] = raw_b[b] >> 8
            body[:, 2] = (np.arange(32) * 7 + b) & 0xFF      # reflectivity: unused by LOAM
            pk[p, o + 4:o + 100] = body.ravel()
    t_us = int(stamp * 1e6)
    for p in range(npk):
        t = (t_us + int(p * 12 * 110.592)) & 0xFFFFFFFF
        pk[p, 1200:1204] = [t & 0xFF, (t >> 8) & 0xFF, (t >> 16) & 0xFF, (t >> 24) & 0xFF]
        pk[p, 1204], pk[p, 1205] = 0x37, 0x22                 # strongest return, VLP-16
    return pk


def hdl64_packets(grid, stamp=0.0):
    """Encodes a preset-2 sweep given as the full firing grid (Synth.sweep_grid: (steps, 64, 3), ring r at
    position r, NaN = no return) as HDL-64E data packets: (n_packets, 1206) uint8.  One azimuth step = one
    firing of the 64 lasers = an upper block (flag 0xEEFF, lasers 32..63) followed by a lower block (0xDDFF,
    lasers 0..31) with the same azimuth; 12 blocks = 6 firings per packet; the last packet is padded with
    empty blocks that continue the azimuth ramp."""
    g = np.asarray(grid, np.float64)
    steps, R = g.shape[0], g.shape[1]
    assert R == 64
    d = np.sqrt((g ** 2).sum(axis=2))
    valid = np.isfinite(d)
    raw = np.where(valid, np.clip(np.rint(np.where(valid, d, 0.0) / 0.002), 1, 65535), 0).astype(np.uint16)
    a = np.arange(steps)
    phi = np.pi - np.pi / steps - 2.0 * np.pi * a / steps          # synth.cpp: clockwise spin, seam behind the sensor
    az = np.rint((np.degrees(-phi) % 360.0) * 100.0).astype(np.int64) % 36000
    nblk = 2 * steps
    npk = (nblk + 11) // 12
    step_az = int((az[1] - az[0]) % 36000) if steps > 1 else 18
    pk = np.zeros((npk, VLP16_PACKET_BYTES), np.uint8)
    for b in range(npk * 12):
        p, k = divmod(b, 12)
        o = 100 * k
        st, upper = divmod(b, 2)
        upper = upper == 0
        pk[p, o], pk[p, o + 1] = 0xFF, (0xEE if upper else 0xDD)
        aa = int(az[st]) if st < steps else int((az[-1] + step_az * (st - steps + 1)) % 36000)
        pk[p, o + 2], pk[p, o + 3] = aa & 0xFF, aa >> 8
        if st < steps:
            rr = raw[st, 32:64] if upper else raw[st, 0:32]
            body = np.zeros((32, 3), np.uint8)
            body[:, 0] = rr & 0xFF
            body[:, 1] = rr >> 8
            body[:, 2] = (np.arange(32) * 5 + b) & 0xFF        # reflectivity: unused by LOAM
            pk[p, o + 4:o + 100] = body.ravel()
    t_us = int(stamp * 1e6)
    for p in range(npk):
        t = (t_us + int(p * 6 * 48.0)) & 0xFFFFFFFF
        pk[p, 1200:1204] = [t & 0xFF, (t >> 8) & 0xFF, (t >> 16) & 0xFF, (t >> 24) & 0xFF]
        pk[p, 1204], pk[p, 1205] = 0x37, 0x21                     # strongest return, HDL-64E
    return pk
