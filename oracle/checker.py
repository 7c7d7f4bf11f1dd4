"""ORACLE (test infrastructure): ctypes wrapper over the checker API (oracle/checker_api.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this.  kind = "port" (CPU restatement), "port-libm" (same with glibc trig), "reference"
(oracle/_ref/libloam_ref.so: the reference's own sources compiled from /root/reference) or
"reference-libm" (the same sources with glibc's trig instead of loam/detmath.h).
"""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class LoamParams(ctypes.Structure):
    _fields_ = [("n_scans", ctypes.c_int), ("ring_model", ctypes.c_int),
                ("ring_lower_deg", ctypes.c_float), ("ring_upper_deg", ctypes.c_float),
                ("max_points", ctypes.c_int), ("system_delay", ctypes.c_int),
                ("lo_max_iter", ctypes.c_int), ("lm_max_iter", ctypes.c_int),
                ("map_arena_points", ctypes.c_int), ("coop_ctas", ctypes.c_int),
                ("dev_flags", ctypes.c_int), ("packet_metadata", ctypes.c_int), ("reserved", ctypes.c_int * 4)]


class SweepResult(ctypes.Structure):
    _fields_ = [("sr_status", ctypes.c_int), ("lo_status", ctypes.c_int), ("lm_status", ctypes.c_int),
                ("n_full", ctypes.c_int), ("n_sharp", ctypes.c_int), ("n_less_sharp", ctypes.c_int),
                ("n_flat", ctypes.c_int), ("n_less_flat", ctypes.c_int),
                ("lo_iters", ctypes.c_int), ("lo_rows", ctypes.c_int), ("lo_degenerate", ctypes.c_int),
                ("lm_iters", ctypes.c_int), ("lm_rows", ctypes.c_int), ("lm_degenerate", ctypes.c_int),
                ("n_map_corner", ctypes.c_int), ("n_map_surf", ctypes.c_int),
                ("n_stack_corner", ctypes.c_int), ("n_stack_surf", ctypes.c_int),
                ("transform", ctypes.c_float * 6), ("transform_sum", ctypes.c_float * 6),
                ("transform_aft_mapped", ctypes.c_float * 6), ("transform_bef_mapped", ctypes.c_float * 6),
                ("transform_mapped", ctypes.c_float * 6), ("imu_trans", ctypes.c_float * 12)]

    def as_dict(self):
        d = {}
        for name, typ in self._fields_:
            v = getattr(self, name)
            d[name] = np.array(v[:], dtype=np.float32) if hasattr(v, "__len__") else v
        return d


def make_params(preset):
    """BASELINE.json configs: 1 = VLP-16 (reference literals), 2 = HDL-64E, 4 = 128-ring."""
    p = LoamParams()
    p.system_delay = 0
    p.lo_max_iter = 25
    p.lm_max_iter = 10
    p.map_arena_points = 0
    if preset == 1:
        p.n_scans, p.ring_model, p.ring_lower_deg, p.ring_upper_deg, p.max_points = 16, 0, -15.0, 15.0, 40000
    elif preset == 2:
        p.n_scans, p.ring_model, p.ring_lower_deg, p.ring_upper_deg, p.max_points = 64, 1, -24.8, 2.0, 160000
    elif preset == 4:
        p.n_scans, p.ring_model, p.ring_lower_deg, p.ring_upper_deg, p.max_points = 128, 1, -25.0, 15.0, 300000
    else:
        raise ValueError(preset)
    return p


def build(verbose=False):
    """compile the CPU restatement (and oracle/_ref when /root/reference is mounted)"""
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if os.path.isdir("/root/reference/src") and os.path.exists(os.path.join(_HERE, "build_ref.sh")):
        r = subprocess.run(["make", "-C", _HERE, "ref"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("oracle/_ref build failed:\n" + r.stdout + r.stderr)


def lib_path(kind):
    if kind.endswith(".so"):   # a library with the checker API built elsewhere (tests/adapter: the ROS adapter nodes)
        return kind
    return {"port": os.path.join(_HERE, "liboracle_port.so"),
            "port-libm": os.path.join(_HERE, "liboracle_port_libm.so"),
            "reference": os.path.join(_HERE, "_ref", "libloam_ref.so"),
            "reference-libm": os.path.join(_HERE, "_ref", "libloam_ref_libm.so")}[kind]


def available(kind):
    return os.path.exists(lib_path(kind))


_libs = {}


def _load(kind):
    if kind in _libs:
        return _libs[kind]
    path = lib_path(kind)
    if not os.path.exists(path):
        if not kind.startswith("reference"):
            build()
        if not os.path.exists(path):
            raise FileNotFoundError(path)
    if kind.endswith(".so") and not os.path.exists(path):
        raise FileNotFoundError(path)
    L = ctypes.CDLL(path)
    L.chk_kind.restype = ctypes.c_char_p
    L.chk_create.restype = ctypes.c_void_p
    L.chk_create.argtypes = [ctypes.POINTER(LoamParams)]
    L.chk_destroy.argtypes = [ctypes.c_void_p]
    L.chk_imu.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p]
    L.chk_push_sweep.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                 ctypes.POINTER(SweepResult)]
    L.chk_get_cloud.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    L.chk_get_ints.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    L.chk_get_floats.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    L.chk_busy_seconds.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    L.chk_lo_set_state.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    if hasattr(L, "chk_lo_push"):   # not in the ROS-adapter driver (tests/adapter)
        L.chk_lo_push.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p, ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_double, ctypes.POINTER(SweepResult)]
    if hasattr(L, "chk_run_pipelined"):   # port and reference builds: the nodes on three threads
        L.chk_run_pipelined.restype = ctypes.c_double
        L.chk_run_pipelined.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    if hasattr(L, "chk_lm_push"):   # the CPU restatement only
        L.chk_lm_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                  ctypes.c_int, ctypes.c_void_p, ctypes.c_double, ctypes.POINTER(SweepResult)]
        L.chk_lm_save_map.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.chk_lm_load_map.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.chk_run_pipelined.restype = ctypes.c_double
        L.chk_run_pipelined.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.chk_decode_packets.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.chk_packet_front_end.argtypes = [ctypes.POINTER(LoamParams), ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.chk_push_packets.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.POINTER(SweepResult)]
        L.chk_solve_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    _libs[kind] = L
    return L


def decode_packets(packets, kind="port"):
    """CPU restatement of the VLP-16 packet decoder (oracle/packet_port.hpp) -> (n_packets * 384, 3) float32"""
    pk = np.ascontiguousarray(packets, dtype=np.uint8)
    assert pk.ndim == 2 and pk.shape[1] == 1206
    out = np.empty((pk.shape[0] * 384, 3), np.float32)
    _load(kind).chk_decode_packets(pk.ctypes.data, pk.shape[0], out.ctypes.data)
    return out


def packet_front_end(params, packets, metadata=False, kind="port"):
    """CPU restatement of the packet front end of the sensor `params` describes -> xyz (n, 3) and, with
    metadata=True, the ring ids and relTime of the packet-metadata SR-1 (oracle/packet_port.hpp)"""
    pk = np.ascontiguousarray(packets, dtype=np.uint8)
    assert pk.ndim == 2 and pk.shape[1] == 1206
    n = pk.shape[0] * 384
    xyz = np.empty((n, 3), np.float32)
    ring = np.empty(n, np.int32) if metadata else None
    rt = np.empty(n, np.float32) if metadata else None
    rc = _load(kind).chk_packet_front_end(ctypes.byref(params), pk.ctypes.data, pk.shape[0], xyz.ctypes.data,
                                          ring.ctypes.data if metadata else None, rt.ctypes.data if metadata else None)
    assert rc == 0
    return (xyz, ring, rt) if metadata else xyz


class Checker:
    """the four reference nodes wired in-process: sr -> lo -> tm -> (lm -> tm)"""

    def __init__(self, params, kind="port"):
        self.L = _load(kind)
        self.kind = kind
        self.params = params
        self.h = self.L.chk_create(ctypes.byref(params))
        self.cap = max(params.max_points, 1 << 16)

    def close(self):
        if self.h:
            self.L.chk_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def imu(self, stamp, quat, acc):
        q = np.ascontiguousarray(quat, dtype=np.float64)
        a = np.ascontiguousarray(acc, dtype=np.float32)
        self.L.chk_imu(self.h, float(stamp), q.ctypes.data, a.ctypes.data)

    def push(self, xyz, stamp):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        res = SweepResult()
        self.L.chk_push_sweep(self.h, xyz.ctypes.data, xyz.shape[0], float(stamp), ctypes.byref(res))
        return res

    def push_packets(self, packets, stamp, metadata=False):
        """one revolution of data packets (VLP-16 / HDL-64E by the handle's sensor); metadata = loam_params.packet_metadata"""
        pk = np.ascontiguousarray(packets, dtype=np.uint8)
        res = SweepResult()
        self.L.chk_push_packets(self.h, pk.ctypes.data, pk.shape[0], 1 if metadata else 0, float(stamp), ctypes.byref(res))
        return res

    def cloud(self, which, cap=None):
        cap = cap or 4 * self.cap
        buf = np.empty((cap, 4), dtype=np.float32)
        n = self.L.chk_get_cloud(self.h, which, buf.ctypes.data, cap)
        if n > cap:
            return self.cloud(which, n)
        return buf[:max(n, 0)].copy()

    def ints(self, which, cap=None):
        cap = cap or 4 * self.cap
        buf = np.empty(cap, dtype=np.int32)
        n = self.L.chk_get_ints(self.h, which, buf.ctypes.data, cap)
        if n > cap:
            return self.ints(which, n)
        return buf[:max(n, 0)].copy()

    def floats(self, which, cap=None):
        cap = cap or 4 * self.cap
        buf = np.empty(cap, dtype=np.float32)
        n = self.L.chk_get_floats(self.h, which, buf.ctypes.data, cap)
        return buf[:max(n, 0)].copy()

    def busy_seconds(self):
        b = np.zeros(4, dtype=np.float64)
        self.L.chk_busy_seconds(self.h, b.ctypes.data)
        return b

    def lo_set_state(self, transform, transform_sum):
        t = np.ascontiguousarray(transform, dtype=np.float32)
        s = np.ascontiguousarray(transform_sum, dtype=np.float32)
        self.L.chk_lo_set_state(self.h, t.ctypes.data, s.ctypes.data)

    def lo_push(self, full, sharp, less_sharp, flat, less_flat, imu_trans=None, stamp=0.0):
        """laserOdometry (and what follows it) fed with host copies of scanRegistration's six topics"""
        cl = [np.ascontiguousarray(c, dtype=np.float32).reshape(-1, 4) for c in (full, sharp, less_sharp, flat, less_flat)]
        it = None if imu_trans is None else np.ascontiguousarray(imu_trans, dtype=np.float32)
        res = SweepResult()
        args = []
        for c in cl:
            args += [c.ctypes.data, c.shape[0]]
        self.L.chk_lo_push(self.h, *args, None if it is None else it.ctypes.data, float(stamp), ctypes.byref(res))
        return res

    # ---- laserMapping alone (kind "port"): recorded laserOdometry output in, map files ----
    def lm_push(self, corner, surf, full, transform_sum, stamp):
        c = np.ascontiguousarray(corner, dtype=np.float32).reshape(-1, 4)
        f = np.ascontiguousarray(surf, dtype=np.float32).reshape(-1, 4)
        a = np.ascontiguousarray(full, dtype=np.float32).reshape(-1, 4)
        t = np.ascontiguousarray(transform_sum, dtype=np.float32)
        res = SweepResult()
        self.L.chk_lm_push(self.h, c.ctypes.data, c.shape[0], f.ctypes.data, f.shape[0], a.ctypes.data, a.shape[0],
                           t.ctypes.data, float(stamp), ctypes.byref(res))
        return res

    def lm_save_map(self, path):
        rc = self.L.chk_lm_save_map(self.h, os.fsencode(path))
        if rc != 0:
            raise RuntimeError(f"chk_lm_save_map({path}) = {rc}")

    def lm_load_map(self, path):
        rc = self.L.chk_lm_load_map(self.h, os.fsencode(path))
        if rc != 0:
            raise RuntimeError(f"chk_lm_load_map({path}) = {rc}")

    def run_pipelined(self, sweeps):
        """kind "port" or "reference": sr / lo / lm on three threads joined by depth-2 queues (the reference's process
        pipeline); sweeps = [(xyz, stamp), ...]; returns (wall seconds, /integrated_to_init poses [n, 6])"""
        xyz = np.ascontiguousarray(np.concatenate([np.asarray(x, np.float32).reshape(-1, 3) for x, _ in sweeps]), np.float32)
        counts = np.array([np.asarray(x).shape[0] for x, _ in sweeps], np.int32)
        stamps = np.array([st for _, st in sweeps], np.float64)
        poses = np.zeros((len(sweeps), 6), np.float32)
        t = self.L.chk_run_pipelined(self.h, xyz.ctypes.data, counts.ctypes.data, stamps.ctypes.data, len(sweeps), poses.ctypes.data)
        return float(t), poses
