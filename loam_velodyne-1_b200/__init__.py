"""loam_velodyne-1_b200: B200-native LOAM per-sweep pipeline (scanRegistration -> laserOdometry ->
laserMapping, transformMaintenance on the host).

This Python layer is test / bench orchestration only: a ctypes binding of the C ABI in
include/loam_b200.h (libloam_b200.so, hand-written sm_100a CUDA kernels).  There is no CPU
fallback: importing works anywhere, but creating a handle without a usable sm_100 device raises.
"""
import ctypes
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LOAM_B200_LIB") or os.path.join(_HERE, "libloam_b200.so")   # LOAM_B200_LIB: a development build (e.g. with -DLOAM_PHASE_CLOCKS)

LOAM_OK, LOAM_NOT_READY = 0, 1

# enum loam_cloud / loam_ints / loam_floats (include/loam_b200.h)
CLOUD_FULL, CLOUD_SHARP, CLOUD_LESS_SHARP, CLOUD_FLAT, CLOUD_LESS_FLAT = 0, 1, 2, 3, 4
CLOUD_CORNER_LAST, CLOUD_SURF_LAST, CLOUD_FULL_END, CLOUD_REGISTERED, CLOUD_SURROUND = 5, 6, 7, 8, 9
CLOUD_MAP_CORNER, CLOUD_MAP_SURF, CLOUD_STACK_CORNER, CLOUD_STACK_SURF = 10, 11, 12, 13
INTS_LABEL, INTS_PICKED, INTS_SORT_IND, INTS_SHARP_IDX, INTS_LESS_SHARP_IDX, INTS_FLAT_IDX = 0, 1, 2, 3, 4, 5
INTS_CORNER_IND1, INTS_CORNER_IND2, INTS_SURF_IND1, INTS_SURF_IND2, INTS_SURF_IND3 = 6, 7, 8, 9, 10
INTS_SCAN_START, INTS_SCAN_END = 11, 12
FLOATS_CURVATURE = 0


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
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = np.array(v[:], dtype=np.float32) if hasattr(v, "__len__") else v
        return d


class LoamError(RuntimeError):
    pass


_lib = None


def build(force=False):
    from . import build as _b
    return _b.build_all(force=force)


def lib():
    """loads libloam_b200.so; fails loudly if it has not been built"""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LoamError("libloam_b200.so is missing: run __graft_entry__.build() (there is no CPU fallback)")
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, cd = ctypes.c_void_p, ctypes.c_int, ctypes.c_double
    L.loam_last_error.restype = ctypes.c_char_p
    L.loam_kernel_launch_count.restype = ctypes.c_longlong
    L.loam_debug_host_wait_ns.restype = ctypes.c_longlong
    L.loam_params_preset.argtypes = [ctypes.POINTER(LoamParams), ci]
    L.loam_profile_enable.argtypes = [ci]
    L.loam_profile_report.argtypes = [ctypes.c_char_p, ci]
    for node in ("sr", "lo", "lm", "pipeline"):
        getattr(L, f"loam_{node}_create").argtypes = [ctypes.POINTER(LoamParams), ci, ctypes.POINTER(vp)]
        getattr(L, f"loam_{node}_destroy").argtypes = [vp]
    L.loam_tm_create.argtypes = [ctypes.POINTER(vp)]
    L.loam_tm_destroy.argtypes = [vp]
    L.loam_sr_imu.argtypes = [vp, cd, vp, vp]
    L.loam_sr_process.argtypes = [vp, vp, ci, ci, cd]
    L.loam_sr_process_packets.argtypes = [vp, vp, ci, cd]
    L.loam_pipeline_push_packets.argtypes = [vp, vp, ci, cd, ctypes.POINTER(SweepResult)]
    L.loam_pipeline_push_packets_async.argtypes = [vp, vp, ci, cd]
    L.loam_debug_decode_packets.argtypes = [vp, ci, vp]
    L.loam_debug_packet_front_end.argtypes = [ctypes.POINTER(LoamParams), vp, ci, vp, vp, vp]
    L.loam_sr_download.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_sr_download_ints.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_sr_download_floats.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_sr_imu_trans.argtypes = [vp, vp]
    L.loam_lo_process.argtypes = [vp, vp, ctypes.POINTER(ci)]
    L.loam_lo_process_host.argtypes = [vp, vp, ci, vp, ci, vp, ci, vp, ci, vp, ci, vp, ctypes.POINTER(ci)]
    L.loam_lo_download.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_lo_download_ints.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_lo_state.argtypes = [vp, vp, vp, vp]
    L.loam_lo_set_state.argtypes = [vp, vp, vp]
    L.loam_lm_imu.argtypes = [vp, cd, vp]
    L.loam_lm_process.argtypes = [vp, vp, cd]
    L.loam_lm_process_host.argtypes = [vp, vp, ci, vp, ci, vp, ci, vp, cd]
    L.loam_lm_process_device.argtypes = [vp, vp, ci, vp, ci, vp, ci, vp, cd]
    L.loam_lm_sync.argtypes = [vp]
    L.loam_lm_stream.restype = vp
    L.loam_lm_stream.argtypes = [vp]
    L.loam_lm_save_map.argtypes = [vp, ctypes.c_char_p]
    L.loam_lm_load_map.argtypes = [vp, ctypes.c_char_p]
    L.loam_lm_export_pcd.argtypes = [vp, ci, ctypes.c_char_p]
    L.loam_lm_download.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_lm_download_ints.argtypes = [vp, ci, vp, ci, ctypes.POINTER(ci)]
    L.loam_lm_state.argtypes = [vp, vp, vp, vp, vp]
    L.loam_tm_on_mapping.argtypes = [vp, vp, vp]
    L.loam_tm_on_odometry.argtypes = [vp, vp, vp]
    L.loam_pipeline_imu.argtypes = [vp, cd, vp, vp]
    L.loam_pipeline_push.argtypes = [vp, vp, ci, ci, cd, ctypes.POINTER(SweepResult)]
    L.loam_pipeline_push_async.argtypes = [vp, vp, ci, ci, cd]
    L.loam_pipeline_push_device.argtypes = [vp, vp, ci, vp, ci, cd]
    L.loam_pipeline_collect.argtypes = [vp, vp, ci, ctypes.POINTER(ci)]
    L.loam_pipeline_stream.restype = vp
    L.loam_pipeline_stream.argtypes = [vp]
    L.loam_pipeline_join.argtypes = [vp]
    L.loam_pipeline_batch_create.argtypes = [ctypes.POINTER(LoamParams), ci, ci, ctypes.POINTER(vp)]
    L.loam_pipeline_batch_destroy.argtypes = [vp]
    L.loam_pipeline_batch_get.restype = vp
    L.loam_pipeline_batch_get.argtypes = [vp, ci]
    L.loam_pipeline_batch_push_async.argtypes = [vp, vp, vp, ci, vp]
    L.loam_pipeline_batch_push_device.argtypes = [vp, vp, vp, vp, vp, vp]
    for f in ("sr", "lo", "lm"):
        fn = getattr(L, f"loam_pipeline_{f}")
        fn.restype = vp
        fn.argtypes = [vp]
    _lib = L
    return L


def _check(rc):
    if rc < 0:
        raise LoamError(f"loam_b200 error {rc}: {lib().loam_last_error().decode()}")
    return rc


def make_params(preset, system_delay=0):
    p = LoamParams()
    _check(lib().loam_params_preset(ctypes.byref(p), preset))
    p.system_delay = system_delay
    return p


def kernel_launch_count():
    return int(lib().loam_kernel_launch_count())


def host_wait_ns():
    return int(lib().loam_debug_host_wait_ns())


def host_node_ns():
    out = (ctypes.c_longlong * 3)()
    lib().loam_debug_host_node_ns(out)
    return [int(x) for x in out]


def profile_enable(on=True):
    lib().loam_profile_enable(1 if on else 0)


def profile_report():
    buf = ctypes.create_string_buffer(1 << 16)
    lib().loam_profile_report(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, calls, ms = line.split()
        out[name] = (int(calls), float(ms))
    return out


def _packets(packets):
    pk = np.ascontiguousarray(packets, dtype=np.uint8)
    if pk.ndim != 2 or pk.shape[1] != 1206:
        raise ValueError("packets must be a (n_packets, 1206) uint8 array")
    return pk


def decode_packets(packets):
    """test hook: the device VLP-16 decoder alone -> (n_packets * 384, 3) float32, NaN = no return"""
    pk = _packets(packets)
    out = np.empty((pk.shape[0] * 384, 3), np.float32)
    _check(lib().loam_debug_decode_packets(pk.ctypes.data, pk.shape[0], out.ctypes.data))
    return out


def packet_front_end(params, packets, metadata=False):
    """test hook: the device packet front end of the sensor `params` describes (VLP-16 or HDL-64E) -> xyz (n, 3)
    and, with metadata=True, ring ids and relTime as loam_params.packet_metadata = 1 derives them"""
    pk = _packets(packets)
    n = pk.shape[0] * 384
    xyz = np.empty((n, 3), np.float32)
    ring = np.empty(n, np.int32) if metadata else None
    rt = np.empty(n, np.float32) if metadata else None
    _check(lib().loam_debug_packet_front_end(ctypes.byref(params), pk.ctypes.data, pk.shape[0], xyz.ctypes.data,
                                             ring.ctypes.data if metadata else None, rt.ctypes.data if metadata else None))
    return (xyz, ring, rt) if metadata else xyz


class _Node:
    _prefix = None

    def __init__(self, params, device=0):
        self.L = lib()
        self.params = params
        self.h = ctypes.c_void_p()
        _check(getattr(self.L, f"loam_{self._prefix}_create")(ctypes.byref(params), device, ctypes.byref(self.h)))
        self.cap = params.max_points

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            getattr(self.L, f"loam_{self._prefix}_destroy")(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dl(self, fn, which, dtype, width, cap=None):
        cap = cap or self.cap
        buf = np.empty((cap, width) if width > 1 else (cap,), dtype=dtype)
        n = ctypes.c_int()
        _check(fn(self.h, which, buf.ctypes.data, cap, ctypes.byref(n)))
        if n.value > cap:
            return self._dl(fn, which, dtype, width, n.value)
        return buf[:n.value].copy()


class ScanRegistration(_Node):
    """scanRegistration node (scanRegistration.cpp): imu() = imuHandler, process() = laserCloudHandler"""
    _prefix = "sr"

    def imu(self, stamp, quat, acc):
        q = np.ascontiguousarray(quat, dtype=np.float64)
        a = np.ascontiguousarray(acc, dtype=np.float32)
        _check(self.L.loam_sr_imu(self.h, float(stamp), q.ctypes.data, a.ctypes.data))

    def process(self, xyz, stamp):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        return _check(self.L.loam_sr_process(self.h, xyz.ctypes.data, xyz.shape[0], xyz.strides[0], float(stamp)))

    def process_packets(self, packets, stamp):
        """packets: uint8 array (n_packets, 1206) of VLP-16 data packets of one revolution"""
        pk = _packets(packets)
        return _check(self.L.loam_sr_process_packets(self.h, pk.ctypes.data, pk.shape[0], float(stamp)))

    def cloud(self, which):
        return self._dl(self.L.loam_sr_download, which, np.float32, 4)

    def ints(self, which):
        return self._dl(self.L.loam_sr_download_ints, which, np.int32, 1)

    def floats(self, which):
        return self._dl(self.L.loam_sr_download_floats, which, np.float32, 1)

    def imu_trans(self):
        out = np.zeros(12, dtype=np.float32)
        _check(self.L.loam_sr_imu_trans(self.h, out.ctypes.data))
        return out


class LaserOdometry(_Node):
    """laserOdometry node (laserOdometry.cpp main loop body)"""
    _prefix = "lo"

    def process(self, sr):
        flag = ctypes.c_int()
        rc = _check(self.L.loam_lo_process(self.h, sr.h, ctypes.byref(flag)))
        return rc, bool(flag.value)

    def process_host(self, full, sharp, less_sharp, flat, less_flat, imu_trans=None):
        """the same pass fed from host copies of scanRegistration's six topics"""
        cl = [np.ascontiguousarray(c, dtype=np.float32).reshape(-1, 4) for c in (full, sharp, less_sharp, flat, less_flat)]
        it = None if imu_trans is None else np.ascontiguousarray(imu_trans, dtype=np.float32)
        flag = ctypes.c_int()
        args = []
        for c in cl:
            args += [c.ctypes.data, c.shape[0]]
        rc = _check(self.L.loam_lo_process_host(self.h, *args, None if it is None else it.ctypes.data, ctypes.byref(flag)))
        return rc, bool(flag.value)

    def cloud(self, which):
        return self._dl(self.L.loam_lo_download, which, np.float32, 4)

    def ints(self, which):
        return self._dl(self.L.loam_lo_download_ints, which, np.int32, 1)

    def state(self):
        t = np.zeros(6, np.float32)
        s = np.zeros(6, np.float32)
        st = np.zeros(4, np.int32)
        _check(self.L.loam_lo_state(self.h, t.ctypes.data, s.ctypes.data, st.ctypes.data))
        return t, s, st

    def set_state(self, transform, transform_sum):
        t = np.ascontiguousarray(transform, np.float32)
        s = np.ascontiguousarray(transform_sum, np.float32)
        _check(self.L.loam_lo_set_state(self.h, t.ctypes.data, s.ctypes.data))


class LaserMapping(_Node):
    """laserMapping node (laserMapping.cpp main loop body)"""
    _prefix = "lm"

    def imu(self, stamp, quat):
        q = np.ascontiguousarray(quat, dtype=np.float64)
        _check(self.L.loam_lm_imu(self.h, float(stamp), q.ctypes.data))

    def process(self, lo, stamp):
        return _check(self.L.loam_lm_process(self.h, lo.h, float(stamp)))

    def process_host(self, corner_last, surf_last, full_res, transform_sum, stamp):
        """the same pass fed from host copies of laserOdometry's output (recorded / replayed)"""
        c = np.ascontiguousarray(corner_last, dtype=np.float32).reshape(-1, 4)
        f = np.ascontiguousarray(surf_last, dtype=np.float32).reshape(-1, 4)
        a = np.ascontiguousarray(full_res, dtype=np.float32).reshape(-1, 4)
        t = np.ascontiguousarray(transform_sum, dtype=np.float32)
        return _check(self.L.loam_lm_process_host(self.h, c.ctypes.data, c.shape[0], f.ctypes.data, f.shape[0],
                                                  a.ctypes.data, a.shape[0], t.ctypes.data, float(stamp)))

    def process_device(self, d_corner, n_corner, d_surf, n_surf, d_full, n_full, transform_sum, stamp):
        """enqueue only; device pointers to float4 clouds that stay unchanged until sync()"""
        t = np.ascontiguousarray(transform_sum, dtype=np.float32)
        return _check(self.L.loam_lm_process_device(self.h, ctypes.c_void_p(d_corner), n_corner, ctypes.c_void_p(d_surf), n_surf,
                                                    ctypes.c_void_p(d_full), n_full, t.ctypes.data, float(stamp)))

    def sync(self):
        return _check(self.L.loam_lm_sync(self.h))

    def stream(self):
        return self.L.loam_lm_stream(self.h)

    def save_map(self, path):
        _check(self.L.loam_lm_save_map(self.h, os.fsencode(path)))

    def load_map(self, path):
        _check(self.L.loam_lm_load_map(self.h, os.fsencode(path)))

    def export_pcd(self, which, path):
        _check(self.L.loam_lm_export_pcd(self.h, which, os.fsencode(path)))

    def cloud(self, which, cap=None):
        return self._dl(self.L.loam_lm_download, which, np.float32, 4, cap=cap)

    def ints(self, which):
        return self._dl(self.L.loam_lm_download_ints, which, np.int32, 1, cap=5 * self.cap)

    def state(self):
        a = np.zeros(6, np.float32)
        b = np.zeros(6, np.float32)
        t = np.zeros(6, np.float32)
        st = np.zeros(8, np.int32)
        _check(self.L.loam_lm_state(self.h, a.ctypes.data, b.ctypes.data, t.ctypes.data, st.ctypes.data))
        return a, b, t, st


class TransformMaintenance:
    """transformMaintenance node (host only)"""

    def __init__(self):
        self.L = lib()
        self.h = ctypes.c_void_p()
        _check(self.L.loam_tm_create(ctypes.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.loam_tm_destroy(self.h)
            self.h = ctypes.c_void_p()

    def on_mapping(self, aft, bef):
        a = np.ascontiguousarray(aft, np.float32)
        b = np.ascontiguousarray(bef, np.float32)
        _check(self.L.loam_tm_on_mapping(self.h, a.ctypes.data, b.ctypes.data))

    def on_odometry(self, transform_sum):
        s = np.ascontiguousarray(transform_sum, np.float32)
        out = np.zeros(6, np.float32)
        _check(self.L.loam_tm_on_odometry(self.h, s.ctypes.data, out.ctypes.data))
        return out


class _Borrowed:
    def __init__(self, h, cap, L):
        self.h, self.cap, self.L = ctypes.c_void_p(h), cap, L


class Pipeline(_Node):
    """sr -> lo -> tm -> (lm -> tm) on one device / one stream"""
    _prefix = "pipeline"

    def __init__(self, params, device=0):
        super().__init__(params, device)
        mk = lambda cls, h: _borrow(cls, h, params.max_points, self.L)
        self.sr = mk(ScanRegistration, self.L.loam_pipeline_sr(self.h))
        self.lo = mk(LaserOdometry, self.L.loam_pipeline_lo(self.h))
        self.lm = mk(LaserMapping, self.L.loam_pipeline_lm(self.h))

    def imu(self, stamp, quat, acc):
        q = np.ascontiguousarray(quat, dtype=np.float64)
        a = np.ascontiguousarray(acc, dtype=np.float32)
        _check(self.L.loam_pipeline_imu(self.h, float(stamp), q.ctypes.data, a.ctypes.data))

    def push(self, xyz, stamp):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        res = SweepResult()
        _check(self.L.loam_pipeline_push(self.h, xyz.ctypes.data, xyz.shape[0], xyz.strides[0], float(stamp), ctypes.byref(res)))
        return res


    def push_packets(self, packets, stamp):
        pk = _packets(packets)
        res = SweepResult()
        _check(self.L.loam_pipeline_push_packets(self.h, pk.ctypes.data, pk.shape[0], float(stamp), ctypes.byref(res)))
        return res

    def push_packets_async(self, packets, stamp):
        pk = _packets(packets)
        return _check(self.L.loam_pipeline_push_packets_async(self.h, pk.ctypes.data, pk.shape[0], float(stamp)))

    def push_async(self, xyz, stamp):
        """pageable xyz is staged before this returns; CUDA-pinned packed xyz is read in place until collect()"""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        return _check(self.L.loam_pipeline_push_async(self.h, xyz.ctypes.data, xyz.shape[0], xyz.strides[0], float(stamp)))

    def push_device(self, d_ptr, n, first_last_xyz, first_index, stamp):
        fl = np.ascontiguousarray(first_last_xyz, dtype=np.float32)
        return _check(self.L.loam_pipeline_push_device(self.h, ctypes.c_void_p(d_ptr), n, fl.ctypes.data, first_index, float(stamp)))

    def collect(self, cap=1024):
        arr = (SweepResult * cap)()
        n = ctypes.c_int()
        _check(self.L.loam_pipeline_collect(self.h, arr, cap, ctypes.byref(n)))
        return [arr[i] for i in range(min(n.value, cap))]

    def stream(self):
        return self.L.loam_pipeline_stream(self.h)

    def join(self):
        """device-side: the stream() waits for the registration / mapping streams"""
        _check(self.L.loam_pipeline_join(self.h))


def _borrow(cls, h, cap, L):
    obj = cls.__new__(cls)
    obj.L = L
    obj.h = ctypes.c_void_p(h)
    obj.cap = cap
    obj.close = lambda: None
    return obj


class PipelineBatch:
    """`count` pipelines of one GPU whose scanRegistration passes go out as ONE launch chain per step
    (loam_pipeline_batch_*, BASELINE.json configs[4]); self.pipes[i] is the i-th sequence's Pipeline view"""

    def __init__(self, params, count, device=0):
        self.L = lib()
        self.h = ctypes.c_void_p()
        _check(self.L.loam_pipeline_batch_create(ctypes.byref(params), device, count, ctypes.byref(self.h)))
        self.count = count
        self.pipes = []
        for i in range(count):
            p = Pipeline.__new__(Pipeline)
            p.L = self.L
            p.params = params
            p.cap = params.max_points
            p.h = ctypes.c_void_p(self.L.loam_pipeline_batch_get(self.h, i))
            p.close = lambda: None
            mk = lambda cls, h: _borrow(cls, h, params.max_points, self.L)
            p.sr = mk(ScanRegistration, self.L.loam_pipeline_sr(p.h))
            p.lo = mk(LaserOdometry, self.L.loam_pipeline_lo(p.h))
            p.lm = mk(LaserMapping, self.L.loam_pipeline_lm(p.h))
            self.pipes.append(p)
        self._ptrs = (ctypes.c_void_p * count)()
        self._n = (ctypes.c_int * count)()
        self._st = (ctypes.c_double * count)()

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.loam_pipeline_batch_destroy(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def push_async(self, sweeps, stamps):
        """sweeps: one (n_i, 3) float32 array per sequence (pageable or CUDA-pinned host memory)"""
        keep = [np.ascontiguousarray(x, dtype=np.float32) for x in sweeps]
        for i, x in enumerate(keep):
            self._ptrs[i] = x.ctypes.data
            self._n[i] = x.shape[0]
            self._st[i] = float(stamps[i])
        return _check(self.L.loam_pipeline_batch_push_async(self.h, self._ptrs, self._n, 12, self._st))

    def push_device(self, d_ptrs, ns, first_last, first_index, stamps):
        """sweeps resident in device memory: device pointers, point counts, per sequence the 6 floats (first xyz,
        last xyz) and the index of the first finite point"""
        fl = [np.ascontiguousarray(a, dtype=np.float32) for a in first_last]
        flp = (ctypes.c_void_p * self.count)(*[a.ctypes.data for a in fl])
        fi = (ctypes.c_int * self.count)(*[int(v) for v in first_index])
        for i in range(self.count):
            self._ptrs[i] = int(d_ptrs[i])
            self._n[i] = int(ns[i])
            self._st[i] = float(stamps[i])
        return _check(self.L.loam_pipeline_batch_push_device(self.h, self._ptrs, self._n, flp, fi, self._st))
