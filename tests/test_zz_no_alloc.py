"""loam_voxel_filter_noalloc (include/loam_b200.h): the allocation-free variant of the one-cluster voxel
filter, scratch in a static __device__ array.  No counterpart in the reference.

Collected last on purpose.  Checked through the C ABI against the
existing entry point loam_debug_voxel_filter_coop (which cudaMallocs its scratch per call) and against
the oracle's PCL VoxelGrid definition, bit for bit -- the tolerance every voxel-filter test here uses --
on two calls in a row that share the scratch array (the second smaller, so stale contents would show).
"""
import ctypes

import numpy as np
import pytest

from conftest import biteq

MAX_POINTS = 131072  # LOAM_NOALLOC_MAX_POINTS


def _cloud(seed, n, nan_every=0):
    rng = np.random.default_rng(seed)
    pts = np.zeros((n, 4), np.float32)
    pts[:, 0] = rng.uniform(-60, 60, n)
    pts[:, 2] = rng.uniform(-40, 40, n)
    pts[:, 1] = np.where(rng.random(n) < 0.6, -1.7 + rng.normal(0, 0.01, n), rng.uniform(-2, 10, n))
    pts[:, 3] = rng.integers(0, 64, n)
    if nan_every:
        pts[::nan_every, 0] = np.nan
    return pts


def _bind(L):
    vp, ci = ctypes.c_void_p, ctypes.c_int
    L.loam_voxel_filter_noalloc.argtypes = [vp, ci, ctypes.c_float, vp, ci, vp, vp]
    L.loam_voxel_filter_noalloc.restype = ci
    L.loam_debug_voxel_filter_coop.argtypes = [vp, ci, ctypes.c_float, vp, ci, ctypes.POINTER(ci)]
    L.loam_last_error.restype = ctypes.c_char_p


def _existing(L, pts, leaf):
    n = pts.shape[0]
    out = np.zeros((max(n, 1), 4), np.float32)
    cnt = ctypes.c_int()
    assert L.loam_debug_voxel_filter_coop(pts.ctypes.data, n, leaf, out.ctypes.data, n, ctypes.byref(cnt)) == 0
    return out[:cnt.value].copy()


@pytest.mark.gpu
def test_noalloc_voxel_filter_equals_existing_entry_point_twice_in_a_row(pkg, checker):
    import torch
    L = pkg.lib()
    _bind(L)
    Lo = checker._load("port")
    Lo.chk_voxel_filter.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int]
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    # caller-owned device buffers, allocated once up front; the calls below allocate nothing
    d_in = torch.empty((MAX_POINTS, 4), dtype=torch.float32, device=dev)
    d_out = torch.empty((MAX_POINTS, 4), dtype=torch.float32, device=dev)
    d_cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    # once-per-device set-up (symbol address, the driver's lazy loading of the two kernels) happens here
    assert L.loam_voxel_filter_noalloc(d_in.data_ptr(), 0, 0.2, d_out.data_ptr(), MAX_POINTS, d_cnt.data_ptr(), None) == 0, L.loam_last_error()
    torch.cuda.synchronize()
    # two calls in a row on the shared scratch: a large cloud (4-pass key range) first, then a smaller
    # one with NaNs and a coarser leaf (fewer radix passes, fewer tiles) -- then the exact maximum
    cases = [(0.2, _cloud(5, 120_000)), (0.4, _cloud(6, 9_000, nan_every=97)), (0.2, _cloud(7, MAX_POINTS)), (0.3, _cloud(8, 1))]
    for leaf, pts in cases:
        n = pts.shape[0]
        want = _existing(L, pts, leaf)
        good = np.ascontiguousarray(pts[np.isfinite(pts[:, 0])])
        ref = np.zeros((n, 4), np.float32)
        m = Lo.chk_voxel_filter(good.ctypes.data, good.shape[0], leaf, ref.ctypes.data, n)
        assert want.shape[0] == m and biteq(want, ref[:m])
        d_in[:n].copy_(torch.from_numpy(pts))
        d_out.fill_(float("nan"))
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = L.loam_voxel_filter_noalloc(d_in.data_ptr(), n, leaf, d_out.data_ptr(), MAX_POINTS, d_cnt.data_ptr(), None)
        assert rc == 0, L.loam_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0, "the call changed the amount of free device memory"
        got_n = int(d_cnt.cpu()[0])
        got = d_out[:got_n].cpu().numpy()
        assert got_n == m, (leaf, n, got_n, m)
        assert biteq(got, want), (leaf, n)
    # empty input: zero output points, no error
    assert L.loam_voxel_filter_noalloc(d_in.data_ptr(), 0, 0.2, d_out.data_ptr(), MAX_POINTS, d_cnt.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert int(d_cnt.cpu()[0]) == 0


@pytest.mark.gpu
def test_noalloc_voxel_filter_rejects_sizes_beyond_the_static_scratch(pkg):
    import torch
    L = pkg.lib()
    _bind(L)
    d = torch.zeros((16, 4), dtype=torch.float32, device="cuda:0")
    c = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    rc = L.loam_voxel_filter_noalloc(d.data_ptr(), MAX_POINTS + 1, 0.2, d.data_ptr(), MAX_POINTS + 1, c.data_ptr(), None)
    assert rc == -3  # LOAM_ERR_CAPACITY
    assert b"LOAM_NOALLOC_MAX_POINTS" in L.loam_last_error()
    # the output buffer must be able to hold one point per input point
    assert L.loam_voxel_filter_noalloc(d.data_ptr(), 16, 0.2, d.data_ptr(), 8, c.data_ptr(), None) == -1
    assert L.loam_voxel_filter_noalloc(None, 16, 0.2, d.data_ptr(), 16, c.data_ptr(), None) == -1
    assert L.loam_voxel_filter_noalloc(d.data_ptr(), 16, 0.0, d.data_ptr(), 16, c.data_ptr(), None) == -1
