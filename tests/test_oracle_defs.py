"""CPU: the third-party definitions the oracle adopts (oracle/defs/*) pinned
against independent evidence: OpenCV 4.13 golden vectors, numpy restatements, brute force."""
import ctypes
import os

import numpy as np
import pytest

from conftest import GOLDEN, biteq


def _lib(checker):
    L = checker._load("port")
    L.chk_math.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int]
    L.chk_linalg.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3
    L.chk_gemm.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3
    L.chk_voxel_filter.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int]
    L.chk_knn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    return L


def linalg(L, op, A, b, m, n):
    A = np.ascontiguousarray(A, np.float32)
    batch = A.shape[0]
    outw = n if op == 0 else (n + n * n if op == 1 else n * n)
    out = np.zeros((batch, outw), np.float32)
    bb = np.ascontiguousarray(b, np.float32) if b is not None else np.zeros((batch, m), np.float32)
    L.chk_linalg(op, A.ctypes.data, bb.ctypes.data, out.ctypes.data, m, n, batch)
    return out


def test_linalg_bitexact_vs_opencv(checker):
    """cv::solve(QR) 6x6 and 5x3, cv::eigen 3x3 and 6x6, Mat::inv, gemm: every bit equal to cv2 4.13"""
    L = _lib(checker)
    G = np.load(os.path.join(GOLDEN, "linalg_cv2.npz"))
    assert biteq(linalg(L, 0, G["qr6_A"], G["qr6_b"], 6, 6), G["qr6_x"])
    assert biteq(linalg(L, 0, G["qr53_A"], -np.ones((200, 5), np.float32), 5, 3), G["qr53_x"])
    for n in (3, 6):
        o = linalg(L, 1, G[f"eig{n}_A"], None, n, n)
        assert biteq(o[:, :n], G[f"eig{n}_w"])
        assert biteq(o[:, n:].reshape(-1, n, n), G[f"eig{n}_v"])
    assert biteq(linalg(L, 2, G["inv6_A"], None, 6, 6).reshape(-1, 6, 6), G["inv6_inv"])
    for A, C in zip(G["gemm_A"], G["gemm_C"]):
        B = np.ascontiguousarray(A.T)
        out = np.zeros((6, 6), np.float32)
        A = np.ascontiguousarray(A)
        L.chk_gemm(A.ctypes.data, B.ctypes.data, out.ctypes.data, 6, A.shape[1], 6)
        assert biteq(out, C)


def test_detmath_close_to_glibc(checker):
    """detmath (the adopted libm definition) is within 1 ulp of glibc everywhere sampled and the
    non-transcendental ops (sqrt, div, unfused mul-add) are bit-identical"""
    Ld = _lib(checker)
    Lm = checker._load("port-libm")
    Lm.chk_math.argtypes = Ld.chk_math.argtypes
    rng = np.random.default_rng(7)
    n = 400000
    x = np.concatenate([rng.uniform(-7, 7, n // 2), rng.normal(0, 0.05, n // 2)]).astype(np.float32)
    y = rng.uniform(-7, 7, n).astype(np.float32)
    u = rng.uniform(-1, 1, n).astype(np.float32)
    for fn, a, b in ((0, x, y), (1, x, y), (2, x, y), (3, x, y), (4, u, y), (5, np.abs(x), y), (6, x, y), (7, x, y)):
        od = np.zeros(n, np.float32)
        om = np.zeros(n, np.float32)
        Ld.chk_math(fn, a.ctypes.data, b.ctypes.data, od.ctypes.data, n)
        Lm.chk_math(fn, a.ctypes.data, b.ctypes.data, om.ctypes.data, n)
        if fn >= 5:
            assert biteq(od, om), fn
        else:
            ulp = np.abs(od.view(np.int32).astype(np.int64) - om.view(np.int32).astype(np.int64))
            same_sign = np.signbit(od) == np.signbit(om)
            assert ulp[same_sign].max() <= 1, (fn, ulp.max())
            assert np.all(same_sign | (np.abs(od) < 1e-30))
            assert (ulp == 0).mean() > 0.8, (fn, (ulp == 0).mean())


def voxel_numpy(pts, leaf):
    """independent numpy restatement of PCL VoxelGrid (App. B.1)"""
    pts = np.asarray(pts, np.float32)
    inv = np.float32(1.0) / np.float32(leaf)
    mn = pts[:, :3].min(0)
    mx = pts[:, :3].max(0)
    minb = np.floor(mn * inv).astype(np.int64)
    maxb = np.floor(mx * inv).astype(np.int64)
    div = maxb - minb + 1
    ijk = (np.floor(pts[:, :3] * inv) - minb.astype(np.float32)).astype(np.int64)
    idx = ijk[:, 0] + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    order = np.lexsort((np.arange(len(pts)), idx))
    out = []
    i = 0
    while i < len(order):
        j = i
        acc = np.zeros(4, np.float32)
        while j < len(order) and idx[order[j]] == idx[order[i]]:
            acc = (acc + pts[order[j]]).astype(np.float32)
            j += 1
        out.append(acc / np.float32(j - i))
        i = j
    return np.array(out, np.float32)


@pytest.mark.parametrize("leaf,n", [(0.2, 3000), (0.4, 3000), (0.2, 1), (0.4, 17)])
def test_voxel_grid_definition(checker, leaf, n):
    L = _lib(checker)
    rng = np.random.default_rng(n)
    pts = np.zeros((n, 4), np.float32)
    pts[:, :3] = rng.uniform(-6, 6, (n, 3))
    pts[: n // 3, 2] = 1.0  # a dense plane: many points per voxel
    pts[:, 3] = rng.uniform(0, 16, n)
    out = np.zeros((n, 4), np.float32)
    m = L.chk_voxel_filter(pts.ctypes.data, n, leaf, out.ctypes.data, n)
    ref = voxel_numpy(pts, leaf)
    assert m == len(ref)
    assert biteq(out[:m], ref)


def test_kdtree_is_exact_with_lowest_index_tiebreak(checker):
    L = _lib(checker)
    rng = np.random.default_rng(3)
    n, nq = 5000, 300
    pts = np.zeros((n, 4), np.float32)
    pts[:, :3] = np.round(rng.uniform(-5, 5, (n, 3)) * 4) / 4  # lattice => many exact ties
    q = np.round(rng.uniform(-5, 5, (nq, 3)) * 4).astype(np.float32) / 4
    for k in (1, 5):
        idx = np.zeros((nq, k), np.int32)
        d2 = np.zeros((nq, k), np.float32)
        L.chk_knn(pts.ctypes.data, n, np.ascontiguousarray(q).ctypes.data, nq, k, idx.ctypes.data, d2.ctypes.data)
        for i in range(nq):
            dx = (q[i, 0] - pts[:, 0]).astype(np.float32)
            dy = (q[i, 1] - pts[:, 1]).astype(np.float32)
            dz = (q[i, 2] - pts[:, 2]).astype(np.float32)
            dd = ((dx * dx).astype(np.float32) + (dy * dy).astype(np.float32)).astype(np.float32)
            dd = (dd + (dz * dz).astype(np.float32)).astype(np.float32)
            order = np.lexsort((np.arange(n), dd))[:k]
            assert np.array_equal(idx[i], order), (k, i)
            assert biteq(d2[i], dd[order])
