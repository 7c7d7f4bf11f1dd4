"""GPU (-m gpu): the sm_100a path, called through the C ABI, against the oracle on identical seeded
sweeps.  Bar: bit-exact -- every cloud, every index array, every pose.  The oracle is only the
checker here; the product never touches it."""
import ctypes
import os

import numpy as np
import pytest

from conftest import GOLDEN, biteq, cloud_checksum, feed_imu

pytestmark = pytest.mark.gpu


def _oracle_lib(checker):
    L = checker._load("port")
    L.chk_math.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int]
    L.chk_voxel_filter.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int]
    return L


def test_device_detmath_equals_host_bit_for_bit(pkg, checker):
    """the definition of sin/cos/atan/atan2/asin (loam/detmath.h) must give the same bits on device"""
    Lo = _oracle_lib(checker)
    Lg = pkg.lib()
    Lg.loam_debug_math.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int]
    rng = np.random.default_rng(11)
    n = 2_000_000
    x = np.concatenate([rng.uniform(-10, 10, n // 2), rng.normal(0, 1e-2, n // 4), rng.uniform(-400, 400, n // 4)]).astype(np.float32)
    y = rng.uniform(-10, 10, n).astype(np.float32)
    u = rng.uniform(-1, 1, n).astype(np.float32)
    for fn, a, b in ((0, x, y), (1, x, y), (2, x, y), (3, x, y), (4, u, y), (5, np.abs(x), y), (6, x, y), (7, x, y)):
        oh = np.zeros(n, np.float32)
        od = np.zeros(n, np.float32)
        Lo.chk_math(fn, a.ctypes.data, b.ctypes.data, oh.ctypes.data, n)
        assert Lg.loam_debug_math(fn, a.ctypes.data, b.ctypes.data, od.ctypes.data, n) == 0
        assert biteq(oh, od), (fn, int((oh.view(np.uint32) != od.view(np.uint32)).sum()))


def test_device_linalg_equals_opencv_golden(pkg):
    """device QR / Jacobi / LU against the cv2 4.13 golden vectors"""
    L = pkg.lib()
    L.loam_debug_linalg.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3
    G = np.load(os.path.join(GOLDEN, "linalg_cv2.npz"))

    def run(op, A, b, m, n):
        A = np.ascontiguousarray(A, np.float32)
        batch = A.shape[0]
        outw = n if op in (0, 3) else (n + n * n if op in (1, 4) else n * n)
        out = np.zeros((batch, outw), np.float32)
        bb = np.ascontiguousarray(b, np.float32) if b is not None else np.zeros((batch, m), np.float32)
        assert L.loam_debug_linalg(op, A.ctypes.data, bb.ctypes.data, out.ctypes.data, m, n, batch) == 0
        return out

    assert biteq(run(0, G["qr6_A"], G["qr6_b"], 6, 6), G["qr6_x"])
    assert biteq(run(0, G["qr53_A"], -np.ones((200, 5), np.float32), 5, 3), G["qr53_x"])
    for n in (3, 6):
        o = run(1, G[f"eig{n}_A"], None, n, n)
        assert biteq(o[:, :n], G[f"eig{n}_w"])
        assert biteq(o[:, n:].reshape(-1, n, n), G[f"eig{n}_v"])
    assert biteq(run(2, G["inv6_A"], None, 6, 6).reshape(-1, 6, 6), G["inv6_inv"])
    # the register-resident specialisations used on the hot paths give the same bits
    assert biteq(run(3, G["qr6_A"], G["qr6_b"], 6, 6), G["qr6_x"])
    assert biteq(run(3, G["qr53_A"], -np.ones((200, 5), np.float32), 5, 3), G["qr53_x"])
    o = run(4, G["eig3_A"], None, 3, 3)
    assert biteq(o[:, :3], G["eig3_w"])
    assert biteq(o[:, 3:].reshape(-1, 3, 3), G["eig3_v"])
    # near-degenerate 3x3 inputs (repeated / zero eigenvalues) exercise the pivot bookkeeping
    rng = np.random.default_rng(7)
    B = rng.standard_normal((500, 3, 1)).astype(np.float32) * rng.uniform(0.01, 3.0, (500, 1, 1)).astype(np.float32)
    A3 = (B @ B.transpose(0, 2, 1)).astype(np.float32)
    A3[::3] += np.eye(3, dtype=np.float32) * 1e-3
    A3[5::7] = 0
    assert biteq(run(4, A3, None, 3, 3), run(1, A3, None, 3, 3))


@pytest.mark.parametrize("n", [0, 1, 31, 4096, 4097, 100_003, 1_500_000])
def test_device_radix_sort_is_stable(pkg, n):
    L = pkg.lib()
    L.loam_debug_sort_pairs.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 1 << 32, max(n, 1), dtype=np.uint64).astype(np.uint32)[:n]
    if n > 10:
        keys[: n // 2] = rng.integers(0, 50, n // 2).astype(np.uint32)  # many duplicates: stability matters
    vals = np.arange(n, dtype=np.uint32)
    k2, v2 = keys.copy(), vals.copy()
    assert L.loam_debug_sort_pairs(k2.ctypes.data, v2.ctypes.data, n) == 0
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k2, keys[order])
    assert np.array_equal(v2, vals[order])


@pytest.mark.parametrize("leaf,n", [(0.2, 1), (0.2, 5000), (0.4, 120_000), (0.2, 700_000)])
def test_device_voxel_filter_equals_pcl_definition(pkg, checker, leaf, n):
    Lo = _oracle_lib(checker)
    L = pkg.lib()
    L.loam_debug_voxel_filter.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    rng = np.random.default_rng(n)
    pts = np.zeros((n, 4), np.float32)
    pts[:, 0] = rng.uniform(-60, 60, n); pts[:, 2] = rng.uniform(-40, 40, n)
    pts[:, 1] = np.where(rng.random(n) < 0.6, -1.7 + rng.normal(0, 0.01, n), rng.uniform(-2, 10, n))
    pts[:, 3] = rng.integers(0, 64, n)
    ref = np.zeros((n, 4), np.float32)
    m = Lo.chk_voxel_filter(pts.ctypes.data, n, leaf, ref.ctypes.data, n)
    out = np.zeros((n, 4), np.float32)
    cnt = ctypes.c_int()
    assert L.loam_debug_voxel_filter(pts.ctypes.data, n, leaf, out.ctypes.data, n, ctypes.byref(cnt)) == 0
    assert cnt.value == m
    assert biteq(out[:m], ref[:m])
    # the one-kernel cooperative variant (laserMapping's incoming-sweep filters) gives the same bits,
    # also with non-finite points in the input (dropped, like removeNaN + VoxelGrid)
    L.loam_debug_voxel_filter_coop.argtypes = L.loam_debug_voxel_filter.argtypes
    outc = np.zeros((n, 4), np.float32)
    cntc = ctypes.c_int()
    assert L.loam_debug_voxel_filter_coop(pts.ctypes.data, n, leaf, outc.ctypes.data, n, ctypes.byref(cntc)) == 0
    assert cntc.value == m and biteq(outc[:m], ref[:m])
    if n > 100:
        bad = pts.copy()
        bad[::97, 0] = np.nan
        for fn in (L.loam_debug_voxel_filter, L.loam_debug_voxel_filter_coop):
            oa = np.zeros((n, 4), np.float32); ca = ctypes.c_int()
            assert fn(bad.ctypes.data, n, leaf, oa.ctypes.data, n, ctypes.byref(ca)) == 0
            good = np.ascontiguousarray(bad[np.isfinite(bad[:, 0])])
            mg = Lo.chk_voxel_filter(good.ctypes.data, good.shape[0], leaf, ref.ctypes.data, n)
            assert ca.value == mg and biteq(oa[:mg], ref[:mg])
    # idempotence on the output is NOT guaranteed by PCL (centroids can cross cell borders), but the
    # output must be sorted by voxel index, i.e. filtering it again never increases the size
    cnt2 = ctypes.c_int()
    out2 = np.zeros((m, 4), np.float32)
    assert L.loam_debug_voxel_filter(out.ctypes.data, m, leaf, out2.ctypes.data, m, ctypes.byref(cnt2)) == 0
    assert cnt2.value <= m


def _sr_compare(C, G, r, k):
    for which in (0, 1, 2, 3, 4):
        assert biteq(C.cloud(which), G.cloud(which)), (k, "cloud", which)
    for which in (0, 1, 2):  # cloudLabel / cloudNeighborPicked / cloudSortInd
        assert np.array_equal(C.ints(which)[5:-6], G.ints(which)[5:-6]), (k, "ints", which)
    for which in (3, 4, 5, 11, 12):  # feature index lists in emission order, scanStartInd / scanEndInd
        assert np.array_equal(C.ints(which), G.ints(which)), (k, "ints", which)
    assert biteq(C.floats(0)[5:-5], G.floats(0)[5:-5]), (k, "curvature")
    assert biteq(np.array(r.imu_trans[:], np.float32), G.imu_trans()), (k, "imu_trans")


@pytest.mark.parametrize("preset,nsw", [(1, 12), (2, 5), (4, 5)])
def test_scan_registration_node(pkg, checker, synth, preset, nsw):
    """SR-1..SR-6: feature index sets bit-exact, incl. IMU deskew (preset 4)"""
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    G = pkg.ScanRegistration(pkg.make_params(preset, system_delay=0))
    for k in range(nsw):
        xyz, st = S.sweep(k)
        feed_imu(S, k, C, G)
        r = C.push(xyz, st)
        rc = G.process(xyz, st)
        assert rc == r.sr_status
        if rc == 0:
            _sr_compare(C, G, r, k)


def test_scan_registration_edge_cases(pkg, checker, synth):
    """NaN points, out-of-range elevations (dropped rings), strided input, system delay, bad input"""
    S = synth.Synth(1)
    P = pkg.make_params(1, system_delay=3)
    G = pkg.ScanRegistration(P)
    Pc = checker.make_params(1)
    Pc.system_delay = 3
    C = checker.Checker(Pc, "port")
    rng = np.random.default_rng(5)
    for k in range(6):
        xyz, st = S.sweep(k)
        xyz = xyz.copy()
        bad = rng.choice(len(xyz), 300, replace=False)
        xyz[bad[:100], 0] = np.nan
        xyz[bad[100:200], 2] = np.inf
        xyz[bad[200:], 2] = 50.0  # elevation far outside +-15 deg: ring id out of range, point dropped
        if k == 4:
            xyz[0] = np.nan  # first raw point invalid: startOri comes from the first finite one
        r = C.push(xyz, st)
        padded = np.zeros((len(xyz), 5), np.float32)  # stride 20 bytes
        padded[:, :3] = xyz
        rc = pkg._check(G.L.loam_sr_process(G.h, padded.ctypes.data, len(xyz), 20, float(st)))
        assert rc == r.sr_status, k  # first 3 sweeps swallowed by systemDelay (:213-219)
        if rc == 0:
            assert r.n_full < len(xyz) - 250
            _sr_compare(C, G, r, k)
    with pytest.raises(pkg.LoamError):
        G.process(np.zeros((P.max_points + 1, 3), np.float32), 1.0)  # capacity error, not a crash
    assert G.L.loam_sr_process(G.h, None, 10, 12, 0.0) < 0
    assert G.process(np.full((100, 3), np.nan, np.float32), 2.0) == 1  # nothing finite: NOT_READY


@pytest.mark.parametrize("preset,nsw", [(1, 40), (2, 10)])
def test_laser_odometry_node(pkg, checker, synth, preset, nsw):
    """LO-1..LO-12 through the node-level API: correspondences, poses and re-projected clouds"""
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    P = pkg.make_params(preset, system_delay=0)
    SR, LO = pkg.ScanRegistration(P), pkg.LaserOdometry(P)
    iters = set()
    for k in range(nsw):
        xyz, st = S.sweep(k)
        r = C.push(xyz, st)
        if SR.process(xyz, st) != 0:
            continue
        rc, pub = LO.process(SR)
        assert rc == r.lo_status, k
        tr, ts, stt = LO.state()
        assert biteq(tr, r.transform[:]) and biteq(ts, r.transform_sum[:]), k
        assert pub == (r.lm_status == 0)
        if rc == 0:
            assert (stt[0], stt[1], stt[2]) == (r.lo_iters, r.lo_rows, r.lo_degenerate), k
            iters.add(int(stt[0]))
            for w in (6, 7, 8, 9, 10):
                assert np.array_equal(C.ints(w), LO.ints(w)), (k, w)
        for w in (5, 6):
            assert biteq(C.cloud(w), LO.cloud(w)), (k, w)
        if pub:
            assert biteq(C.cloud(7), LO.cloud(7)), k
    assert max(iters) >= 6 and min(iters) <= 2  # both the weighted (iter >= 5) and early-exit paths ran


def test_laser_odometry_teacher_forced(pkg, checker, synth):
    """teacher forcing: overwrite transform / transformSum on both sides with the same arbitrary
    state before a sweep; the next outputs must still agree bit for bit (large motion prior =>
    25 iterations, NN refresh at 5/10/15/20, weights)"""
    S = synth.Synth(1)
    C = checker.Checker(checker.make_params(1), "port")
    P = pkg.make_params(1, system_delay=0)
    SR, LO = pkg.ScanRegistration(P), pkg.LaserOdometry(P)
    for k in range(8):
        xyz, st = S.sweep(k)
        if k == 5:
            t = np.array([0.01, -0.02, 0.015, 0.05, -0.03, 0.2], np.float32)
            s = np.array([0.02, 0.3, -0.01, 1.0, 0.1, 4.0], np.float32)
            C.lo_set_state(t, s)
            LO.set_state(t, s)
        r = C.push(xyz, st)
        if SR.process(xyz, st) != 0:
            continue
        rc, pub = LO.process(SR)
        tr, ts, stt = LO.state()
        assert biteq(tr, r.transform[:]) and biteq(ts, r.transform_sum[:]), k
        if k == 5:
            assert stt[0] == r.lo_iters and r.lo_iters >= 10


@pytest.mark.parametrize("preset,nsw", [(1, 60), (2, 16), (4, 8)])
def test_full_pipeline_free_running(pkg, checker, synth, preset, nsw):
    """sr -> lo -> lm -> tm free running: every pose, the local map, the 5-NN indices, the
    registered cloud -- bit-exact against the oracle, and against the reference golden fixture"""
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    G = pkg.Pipeline(pkg.make_params(preset, system_delay=0))
    gold = np.load(os.path.join(GOLDEN, "ref_trajectories.npz"))
    gposes = gold[f"cfg{preset}_poses"]
    mapped = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        feed_imu(S, k, C, G)
        r = C.push(xyz, st)
        g = G.push(xyz, st)
        da, db = r.as_dict(), g.as_dict()
        for key in da:
            if isinstance(da[key], np.ndarray):
                assert biteq(da[key], db[key]), (k, key, da[key], db[key])
            else:
                assert da[key] == db[key], (k, key, da[key], db[key])
        if k < len(gposes):  # the reference's own code produced these
            mine = np.concatenate([db[key] for key in ("transform", "transform_sum", "transform_aft_mapped", "transform_bef_mapped", "transform_mapped", "imu_trans")])
            assert biteq(mine, gposes[k]), k
        if g.lm_status == 0:
            mapped += 1
            for w in (10, 11, 12, 13, 8, 9):   # 9 = /laser_cloud_surround (refreshed every 5th mapped sweep)
                assert biteq(C.cloud(w), G.lm.cloud(w, cap=max(200000, G.lm.cap))), (k, w)
            for w in (100, 101):
                assert np.array_equal(C.ints(w), G.lm.ints(w)), (k, w)
    assert mapped >= nsw // 2 - 1


def test_map_arena_compaction_keeps_parity(pkg, checker, synth):
    """a tiny arena forces the log-structured map to compact many times; results must not change"""
    S = synth.Synth(1)
    C = checker.Checker(checker.make_params(1), "port")
    P = pkg.make_params(1, system_delay=0)
    P.map_arena_points = 1 << 16
    G = pkg.Pipeline(P)
    for k in range(50):
        xyz, st = S.sweep(k)
        r = C.push(xyz, st)
        g = G.push(xyz, st)
        assert biteq(np.array(r.transform_aft_mapped[:]), np.array(g.transform_aft_mapped[:])), k
        assert (r.n_map_corner, r.n_map_surf) == (g.n_map_corner, g.n_map_surf), k


def test_async_and_resident_paths_equal_sync_path(pkg, synth):
    """push_async / push_device (what bench.py times) produce exactly what push() produces, and the
    whole run is deterministic run to run"""
    import torch
    S = synth.Synth(1)
    sweeps = [S.sweep(k) for k in range(30)]
    P = pkg.make_params(1, system_delay=0)
    A = pkg.Pipeline(P)
    ref = [A.push(x, st).as_dict() for x, st in sweeps]
    B = pkg.Pipeline(P)
    for x, st in sweeps:
        B.push_async(x, st)
    got = [r.as_dict() for r in B.collect()]
    D = pkg.Pipeline(P)
    keep = []
    for x, st in sweeps:
        t = torch.from_numpy(x).cuda()
        keep.append(t)
        fin = np.nonzero(np.isfinite(x).all(axis=1))[0]
        D.push_device(t.data_ptr(), len(x), np.concatenate([x[fin[0]], x[fin[-1]]]), int(fin[0]), st)
    got2 = [r.as_dict() for r in D.collect()]
    assert len(got) == len(ref) == len(got2)
    for a, b, c in zip(ref, got, got2):
        for key in a:
            if isinstance(a[key], np.ndarray):
                assert biteq(a[key], b[key]) and biteq(a[key], c[key]), key
            else:
                assert a[key] == b[key] == c[key], key


@pytest.mark.parametrize("preset,nsw", [(1, 60), (2, 24)])
def test_overlapped_streams_equal_sync_path(pkg, synth, preset, nsw):
    """the free-running pipeline overlaps scanRegistration(k+1), laserOdometry(k) and
    laserMapping(k-1) on three streams (banked hand-over buffers, loam_sr.cuh / loam_lm.cuh) and
    reads CUDA-pinned input in place: with IMU data, many sweeps queued at once and the 8-CTA
    cluster map update, every pose and the final map clouds equal the synchronous run bit for bit"""
    import torch
    S = synth.Synth(preset)
    sweeps = [S.sweep(k) for k in range(nsw)]
    P = pkg.make_params(preset, system_delay=0)
    A = pkg.Pipeline(P)
    ref = []
    for k, (x, st) in enumerate(sweeps):
        feed_imu(S, k, A)
        ref.append(A.push(x, st).as_dict())
    B = pkg.Pipeline(P)
    pinned = [torch.from_numpy(x).pin_memory() for x, _ in sweeps]   # read in place by the copy engine
    for k, (x, st) in enumerate(sweeps):
        feed_imu(S, k, B)
        B.push_async(pinned[k].numpy(), st)
        if k == nsw // 2:
            B.join()          # device-side join in the middle of the queue must not change anything
    got = [r.as_dict() for r in B.collect()]
    assert len(got) == len(ref)
    for k, (a, b) in enumerate(zip(ref, got)):
        for key in a:
            if isinstance(a[key], np.ndarray):
                assert biteq(a[key], b[key]), (k, key)
            else:
                assert a[key] == b[key], (k, key)
    cap = max(200000, A.lm.cap)
    for w in (10, 11, 12, 13, 8, 9):
        assert biteq(A.lm.cloud(w, cap=cap), B.lm.cloud(w, cap=cap)), w
    for w in (0, 1, 2, 3, 4):
        assert biteq(A.sr.cloud(w), B.sr.cloud(w)), w
