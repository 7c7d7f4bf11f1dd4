"""GPU (-m gpu): the guard / degeneracy branches of the two LM loops, which the BASELINE.json scenes never
take:

  laserOdometry.cpp:465      cornerLast > 10 && surfLast > 100 gate
  laserOdometry.cpp:697-700  fewer than 10 selected rows -> `continue`
  laserOdometry.cpp:770-797  cv::eigen threshold scan, isDegenerate, matP, X = matP * X
  laserOdometry.cpp:799-811  NaN guard (the fork's one functional change)
  laserMapping.cpp:886-889   fewer than 50 selected rows -> `continue`
  laserMapping.cpp:927-954   the same degeneracy analysis, threshold 100

Every test asserts that the ORACLE takes the branch (counters non-zero on its side) and that the CUDA
path then agrees with it bit for bit: poses, iteration counts, row counts, isDegenerate and matP.
"""
import ctypes
import os

import numpy as np
import pytest

from conftest import GOLDEN, biteq, inside_reference_domain

pytestmark = pytest.mark.gpu

POSE_KEYS = ("transform", "transform_sum", "transform_aft_mapped", "transform_bef_mapped", "transform_mapped", "imu_trans")


def _matp_gpu(pkg, node, which):
    L = pkg.lib()
    L.loam_debug_matp.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    out = np.zeros(36, np.float32)
    assert L.loam_debug_matp(node.h, which, out.ctypes.data) == 0
    return out


def _same_result(r, g, k):
    da, db = r.as_dict(), g.as_dict()
    for key in da:
        if isinstance(da[key], np.ndarray):
            assert biteq(da[key], db[key]), (k, key, da[key], db[key])
        else:
            assert da[key] == db[key], (k, key, da[key], db[key])


@pytest.mark.parametrize("scene,nsw", [(1, 40), (2, 40), (3, 60)])
def test_degenerate_scenes_free_running(pkg, checker, synth, scene, nsw):
    """tunnel / ribbed tunnel / sparse scenes (synth.cpp): isDegenerate is raised in laserMapping (all
    three) and laserOdometry (sparse), the update is projected with matP, runs of 25 iterations that never
    converge -- every sweep bit-equal to the oracle and to the fixture made by the reference's own sources"""
    S = synth.Synth(1, scene=scene)
    C = checker.Checker(checker.make_params(1), "port")
    G = pkg.Pipeline(pkg.make_params(1, system_delay=0))
    gold = np.load(os.path.join(GOLDEN, "ref_trajectories.npz"))
    gposes = gold[f"cfg1s{scene}_poses"]
    lo_deg = lm_deg = lo_deg_matp = lm_deg_matp = 0
    lo_iters, lm_iters = set(), set()
    prev = None
    for k in range(nsw):
        xyz, st = S.sweep(k)
        r = C.push(xyz, st)
        assert inside_reference_domain(prev, r), k
        prev = r
        g = G.push(xyz, st)
        _same_result(r, g, k)
        mine = np.concatenate([g.as_dict()[key] for key in POSE_KEYS])
        assert biteq(mine, gposes[k]), (k, "reference fixture")
        if r.lo_status == 0 and r.lo_iters > 0:
            lo_iters.add(r.lo_iters)
            if r.lo_degenerate:
                lo_deg += 1
                assert biteq(C.floats(1), _matp_gpu(pkg, G.lo, 0)), (k, "lo matP")
                lo_deg_matp += 1
        if r.lm_status == 0:
            lm_iters.add(r.lm_iters)
            if r.lm_degenerate:
                lm_deg += 1
                assert biteq(C.floats(2), _matp_gpu(pkg, G.lm, 1)), (k, "lm matP")
                lm_deg_matp += 1
            for w in (10, 11, 12, 13, 8):
                assert biteq(C.cloud(w), G.lm.cloud(w, cap=200000)), (k, w)
            for w in (100, 101):
                assert np.array_equal(C.ints(w), G.lm.ints(w)), (k, w)
    # the oracle itself took the branches
    assert lm_deg >= 1 and lm_deg_matp == lm_deg, (lm_deg, lm_iters)
    if scene == 1:
        assert lm_deg >= 10
    if scene == 3:
        assert lm_deg >= 20 and lo_deg >= 5 and lo_deg_matp == lo_deg, lo_deg
    if scene != 2:
        assert 25 in lo_iters        # runs that never converge
    assert len(lm_iters) >= 2


def _crafted_plane_features(shift, reltime=0.0, patches=False):
    """features in laserOdometry's frame (y up) that make A^T A exactly singular: every flat point and
    every point of the last surf cloud lie on planes y = const with bit-exact coordinates, the motion
    prior is zero, so the columns of rotation about y and of translation along x and z are exactly 0.
    cv::solve(DECOMP_QR) then divides 0 by 0 (hal::QR has no pivoting): matX is NaN, the eigenvalues
    are 0 (isDegenerate), matP * NaN stays NaN, and the NaN guard of laserOdometry.cpp:799-811 skips the
    update 25 times per sweep.  The corner cloud sits inside one 0.2 m voxel, so laserMapping never
    optimises (its own loop has no NaN guard and would poison the map pose in the reference too).
    patches=True adds two small wall patches (normals along x and z) with a handful of flat points on
    them: the system becomes solvable but three eigenvalues stay far below 10, so the update is
    projected with matP (laserOdometry.cpp:793-797) instead of being dropped."""
    R = 16
    lf, fl = [], []
    for j in range(R):
        for i in range(40):
            lf.append((i * 0.5 - 10.0, -1.5, 2.0 + j * 0.5, float(j)))
        if patches:
            for i in range(3):
                lf.append((8.0, -1.5 + 0.1 * j + 0.03125, 4.0 + 0.3 * i, float(j)))
            for i in range(3):
                lf.append((-2.0 + 0.3 * i, -1.4 + 0.1 * j, 12.0, float(j)))
        for i in range(4):
            fl.append((i * 2.0 - 3.75 + shift, -1.5 + 0.015625, 2.0 + j * 0.5 + 0.125, float(j) + 0.1 * reltime * (i + 1) / 4))
        if patches and j in (3, 7, 11):
            fl.append((8.0 - 0.02, -1.5 + 0.1 * j + 0.08, 4.15 + shift, float(j) + 0.1 * reltime * 0.5))
        if patches and j in (5, 9):
            fl.append((-1.85 + shift, -1.4 + 0.1 * j + 0.05, 12.0 - 0.02, float(j) + 0.1 * reltime * 0.25))
    ls = [(3.0, -1.0 + 0.01 * j, 6.0, float(j)) for j in range(R)]
    lf, fl, ls = np.array(lf, np.float32), np.array(fl, np.float32), np.array(ls, np.float32)
    return lf.copy(), np.zeros((0, 4), np.float32), ls, fl, lf


def test_lo_nan_guard_with_crafted_features(pkg, checker):
    P = pkg.make_params(1, system_delay=0)
    C = checker.Checker(checker.make_params(1), "port")
    LO = pkg.LaserOdometry(P)
    fired = 0
    for k in range(8):
        feats = _crafted_plane_features(0.015625 * k)
        if k == 4:   # a non-trivial accumulated pose for the integration step; the motion prior stays exactly zero
            z = np.zeros(6, np.float32)
            s = np.array([0.02, 0.3, -0.01, 1.0, 0.1, 4.0], np.float32)
            C.lo_set_state(z, s)
            LO.set_state(z, s)
        r = C.lo_push(*feats, stamp=0.1 * k)
        rc, pub = LO.process_host(*feats)
        assert rc == r.lo_status and pub == (r.lm_status == 0), k
        t, s, st = LO.state()
        assert biteq(t, r.transform[:]) and biteq(s, r.transform_sum[:]), k
        if r.lo_iters:
            # 25 iterations without convergence (deltaR is NaN), 64 rows more per iteration, update never applied
            assert (int(st[0]), int(st[1]), int(st[2])) == (r.lo_iters, r.lo_rows, r.lo_degenerate) == (25, 1600, 1), (k, st)
            assert not np.any(np.array(r.transform[:])), k
            fired += 1
    assert fired >= 5


DEG_PRIOR = np.array([0.03, -0.05, 0.04, 0.2, -0.15, 0.1], np.float32)


def test_lo_degenerate_projection_with_crafted_features(pkg, checker):
    """one dominant plane + two tiny wall patches: solvable, but three eigenvalues of A^T A stay below 10, so
    every update of laserOdometry is X = matP * X (laserOdometry.cpp:770-797); teacher-forced motion prior
    so that several iterations run; poses, counters and matP bit-equal"""
    P = pkg.make_params(1, system_delay=0)
    C = checker.Checker(checker.make_params(1), "port")
    LO = pkg.LaserOdometry(P)
    projected, iters = 0, set()
    for k in range(10):
        feats = _crafted_plane_features(0.015625 * k, reltime=0.9, patches=True)
        if k >= 2:
            s = LO.state()[1]
            C.lo_set_state(DEG_PRIOR * (1 + 0.25 * (k % 3)), s)
            LO.set_state(DEG_PRIOR * (1 + 0.25 * (k % 3)), s)
        r = C.lo_push(*feats, stamp=0.1 * k)
        rc, pub = LO.process_host(*feats)
        assert rc == r.lo_status and pub == (r.lm_status == 0), k
        t, s, st = LO.state()
        assert biteq(t, r.transform[:]) and biteq(s, r.transform_sum[:]), k
        if r.lo_iters:
            assert (int(st[0]), int(st[1]), int(st[2])) == (r.lo_iters, r.lo_rows, r.lo_degenerate), (k, st)
            assert r.lo_degenerate == 1 and np.isfinite(np.array(r.transform[:])).all(), k
            assert biteq(C.floats(1), _matp_gpu(pkg, LO, 0)), k
            iters.add(r.lo_iters)
            if np.abs(np.array(r.transform[:]) - DEG_PRIOR * (1 + 0.25 * (k % 3))).max() > 1e-5:
                projected += 1          # a projected, finite, non-zero update was applied
            for w in (8, 9, 10):
                assert np.array_equal(C.ints(w), LO.ints(w)), (k, w)
    assert projected >= 4 and max(iters) >= 2, (projected, iters)


def test_lo_min_rows_gate_and_persisting_matp(pkg, checker, synth):
    """teacher forcing in the sparse scene: a motion prior far off (300 m, later 3000 m along the sweep)
    leaves no correspondence but for the points fired in the first instants of the sweep: whole sweeps
    of 25 iterations that all `continue` with fewer than 10 rows (:697-700), and a sweep whose rows --
    never cleared between iterations in this fork -- creep up by one per iteration until the 11th
    iteration is the first to solve, with the isDegenerate / matP left over from EARLIER sweeps because
    iteration 0 never reached the eigen analysis"""
    S = synth.Synth(1, scene=3)
    C = checker.Checker(checker.make_params(1), "port")
    P = pkg.make_params(1, system_delay=0)
    SR, LO = pkg.ScanRegistration(P), pkg.LaserOdometry(P)
    prior = {6: 300.0, 9: 0.0, 12: 3000.0}
    no_rows = stale_matp = 0
    for k in range(16):
        xyz, st = S.sweep(k)
        if k in prior:
            t = np.array([0.0, 0.0, 0.0, 0.0, 0.0, prior[k]], np.float32)
            s = LO.state()[1]
            C.lo_set_state(t, s)
            LO.set_state(t, s)
        r = C.push(xyz, st)
        assert SR.process(xyz, st) == r.sr_status
        if r.sr_status != 0:
            continue
        rc, pub = LO.process(SR)
        assert rc == r.lo_status, k
        tr, ts, stt = LO.state()
        assert biteq(tr, r.transform[:]) and biteq(ts, r.transform_sum[:]), k
        if rc == 0:
            assert (int(stt[0]), int(stt[1]), int(stt[2])) == (r.lo_iters, r.lo_rows, r.lo_degenerate), (k, stt)
            if r.lo_degenerate:
                assert biteq(C.floats(1), _matp_gpu(pkg, LO, 0)), k
            if r.lo_iters == 25 and r.lo_rows < 10:
                no_rows += 1
            if 1 < r.lo_iters < 25 and 10 <= r.lo_rows < 2 * r.lo_iters and r.lo_degenerate:
                stale_matp += 1          # the first solve happened in a late iteration, projected with the old matP
            for w in (6, 7, 8, 9, 10):
                assert np.array_equal(C.ints(w), LO.ints(w)), (k, w)
    assert no_rows >= 4 and stale_matp >= 1, (no_rows, stale_matp)


def test_lm_min_rows_gate(pkg, checker, synth):
    """laserMapping fed with a pose 4 m above the corridor map: the local map is there (> 10 corner, > 100
    surf points) but no stack point finds five neighbours within 1 m, so all ten iterations `continue`
    with fewer than 50 rows (laserMapping.cpp:886-889) and the pose comes out as predicted"""
    S = synth.Synth(1)
    C = checker.Checker(checker.make_params(1), "port")
    G = pkg.Pipeline(pkg.make_params(1, system_delay=0))
    rec = []
    for k in range(24):
        xyz, st = S.sweep(k)
        r = C.push(xyz, st)
        g = G.push(xyz, st)
        _same_result(r, g, k)
        if r.lm_status == 0:
            rec.append((C.cloud(5), C.cloud(6), C.cloud(7), np.array(r.transform_sum[:], np.float32), st))
    LM = G.lm
    skipped = 0
    for i, (corner, surf, full, tsum, st) in enumerate(rec[-4:]):
        t = tsum.copy()
        if i % 2 == 1:
            t[4] += 4.0 if i == 1 else -4.0   # y = up in laserMapping's frame; the first offset pass leaves points up there
        r = C.lm_push(corner, surf, full, t, st + 10.0)
        rc = LM.process_host(corner, surf, full, t, st + 10.0)
        assert rc == r.lm_status == 0
        aft, bef, tobe, stt = LM.state()
        assert biteq(aft, r.transform_aft_mapped[:]) and biteq(bef, r.transform_bef_mapped[:]), i
        assert (int(stt[0]), int(stt[1]), int(stt[2])) == (r.lm_iters, r.lm_rows, r.lm_degenerate), (i, stt)
        assert (int(stt[3]), int(stt[4])) == (r.n_map_corner, r.n_map_surf), i
        if i % 2 == 1:
            assert r.n_map_corner > 10 and r.n_map_surf > 100 and r.lm_iters == 10 and r.lm_rows < 50, (r.lm_iters, r.lm_rows)
            skipped += 1
        for w in (100, 101):
            assert np.array_equal(C.ints(w), LM.ints(w)), (i, w)
    assert skipped == 2


def _near_threshold_systems(rng, n, thr):
    """symmetric positive definite 6x6 float matrices whose 1..3 smallest eigenvalues lie within
    +-1e-5 * trace of the threshold (where an fp32 Jacobi sweep can land on either side)"""
    A = np.zeros((n, 6, 6), np.float32)
    b = rng.normal(0, 50, (n, 6)).astype(np.float32)
    for t in range(n):
        q, _ = np.linalg.qr(rng.normal(size=(6, 6)))
        lam = np.sort(10.0 ** rng.uniform(np.log10(thr) + 0.05, 6.0, 6))[::-1]
        tr = lam.sum()
        m = 1 + (t % 3)
        lam[6 - m:] = thr + rng.uniform(-1e-5, 1e-5, m) * tr
        if t % 7 == 0:
            lam[5] = thr * rng.uniform(0.0, 0.9)      # clearly degenerate
        if t % 11 == 0:
            lam[6 - m:] = thr * rng.uniform(1.2, 3.0, m)  # clearly not
        M = (q * lam) @ q.T
        A[t] = ((M + M.T) * 0.5).astype(np.float32)
        A[t] = np.triu(A[t]) + np.triu(A[t], 1).T
    return A, b


@pytest.mark.parametrize("thr", [10.0, 100.0])
def test_solve_step_near_the_eigen_threshold(pkg, checker, thr):
    """lm_solve_step's positive-definiteness pre-test (skips cv::eigen when no eigenvalue can be below
    the threshold) must never change isDegenerate, X or -- where it is read -- matP: 10^4 systems with
    eigenvalues within 1e-5 * trace of the threshold against the always-decompose path on the device
    and against the oracle's restatement of laserOdometry.cpp:765-797"""
    L = pkg.lib()
    vp, ci = ctypes.c_void_p, ctypes.c_int
    L.loam_debug_solve_step.argtypes = [vp, vp, ci, ctypes.c_float, ci, vp, vp, vp]
    Lo = checker._load("port")
    n = 10000
    A, b = _near_threshold_systems(np.random.default_rng(int(thr)), n, thr)
    res = []
    for mode in (0, 1):
        deg = np.zeros(n, np.int32); P = np.zeros((n, 36), np.float32); X = np.zeros((n, 6), np.float32)
        assert L.loam_debug_solve_step(A.ctypes.data, b.ctypes.data, n, thr, mode, deg.ctypes.data, P.ctypes.data, X.ctypes.data) == 0
        res.append((deg, P, X))
    dego = np.zeros(n, np.int32); Po = np.zeros((n, 36), np.float32); Xo = np.zeros((n, 6), np.float32)
    Lo.chk_solve_step(A.ctypes.data, b.ctypes.data, n, thr, dego.ctypes.data, Po.ctypes.data, Xo.ctypes.data)
    (d0, P0, X0), (d1, P1, X1) = res
    assert 0.2 * n < dego.sum() < 0.8 * n, dego.sum()           # both outcomes are well represented
    assert np.array_equal(d1, dego) and biteq(X1, Xo) and biteq(P1, Po)      # always-decompose == oracle
    assert np.array_equal(d0, dego), int((d0 != dego).sum())               # the shortcut never flips the flag
    assert biteq(X0, Xo)
    m = dego == 1
    assert biteq(P0[m], Po[m])                                             # matP is only read while isDegenerate


def test_documented_limits_are_refused_with_a_specific_message(pkg, synth):
    """include/loam_b200.h "Hard limits": a parameter block or a sweep beyond a compile-time capacity of the kernels is
    refused with an error text that names the limit -- never truncated silently, never a crash"""
    P = pkg.make_params(1, system_delay=0)
    P.lo_max_iter = 33
    with pytest.raises(pkg.LoamError, match=r"lo_max_iter = 33 is outside 1 \.\. 32"):
        pkg.Pipeline(P)
    P = pkg.make_params(1, system_delay=0)
    P.lm_max_iter = 0
    with pytest.raises(pkg.LoamError, match="lm_max_iter"):
        pkg.LaserMapping(P)
    P = pkg.make_params(1, system_delay=0)
    P.n_scans = 129
    with pytest.raises(pkg.LoamError, match="LOAM_LIMIT_RINGS"):
        pkg.ScanRegistration(P)
    P = pkg.make_params(1, system_delay=0)
    P.map_arena_points = 100
    with pytest.raises(pkg.LoamError, match="map_arena_points"):
        pkg.LaserMapping(P)
    # one ring with far more points than its buffers hold: 9000 returns at -24 deg elevation (all on ring 2 of the HDL-64 model)
    P = pkg.make_params(2, system_delay=0)
    sr = pkg.ScanRegistration(P)
    az = np.linspace(0.0, -2.0 * np.pi, 9000, endpoint=False)
    rng = 10.0 + 0.5 * np.sin(7 * az)
    xyz = np.stack([rng * np.cos(az), rng * np.sin(az), np.tan(np.radians(-24.0)) * rng], axis=1).astype(np.float32)
    with pytest.raises(pkg.LoamError, match="LOAM_LIMIT_(SECTOR|RING)_POINTS"):
        assert sr.process(xyz, 0.0) == 1     # the first sweep is swallowed (scanRegistration.cpp:213-219)
        sr.process(xyz, 0.1)
