"""CPU: the oracle's restatement (oracle/loam_port.hpp) against the reference itself.

Two pins: (1) tests/golden/ref_trajectories.npz, produced in the build container by oracle/_ref =
the reference's own src/*.cpp compiled from /root/reference (tests/golden/make_ref_golden.py);
(2) when oracle/_ref is present next to the repo, a live bit-for-bit comparison."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, biteq, cloud_checksum, feed_imu, inside_reference_domain

G = np.load(os.path.join(GOLDEN, "ref_trajectories.npz"))
INT_KEYS = [str(k) for k in G["int_keys"]]
POSE_KEYS = [str(k) for k in G["pose_keys"]]


@pytest.mark.parametrize("preset,nsw", [(1, 100), (2, 12), (4, 8)])
def test_port_reproduces_reference_golden(checker, synth, preset, nsw):
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    ints, poses, sums = G[f"cfg{preset}_ints"], G[f"cfg{preset}_poses"], G[f"cfg{preset}_sums"]
    for k in range(nsw):
        xyz, st = S.sweep(k)
        feed_imu(S, k, C)
        r = C.push(xyz, st)
        d = r.as_dict()
        assert [d[key] for key in INT_KEYS] == ints[k].tolist(), k
        assert biteq(np.concatenate([d[key] for key in POSE_KEYS]), poses[k]), k
        if r.sr_status == 0:
            for w in (0, 1, 2, 3, 4):
                assert cloud_checksum(C.cloud(w)) == sums[k, w], (k, w)
        if r.lm_status == 0:
            for w in (5, 6, 7, 8, 10, 11, 12, 13):
                assert cloud_checksum(C.cloud(w)) == sums[k, w], (k, w)


@pytest.mark.parametrize("scene,nsw", [(1, 24), (2, 24), (3, 40)])
def test_port_reproduces_reference_golden_on_guard_branch_scenes(checker, synth, scene, nsw):
    """tunnel / ribbed tunnel / sparse scenes (synth.cpp): the restatement's isDegenerate / matP projection,
    small-row-count and small-corner-cloud paths against what the reference's own sources produced"""
    S = synth.Synth(1, scene=scene)
    C = checker.Checker(checker.make_params(1), "port")
    ints, poses, sums = G[f"cfg1s{scene}_ints"], G[f"cfg1s{scene}_poses"], G[f"cfg1s{scene}_sums"]
    deg = 0
    prev = None
    for k in range(nsw):
        xyz, st = S.sweep(k)
        r = C.push(xyz, st)
        assert inside_reference_domain(prev, r), k
        prev = r
        d = r.as_dict()
        assert [d[key] for key in INT_KEYS] == ints[k].tolist(), k
        assert biteq(np.concatenate([d[key] for key in POSE_KEYS]), poses[k]), k
        deg += r.lm_degenerate + r.lo_degenerate
        if r.sr_status == 0:
            for w in (0, 1, 2, 3, 4):
                assert cloud_checksum(C.cloud(w)) == sums[k, w], (k, w)
        if r.lm_status == 0:
            for w in (5, 6, 7, 8, 10, 11, 12, 13):
                assert cloud_checksum(C.cloud(w)) == sums[k, w], (k, w)
    assert deg >= 1          # the branch was taken


def test_nan_guard_port_equals_live_reference(checker):
    """crafted features that make cv::solve return NaN (tests/test_gpu_guard_branches.py): the reference
    prints its [USER WARN] 25 times per sweep and ignores the step; so must the restatement"""
    if not checker.available("reference"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from test_gpu_guard_branches import _crafted_plane_features
    A = checker.Checker(checker.make_params(1), "port")
    B = checker.Checker(checker.make_params(1), "reference")
    fired = 0
    for k in range(6):
        f = _crafted_plane_features(0.015625 * k)
        ra, rb = A.lo_push(*f, stamp=0.1 * k), B.lo_push(*f, stamp=0.1 * k)
        da, db = ra.as_dict(), rb.as_dict()
        for key in da:
            if isinstance(da[key], np.ndarray):
                assert biteq(da[key], db[key]), (k, key)
            elif not key.startswith(("lo_iters", "lo_rows", "lo_deg", "lm_iters", "lm_rows", "lm_deg")):
                assert da[key] == db[key], (k, key)
        fired += int(ra.lo_iters == 25 and ra.lo_rows == 1600 and ra.lo_degenerate == 1)
        for w in (5, 6):
            assert biteq(A.cloud(w), B.cloud(w)), (k, w)
    assert fired >= 3


def test_degenerate_projection_port_equals_live_reference(checker):
    """crafted features whose A^T A is solvable but has three eigenvalues below the threshold: the update
    is projected with matP in the reference (laserOdometry.cpp:793-797) and in the restatement alike"""
    if not checker.available("reference"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from test_gpu_guard_branches import DEG_PRIOR, _crafted_plane_features
    A = checker.Checker(checker.make_params(1), "port")
    B = checker.Checker(checker.make_params(1), "reference")
    deg, iters = 0, set()
    for k in range(8):
        f = _crafted_plane_features(0.015625 * k, reltime=0.9, patches=True)
        if k >= 2:
            s = np.array(ra.transform_sum[:], np.float32)
            A.lo_set_state(DEG_PRIOR * (1 + 0.25 * (k % 3)), s)
            B.lo_set_state(DEG_PRIOR * (1 + 0.25 * (k % 3)), s)
        ra, rb = A.lo_push(*f, stamp=0.1 * k), B.lo_push(*f, stamp=0.1 * k)
        assert biteq(np.array(ra.transform[:]), np.array(rb.transform[:])), k
        assert biteq(np.array(ra.transform_sum[:]), np.array(rb.transform_sum[:])), k
        deg += ra.lo_degenerate
        iters.add(ra.lo_iters)
        if ra.lo_iters:
            for w in (8, 9, 10):
                assert np.array_equal(A.ints(w), B.ints(w)), (k, w)
    assert deg >= 5 and max(iters) >= 2, (deg, iters)


@pytest.mark.parametrize("preset,scene,nsw", [(1, 0, 30), (2, 0, 6), (1, 1, 12), (1, 3, 16)])
def test_port_equals_live_reference(checker, synth, preset, scene, nsw):
    if not checker.available("reference"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    S = synth.Synth(preset, scene=scene)
    A = checker.Checker(checker.make_params(preset), "port")
    B = checker.Checker(checker.make_params(preset), "reference")
    for k in range(nsw):
        xyz, st = S.sweep(k)
        ra, rb = A.push(xyz, st), B.push(xyz, st)
        da, db = ra.as_dict(), rb.as_dict()
        for key in da:
            if isinstance(da[key], np.ndarray):
                assert biteq(da[key], db[key]), (k, key)
            elif not key.startswith(("lo_iters", "lo_rows", "lo_deg", "lm_iters", "lm_rows", "lm_deg")):
                assert da[key] == db[key], (k, key)
        if ra.sr_status == 0:
            for w in (0, 1, 2, 3, 4):
                assert biteq(A.cloud(w), B.cloud(w)), (k, w)
            for w in (0, 1, 2):  # cloudLabel / cloudNeighborPicked / cloudSortInd inside the processed range
                assert np.array_equal(A.ints(w)[5:-6], B.ints(w)[5:-6]), (k, w)
            assert biteq(A.floats(0)[5:-5], B.floats(0)[5:-5])
        if ra.lo_status == 0:
            for w in (6, 7, 8, 9, 10):  # pointSearchCornerInd1/2, pointSearchSurfInd1/2/3
                assert np.array_equal(A.ints(w), B.ints(w)), (k, w)
        if ra.lm_status == 0:
            for w in (5, 6, 7, 8, 10, 11, 12, 13):
                assert biteq(A.cloud(w), B.cloud(w)), (k, w)


def test_glibc_vs_detmath_oracle_within_north_star_tolerance(checker, synth):
    """what substituting detmath for glibc costs: poses stay within 1e-4 rad / 1e-3 m per sweep and
    the feature sets are (nearly) the same -- measured, not assumed"""
    S = synth.Synth(1)
    A = checker.Checker(checker.make_params(1), "port")
    B = checker.Checker(checker.make_params(1), "port-libm")
    same_sets = 0
    n_sets = 0
    for k in range(40):
        xyz, st = S.sweep(k)
        ra, rb = A.push(xyz, st), B.push(xyz, st)
        if ra.sr_status != 0:
            continue
        for w in (3, 4, 5):
            n_sets += 1
            same_sets += int(np.array_equal(A.ints(w), B.ints(w)))
        ta, tb = np.array(ra.transform[:]), np.array(rb.transform[:])
        assert np.abs(ta[:3] - tb[:3]).max() < 1e-4 and np.abs(ta[3:] - tb[3:]).max() < 1e-3, k
        ma, mb = np.array(ra.transform_mapped[:]), np.array(rb.transform_mapped[:])
        assert np.abs(ma[:3] - mb[:3]).max() < 1e-4 * (k + 1) and np.abs(ma[3:] - mb[3:]).max() < 1e-3 * (k + 1), k
    assert same_sets / n_sets > 0.95, (same_sets, n_sets)
