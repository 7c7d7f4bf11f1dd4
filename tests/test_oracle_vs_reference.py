"""CPU: the oracle's restatement (oracle/loam_port.hpp) against the reference itself.

Two pins: (1) tests/golden/ref_trajectories.npz, produced in the build container by oracle/_ref =
the reference's own src/*.cpp compiled from /root/reference (tests/golden/make_ref_golden.py);
(2) when oracle/_ref is present next to the repo, a live bit-for-bit comparison."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, biteq, cloud_checksum, feed_imu

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


@pytest.mark.parametrize("preset,nsw", [(1, 30), (2, 6)])
def test_port_equals_live_reference(checker, synth, preset, nsw):
    if not checker.available("reference"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    S = synth.Synth(preset)
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
