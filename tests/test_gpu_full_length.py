"""GPU (-m gpu): the BASELINE.json configs at their FULL lengths -- 200 VLP-16 sweeps (configs[0]),
500 HDL-64 sweeps (configs[1]), 100 sweeps of the 128-ring + IMU sensor (configs[3]) -- free running
through the C ABI against tests/golden/ref_trajectories.npz, which the reference's OWN sources
(oracle/_ref, compiled from /root/reference/src/*.cpp by oracle/build_ref.sh) produced with
tests/golden/make_ref_golden.py: every status and count, every pose bit for bit, an order-sensitive
checksum of every cloud of every sweep.

North-star gates, stated explicitly:
  * per-sweep pose within 1e-4 rad / 1e-3 m and end-of-trajectory drift within 0.1 % of the path
    length versus the reference CPU path -- met with margin 0 (bit-equal);
  * the same gates against the reference built with glibc's trig instead of loam/detmath.h (the
    `*_libm_poses` fixtures), i.e. what the deterministic-trig substitution costs end to end.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, biteq, cloud_checksum, feed_imu

pytestmark = pytest.mark.gpu

POSE_KEYS = ("transform", "transform_sum", "transform_aft_mapped", "transform_bef_mapped", "transform_mapped", "imu_trans")


def _path_length(poses_mapped):
    return float(np.linalg.norm(np.diff(poses_mapped[:, 3:6], axis=0), axis=1).sum())


@pytest.mark.parametrize("preset,nsw", [(1, 200), (2, 500), (4, 100)])
def test_full_length_config_equals_reference_fixture(pkg, synth, preset, nsw):
    gold = np.load(os.path.join(GOLDEN, "ref_trajectories.npz"))
    int_keys = [str(k) for k in gold["int_keys"]]
    ints, poses, sums = gold[f"cfg{preset}_ints"], gold[f"cfg{preset}_poses"], gold[f"cfg{preset}_sums"]
    assert len(poses) >= nsw
    S = synth.Synth(preset)
    G = pkg.Pipeline(pkg.make_params(preset, system_delay=0))
    mine = np.zeros((nsw, 42), np.float32)
    mapped = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        feed_imu(S, k, G)
        g = G.push(xyz, st)
        d = g.as_dict()
        assert [d[key] for key in int_keys] == ints[k].tolist(), k
        mine[k] = np.concatenate([d[key] for key in POSE_KEYS])
        assert biteq(mine[k], poses[k]), k
        if g.sr_status == 0:
            for w in (0, 1, 2, 3, 4):
                assert cloud_checksum(G.sr.cloud(w)) == sums[k, w], (k, w)
        if g.lm_status == 0:
            mapped += 1
            for w in (5, 6, 7):
                assert cloud_checksum(G.lo.cloud(w)) == sums[k, w], (k, w)
            for w in (8, 10, 11, 12, 13):
                assert cloud_checksum(G.lm.cloud(w, cap=max(400000, G.lm.cap))) == sums[k, w], (k, w)
    assert mapped >= nsw // 2 - 2
    # the north-star gates versus the reference CPU path (trivially met: the arrays are bit-equal)
    ref_m, got_m = poses[:nsw, 24:30], mine[:, 24:30]       # transform_mapped = /integrated_to_init
    L = _path_length(ref_m)
    assert L > (15.0 if preset == 1 else 50.0)
    assert np.abs(got_m[:, :3] - ref_m[:, :3]).max() <= 1e-4 and np.abs(got_m[:, 3:] - ref_m[:, 3:]).max() <= 1e-3
    assert np.linalg.norm(got_m[-1, 3:] - ref_m[-1, 3:]) <= 1e-3 * L


@pytest.mark.parametrize("preset,nsw", [(1, 200), (2, 500)])
def test_full_length_config_within_north_star_tolerance_of_glibc_reference(pkg, synth, preset, nsw):
    """against the reference built with glibc's sin/cos/atan2 (its deployed arithmetic): the sweep-to-sweep
    motion (`transform`, laserOdometry.cpp:93) of every sweep within 1e-4 rad / 1e-3 m, and the end of the
    /integrated_to_init trajectory within 0.1 % of the path length"""
    gold = np.load(os.path.join(GOLDEN, "ref_trajectories.npz"))
    ref = gold[f"cfg{preset}_libm_poses"]
    S = synth.Synth(preset)
    G = pkg.Pipeline(pkg.make_params(preset, system_delay=0))
    mine = np.zeros((nsw, 42), np.float32)
    for k in range(nsw):
        xyz, st = S.sweep(k)
        d = G.push(xyz, st).as_dict()
        mine[k] = np.concatenate([d[key] for key in POSE_KEYS])
    dr = np.abs(mine[:, 0:3] - ref[:nsw, 0:3]).max(axis=1)
    dt = np.abs(mine[:, 3:6] - ref[:nsw, 3:6]).max(axis=1)
    assert dr.max() < 1e-4 and dt.max() < 1e-3, (float(dr.max()), float(dt.max()), int(dr.argmax()), int(dt.argmax()))
    L = _path_length(ref[:nsw, 24:30])
    end = float(np.linalg.norm(mine[-1, 27:30] - ref[nsw - 1, 27:30]))
    assert end <= 1e-3 * L, (end, L)
