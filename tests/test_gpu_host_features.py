"""GPU (-m gpu): laserOdometry fed from HOST copies of scanRegistration's six topics
(loam_lo_process_host, SURVEY.md §8b `LaserOdometry::process(const Features&, OdomResult*)`)."""
import numpy as np
import pytest

from conftest import biteq, feed_imu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("preset,nsw", [(1, 30), (2, 8), (4, 6)])
def test_lo_process_host_equals_oracle_and_device_path(pkg, checker, synth, preset, nsw):
    S = synth.Synth(preset)
    P = pkg.make_params(preset, system_delay=0)
    C = checker.Checker(checker.make_params(preset), "port")
    sr, lo_dev = pkg.ScanRegistration(P), pkg.LaserOdometry(P)
    lo_host = pkg.LaserOdometry(P)
    published = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        feed_imu(S, k, C, sr)
        r = C.push(xyz, st)
        g = sr.process(xyz, st)
        assert g == r.sr_status
        if r.sr_status != 0:
            continue
        feats = [C.cloud(w) for w in (0, 1, 2, 3, 4)]           # the oracle's features, held on the host
        imu_trans = np.array(r.imu_trans[:], np.float32)
        rc_h, pub_h = lo_host.process_host(*feats, imu_trans)
        rc_d, pub_d = lo_dev.process(sr)
        assert (rc_h, pub_h) == (rc_d, pub_d) == (r.lo_status, r.lm_status == 0), k
        th, sh, ih = lo_host.state()
        td, sd, idv = lo_dev.state()
        assert biteq(th, td) and biteq(sh, sd) and np.array_equal(ih[:3], idv[:3]), k
        assert biteq(th, r.transform[:]) and biteq(sh, r.transform_sum[:]), k
        if rc_h == 0:
            for which in (5, 6):
                assert np.array_equal(lo_host.cloud(which).view(np.uint32), C.cloud(which).view(np.uint32)), (k, which)
            if pub_h:
                published += 1
                assert np.array_equal(lo_host.cloud(7).view(np.uint32), C.cloud(7).view(np.uint32)), k
    assert published >= 2


def test_lo_process_host_rejects_unordered_features(pkg):
    P = pkg.make_params(1, system_delay=0)
    lo = pkg.LaserOdometry(P)
    pts = np.zeros((4, 4), np.float32)
    pts[:, 3] = [0.0, 2.0, 1.0, 3.0]          # rings out of order
    with pytest.raises(pkg.LoamError):
        lo.process_host(pts, pts[:1], pts, pts[:1], pts)
    pts[:, 3] = [0.0, 1.0, 2.0, 99.0]         # ring outside the sensor
    with pytest.raises(pkg.LoamError):
        lo.process_host(pts, pts[:1], pts, pts[:1], pts)


def test_lo_process_host_rejects_feature_clouds_beyond_the_banks(pkg):
    """sharp / less-sharp / flat clouds larger than what scanRegistration can emit (12 / 120 / 24 points per
    ring) but below max_points must be refused, not copied past the device banks"""
    P = pkg.make_params(1, system_delay=0)
    lo = pkg.LaserOdometry(P)
    def cloud(n):
        c = np.zeros((n, 4), np.float32)
        c[:, 3] = np.minimum(np.arange(n) * 16 // max(n, 1), 15)
        return c
    ok = cloud(100)
    for which, n in ((1, 16 * 12 + 1), (2, 16 * 120 + 1), (3, 16 * 24 + 1)):
        args = [cloud(2000), ok[:10], ok, ok[:10], cloud(2000)]
        args[which] = cloud(n)
        with pytest.raises(pkg.LoamError, match="exceeds the bank"):
            lo.process_host(*args)
    rc, _ = lo.process_host(cloud(2000), cloud(16 * 12), cloud(16 * 120), cloud(16 * 24), cloud(2000))   # exactly full banks are fine
    assert rc in (0, 1)
