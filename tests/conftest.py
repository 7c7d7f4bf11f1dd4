import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "loam_velodyne-1_b200"
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


@pytest.fixture(scope="session")
def checker():
    from oracle import checker as c
    c.build()
    return c


@pytest.fixture(scope="session")
def synth():
    m = importlib.import_module(PKG + ".synth")
    m._load()
    return m


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module(PKG)


def cloud_checksum(c):
    u = np.ascontiguousarray(c, np.float32).view(np.uint32).astype(np.uint64).ravel()
    w = (np.arange(u.size, dtype=np.uint64) * np.uint64(2654435761) + np.uint64(1)) & np.uint64(0xFFFFFFFF)
    return np.uint64((u * w).sum(dtype=np.uint64)) ^ np.uint64(u.size)


def biteq(a, b):
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def feed_imu(S, k, *sinks):
    """the 20 IMU samples (200 Hz) that precede sweep k+1"""
    if S.has_imu:
        for j in range(k * 20, (k + 1) * 20):
            t, q, a = S.imu(j)
            for s in sinks:
                s.imu(t, q, a)


def inside_reference_domain(prev, r):
    """laserOdometry.cpp:486 / :592 bound the forward scans over the LAST corner / surf cloud by THIS sweep's
    sharp / flat count; when that exceeds the last cloud's size the reference reads past the end of the
    cloud (undefined behaviour, nothing to be equal to).  True when sweep `r` (after `prev`) stays clear of
    that: every guard-branch scene is checked for it, so parity is only claimed where the reference defines it."""
    if prev is None or prev.sr_status != 0 or r.lo_iters == 0:
        return True
    return r.n_sharp <= prev.n_less_sharp and r.n_flat <= prev.n_less_flat
