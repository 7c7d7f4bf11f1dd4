"""GPU (-m gpu): ROS-facing adapter.  The four nodes of adapters/ros/ are
compiled against the in-process ROS stand-ins of oracle/shims/ and driven by the same topic-graph
driver (oracle/ref_capi.cpp) that drives the reference's own nodes in oracle/_ref: the same
messages on the same topics must come out, bit for bit."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, biteq, feed_imu

pytestmark = pytest.mark.gpu

ADAPTER = os.path.join(ROOT, "tests", "adapter", "_build", "libloam_ref.so")
KEYS = ("transform", "transform_sum", "transform_aft_mapped", "transform_bef_mapped", "transform_mapped", "imu_trans")
INTS = ("sr_status", "lo_status", "lm_status", "n_full", "n_sharp", "n_less_sharp", "n_flat", "n_less_flat",
        "n_map_corner", "n_map_surf", "n_stack_corner", "n_stack_surf")


@pytest.fixture(scope="module")
def adapter_lib(pkg):
    assert os.path.exists(pkg.LIB_PATH)
    r = subprocess.run(["bash", os.path.join(ROOT, "tests", "adapter", "build_adapter.sh")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return ADAPTER


@pytest.mark.parametrize("preset,nsw", [(1, 26), (2, 8), (4, 6)])
def test_adapter_nodes_publish_what_the_reference_nodes_publish(adapter_lib, checker, synth, preset, nsw):
    os.environ["LOAM_B200_PRESET"] = str(preset)
    S = synth.Synth(preset)
    P = checker.make_params(preset)
    A = checker.Checker(P, adapter_lib)
    R = checker.Checker(P, "reference" if checker.available("reference") else "port")
    mapped = surround = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        feed_imu(S, k, A, R)
        a, r = A.push(xyz, st), R.push(xyz, st)
        for key in INTS:
            assert getattr(a, key) == getattr(r, key), (k, key, getattr(a, key), getattr(r, key))
        for key in KEYS:
            assert biteq(getattr(a, key)[:], getattr(r, key)[:]), (k, key)
        if a.sr_status == 0:
            for which in (0, 1, 2, 3, 4):           # the five clouds scanRegistration publishes
                assert np.array_equal(A.cloud(which).view(np.uint32), R.cloud(which).view(np.uint32)), (k, which)
        if a.lo_status == 0 or k == 1:
            for which in (5, 6):                    # /laser_cloud_corner_last, /laser_cloud_surf_last
                assert np.array_equal(A.cloud(which).view(np.uint32), R.cloud(which).view(np.uint32)), (k, which)
        if a.lm_status == 0:
            mapped += 1
            for which in (7, 8, 9):                 # /velodyne_cloud_3, /velodyne_cloud_registered, /laser_cloud_surround
                ca, cr = A.cloud(which), R.cloud(which)
                assert np.array_equal(ca.view(np.uint32), cr.view(np.uint32)), (k, which)
                surround += which == 9 and ca.shape[0] > 0
    assert mapped >= 2 and surround >= 1
    A.close()
    R.close()
