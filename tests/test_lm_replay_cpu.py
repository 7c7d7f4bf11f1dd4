"""CPU: the oracle's laserMapping alone on recorded laserOdometry output (BASELINE.json configs[2])
and the oracle's reader / writer of the LOAMMAP1 map file."""
import os
import struct

import numpy as np

from conftest import biteq
from lm_replay import record_lm_inputs_oracle, shifted


def test_oracle_lm_alone_equals_free_running_pipeline(checker, synth):
    rec = record_lm_inputs_oracle(checker, synth, 1, 40)
    assert len(rec) >= 18
    C = checker.Checker(checker.make_params(1), "port")
    for r in rec:
        o = C.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
        assert o.lm_status == 0 and o.lm_iters == r["iters"]
        assert biteq(o.transform_aft_mapped[:], r["aft"]) and biteq(o.transform_bef_mapped[:], r["bef"])
        assert (o.n_map_corner, o.n_map_surf) == r["n_map"]


def test_oracle_map_file_roundtrip_resumes_bit_identically(checker, synth, tmp_path):
    rec = record_lm_inputs_oracle(checker, synth, 1, 40)
    half = len(rec) // 2
    path = str(tmp_path / "half.loammap")
    A = checker.Checker(checker.make_params(1), "port")
    for r in rec[:half]:
        A.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
    A.lm_save_map(path)
    # header as documented in csrc/lm_persist.cuh
    raw = open(path, "rb").read()
    magic, version, w, h, d = struct.unpack_from("<8s4i", raw, 0)
    assert magic == b"LOAMMAP1" and (version, w, h, d) == (1, 21, 11, 21)
    B = checker.Checker(checker.make_params(1), "port")
    B.lm_load_map(path)
    for r in rec[half:]:
        a = A.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
        b = B.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
        assert biteq(a.transform_aft_mapped[:], b.transform_aft_mapped[:]) and biteq(a.transform_aft_mapped[:], r["aft"])
        assert (a.n_map_corner, a.n_map_surf, a.lm_iters) == (b.n_map_corner, b.n_map_surf, b.lm_iters)
        assert np.array_equal(A.cloud(10).view(np.uint32), B.cloud(10).view(np.uint32))
    # a second save of the resumed run is byte-identical to a save of the uninterrupted one
    pa, pb = str(tmp_path / "a.loammap"), str(tmp_path / "b.loammap")
    A.lm_save_map(pa)
    B.lm_save_map(pb)
    assert open(pa, "rb").read() == open(pb, "rb").read()


def test_oracle_lateral_replicas_grow_the_local_map(checker, synth):
    """the pre-population scheme of the mapping-only configuration: the same street replayed a few
    metres to the side adds its points to the cube map instead of merging into the first one"""
    rec = record_lm_inputs_oracle(checker, synth, 1, 16)
    C = checker.Checker(checker.make_params(1), "port")
    sizes = []
    for dx in (0.0, 9.0, -9.0):
        for r in rec:
            o = C.lm_push(*shifted(r, dx))
        sizes.append(o.n_map_corner + o.n_map_surf)
    assert sizes[1] > 1.5 * sizes[0] and sizes[2] > 2.2 * sizes[0]
    assert np.isfinite(np.array(o.transform_aft_mapped[:])).all()


def test_oracle_three_thread_pipeline_builds_the_same_map(checker, synth):
    """the measured 'reference deployment' baseline of bench.py (one thread per node, depth-2 queues):
    odometry and mapping do not depend on timing, so the map it builds equals the sequential run's
    (only /integrated_to_init depends on when the mapping correction arrives, as in the reference)"""
    S = synth.Synth(1)
    sw = [S.sweep(k) for k in range(24)]
    A = checker.Checker(checker.make_params(1), "port")
    for x, st in sw:
        A.push(x, st)
    B = checker.Checker(checker.make_params(1), "port")
    wall, poses = B.run_pipelined(sw)
    assert wall > 0 and poses.shape == (24, 6) and np.isfinite(poses).all()
    for which in (10, 11, 5, 6):
        assert np.array_equal(A.cloud(which).view(np.uint32), B.cloud(which).view(np.uint32)), which
    assert np.abs(poses[-1] - np.array(A.push(*S.sweep(24)).transform_mapped[:])).max() < 0.5
