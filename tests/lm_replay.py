"""Shared by the mapping-only tests (BASELINE.json configs[2]): the oracle-side
recording of laserOdometry's output; the GPU-side recorder, the replica replay and the map-file parser
live in the package (loam_velodyne-1_b200/replay.py) because bench.py uses them too."""
import importlib

import numpy as np

_replay = importlib.import_module("loam_velodyne-1_b200.replay")
record_lm_inputs_gpu = _replay.record_lm_inputs_gpu
shifted = _replay.shifted
replica_schedule = _replay.replica_schedule
read_map_file = _replay.read_map_file


def record_lm_inputs_oracle(checker, synth, preset, nsw):
    """the four topics laserMapping receives, recorded from the oracle's own sr -> lo, plus what
    the oracle's laserMapping made of them (free-running)"""
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    rec = []
    for k in range(nsw):
        xyz, st = S.sweep(k)
        r = C.push(xyz, st)
        if r.lm_status == 0:
            rec.append(dict(corner=C.cloud(5), surf=C.cloud(6), full=C.cloud(7), sum=np.array(r.transform_sum[:], np.float32), stamp=st,
                            aft=np.array(r.transform_aft_mapped[:], np.float32), bef=np.array(r.transform_bef_mapped[:], np.float32),
                            iters=r.lm_iters, n_map=(r.n_map_corner, r.n_map_surf)))
    C.close()
    return rec
