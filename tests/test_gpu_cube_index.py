"""GPU (-m gpu): the persistent per-cube index (csrc/lm_cube_index.cuh) against the
scheme it replaces (index over the whole gathered local map, every pass: laserMapping.cpp:707-708) and against
the oracle: same 5-NN lists, same poses, same maps -- while only the cubes a sweep changed are re-indexed."""
import numpy as np
import pytest

from conftest import biteq
from lm_replay import record_lm_inputs_gpu, record_lm_inputs_oracle, shifted

pytestmark = pytest.mark.gpu

PER_PASS_INDEX = 32      # loam_params.dev_flags: rebuild the index over the gathered local map every pass (round-1 scheme)
SMALLEST_ARENA = 64      # loam_params.dev_flags: index arena just large enough for one local map (test hook: resets)


def _node(pkg, preset, flags=0):
    P = pkg.make_params(preset, system_delay=0)
    P.dev_flags = flags
    return pkg.LaserMapping(P)


def _same_pass(A, B, i):
    sa, sb = A.state(), B.state()
    for x, y in zip(sa[:3], sb[:3]):
        assert biteq(x, y), i
    assert np.array_equal(sa[3], sb[3]), i
    assert np.array_equal(A.ints(100), B.ints(100)) and np.array_equal(A.ints(101), B.ints(101)), i


def test_per_cube_index_reindexes_only_what_changed(pkg, synth):
    """HDL-64 laserOdometry output replayed on five parallel streets (a map of several hundred thousand points
    spread over many cubes): per pass, the incremental node and the per-pass-index node agree bit for bit on
    poses, counters and the 5-NN index lists; and once the streets are laid down, a pass re-indexes a fraction
    of the local map"""
    rec = record_lm_inputs_gpu(pkg, synth, 2, 60)
    A, B = _node(pkg, 2), _node(pkg, 2, PER_PASS_INDEX)
    assert A.ints(110).size == 8 and B.ints(110).size == 0
    t = 0.0
    i = 0
    frac = []
    for dx in (0.0, 18.0, -18.0, 36.0, -36.0, 54.0, -54.0):
        for r in rec:
            c, s, f, tsum, st = shifted(r, dx, t)
            assert A.process_host(c, s, f, tsum, st) == 0 and B.process_host(c, s, f, tsum, st) == 0
            if i % 7 == 0 or dx == -54.0:
                _same_pass(A, B, i)
            st_a = A.state()[3]
            ci = A.ints(110)
            n_map = int(st_a[3] + st_a[4])
            if n_map > 0 and dx in (54.0, -54.0):
                frac.append(ci[4] / n_map)
            i += 1
        t += 100.0
    for w in (10, 11):
        assert np.array_equal(A.cloud(w, cap=1 << 21).view(np.uint32), B.cloud(w, cap=1 << 21).view(np.uint32)), w
    ci = A.ints(110)
    n_map = int(A.state()[3][3] + A.state()[3][4])
    assert n_map > 250_000, n_map
    # a sweep touches the cubes around the sensor; the other streets' cubes keep their index
    assert np.median(frac) < 0.7, np.median(frac)
    assert ci[1] < 0.8 * i * n_map        # far fewer points indexed than "the whole local map, every pass"
    assert ci[5] <= ci[7] and ci[6] <= ci[7]


@pytest.mark.parametrize("preset,nsw", [(1, 70), (2, 16)])
def test_per_cube_index_with_the_smallest_arena_equals_oracle(pkg, checker, synth, preset, nsw):
    """index arena sized for exactly one local map: regions outgrown by their cubes exhaust it every few passes and
    the whole table starts over; every pass still equals the oracle (poses, counters, 5-NN lists, local map)"""
    rec = record_lm_inputs_oracle(checker, synth, preset, nsw)
    C = checker.Checker(checker.make_params(preset), "port")
    M = _node(pkg, preset, SMALLEST_ARENA)
    for i, r in enumerate(rec):
        o = C.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
        assert M.process_host(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"]) == 0
        aft, bef, tobe, st = M.state()
        assert biteq(aft, o.transform_aft_mapped[:]) and biteq(bef, o.transform_bef_mapped[:]), i
        assert (st[0], st[1], st[3], st[4]) == (o.lm_iters, o.lm_rows, o.n_map_corner, o.n_map_surf), i
        assert np.array_equal(M.ints(100), C.ints(100)) and np.array_equal(M.ints(101), C.ints(101)), i
        if i % 5 == 0:
            for w in (10, 11):
                assert np.array_equal(M.cloud(w, cap=1 << 20).view(np.uint32), C.cloud(w).view(np.uint32)), (i, w)
    ci = M.ints(110)
    assert ci[0] > 0 and ci[1] > 0


def test_per_cube_index_survives_window_shifts_and_map_reload(pkg, checker, synth, tmp_path):
    """odometry jumps move the 5x5x5 window by several cubes (slots change owner, cubes enter and leave), a map
    file is loaded over a used node (every slot forgets its cube): the incremental node keeps equalling the
    per-pass-index node, and the oracle at the end"""
    rec = record_lm_inputs_oracle(checker, synth, 1, 30)
    A, B = _node(pkg, 1), _node(pkg, 1, PER_PASS_INDEX)
    C = checker.Checker(checker.make_params(1), "port")
    plan = [(0, 0.0), (0, 60.0), (0, 10.0), (2, 130.0), (1, 55.0), (1, -10.0), (2, -240.0), (0, -170.0), (0, 0.0), (0, 45.0), (0, 0.0)]
    t = 0.0
    i = 0
    for axis, off in plan:
        for r in rec[:10]:
            tsum = r["sum"].copy()
            tsum[3 + axis] += off
            st = r["stamp"] + t
            o = C.lm_push(r["corner"], r["surf"], r["full"], tsum, st)
            assert A.process_host(r["corner"], r["surf"], r["full"], tsum, st) == 0
            assert B.process_host(r["corner"], r["surf"], r["full"], tsum, st) == 0
            _same_pass(A, B, i)
            assert biteq(A.state()[0], o.transform_aft_mapped[:]), i
            assert np.array_equal(A.ints(100), C.ints(100)) and np.array_equal(A.ints(101), C.ints(101)), i
            i += 1
        t += 10.0
    path = str(tmp_path / "a.loammap")
    A.save_map(path)
    built = int(A.ints(110)[0])
    A.load_map(path)
    B.load_map(path)
    assert int(A.ints(110)[0]) == 0      # counters and slots start over
    for r in rec[10:16]:
        st = r["stamp"] + t
        assert A.process_host(r["corner"], r["surf"], r["full"], r["sum"], st) == 0
        assert B.process_host(r["corner"], r["surf"], r["full"], r["sum"], st) == 0
        _same_pass(A, B, i)
        i += 1
    assert built > 0 and int(A.ints(110)[0]) > 0
