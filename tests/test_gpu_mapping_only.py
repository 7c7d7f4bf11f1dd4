"""GPU (-m gpu): laserMapping alone on recorded laserOdometry output (BASELINE.json configs[2],
SURVEY.md §8d config 3), map checkpoint / resume through LOAMMAP1 files exchanged with the oracle
in both directions, PCD export."""
import numpy as np
import pytest

from conftest import biteq
from lm_replay import read_map_file, record_lm_inputs_gpu, record_lm_inputs_oracle, shifted

pytestmark = pytest.mark.gpu


def _gpu_pass(M, r, dx=0.0, dstamp=0.0):
    c, s, f, tsum, st = shifted(r, dx, dstamp)
    assert M.process_host(c, s, f, tsum, st) == 0
    return M.state()


@pytest.mark.parametrize("preset,nsw", [(1, 40), (2, 14)])
def test_lm_process_host_equals_oracle(pkg, checker, synth, preset, nsw):
    rec = record_lm_inputs_oracle(checker, synth, preset, nsw)
    assert len(rec) >= 6
    C = checker.Checker(checker.make_params(preset), "port")
    M = pkg.LaserMapping(pkg.make_params(preset, system_delay=0))
    for i, r in enumerate(rec):
        o = C.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
        aft, bef, tobe, st = _gpu_pass(M, r)
        assert biteq(aft, r["aft"]) and biteq(bef, r["bef"]), i       # equals the free-running oracle pipeline
        assert biteq(aft, o.transform_aft_mapped[:]), i              # and the oracle's laserMapping alone
        assert (st[0], st[3], st[4], st[5], st[6]) == (o.lm_iters, o.n_map_corner, o.n_map_surf, o.n_stack_corner, o.n_stack_surf), i
        for which in (10, 11, 12, 13, 8):
            assert np.array_equal(M.cloud(which, cap=1 << 20).view(np.uint32), C.cloud(which).view(np.uint32)), (i, which)
        assert np.array_equal(M.ints(100), C.ints(100)) and np.array_equal(M.ints(101), C.ints(101)), i


def test_lm_process_host_rejects_bad_arguments(pkg):
    P = pkg.make_params(1, system_delay=0)
    M = pkg.LaserMapping(P)
    big = np.zeros((P.max_points + 1, 4), np.float32)
    one = np.zeros((1, 4), np.float32)
    with pytest.raises(pkg.LoamError):
        M.process_host(big, one, one, np.zeros(6, np.float32), 0.0)
    # empty clouds are legal: the reference runs the pass with nothing to match (:706 guards the LM loop)
    e = np.zeros((0, 4), np.float32)
    assert M.process_host(e, e, e, np.zeros(6, np.float32), 0.0) == 0
    assert M.state()[3][0] == 0
    with pytest.raises(pkg.LoamError):
        M.load_map("/nonexistent/dir/x.loammap")


def test_map_checkpoint_resume_gpu_and_oracle_both_directions(pkg, checker, synth, tmp_path):
    rec = record_lm_inputs_oracle(checker, synth, 1, 48)
    half = len(rec) // 2
    P = pkg.make_params(1, system_delay=0)
    A = pkg.LaserMapping(P)
    C = checker.Checker(checker.make_params(1), "port")
    for r in rec[:half]:
        _gpu_pass(A, r)
        C.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
    g_path, o_path = str(tmp_path / "gpu.loammap"), str(tmp_path / "oracle.loammap")
    A.save_map(g_path)
    C.lm_save_map(o_path)
    fg, fo = read_map_file(g_path), read_map_file(o_path)
    for T in range(2):   # same cubes, same points, same order; only the dirty hints may differ
        assert np.array_equal(fg["cnt"][T], fo["cnt"][T]) and np.array_equal(fg["pts"][T].view(np.uint32), fo["pts"][T].view(np.uint32))
    for name in ("cen", "frameCount", "mapFrameCount", "tf", "total"):
        assert np.array_equal(fg["header"][name], fo["header"][name]), name
    B = pkg.LaserMapping(P)          # GPU resumes from its own file
    B.load_map(g_path)
    D = pkg.LaserMapping(P)          # GPU resumes from the oracle's file (every cube flagged dirty)
    D.load_map(o_path)
    C2 = checker.Checker(checker.make_params(1), "port")   # oracle resumes from the GPU's file
    C2.lm_load_map(g_path)
    for i, r in enumerate(rec[half:]):
        a = _gpu_pass(A, r)
        b = _gpu_pass(B, r)
        d = _gpu_pass(D, r)
        o = C2.lm_push(r["corner"], r["surf"], r["full"], r["sum"], r["stamp"])
        for x in (b, d):
            assert biteq(a[0], x[0]) and biteq(a[1], x[1]) and np.array_equal(a[3], x[3]), i
        assert biteq(a[0], o.transform_aft_mapped[:]) and biteq(a[0], r["aft"]), i
        assert np.array_equal(A.cloud(10).view(np.uint32), D.cloud(10).view(np.uint32))
        assert np.array_equal(A.cloud(11).view(np.uint32), C2.cloud(11).view(np.uint32))
    pa, pb = str(tmp_path / "a.loammap"), str(tmp_path / "d.loammap")
    A.save_map(pa)
    D.save_map(pb)
    fa, fd = read_map_file(pa), read_map_file(pb)
    for T in range(2):
        assert np.array_equal(fa["cnt"][T], fd["cnt"][T]) and np.array_equal(fa["pts"][T].view(np.uint32), fd["pts"][T].view(np.uint32))


def test_pcd_export(pkg, checker, synth, tmp_path):
    rec = record_lm_inputs_oracle(checker, synth, 1, 24)
    M = pkg.LaserMapping(pkg.make_params(1, system_delay=0))
    for r in rec:
        _gpu_pass(M, r)
    for which in (10, 9, 8, 20, 21):
        path = str(tmp_path / f"c{which}.pcd")
        M.export_pcd(which, path)
        raw = open(path, "rb").read()
        head, _, body = raw.partition(b"DATA binary\n")
        fields = dict(l.split(" ", 1) for l in head.decode().splitlines() if not l.startswith("#"))
        assert fields["FIELDS"] == "x y z intensity" and fields["SIZE"] == "4 4 4 4" and fields["TYPE"] == "F F F F"
        n = int(fields["POINTS"])
        assert int(fields["WIDTH"]) == n and len(body) == 16 * n and n > 0
        pts = np.frombuffer(body, "<f4").reshape(n, 4)
        if which < 20:
            assert np.array_equal(pts.view(np.uint32), M.cloud(which, cap=1 << 20).view(np.uint32))
        else:
            mp = str(tmp_path / "m.loammap")
            M.save_map(mp)
            assert np.array_equal(pts.view(np.uint32), read_map_file(mp)["pts"][which - 20].view(np.uint32))


def test_mapping_only_stress_prepopulated_map_parity(pkg, checker, synth, tmp_path):
    """config 3 at a size the oracle finishes in seconds: HDL-64 laserOdometry output replayed on
    five parallel streets, the map handed to the oracle as a file, then the same passes on both"""
    rec = record_lm_inputs_gpu(pkg, synth, 2, 42)
    assert len(rec) >= 18
    P = pkg.make_params(2, system_delay=0)
    M = pkg.LaserMapping(P)
    sizes = []
    t = 0.0
    for dx in (0.0, 18.0, -18.0, 36.0, -36.0):
        for r in rec:
            st = _gpu_pass(M, r, dx, t)[3]
        t += 10.0
        sizes.append(int(st[3] + st[4]))
    assert sizes[-1] > 3 * sizes[0]
    path = str(tmp_path / "stress.loammap")
    M.save_map(path)
    C = checker.Checker(checker.make_params(2), "port")
    C.lm_load_map(path)
    for i, r in enumerate(rec[8:12]):
        c, s, f, tsum, stamp = shifted(r, 0.0, t)
        o = C.lm_push(c, s, f, tsum, stamp)
        aft, bef, tobe, st = _gpu_pass(M, r, 0.0, t)
        assert o.lm_iters > 0
        assert biteq(aft, o.transform_aft_mapped[:]) and biteq(bef, o.transform_bef_mapped[:]), i
        assert (st[0], st[3], st[4]) == (o.lm_iters, o.n_map_corner, o.n_map_surf), i
        assert np.array_equal(M.ints(100), C.ints(100)) and np.array_equal(M.ints(101), C.ints(101)), i
        assert np.array_equal(M.cloud(10, cap=1 << 21).view(np.uint32), C.cloud(10, cap=1 << 21).view(np.uint32)), i


def test_mapping_only_at_the_benchmarked_map_size(pkg, checker, synth, tmp_path):
    """BASELINE.json configs[2] at the size bench.py --workload mapstress times: 260 HDL-64 sweeps of recorded
    laserOdometry output replayed on eight parallel streets (~1.4 M local-map points), the map handed to the
    oracle as a file, then six further passes on both sides: poses, iteration / row counts, the 5-NN index
    lists of the last iteration, the gathered local map and the stack clouds bit-equal; at the end the two
    map files agree cube by cube.  (The 8-CTA-cluster refilter, the arena compaction and the bounded 5-NN
    re-search are the parts most likely to diverge late -- this is where they would.)"""
    rec = record_lm_inputs_gpu(pkg, synth, 2, 260)
    P = pkg.make_params(2, system_delay=0)
    M = pkg.LaserMapping(P)
    t = 0.0
    offsets = [((j + 1) // 2) * 18.0 * (1.0 if j % 2 else -1.0) for j in range(8)]     # bench._lateral_offsets(8)
    for dx in offsets:
        for r in rec:
            c, s, f, tsum, st = shifted(r, dx, t)
            assert M.process_host(c, s, f, tsum, st) == 0
        t += 100.0
    path = str(tmp_path / "prepopulated.loammap")
    M.save_map(path)
    C = checker.Checker(checker.make_params(2), "port")
    C.lm_load_map(path)
    n_map = 0
    for i, r in enumerate(rec[30:36]):
        c, s, f, tsum, stamp = shifted(r, 0.0, t + 0.2 * i)
        o = C.lm_push(c, s, f, tsum, stamp)
        assert M.process_host(c, s, f, tsum, stamp) == 0
        aft, bef, tobe, st = M.state()
        assert o.lm_iters > 0
        assert biteq(aft, o.transform_aft_mapped[:]) and biteq(bef, o.transform_bef_mapped[:]), i
        assert (st[0], st[1], st[3], st[4], st[5], st[6]) == (o.lm_iters, o.lm_rows, o.n_map_corner, o.n_map_surf, o.n_stack_corner, o.n_stack_surf), i
        n_map = o.n_map_corner + o.n_map_surf
        assert np.array_equal(M.ints(100), C.ints(100)) and np.array_equal(M.ints(101), C.ints(101)), i
        for w in (10, 11, 12, 13):
            assert np.array_equal(M.cloud(w, cap=1 << 21).view(np.uint32), C.cloud(w, cap=1 << 21).view(np.uint32)), (i, w)
    assert n_map > 1_000_000, n_map
    g_path, o_path = str(tmp_path / "g.loammap"), str(tmp_path / "o.loammap")
    M.save_map(g_path)
    C.lm_save_map(o_path)
    fg, fo = read_map_file(g_path), read_map_file(o_path)
    for T in range(2):
        assert np.array_equal(fg["cnt"][T], fo["cnt"][T]), T
        assert np.array_equal(fg["pts"][T].view(np.uint32), fo["pts"][T].view(np.uint32)), T


def test_map_window_shifts_keep_parity(pkg, checker, synth, tmp_path):
    """the 21x11x21 cube window follows the pose (laserMapping.cpp:454-614): odometry jumps of several
    hundred metres along every axis, in both directions, shift the descriptor table by up to eight cubes
    at once and clear the wrapped slabs; every pass must stay bit-equal to the oracle, and so must the map"""
    rec = record_lm_inputs_oracle(checker, synth, 1, 26)
    P = pkg.make_params(1, system_delay=0)
    M = pkg.LaserMapping(P)
    C = checker.Checker(checker.make_params(1), "port")
    plan = [(0, 0.0), (0, 420.0), (0, -30.0), (2, 390.0), (1, 160.0), (1, -10.0), (2, -700.0), (0, -460.0), (0, 0.0)]
    t = 0.0
    cens = set()
    for axis, off in plan:
        for i, r in enumerate(rec[:8]):
            c, s, f, tsum, st = shifted(r, off, t, axis)
            o = C.lm_push(c, s, f, tsum, st)
            assert M.process_host(c, s, f, tsum, st) == 0
            aft, bef, tobe, stt = M.state()
            assert biteq(aft, o.transform_aft_mapped[:]) and biteq(bef, o.transform_bef_mapped[:]), (axis, off, i)
            assert (stt[0], stt[3], stt[4]) == (o.lm_iters, o.n_map_corner, o.n_map_surf), (axis, off, i)
            assert np.array_equal(M.cloud(10, cap=1 << 20).view(np.uint32), C.cloud(10).view(np.uint32)), (axis, off, i)
        t += 10.0
        g_path, o_path = str(tmp_path / "g.loammap"), str(tmp_path / "o.loammap")
        M.save_map(g_path)
        C.lm_save_map(o_path)
        fg, fo = read_map_file(g_path), read_map_file(o_path)
        assert np.array_equal(fg["header"]["cen"], fo["header"]["cen"])
        cens.add(tuple(int(v) for v in fg["header"]["cen"]))
        for T in range(2):
            assert np.array_equal(fg["cnt"][T], fo["cnt"][T]) and np.array_equal(fg["pts"][T].view(np.uint32), fo["pts"][T].view(np.uint32)), (axis, off, T)
    assert len(cens) >= 5     # the window really moved


def test_load_map_refuses_corrupt_and_foreign_files(pkg, checker, synth, tmp_path):
    """a LOAMMAP1 file is input from outside: header fields that would index the cube array or the IMU ring out of
    range, negative counts, truncation and a wrong magic are refused with an error -- and the node that refused a file
    keeps working"""
    rec = record_lm_inputs_oracle(checker, synth, 1, 14)
    P = pkg.make_params(1, system_delay=0)
    M = pkg.LaserMapping(P)
    for r in rec[:4]:
        _gpu_pass(M, r)
    good = str(tmp_path / "good.loammap")
    M.save_map(good)
    raw = bytearray(open(good, "rb").read())
    hdr = np.dtype([("magic", "S8"), ("version", "<i4"), ("dims", "<i4", 3), ("cen", "<i4", 3), ("frameCount", "<i4"),
                    ("mapFrameCount", "<i4"), ("isDegenerate", "<i4"), ("tf", "<f4", (5, 6)), ("imuPointerFront", "<i4"),
                    ("imuPointerLast", "<i4"), ("hasImu", "<i4"), ("imuQueLength", "<i4"), ("total", "<i8", 2)])

    def variant(name, edit):
        b = bytearray(raw)
        H = np.frombuffer(b, hdr, 1)     # writable view of the header
        edit(H[0], b)
        p = str(tmp_path / (name + ".loammap"))
        open(p, "wb").write(bytes(b))
        return p

    def set_field(field, value, idx=None):
        def edit(H, b):
            if idx is None:
                H[field] = value
            else:
                H[field][idx] = value
        return edit

    bad = [variant("cen_hi", set_field("cen", 21, 0)), variant("cen_neg", set_field("cen", -1, 2)),
           variant("imu_front", set_field("imuPointerFront", 100000)), variant("imu_last", set_field("imuPointerLast", -2)),
           variant("frames", set_field("frameCount", -5)), variant("magic", set_field("magic", b"LOAMMAP2")),
           variant("dims", set_field("dims", 22, 0)), variant("total", set_field("total", -1, 0))]
    cnt_off = hdr.itemsize + 200 * 16      # IMU ring: 200 x (double + 2 floats)
    assert int(np.frombuffer(raw, hdr, 1)[0]["imuQueLength"]) == 200
    neg = bytearray(raw)
    np.frombuffer(neg, "<i4", 1, cnt_off)[0] = -3
    p = str(tmp_path / "neg_count.loammap"); open(p, "wb").write(bytes(neg)); bad.append(p)
    p = str(tmp_path / "short.loammap"); open(p, "wb").write(bytes(raw[:len(raw) - 40])); bad.append(p)
    for path in bad:
        with pytest.raises(pkg.LoamError):
            M.load_map(path)
    # the node is still usable and still agrees with a node that never saw the bad files
    N = pkg.LaserMapping(P)
    N.load_map(good)
    M.load_map(good)
    for r in rec[4:8]:
        a, b = _gpu_pass(M, r), _gpu_pass(N, r)
        assert biteq(a[0], b[0]) and np.array_equal(a[3], b[3])
