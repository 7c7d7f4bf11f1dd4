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


def _cube_of(p, cen=(10, 5, 10)):
    """laserMapping.cpp:983-989: cube indices of a map point (double arithmetic, truncation + fix-up for negatives)"""
    out = []
    for a in range(3):
        v = float(p[a]) + 25.0
        c = int(v / 50.0) + cen[a]
        if v < 0:
            c -= 1
        out.append(c)
    return out[0] + 21 * out[1] + 21 * 11 * out[2]


def _write_map_file(path, clouds):
    """a LOAMMAP1 file (csrc/lm_persist.cuh) holding the two given clouds, every cube flagged dirty (= "written by
    someone else": the first refilter of a cube sorts it)"""
    hdr = np.zeros(1, np.dtype([("magic", "S8"), ("version", "<i4"), ("dims", "<i4", 3), ("cen", "<i4", 3), ("frameCount", "<i4"),
                                ("mapFrameCount", "<i4"), ("isDegenerate", "<i4"), ("tf", "<f4", (5, 6)), ("imuPointerFront", "<i4"),
                                ("imuPointerLast", "<i4"), ("hasImu", "<i4"), ("imuQueLength", "<i4"), ("total", "<i8", 2)]))
    hdr["magic"] = b"LOAMMAP1"; hdr["version"] = 1; hdr["dims"] = (21, 11, 21); hdr["cen"] = (10, 5, 10)
    hdr["mapFrameCount"] = 4; hdr["imuPointerLast"] = -1; hdr["imuQueLength"] = 200
    body = []
    for T, pts in enumerate(clouds):
        cube = np.array([_cube_of(p) for p in pts], np.int64)
        order = np.argsort(cube, kind="stable")
        cnt = np.bincount(cube, minlength=21 * 11 * 21).astype("<i4")
        hdr["total"][0][T] = len(pts)
        body += [cnt.tobytes(), (cnt > 0).astype(np.uint8).tobytes(), np.ascontiguousarray(pts[order], "<f4").tobytes()]
    with open(path, "wb") as f:
        f.write(hdr.tobytes())
        f.write(np.zeros(200, "<f8").tobytes() + np.zeros(400, "<f4").tobytes())
        for b in body:
            f.write(b)


def test_per_cube_index_at_cube_boundaries(pkg, checker, tmp_path):
    """the one place where the per-cube index differs structurally from an index over the gathered map: neighbourhoods
    that straddle cube faces, edges and corners (a row of three lattice cells split over two cubes, neighbours in up to
    eight cubes, points stored a rounding error outside their cube), negative coordinates, cubes the FOV test leaves
    out.  A hand-made map of line- and plane-shaped clusters centred ON the cube boundaries, queried by stack points
    next to every cluster, over three passes between which the pose jumps by a cube: incremental node == per-pass-index
    node == oracle in poses, counters and every 5-NN index list"""
    rng = np.random.default_rng(77)
    edges = np.array([-75.0, -25.0, 25.0, 75.0])

    def centres(n):
        c = np.stack([rng.uniform(-70, 70, n), rng.uniform(-28, 28, n), rng.uniform(-70, 70, n)], axis=1)
        k = n // 6
        c[:k, 0] = rng.choice(edges, k) + rng.uniform(-0.3, 0.3, k)                       # on a face
        c[k:2 * k, 2] = rng.choice(edges, k) + rng.uniform(-0.3, 0.3, k)
        c[2 * k:3 * k, 1] = rng.choice([-25.0, 25.0], k) + rng.uniform(-0.3, 0.3, k)
        c[3 * k:4 * k, 0] = rng.choice(edges, k); c[3 * k:4 * k, 2] = rng.choice(edges, k)   # exactly on an edge
        c[4 * k:5 * k, 0] = rng.choice(edges, k) + rng.uniform(-0.05, 0.05, k)              # around a corner
        c[4 * k:5 * k, 1] = rng.choice([-25.0, 25.0], k) + rng.uniform(-0.05, 0.05, k)
        c[4 * k:5 * k, 2] = rng.choice(edges, k) + rng.uniform(-0.05, 0.05, k)
        return c

    def corner_map(c):      # 30 points along a random line through every centre
        d = rng.normal(size=(len(c), 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
        t = rng.uniform(-0.8, 0.8, (len(c), 30, 1))
        return (c[:, None, :] + t * d[:, None, :] + rng.normal(0, 0.01, (len(c), 30, 3))).reshape(-1, 3)

    def surf_map(c):        # 40 points on a random disc around every centre
        n = rng.normal(size=(len(c), 3)); n /= np.linalg.norm(n, axis=1, keepdims=True)
        u = np.cross(n, [0.3, 0.5, 0.8]); u /= np.linalg.norm(u, axis=1, keepdims=True)
        v = np.cross(n, u)
        a, b = rng.uniform(-0.8, 0.8, (2, len(c), 40, 1))
        return (c[:, None, :] + a * u[:, None, :] + b * v[:, None, :] + rng.normal(0, 0.005, (len(c), 40, 3))).reshape(-1, 3)

    cc, cs = centres(360), centres(360)
    maps = [np.concatenate([m, np.zeros((len(m), 1))], axis=1).astype(np.float32) for m in (corner_map(cc), surf_map(cs))]
    path = str(tmp_path / "boundaries.loammap")
    _write_map_file(path, maps)
    A, B = _node(pkg, 1), _node(pkg, 1, PER_PASS_INDEX)
    C = checker.Checker(checker.make_params(1), "port")
    A.load_map(path); B.load_map(path); C.lm_load_map(path)
    found = 0
    for i, jump in enumerate([(0.0, 0.0, 0.0), (52.0, 0.0, 0.0), (-49.5, 3.0, 51.0), (0.0, -2.0, -52.0)]):
        tsum = np.zeros(6, np.float32); tsum[3:] = jump
        def stack(c, k):
            q = np.repeat(c, k, axis=0) + rng.uniform(-0.35, 0.35, (len(c) * k, 3)) - np.asarray(jump)
            return np.concatenate([q, np.zeros((len(q), 1))], axis=1).astype(np.float32)
        corner, surf = stack(cc, 2), stack(cs, 2)
        full = surf[:64].copy()
        o = C.lm_push(corner, surf, full, tsum, 10.0 + i)
        assert A.process_host(corner, surf, full, tsum, 10.0 + i) == 0 and B.process_host(corner, surf, full, tsum, 10.0 + i) == 0
        _same_pass(A, B, i)
        aft, bef, tobe, st = A.state()
        assert biteq(aft, o.transform_aft_mapped[:]) and (st[0], st[1], st[3], st[4]) == (o.lm_iters, o.lm_rows, o.n_map_corner, o.n_map_surf), i
        nn_c, nn_s = A.ints(100), A.ints(101)
        assert np.array_equal(nn_c, C.ints(100)) and np.array_equal(nn_s, C.ints(101)), i
        found += int((nn_c.reshape(-1, 5)[:, 0] >= 0).sum() + (nn_s.reshape(-1, 5)[:, 0] >= 0).sum())
        assert o.lm_iters >= 1
    assert found > 1500, found         # the comparison is about real neighbour lists, not about "nothing found"
    for w in (10, 11):
        assert np.array_equal(A.cloud(w, cap=1 << 20).view(np.uint32), C.cloud(w).view(np.uint32)), w
