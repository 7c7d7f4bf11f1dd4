"""Host-side helpers for the mapping-only configuration (BASELINE.json configs[2]): recording of
laserOdometry's output from the sm_100a nodes, the lateral-replica replay that pre-populates the cube
map, and a parser of the LOAMMAP1 map file (csrc/lm_persist.cuh).  Test / bench orchestration only."""
import numpy as np


def record_lm_inputs_gpu(pkg, synth, preset, nsw, seed=None):
    """same recording from the sm_100a scanRegistration -> laserOdometry nodes (bit-identical to the
    oracle's by the node parity tests, and fast enough for the full-size configurations)"""
    S = synth.Synth(preset) if seed is None else synth.Synth(preset, seed=seed)
    P = pkg.make_params(preset, system_delay=0)
    sr, lo = pkg.ScanRegistration(P), pkg.LaserOdometry(P)
    rec = []
    for k in range(nsw):
        xyz, st = S.sweep(k)
        if sr.process(xyz, st) != 0:
            continue
        rc, pub = lo.process(sr)
        if rc == 0 and pub:
            _, tsum, _ = lo.state()
            rec.append(dict(corner=lo.cloud(5), surf=lo.cloud(6), full=lo.cloud(7), sum=tsum.copy(), stamp=st))
    return rec


def shifted(rec_item, dx, dstamp=0.0, axis=0):
    """the same laserOdometry output on a street `dx` metres to the left (LOAM x = left; axis 1 = up, 2 = forward)"""
    s = rec_item["sum"].copy()
    s[3 + axis] = np.float32(s[3 + axis] + np.float32(dx))
    return rec_item["corner"], rec_item["surf"], rec_item["full"], s, rec_item["stamp"] + dstamp


def replica_schedule(n_inputs, offsets):
    """pre-population order: every recorded input once per lateral offset"""
    out = []
    for j, dx in enumerate(offsets):
        for i in range(n_inputs):
            out.append((i, dx, j))
    return out


def read_map_file(path):
    """parser of the LOAMMAP1 layout (csrc/lm_persist.cuh) for the tests"""
    raw = open(path, "rb").read()
    hdr = np.dtype([("magic", "S8"), ("version", "<i4"), ("dims", "<i4", 3), ("cen", "<i4", 3), ("frameCount", "<i4"),
                    ("mapFrameCount", "<i4"), ("isDegenerate", "<i4"), ("tf", "<f4", (5, 6)), ("imuPointerFront", "<i4"),
                    ("imuPointerLast", "<i4"), ("hasImu", "<i4"), ("imuQueLength", "<i4"), ("total", "<i8", 2)])
    H = np.frombuffer(raw, hdr, 1)[0]
    off = hdr.itemsize
    q = int(H["imuQueLength"])
    off += q * 8 + q * 4 + q * 4
    ncube = int(np.prod(H["dims"]))
    out = {"header": H, "cnt": [], "dirty": [], "pts": []}
    for T in range(2):
        cnt = np.frombuffer(raw, "<i4", ncube, off); off += 4 * ncube
        dirty = np.frombuffer(raw, np.uint8, ncube, off); off += ncube
        n = int(cnt.sum())
        pts = np.frombuffer(raw, "<f4", 4 * n, off).reshape(n, 4); off += 16 * n
        out["cnt"].append(cnt); out["dirty"].append(dirty); out["pts"].append(pts)
    assert off == len(raw), (off, len(raw))
    return out
