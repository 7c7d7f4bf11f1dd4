"""dev script: scanRegistration+laserOdometry CUDA vs oracle, free running (run on a GPU box)."""
import sys, os, importlib, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import checker
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")

def biteq(a, b):
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))

def run(preset, nsw):
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    P = pkg.make_params(preset, system_delay=0)
    G = pkg.ScanRegistration(P); LO = pkg.LaserOdometry(P)
    bad = 0; tg = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        if S.has_imu:
            for j in range(k * 20, (k + 1) * 20):
                t, q, a = S.imu(j); C.imu(t, q, a); G.imu(t, q, a)
        r = C.push(xyz, st)
        t0 = time.time()
        rc = G.process(xyz, st)
        if rc != 0: continue
        lrc, pub = LO.process(G)
        tg += time.time() - t0
        msgs = []
        if lrc != r.lo_status: msgs.append(("lo_status", lrc, r.lo_status))
        tr, ts, stt = LO.state()
        if not biteq(tr, r.transform[:]): msgs.append(("transform", tr, np.array(r.transform[:])))
        if not biteq(ts, r.transform_sum[:]): msgs.append(("transform_sum", ts, np.array(r.transform_sum[:])))
        if lrc == 0:
            if stt[0] != r.lo_iters or stt[1] != r.lo_rows: msgs.append(("iters/rows", stt[:3].tolist(), r.lo_iters, r.lo_rows, r.lo_degenerate))
            for w in (6, 7, 8, 9, 10):
                a = C.ints(w); b = LO.ints(w)
                if a.shape != b.shape or not np.array_equal(a, b):
                    msgs.append(("corr", w, a.shape, b.shape, int((a != b).sum()) if a.shape == b.shape else None))
        for w in (5, 6):
            a = C.cloud(w); b = LO.cloud(w)
            if not biteq(a, b): msgs.append(("cloud", w, a.shape, b.shape))
        if pub:
            a = C.cloud(7); b = LO.cloud(7)
            if not biteq(a, b): msgs.append(("cloud", 7, a.shape, b.shape))
        if pub != (r.lm_status == 0): msgs.append(("pub", pub, r.lm_status))
        if msgs:
            bad += 1
            if bad < 5: print("sweep", k, msgs[:5])
    print("preset", preset, "sweeps", nsw, "mismatching", bad, "gpu wall s", round(tg, 3), "launches", pkg.kernel_launch_count())
    return bad

if __name__ == "__main__":
    tot = 0
    cfg = ((1, 60), (2, 30), (4, 12)) if len(sys.argv) < 2 else ((int(sys.argv[1]), int(sys.argv[2])),)
    for preset, nsw in cfg:
        tot += run(preset, nsw)
    print("TOTAL MISMATCH", tot)
    sys.exit(1 if tot else 0)
