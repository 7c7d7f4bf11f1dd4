"""dev script: full pipeline (sr -> lo -> lm -> tm) CUDA vs oracle, free running (GPU box)."""
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

def run(preset, nsw, kind="port"):
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), kind)
    P = pkg.make_params(preset, system_delay=0)
    G = pkg.Pipeline(P)
    bad = 0; tg = 0; tc = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        if S.has_imu:
            for j in range(k * 20, (k + 1) * 20):
                t, q, a = S.imu(j); C.imu(t, q, a); G.imu(t, q, a)
        t0 = time.time(); r = C.push(xyz, st); tc += time.time() - t0
        t0 = time.time(); g = G.push(xyz, st); tg += time.time() - t0
        da, db = r.as_dict(), g.as_dict()
        msgs = []
        for key in da:
            va, vb = da[key], db[key]
            if isinstance(va, np.ndarray):
                if not biteq(va, vb): msgs.append((key, va, vb))
            elif va != vb: msgs.append((key, va, vb))
        if g.lm_status == 0 and r.lm_status == 0:
            for w in (10, 11, 12, 13, 8):
                a = C.cloud(w); b = G.lm.cloud(w)
                if not biteq(a, b):
                    msgs.append(("cloud", w, a.shape, b.shape, float(np.abs(a - b).max()) if a.shape == b.shape else None))
            if kind == "port":
                for w in (100, 101):
                    a = C.ints(w); b = G.lm.ints(w)
                    if a.shape != b.shape or not np.array_equal(a, b):
                        msgs.append(("nn", w, a.shape, b.shape, int((a != b).sum()) if a.shape == b.shape else None))
        if msgs:
            bad += 1
            if bad < 4: print("sweep", k, msgs[:6])
    print("preset", preset, "sweeps", nsw, "mismatching", bad, "gpu wall s", round(tg, 3), "cpu s", round(tc, 3), "launches", pkg.kernel_launch_count())
    return bad

if __name__ == "__main__":
    tot = 0
    cfg = ((1, 60), (2, 30), (4, 12)) if len(sys.argv) < 3 else ((int(sys.argv[1]), int(sys.argv[2])),)
    for preset, nsw in cfg:
        tot += run(preset, nsw)
    print("TOTAL MISMATCH", tot)
    sys.exit(1 if tot else 0)
