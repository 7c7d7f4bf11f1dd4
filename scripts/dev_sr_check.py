"""dev script: scanRegistration CUDA vs oracle on a few sweeps (run on a GPU box)."""
import sys, os, importlib, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import checker
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")

def run(preset, nsw):
    S = synth.Synth(preset)
    C = checker.Checker(checker.make_params(preset), "port")
    P = pkg.make_params(preset, system_delay=0)
    G = pkg.ScanRegistration(P)
    bad = 0
    for k in range(nsw):
        xyz, st = S.sweep(k)
        if S.has_imu:
            for j in range(k * 20, (k + 1) * 20):
                t, q, a = S.imu(j); C.imu(t, q, a); G.imu(t, q, a)
        r = C.push(xyz, st)
        rc = G.process(xyz, st)
        if rc != r.sr_status:
            print("status mismatch", k, rc, r.sr_status); bad += 1; continue
        if rc != 0: continue
        msgs = []
        for which in (0, 1, 2, 3, 4):
            a = C.cloud(which); b = G.cloud(which)
            if a.shape != b.shape or not np.array_equal(a.view(np.uint32), b.view(np.uint32)):
                d = None
                if a.shape == b.shape:
                    neq = np.nonzero((a.view(np.uint32) != b.view(np.uint32)).any(axis=1))[0]
                    d = (len(neq), neq[:5].tolist(), a[neq[:2]].tolist(), b[neq[:2]].tolist())
                msgs.append(("cloud", which, a.shape, b.shape, d))
        for which in (0, 1, 2):
            a = C.ints(which)[5:-6]; b = G.ints(which)[5:-6]
            if a.shape != b.shape or not np.array_equal(a, b):
                msgs.append(("ints", which, a.shape, b.shape, int((a != b).sum()) if a.shape == b.shape else None))
        for which in (3, 4, 5, 11, 12):
            a = C.ints(which); b = G.ints(which)
            if a.shape != b.shape or not np.array_equal(a, b): msgs.append(("ints", which, a.shape, b.shape))
        a = C.floats(0)[5:-5]; b = G.floats(0)[5:-5]
        if a.shape != b.shape or not np.array_equal(a.view(np.uint32), b.view(np.uint32)): msgs.append(("curv",))
        it = np.array(r.imu_trans[:], np.float32)
        if not np.array_equal(it.view(np.uint32), G.imu_trans().view(np.uint32)): msgs.append(("imu_trans", it, G.imu_trans()))
        if msgs:
            bad += 1
            print("sweep", k, msgs[:4])
    print("preset", preset, "sweeps", nsw, "mismatching", bad, "launches", pkg.kernel_launch_count())
    return bad

if __name__ == "__main__":
    tot = 0
    for preset, nsw in ((1, 6), (2, 4), (4, 4)):
        tot += run(preset, nsw)
    print("TOTAL MISMATCH", tot)
    sys.exit(1 if tot else 0)
