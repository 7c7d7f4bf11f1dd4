"""dev tool: a short run that touches every kernel of the three nodes (VLP-16, 14 sweeps free-running, then two
window jumps), meant to be run under `compute-sanitizer --tool memcheck` where that tool is available (it is closed on this pool)"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")
S = synth.Synth(1)
P = pkg.make_params(1, system_delay=0)
G = pkg.Pipeline(P, device=0)
res = [G.push(*S.sweep(k)) for k in range(14)]
print("pipeline", [(r.lo_iters, r.lm_iters) for r in res[-4:]])
M = pkg.LaserMapping(P)
r = None
for i in range(6):
    c, s, f = G.lo.cloud(5), G.lo.cloud(6), G.lo.cloud(7)
    t = np.array(res[-1].transform_sum[:], np.float32)
    t[3] += (0.0, 130.0, -60.0)[i % 3]
    M.process_host(c, s, f, t, 100.0 + i)
print("mapping", M.state()[3][:5], M.ints(110))
