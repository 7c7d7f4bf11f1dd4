"""dev: how large the local map gets when HDL-64 laserOdometry output is replayed on parallel streets"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")
_r = importlib.import_module("loam_velodyne-1_b200.replay"); record_lm_inputs_gpu, shifted = _r.record_lm_inputs_gpu, _r.shifted
nsw = int(sys.argv[1]) if len(sys.argv) > 1 else 260
spacing = float(sys.argv[2]) if len(sys.argv) > 2 else 18.0
t0 = time.time()
rec = record_lm_inputs_gpu(pkg, synth, 2, nsw)
print("recorded", len(rec), "inputs in", round(time.time() - t0, 1), "s", flush=True)
P = pkg.make_params(2, system_delay=0)
M = pkg.LaserMapping(P)
t = 0.0
for j in range(16):
    dx = ((j + 1) // 2) * spacing * (1 if j % 2 else -1)
    t1 = time.time()
    for r in rec:
        c, s, f, ts, st = shifted(r, dx, t)
        M.process_host(c, s, f, ts, st)
    t += 100.0
    stt = M.state()[3]
    print(j, dx, "local map", stt[3], stt[4], "stack", stt[5], stt[6], "iters", stt[0], "ms/pass", round(1e3 * (time.time() - t1) / len(rec), 2), flush=True)
    if stt[3] + stt[4] > 1_200_000:
        break
