"""dev tool: per-phase SM-cycle breakdown of lo_optimize / lm_optimize (CTA 0) on a steady-state HDL-64 sweep"""
import ctypes, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")
L = pkg.lib()
L.loam_debug_phase_clocks.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
preset = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nsw = int(sys.argv[2]) if len(sys.argv) > 2 else 24
S = synth.Synth(preset)
G = pkg.Pipeline(pkg.make_params(preset, system_delay=0), device=0)
L.loam_debug_phase_enable(1)
names = {0: ["start", "feat", "rows", "gsync", "psum", "solve"], 1: ["start", "trig", "knn", "fit", "csum", "ctot", "gsync", "psum", "solve"]}
for k in range(nsw):
    xyz, st = S.sweep(k)
    r = G.push(xyz, st)
    if k < nsw - 4:
        continue
    for which, tag in ((0, "LO"), (1, "LM")):
        if which == 1 and r.lm_status != 0:
            continue
        buf = np.zeros(512, np.int64)
        L.loam_debug_phase_clocks(which, buf.ctypes.data, 512)
        n = int(buf[0])
        t = buf[1:1 + n]
        d = np.diff(t) / 1.965e3  # us at 1965 MHz
        print(f"sweep {k} {tag}: iters={r.lo_iters if which == 0 else r.lm_iters} stamps={n} total={(t[-1]-t[0])/1.965e3:.1f}us")
        per = len(names[which]) - 1
        for it in range(0, len(d), per):
            print("   it%-2d " % (it // per) + " ".join(f"{nm}={v:6.1f}" for nm, v in zip(names[which][1:], d[it:it + per])))
