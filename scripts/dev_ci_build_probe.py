"""dev tool (LOAM_PHASE_CLOCKS build): phase stamps of ci_build for the largest cube slot of a pass on a multi-street map"""
import ctypes, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")
_r = importlib.import_module("loam_velodyne-1_b200.replay"); record_lm_inputs_gpu, shifted = _r.record_lm_inputs_gpu, _r.shifted
L = pkg.lib()
L.loam_debug_phase_clocks.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
nsw = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rec = record_lm_inputs_gpu(pkg, synth, 2, nsw)
M = pkg.LaserMapping(pkg.make_params(2, system_delay=0))
t = 0.0
for j in range(8):
    dx = ((j + 1) // 2) * 18.0 * (1 if j % 2 else -1)
    for r in rec:
        c, s, f, ts, st = shifted(r, dx, t)
        M.process_host(c, s, f, ts, st)
    t += 100.0
L.loam_debug_phase_enable(1)
for r in rec[30:40]:
    c, s, f, ts, st = shifted(r, 0.0, t)
    M.process_host(c, s, f, ts, st)
    dc = np.zeros(1024, np.int64)
    L.loam_debug_phase_clocks(3, dc.ctypes.data, 1024)
    d = np.diff(dc[960:965]) / 1.965e3
    print(f"largest slot {dc[967]} points, {dc[968]} slots: zero {d[0]:.1f}  count {d[1]:.1f}  alloc {d[2]:.1f}  scatter {d[3]:.1f} us   map {M.state()[3][3:5]}")
