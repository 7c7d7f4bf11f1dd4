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
    if r.lm_status == 0:
        cnt = np.zeros(16, np.int32)
        L.loam_debug_lm_counters.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.loam_debug_lm_counters(G.lm.h, cnt.ctypes.data, 16)
        print(f"sweep {k} refilter corner {cnt[:8].tolist()} surf {cnt[8:].tolist()}  map {r.n_map_corner}/{r.n_map_surf} stack {r.n_stack_corner}/{r.n_stack_surf}")
    if r.lm_status == 0:
        rf = np.zeros(2 * 250 * 8, np.int64)
        L.loam_debug_phase_clocks(2, rf.ctypes.data, rf.size)
        rf = rf.reshape(2, 250, 8)
        for T in range(2):
            for rr in range(250):
                t = rf[T, rr]
                if t[0] == 0 or t[4] == 0:
                    continue
                n_old, n_new = int(t[7] >> 32), int(t[7] & 0xffffffff)
                if n_old < 2000:
                    continue
                st_ = [(t[k] - t[0]) / 1.965e3 if t[k] else -1 for k in range(1, 7)]
                print(f"   rf T{T} r{rr} n_old={n_old} n_new={n_new} stage={st_[4]:.1f} passA={st_[0]:.1f} nolist={st_[5]:.1f} scan+alloc={st_[1]:.1f} passB={st_[2]:.1f} end={st_[3]:.1f} us")
    if r.lm_status == 0:
        dc = np.zeros(64, np.int64)
        L.loam_debug_phase_clocks(3, dc.ctypes.data, 64)
        us = lambda a, b: (dc[b] - dc[a]) / 1.965e3
        print(f"   sort_new: build={us(0,1):.1f} sort={us(1,2):.1f}  pass0: walk1={us(8,9):.1f} prefix={us(9,10):.1f} walk2={us(10,11):.1f}")
    dc = np.zeros(1024, np.int64)
    L.loam_debug_phase_clocks(3, dc.ctypes.data, 1024)
    ft = dc[64:64 + 148] / 1.965e3
    for b in (0, 20, 60, 80, 100, 120, 130, 140):
        t0_, t1_, t2_, t3_ = dc[256 + b], dc[832 + b], dc[448 + b], dc[640 + b]
        print(f"   LO last-search CTA{b} warp0: transform={(t1_-t0_)/1.965e3:.1f} nn1={(t2_-t1_)/1.965e3:.1f} window+residual={(t3_-t2_)/1.965e3:.1f} us")
    print(f"   sr_select ring0: stage={(dc[26]-dc[25])/1.965e3:.1f} six sectors={(dc[27]-dc[26])/1.965e3:.1f} accept/redo={(dc[28]-dc[27])/1.965e3:.1f} emit={(dc[29]-dc[28])/1.965e3:.1f} us")
    print(f"   sr_lessflat ring0: compact={(dc[21]-dc[20])/1.965e3:.1f} keys={(dc[22]-dc[21])/1.965e3:.1f} sort={(dc[23]-dc[22])/1.965e3:.1f} centroids={(dc[24]-dc[23])/1.965e3:.1f} us")
    print("   LO it0 per-CTA feature phase us: " + " ".join(f"{v:.0f}" for v in ft))
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
