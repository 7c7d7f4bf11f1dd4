"""run n sweeps of a synthetic sequence through the free-running pipeline (profiling target for ncu)

    python scripts/run_sequence.py [preset] [n_sweeps]
"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("loam_velodyne-1_b200")
synth = importlib.import_module("loam_velodyne-1_b200.synth")
preset = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nsw = int(sys.argv[2]) if len(sys.argv) > 2 else 40
S = synth.Synth(preset)
G = pkg.Pipeline(pkg.make_params(preset, system_delay=0), device=0)
for k in range(nsw):
    G.push_async(*S.sweep(k))
res = G.collect()
print("sweeps", len(res), "mapped", sum(1 for r in res if r.lm_status == 0), "lo_iters", [r.lo_iters for r in res[-6:]], "lm_iters", [r.lm_iters for r in res[-6:]])
