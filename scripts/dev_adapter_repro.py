"""dev: the ROS-adapter test body as a plain script (tests/test_gpu_ros_adapter.py)"""
import sys, os, faulthandler, importlib
faulthandler.enable()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle import checker
synth = importlib.import_module("loam_velodyne-1_b200.synth")
preset = int(sys.argv[1]) if len(sys.argv) > 1 else 1
os.environ["LOAM_B200_PRESET"] = str(preset)
lib = os.path.join(ROOT, "tests/adapter/_build/libloam_ref.so")
P = checker.make_params(preset)
print("step 0", flush=True)
if len(sys.argv) > 2:
    R = checker.Checker(P, "reference" if checker.available("reference") else "port")
    print("ref created", flush=True)
A = checker.Checker(P, lib)
print("adapter created", flush=True)
if len(sys.argv) <= 2:
    R = checker.Checker(P, "reference" if checker.available("reference") else "port")
print("created", A.h, R.kind, flush=True)
S = synth.Synth(preset)
for k in range(8):
    xyz, st = S.sweep(k)
    a, r = A.push(xyz, st), R.push(xyz, st)
    print(k, a.sr_status, a.lo_status, a.lm_status, r.sr_status, r.lo_status, r.lm_status,
          bytes(a.transform_sum) == bytes(r.transform_sum), bytes(a.transform_aft_mapped) == bytes(r.transform_aft_mapped), bytes(a.transform_mapped) == bytes(r.transform_mapped), flush=True)
