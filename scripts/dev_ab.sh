for fl in 0 16 8; do
  python bench.py --no-cpu-baseline --no-profile --dev-flags $fl > gpurun_out/ab.json 2>gpurun_out/ab.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab.json").read().strip().split("\n")[-1])
print("flags $fl", round(d["value"]), round(d["e2e"]["value"]), round(d["e2e"]["sync"]["p50_ms_per_sweep"],3), d["gpu_launches"], d["e2e"]["poses_identical_to_resident_arm"])
PY
done
