for lib in libloam_b200.so libloam_b200_r112.so libloam_b200_r96.so; do
  for fl in 0 4; do
    LOAM_B200_LIB=$PWD/loam_velodyne-1_b200/$lib python bench.py --no-cpu-baseline --no-profile --dev-flags $fl > gpurun_out/ab.json 2>gpurun_out/ab.err
    python - <<PY
import json
d=json.loads(open("gpurun_out/ab.json").read().strip().split("\n")[-1])
print("$lib flags $fl", round(d["value"]), round(d["e2e"]["value"]), round(d["e2e"]["sync"]["p50_ms_per_sweep"],3), d["gpu_launches"], d["e2e"]["poses_identical_to_resident_arm"])
PY
  done
done
