# A/B of build variants (build_ab/*.so, made with LOAM_LIB_OUT=... python loam_velodyne-1_b200/build.py --force) on one box:
#   bash scripts/dev_ab2.sh <workload> <steps> <variant> [<variant> ...]     (variant "base" = the in-tree library)
wl=$1; steps=$2; shift 2
for v in "$@"; do
  if [ "$v" = base ]; then unset LOAM_B200_LIB; else export LOAM_B200_LIB=$PWD/build_ab/$v.so; fi
  python bench.py --workload $wl --steps $steps --warmup 10 --no-cpu-baseline > gpurun_out/ab_${wl}_$v.json 2> gpurun_out/ab_${wl}_$v.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ab_${wl}_$v.json").read().strip().split("\n")[-1])
    print("$wl $v", round(d["value"]), round(d["e2e"]["value"]), [(k["kernel"], k["avg_launch_us"]) for k in (d.get("top_kernels") or [])[:9]])
except Exception as e:
    print("$wl $v ERR", e, open("gpurun_out/ab_${wl}_$v.err").read()[-400:])
PY
done
