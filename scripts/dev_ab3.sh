# A/B of build variants at the driver's settings and over 200 sweeps: bash scripts/dev_ab3.sh <variant> ...
for v in "$@"; do
  if [ "$v" = base ]; then unset LOAM_B200_LIB; else export LOAM_B200_LIB=$PWD/build_ab/$v.so; fi
  for st in 20 200; do
    python bench.py --steps $st --warmup 5 --no-cpu-baseline --no-profile 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v steps $st', round(d['value']), round(d['e2e']['value']), d['e2e']['poses_identical_to_resident_arm'])"
  done
done
