timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_pytest_gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_driver.json 2> gpurun_out/r02_bench_driver.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err
python bench.py --steps 200 --warmup 10 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err
for wl in vlp16 ring128 seq64 mapstress; do python bench.py --workload $wl --steps 200 --warmup 10 > gpurun_out/r02_bench_$wl.json 2> gpurun_out/r02_bench_$wl.err; done
python - <<PY
import json
for f in ("_driver","_reference","","_vlp16","_ring128","_seq64","_mapstress"):
    n="gpurun_out/r02_bench"+f+".json"
    try:
        d=json.loads(open(n).read().strip().splitlines()[-1]); print(n, round(d["value"],1), d.get("e2e",{}).get("value"), (d.get("roofline") or {}).get("frac"), (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(n,"ERR",e)
PY
