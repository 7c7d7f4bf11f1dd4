# N-GPU lines of round 2: bash scripts/r02_scale.sh N   (under gpurun --gpus N)
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_${N}gpu_driver.json 2> gpurun_out/r02_bench_${N}gpu_driver.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload seq64 --steps 200 --warmup 10 > gpurun_out/r02_bench_seq64_${N}gpu.json 2> gpurun_out/r02_bench_seq64_${N}gpu.err
python - <<PY
import json
for f in ("${N}gpu_driver","${N}gpu","seq64_${N}gpu"):
    n="gpurun_out/r02_bench_"+f+".json"
    try:
        d=json.loads(open(n).read().strip().splitlines()[-1]); print(n, round(d["value"],1), round(d["e2e"]["value"],1), d.get("per_rank_ms"))
    except Exception as e: print(n,"ERR",e, open(n.replace(".json",".err")).read()[-600:])
PY
