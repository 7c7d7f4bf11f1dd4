# ncu evidence of round 2 (one gpurun call): launch list of the driver's bench command, then a full capture of one
# steady-state period of the overlapped pipeline.  Each command first runs plain and must exit 0.
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 460 --csv --log-file gpurun_out/r02_ncu_launches.csv \
    python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-profile > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python scripts/run_sequence.py 2 50 > gpurun_out/ncu_plain_seq.log 2>&1 &&
ncu --set full --clock-control none --import-source on --launch-skip 900 -c 48 -o gpurun_out/r02_ncu_full \
    python scripts/run_sequence.py 2 50 > gpurun_out/ncu_seq.log 2>&1
echo "full capture rc=$?"
ncu -i gpurun_out/r02_ncu_full.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_raw.csv 2>/dev/null
python scripts/dev_ncu_summary.py gpurun_out/r02_ncu_full_raw.csv gpurun_out/r02_ncu_full_summary.json > gpurun_out/r02_ncu_full_summary.txt 2>&1
rm -f gpurun_out/r02_ncu_full.ncu-rep     # 100+ MB: only the exported pages travel back
ls -la gpurun_out | head -30
tail -40 gpurun_out/r02_ncu_full_summary.txt
