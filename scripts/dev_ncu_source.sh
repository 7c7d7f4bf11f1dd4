# per-line stall sampling of one kernel: bash scripts/dev_ncu_source.sh <kernel regex> <skip> ; output gpurun_out/src_<name>.csv
K=$1; S=${2:-10}
python scripts/run_sequence.py 2 40 > gpurun_out/ncu_plain_seq.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -o gpurun_out/src_cap python scripts/run_sequence.py 2 40 > gpurun_out/ncu_src.log 2>&1
echo "rc=$?"
ncu -i gpurun_out/src_cap.ncu-rep --page source --csv > gpurun_out/src_$K.csv 2>/dev/null
ncu -i gpurun_out/src_cap.ncu-rep --page raw --csv > gpurun_out/src_${K}_raw.csv 2>/dev/null
rm -f gpurun_out/src_cap.ncu-rep
ls -la gpurun_out/src_* | head
