#!/bin/bash
# Round 2: parity, instruction counts per kernel (ncu metrics pass), rate vs grid / slots.
set -u
out=gpurun_out/r2b; mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu > $out/parity.txt 2>&1
echo "parity rc=$?" | tee -a $out/parity.txt; tail -4 $out/parity.txt
timeout 600 python -m pytest tests -q -m gpu --deselect tests/test_parity_gpu.py --deselect tests/test_random_sweep_gpu.py > $out/rest.txt 2>&1
echo "rest rc=$?" | tee -a $out/rest.txt; tail -4 $out/rest.txt
for cfg in "1 12 12 2" "1 12 12 1" "1 12 12 4" "1 24 24 1" "4 12 3 1" "2 16 8 1"; do
  set -- $cfg
  timeout 300 python tools/rate.py $1 $2 $3 $4 2>&1 | grep "mb/s"
done
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct \
  --clock-control none --profile-from-start off --csv --log-file $out/metrics.csv python tools/prof_one.py > $out/ncu.log 2>&1
echo "ncu rc=$?"
