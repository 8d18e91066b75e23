#!/bin/bash
set -u
out=gpurun_out/r2f; mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu > $out/parity.txt 2>&1; echo "parity rc=$?"; tail -2 $out/parity.txt
for cfg in "1 8 8 2" "2 8 4 2"; do set -- $cfg; timeout 300 python tools/rate.py $1 $2 $3 $4 2>&1 | grep "mb/s"; done
V=$PWD/occ-gnn_b200/cslicer/variants
echo "variant r5"; GSPLIT_SLICER_LIB=$V/lib_r5.so timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,launch__grid_size,launch__registers_per_thread \
  --clock-control none --profile-from-start off --csv --log-file $out/metrics.csv python tools/prof_one.py > $out/ncu.log 2>&1
echo "ncu rc=$?"
