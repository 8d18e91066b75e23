#!/bin/bash
set -u
out=gpurun_out/r2d; mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu > $out/parity.txt 2>&1
echo "parity rc=$?" | tee -a $out/parity.txt; tail -3 $out/parity.txt
for cfg in "1 1 1 2" "1 3 3 2" "1 6 6 2" "1 12 12 2" "1 16 16 2" "4 12 3 2" "2 12 6 2"; do
  set -- $cfg
  timeout 300 python tools/rate.py $1 $2 $3 $4 2>&1 | grep "mb/s"
done
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -o $out/one_full python tools/prof_one.py > $out/ncu.log 2>&1
echo "ncu rc=$?"
