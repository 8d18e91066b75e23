#!/bin/bash
set -u
out=gpurun_out/r2e; mkdir -p $out
python tools/probe_sectors.py 2>&1 | tee $out/sector_probe.txt
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -m gpu > $out/parity.txt 2>&1; echo "parity rc=$?"; tail -2 $out/parity.txt
for cfg in "1 6 6 2" "1 12 12 2" "2 12 6 2"; do set -- $cfg; timeout 300 python tools/rate.py $1 $2 $3 $4 2>&1 | grep "mb/s"; done
