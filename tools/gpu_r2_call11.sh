#!/bin/bash
set -u
out=gpurun_out/r2i; mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu > $out/parity.txt 2>&1; echo "parity rc=$?"; tail -2 $out/parity.txt
for cfg in "1 8 8 2" "1 12 12 2" "1 8 8 3"; do set -- $cfg; timeout 300 python tools/rate.py $1 $2 $3 $4 2>&1 | grep "mb/s"; done
for r in 1 2 3; do echo "roles_off=$r"; GSPLIT_DIAG_ROLES=$r timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"; done
