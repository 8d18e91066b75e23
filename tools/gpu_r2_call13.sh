#!/bin/bash
set -u
out=gpurun_out/r2j; mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu > $out/parity.txt 2>&1; echo "parity rc=$?"; tail -2 $out/parity.txt
timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"
for r in "$@"; do echo "roles_off=$r"; GSPLIT_DIAG_ROLES=$r timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"; done
