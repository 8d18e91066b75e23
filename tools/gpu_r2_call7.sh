#!/bin/bash
# no-alloc entry point: alone in a fresh process, then after the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/c7_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/c7_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/c7_gpu_suite.log 2>&1; echo "suite rc=$?" >> gpurun_out/c7_gpu_suite.log
timeout 300 python tools/rate.py 1 8 8 2 > gpurun_out/c7_rate.log 2>&1
tail -3 gpurun_out/c7_alone.log gpurun_out/c7_gpu_suite.log gpurun_out/c7_rate.log
