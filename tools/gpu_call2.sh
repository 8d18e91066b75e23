#!/bin/bash
set -u
out=gpurun_out/c2; mkdir -p $out
V=$PWD/occ-gnn_b200/cslicer/variants
GSPLIT_SLICER_LIB=$V/lib_stage1.so timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu 2>&1 | tail -5 | tee $out/parity_stage1.txt
for v in base stage1; do
  GSPLIT_SLICER_LIB=$V/lib_$v.so POINTS="8,8,2;12,12,2" timeout 300 python tools/variant_rate.py 2>&1 | grep "mb/s" | tee -a $out/variants.txt
done
