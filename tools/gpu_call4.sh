#!/bin/bash
set -u
out=gpurun_out/c4; mkdir -p $out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | tee $out/tests_tags.txt
V=$PWD/occ-gnn_b200/cslicer/variants
GSPLIT_SLICER_LIB=$V/lib_notags.so timeout 300 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "tagged or philox" 2>&1 | tail -3 | tee $out/tests_notags.txt
for v in notags default; do
  if [ $v = default ]; then unset GSPLIT_SLICER_LIB; else export GSPLIT_SLICER_LIB=$V/lib_$v.so; fi
  POINTS="8,8,2;12,12,2" timeout 300 python tools/variant_rate.py 2>&1 | grep "mb/s" | tee -a $out/variants.txt
done
