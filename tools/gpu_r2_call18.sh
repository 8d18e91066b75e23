#!/bin/bash
V=$PWD/occ-gnn_b200/cslicer/variants
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu 2>&1 | tail -1
timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"
for v in "$@"; do echo "variant $v"; GSPLIT_SLICER_LIB=$V/lib_$v.so timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"; done
