#!/bin/bash
set -u
V=$PWD/occ-gnn_b200/cslicer/variants
for v in e6 e6c6 r5e6c6 e8c8; do
  echo "variant $v"; GSPLIT_SLICER_LIB=$V/lib_$v.so timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"
done
for cfg in "1 8 8 1" "1 8 8 3" "1 8 8 4"; do set -- $cfg; timeout 300 python tools/rate.py $1 $2 $3 $4 2>&1 | grep "mb/s"; done
