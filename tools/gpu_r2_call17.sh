#!/bin/bash
V=$PWD/occ-gnn_b200/cslicer/variants
timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"
timeout 300 python tools/rate.py 1 6 6 2 2>&1 | grep "mb/s"
for v in "$@"; do echo "variant $v"; GSPLIT_SLICER_LIB=$V/lib_$v.so timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"; done
