#!/bin/bash
# Compile-time variants of libgsplit_slicer.so for A/B runs on the GPU box
# (built here with nvcc, shipped by gpurun; selected with GSPLIT_SLICER_LIB).
#   bash tools/build_variants.sh name1:"-DFLAG=1 ..." name2:"..."
set -eu
cd "$(dirname "$0")/.."
out=occ-gnn_b200/cslicer/variants
mkdir -p $out
src="occ-gnn_b200/csrc/slicer_kernels.cu occ-gnn_b200/csrc/slicer_api.cu occ-gnn_b200/csrc/pslicer_kernels.cu occ-gnn_b200/csrc/pslicer_api.cu occ-gnn_b200/csrc/noalloc.cu occ-gnn_b200/csrc/epart.cu occ-gnn_b200/csrc/flags.cu"
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared \
    -Iinclude $flags $src -o $out/lib_$name.so &
done
wait
ls -la $out
