#!/bin/bash
set -u
out=gpurun_out/r2p; mkdir -p $out
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -o $out/one_full python tools/prof_one.py > $out/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 $out/ncu.log; ls -la $out
