#!/bin/bash
# Round evidence on one B200: bench line, reference arm, ncu launch list, one
# `ncu --set full` capture of the dominant kernel (k_expand, layer 2).
# Run on the GPU box from the repo root: bash tools/capture_evidence.sh
set -u
out=gpurun_out/ev
mkdir -p $out
python bench.py > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_reference.json 2> $out/bench_reference.err; echo "reference rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 600 --csv \
  --log-file $out/launches.csv python bench.py --steps 1 --warmup 3 --mb-per-step 8 --no-cpu-baseline \
  > $out/launches.log 2>&1; echo "launch list rc=$?"
SLOTS=1 NMB=1 timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:k_expand -o $out/expand_full python tools/prof_range.py > $out/expand_full.log 2>&1; echo "full capture rc=$?"
ncu -i $out/expand_full.ncu-rep --page raw --csv > $out/expand_full_raw.csv 2>/dev/null; echo "export rc=$?"
ls -la $out
