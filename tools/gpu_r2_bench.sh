#!/bin/bash
# default bench (both arms) + full GPU suite on one B200
set -u
out=gpurun_out/r2b; mkdir -p $out
timeout 1500 python -m pytest tests -x -q -m gpu > $out/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -2 $out/gpu_suite.log
timeout 900 python bench.py > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"; tail -3 $out/bench_n1.err; cut -c1-600 $out/bench_n1.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_ref.json 2> $out/bench_ref.err; echo "ref rc=$?"; cut -c1-600 $out/bench_ref.json
