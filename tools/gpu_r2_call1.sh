#!/bin/bash
# Round 2, first GPU call: parity of the fused chain, then first rates.
set -u
out=gpurun_out/r2a; mkdir -p $out
nvidia-smi --query-gpu=name,memory.total --format=csv > $out/gpu.txt 2>&1
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_random_sweep_gpu.py -x -q -m gpu > $out/parity.txt 2>&1
echo "parity rc=$?" | tee -a $out/parity.txt
tail -15 $out/parity.txt
timeout 600 python -m pytest tests -q -m gpu --deselect tests/test_parity_gpu.py --deselect tests/test_random_sweep_gpu.py > $out/rest.txt 2>&1
echo "rest rc=$?" | tee -a $out/rest.txt
tail -8 $out/rest.txt
for cfg in "1 12 12" "2 12 6" "4 12 3" "1 6 6"; do
  set -- $cfg
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --group $1 --slots $2 --streams $3 > $out/bench_g$1_s$2_t$3.json 2> $out/bench_g$1_s$2_t$3.err
  echo "bench group=$1 slots=$2 streams=$3 rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("$out/bench_g$1_s$2_t$3.json"))
    print("  value", d["value"], "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"], "kernels", d["roofline"]["kernels_ms"])
except Exception as e:
    print("  no line:", e)
PY
done
