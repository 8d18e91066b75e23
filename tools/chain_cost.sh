#!/bin/bash
# Machine time of every launch of the chain WITH ALL SLOTS IN FLIGHT: the pipelined rate of the
# chain cut after k launches (GSPLIT_DIAG_STOP, diagnostic switch of enqueue_group), k = 1..4L+1.
# 1e6 / rate(k) - 1e6 / rate(k-1) = microseconds of the whole GPU that launch k costs per minibatch.
out=${1:-gpurun_out/chain_cost.txt}; : > $out
for k in 0 1 2 3 4 5 6 7 8 9 10 11 12 13; do
  r=$(GSPLIT_DIAG_STOP=$k timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s" | sed 's/.*: *//; s/ mb.*//')
  echo "$k $r" | tee -a $out
done
python - "$out" <<'PY'
import sys
names = ["housekeeping"] + [f"{n}@{l}" for l in range(3) for n in ("rows", "expand", "cand_flags", "cand_apply")] + ["tail"]
prev = 0.0
for line in open(sys.argv[1]):
    k, r = line.split(); k = int(k); t = 1e6 / float(r)
    print(f"{names[k]:16s} cumulative {t:7.1f} us   this launch {t - prev:6.1f} us")
    prev = t
PY
