#!/bin/bash
# 2 GPUs: the full GPU suite (incl. the real 2-rank tests, both hand-over copy paths) and the N=2 bench, both arms
set -u
out=gpurun_out/r2n2; mkdir -p $out
nvidia-smi -L > $out/gpus.txt
for tma in 0 1; do
  GSPLIT_PACK_TMA=$tma timeout 900 python -m pytest tests/test_handover.py tests/test_partitioned.py -x -q -m gpu -rs > $out/two_rank_tests_tma$tma.log 2>&1; echo "2-rank tests (GSPLIT_PACK_TMA=$tma) rc=$?"; tail -3 $out/two_rank_tests_tma$tma.log
done
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $out/bench_n2.json 2> $out/bench_n2.err; echo "bench n2 rc=$?"; tail -5 $out/bench_n2.err; cut -c1-1500 $out/bench_n2.json
