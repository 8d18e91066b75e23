#!/bin/bash
# 8 GPUs: the N=8 bench (hand-over headline, replicas, partitioned), products shape
set -u
out=gpurun_out/r2n8; mkdir -p $out
nvidia-smi -L > $out/gpus.txt
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 10 --warmup 3 > $out/bench_n8.json 2> $out/bench_n8.err; echo "bench n8 rc=$?"; tail -3 $out/bench_n8.err; cut -c1-300 $out/bench_n8.json
