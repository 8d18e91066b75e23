#!/bin/bash
set -u
out=gpurun_out/r2n8; mkdir -p $out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 10 --warmup 3 --no-partitioned > $out/bench_n8b.json 2> $out/bench_n8b.err; echo "bench n8 rc=$?"; grep -v "^\[bench\] gen\|OMP\|^\*\*\*" $out/bench_n8b.err | tail -5
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2n8/bench_n8b.json').read().strip().splitlines()[-1])
h=d['hand_over']
print('headline',h['headline_exchange'],h['value'],h['parity'],'e2e',h['e2e']['value'])
for k in ('staged','direct','peer_stores'): print(k,h[k]['value'],h[k]['parity'],h[k]['e2e'])
print('nccl',h['same_with_nccl_all_to_all']['value'],'replicas',d['replicas']['value'])
PY
