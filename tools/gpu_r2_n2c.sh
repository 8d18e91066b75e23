#!/bin/bash
set -u
out=gpurun_out/r2n2; mkdir -p $out
timeout 600 python -m pytest tests/test_handover.py -x -q -m gpu -rs > $out/two_rank_handover.log 2>&1; echo "handover tests rc=$?"; tail -4 $out/two_rank_handover.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-partitioned > $out/bench_n2.json 2> $out/bench_n2.err; echo "bench n2 rc=$?"; grep -v "^\[bench\] gen\|OMP\|^\*\*\*" $out/bench_n2.err | tail -5
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2n2/bench_n2.json').read().strip().splitlines()[-1])
h=d['hand_over']
print('headline',h['headline_exchange'],h['value'],h['parity'],'e2e',h['e2e']['value'])
for k in ('staged','direct','peer_stores'): print(k,h[k]['value'],h[k]['parity'],h[k]['e2e'])
print('nccl',h['same_with_nccl_all_to_all']['value'],'replicas',d['replicas']['value'])
PY
