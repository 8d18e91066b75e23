#!/bin/bash
set -u
out=gpurun_out/r2n2; mkdir -p $out
timeout 600 python -m pytest tests/test_handover.py -x -q -m gpu -rs > $out/two_rank_handover.log 2>&1; echo "handover tests rc=$?"; tail -3 $out/two_rank_handover.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/diag_handover.py 2>&1 | grep "minibatches/s\|Error\|error" | tail -8
