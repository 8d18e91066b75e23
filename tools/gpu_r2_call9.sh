#!/bin/bash
# per-kernel counters of one warm minibatch (serialised by the profiler)
set -u
out=gpurun_out/r2h; mkdir -p $out
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 || { tail -5 $out/plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_atom.sum,launch__grid_size,launch__registers_per_thread \
  --clock-control none --cache-control none --profile-from-start off --csv --log-file $out/metrics.csv python tools/prof_one.py > $out/ncu.log 2>&1
echo "ncu rc=$?"
