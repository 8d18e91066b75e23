#!/bin/bash
# what the pipelined chain moves per minibatch: ncu --replay-mode range over 64 minibatches on 8 slots
set -u
out=gpurun_out/r2g; mkdir -p $out
M=gpu__time_duration.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_issued.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sectors.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,lts__t_sector_hit_rate.pct,lts__t_tag_requests.sum,lts__d_sectors.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,l1tex__m_xbar2l1tex_read_sectors.sum,l1tex__m_l1tex2xbar_write_sectors.sum
SLOTS=8 NMB=64 timeout 300 python tools/prof_range.py > $out/range_plain.log 2>&1 || { tail -5 $out/range_plain.log; exit 1; }
tail -1 $out/range_plain.log
SLOTS=8 NMB=64 timeout 900 ncu --replay-mode range --clock-control none --cache-control none --metrics $M --csv \
  --log-file $out/range.csv python tools/prof_range.py > $out/range.log 2>&1; echo "range rc=$?"
tail -2 $out/range.log
