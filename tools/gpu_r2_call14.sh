#!/bin/bash
# per-kernel L1 / L2 transaction counts of one warm minibatch (serialised by the profiler)
set -u
out=gpurun_out/r2k; mkdir -p $out
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 || { tail -5 $out/plain.log; exit 1; }
M=gpu__time_duration.sum,smsp__inst_executed.sum,l1tex__t_requests.sum,l1tex__t_sectors.sum,l1tex__data_pipe_lsu_wavefronts.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__m_xbar2l1tex_read_sectors.sum,l1tex__m_l1tex2xbar_write_sectors.sum,l1tex__m_l1tex2xbar_req_cycles_active.sum,l1tex__lsu_writeback_active.sum,l1tex__lsuin_requests.sum,lts__t_requests.sum,lts__t_sectors.sum,smsp__inst_executed_pipe_lsu.sum,sm__cycles_active.sum,l1tex__throughput.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed
timeout 900 ncu --metrics $M --clock-control none --cache-control none --profile-from-start off --csv --log-file $out/metrics.csv python tools/prof_one.py > $out/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 $out/ncu.log
