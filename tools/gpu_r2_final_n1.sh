#!/bin/bash
# final single-GPU evidence: suite, bench (both arms), launch list, full capture of the top kernel, range replay
set -u
out=gpurun_out/r2f1; mkdir -p $out
timeout 900 python -m pytest tests -x -q -m gpu > $out/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -2 $out/gpu_suite.log
timeout 600 python bench.py > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"; cut -c1-160 $out/bench_n1.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_ref.json 2> $out/bench_ref.err; echo "ref rc=$?"; cut -c1-200 $out/bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --mb-per-step 24 > $out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 300 python tools/prof_one.py > $out/plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_expand_rows -o $out/expand_rows_full python tools/prof_one.py > $out/full.log 2>&1; echo "full capture rc=$?"
ncu -i $out/expand_rows_full.ncu-rep --page raw --csv > $out/expand_rows_full_raw.csv 2>/dev/null; echo "export rc=$?"
M=gpu__time_duration.sum,lts__t_sectors.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_atom.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,l1tex__t_sectors.sum,l1tex__m_xbar2l1tex_read_sectors.sum,l1tex__m_l1tex2xbar_write_sectors.sum
SLOTS=8 NMB=64 timeout 600 ncu --replay-mode range --clock-control none --cache-control none --metrics $M --csv --log-file $out/range.csv python tools/prof_range.py > $out/range.log 2>&1; echo "range rc=$?"
ls -la $out | head -20
