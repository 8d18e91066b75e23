#!/bin/bash
set -u
out=gpurun_out/c3; mkdir -p $out
P=8 timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off --csv --log-file $out/part_p8.csv python tools/prof_partitioned.py > $out/part_p8.log 2>&1; echo rc=$?
tail -2 $out/part_p8.log
