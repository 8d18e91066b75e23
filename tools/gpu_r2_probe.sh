#!/bin/bash
# the random-sector yardstick kernel (k_sector_probe) alone and under ncu --set full
out=gpurun_out/r2probe; mkdir -p $out
timeout 300 python tools/probe_sectors.py > $out/sector_probe.txt 2>&1; cat $out/sector_probe.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sector_probe -c 6 -o $out/sector_probe_full python tools/probe_sectors.py > $out/ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $out/sector_probe_full.ncu-rep --page raw --csv > $out/sector_probe_full_raw.csv 2>/dev/null; echo "export rc=$?"
