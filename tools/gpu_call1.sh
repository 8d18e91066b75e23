#!/bin/bash
# GPU call: A/B of register-cap variants + what the pipelined chain saturates (range replay).
# Build the variants first, in the build container (nvcc cross-compiles, the .so files travel):
#   bash tools/build_variants.sh base:"" minb6:"-DGS_MINB=6" minb8:"-DGS_MINB=8"
# Other A/B switches: -DGS_TABLE_TAGS=0 (untagged dedup table), -DGS_PACK_TMA=0 (hand-over by plain stores).
set -u
out=gpurun_out/c1; mkdir -p $out
V=occ-gnn_b200/cslicer/variants
for v in base minb6 minb8; do
  GSPLIT_SLICER_LIB=$PWD/$V/lib_$v.so timeout 300 python tools/variant_rate.py 2>&1 | grep "mb/s" | tee -a $out/variants.txt
done
M=gpu__time_duration.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_issued.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sectors.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,lts__t_sector_hit_rate.pct,lts__t_tag_requests.sum,lts__d_sectors.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,lts__t_sectors.avg.pct_of_peak_sustained_elapsed,lts__t_tag_requests.avg.pct_of_peak_sustained_elapsed,lts__d_sectors.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_sectors.sum,l1tex__m_l1tex2xbar_write_sectors.sum,lts__xbar2lts_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__lts2xbar_cycles_active.avg.pct_of_peak_sustained_elapsed
SLOTS=8 NMB=64 timeout 600 ncu --replay-mode range --clock-control none --cache-control none --metrics $M --csv \
  --log-file $out/range_base.csv python tools/prof_range.py > $out/range_base.log 2>&1; echo "range rc=$?"
tail -3 $out/range_base.log
