"""Throughput of the pipelined slicer at one (group, slots, streams, CTAs/SM) point.
Diagnostic only (not a bench line).  usage: rate.py GROUP SLOTS STREAMS GRID_CTAS"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from cslicer.engine import Graph, Slicer

group, slots, streams, gc = (int(x) for x in sys.argv[1:5])
parts = int(os.environ.get("PARTS", "4"))
ip, ix = bench.load_graph("products", "cuda:0")
g = Graph(ip, ix)
batches = bench.seed_batches(ip.size - 1, 64, 0, 1)
sl = Slicer(g, bench.FANOUTS, parts, max_batch=4096, num_slots=slots, num_streams=streams, seed=7,
            host_output=False, grid_ctas=gc, group_size=group)
bench.run_pipeline(sl, batches, 3, 96, torch)
best = 0.0
for _ in range(3):
    ms, _, _, _ = bench.run_pipeline(sl, batches, 4, 96, torch)
    best = max(best, 384 / (ms * 1e-3))
print(f"group={group} slots={slots} streams={streams} ctas/sm={gc} parts={parts}: {best:8.1f} mb/s", flush=True)
sl.close(); g.close()
