"""Throughput of the pipelined slicer for the library named by GSPLIT_SLICER_LIB
at a few (slots, streams, CTAs/SM) points.  Diagnostic only (not a bench line)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from cslicer.engine import Graph, Slicer

pts = [tuple(map(int, p.split(","))) for p in os.environ.get("POINTS", "8,8,2;12,12,2;16,16,2;8,8,3").split(";")]
ip, ix = bench.load_graph("products", "cuda:0")
g = Graph(ip, ix)
batches = bench.seed_batches(ip.size - 1, 64, 0, 1)
tag = os.path.basename(os.environ.get("GSPLIT_SLICER_LIB", "default"))
for slots, streams, gc in pts:
    sl = Slicer(g, bench.FANOUTS, 4, max_batch=4096, num_slots=slots, num_streams=streams, seed=7,
                host_output=False, grid_ctas=gc)
    bench.run_pipeline(sl, batches, 3, 64, torch)
    best = 0.0
    for _ in range(3):
        ms, _, _, _ = bench.run_pipeline(sl, batches, 8, 64, torch)
        best = max(best, 512 / (ms * 1e-3))
    print(f"{tag}: slots={slots} streams={streams} ctas/sm={gc}: {best:8.1f} mb/s", flush=True)
    sl.close()
g.close()
