"""Steady-state DRAM / L2 counters of the PIPELINED slicer (all slots in flight):
run under `ncu --replay-mode range` so concurrent kernels are measured together.
  ncu --replay-mode range --metrics dram__bytes_read.sum,dram__bytes_write.sum,... python tools/prof_range.py
Diagnostic only (numbers under a profiler are not bench values)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from cslicer.engine import Graph, Slicer

slots = int(os.environ.get("SLOTS", "8"))
nmb = int(os.environ.get("NMB", "64"))
ip, ix = bench.load_graph("products", "cuda:0")
g = Graph(ip, ix)
batches = bench.seed_batches(ip.size - 1, 64, 0, 1)
sl = Slicer(g, bench.FANOUTS, 4, max_batch=4096, num_slots=slots, num_streams=slots, seed=7,
            host_output=False, grid_ctas=2, use_graph=False)
bench.run_pipeline(sl, batches, 3, 64, torch)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
ms, _, edges, _ = bench.run_pipeline(sl, batches, 1, nmb, torch)
torch.cuda.cudart().cudaProfilerStop()
print(f"slots={slots} minibatches={nmb} ms={ms:.3f} ({nmb / ms * 1e3:.0f} mb/s under the profiler)")
sl.close(); g.close()
