"""One warm minibatch of the products workload for `ncu --set full` (all launches of
one sample, eager, single slot).  Diagnostic only."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from cslicer.engine import Graph, Slicer

group = int(os.environ.get("GROUP", "1"))
gc = int(os.environ.get("GRID_CTAS", "2"))
ip, ix = bench.load_graph("products", "cuda:0")
g = Graph(ip, ix)
batches = bench.seed_batches(ip.size - 1, 64, 0, 1)
sl = Slicer(g, bench.FANOUTS, 4, max_batch=4096, num_slots=group, num_streams=1, seed=7,
            host_output=False, grid_ctas=gc, use_graph=False, group_size=group)
bench.run_pipeline(sl, batches, 2, 4 * group, torch)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
ms, _, edges, _ = bench.run_pipeline(sl, batches, 1, group, torch)
torch.cuda.cudart().cudaProfilerStop()
print(f"group={group} ms={ms:.3f}")
sl.close(); g.close()
