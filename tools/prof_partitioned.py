"""Per-kernel times of the partitioned slicer, P ranks emulated on one device
(LocalGroup).  Run under `ncu --metrics gpu__time_duration.sum --profile-from-start off`.
Diagnostic only."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from cslicer.partitioned import LocalGroup

P = int(os.environ.get("P", "8"))
ip, ix = bench.load_graph("products", "cuda:0")
grp = LocalGroup(ip, ix, P, bench.FANOUTS, max_batch=4096, num_slots=2, seed=7, host_output=False)
batches = bench.seed_batches(ip.size - 1, 8, 0, 1)
for i in range(3):
    for v in grp.slice(batches[i][1], batches[i][0]):
        v.release()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
for v in grp.slice(batches[3][1], batches[3][0]):
    v.release()
torch.cuda.cudart().cudaProfilerStop()
grp.close()
