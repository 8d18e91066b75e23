"""Where does a pipelined minibatch go?  CPU cost of submit/wait, and throughput
vs slots / streams / grid size.  Diagnostic only (not a bench line)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from cslicer.engine import Graph, Slicer

ip, ix = bench.load_graph("products", "cuda:0")
n = ip.size - 1
g = Graph(ip, ix)
batches = bench.seed_batches(n, 64, 0, 1)

def rate(sl, nmb=256):
    bench.run_pipeline(sl, batches, 2, 64, torch)
    ms, wall, e, _ = bench.run_pipeline(sl, batches, nmb // 64, 64, torch)
    return nmb / (ms * 1e-3)

import itertools
for (slots, streams, gc), hints in itertools.product(
        [(1, 1, 4), (2, 2, 4), (3, 3, 4), (4, 4, 4), (5, 5, 4), (6, 6, 4), (8, 8, 4), (12, 12, 4), (16, 16, 4),
         (4, 4, 2), (4, 4, 8), (8, 8, 2)], [False, True]):
    sl = Slicer(g, bench.FANOUTS, 4, max_batch=4096, num_slots=slots, num_streams=streams, seed=7,
                host_output=False, grid_ctas=gc, l2_hints=hints)
    r = rate(sl)
    # CPU cost of submit alone: fill all slots without waiting
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ss = [sl.submit(batches[i][1], batches[i][0]) for i in range(slots)]
    t1 = time.perf_counter()
    for s_ in ss:
        sl.wait_header(s_); sl.release(s_)
    t2 = time.perf_counter()
    print(f"hints={int(hints)} slots={slots:2d} streams={streams:2d} ctas/sm={gc}: {r:8.1f} mb/s | submit {1e6*(t1-t0)/slots:6.1f} us/call, "
          f"drain {1e6*(t2-t1)/slots:6.1f} us/mb", flush=True)
    sl.close()
