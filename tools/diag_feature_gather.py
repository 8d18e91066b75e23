"""Feature-row gather rate (gs_features_gather, one warp per row) against this GPU's
plain copy rate and torch.index_select, for several row widths.  Diagnostic only."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from cslicer import _native as N
from cslicer.engine import Features
import ctypes as C

def timeit(fn, reps=8):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

n_rows, n_ids = 2_449_029, 1_265_000
ids = torch.randint(0, n_rows, (n_ids,), device="cuda", dtype=torch.int32)
ids64 = ids.long()
for dim in (128, 600, 602, 604, 608):
    rows = np.zeros((n_rows, dim), dtype=np.float32)
    F = Features(rows)
    tab = torch.zeros((n_rows, dim), device="cuda")
    out = torch.empty((n_ids, dim), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ms = timeit(lambda: N.check(N.lib().gs_features_gather(F._h, ids.data_ptr(), n_ids, out.data_ptr(), st)))
    ms_t = timeit(lambda: torch.index_select(tab, 0, ids64, out=out))
    src = torch.empty((n_ids, dim), device="cuda")
    ms_c = timeit(lambda: out.copy_(src))
    b = 2 * n_ids * dim * 4
    print(f"dim {dim:4d}: gather {b/ms/1e6:7.1f} GB/s   torch.index_select {b/ms_t/1e6:7.1f} GB/s   "
          f"plain copy {b/ms_c/1e6:7.1f} GB/s", flush=True)
    F.close(); del tab, out, src, rows
