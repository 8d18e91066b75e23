"""Generates tests/golden/cache_policy.npz by RUNNING THE REFERENCE'S OWN
MemoryManager (python/utils/memory_manager.py, imported unmodified from
/root/reference) on a small graph.  Run in the build container only:
    python -O tests/golden/make_cache_policy_golden.py     (-O: the module asserts tensor.device == cuda:i)

MemoryManager needs a DGL graph, four CUDA devices and pinned memory; none of
that touches the policy, so the script hands it (a) a stand-in graph object with
the two methods it calls (num_nodes, out_degree), (b) torch.zeros with the
`device=` argument dropped, (c) a no-op Tensor.pin_memory.  What is recorded is
`local_to_global_id[g]` -- the nodes GPU g caches, in slot order -- for cache
fractions below, at and above 1/4, plus refresh_cache's per-batch additions.
Out-degrees are made distinct so that torch.sort's order among ties cannot matter.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/python/utils/memory_manager.py"


class FakeGraph:
    def __init__(self, out_deg):
        self._d = torch.as_tensor(out_deg, dtype=torch.long)

    def num_nodes(self):
        return int(self._d.numel())

    def num_edges(self):
        return int(self._d.sum())

    def out_degree(self, nds):
        return self._d[nds]


def main():
    spec = importlib.util.spec_from_file_location("ref_memory_manager", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    real_zeros = torch.zeros
    torch.zeros = lambda *a, **k: real_zeros(*a, **{x: y for x, y in k.items() if x != "device"})
    torch.Tensor.pin_memory = lambda self, *a, **k: self
    rng = np.random.default_rng(7)
    n = 400
    out_deg = rng.permutation(n).astype(np.int64) + 1          # distinct degrees 1..n
    part = rng.integers(0, 4, n).astype(np.int64)              # METIS stand-in
    feats = torch.arange(n, dtype=torch.float32).reshape(n, 1)
    out = {"out_degree": out_deg, "partition_map": part}
    import contextlib
    import io
    for frac in (0.05, 0.25, 0.5, 1.0):
        with contextlib.redirect_stdout(io.StringIO()):
            mm = mod.MemoryManager(FakeGraph(out_deg), feats, frac, [10, 10, 10], 64,
                                   torch.as_tensor(part))
        for g in range(4):
            out[f"cached_{frac}_{g}"] = mm.local_to_global_id[g].numpy().astype(np.int64)
        if frac <= 0.25:
            last = torch.as_tensor(rng.permutation(n)[:120])
            mm.refresh_cache(last)
            out[f"last_layer_{frac}"] = last.numpy().astype(np.int64)
            for g in range(4):
                out[f"refreshed_{frac}_{g}"] = mm.local_to_global_id[g].numpy().astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "cache_policy.npz"), **out)
    print("wrote cache_policy.npz:", {k: v.shape for k, v in out.items() if k.startswith("cached")})


if __name__ == "__main__":
    main()
