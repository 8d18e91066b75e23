"""Which feature rows a GPU keeps in HBM.

Host-side restatement of MemoryManager.initialize's policy
(/root/reference python/utils/memory_manager.py:40-74): with a cache fraction c
of all N nodes per GPU,
  * c <= 1/P : the c*N highest-degree nodes of the GPU's own partition;
  * c  > 1/P : the whole partition, then the highest-degree nodes of the rest
               until c*N rows are held.
The reference orders by `torch.sort(descending=True)`, which leaves the order of
equal degrees unspecified; here ties go to the smaller node id (a stable sort),
so the choice is deterministic.  Pinned against the reference's own code:
tests/golden/cache_policy.npz holds `MemoryManager.local_to_global_id` as the
unmodified module computes it (tests/golden/make_cache_policy_golden.py).  The rows chosen feed
`engine.Features(rows, cache_rows=...)`; nothing else of MemoryManager is needed:
a miss is read from host memory by the gather kernel, so there is no
refresh_cache / clean_up cycle (memory_manager.py:75-104).
"""
import numpy as np


def degree_ordered_cache(indptr, fraction, owner=None, part=0, num_parts=4, indices=None, degrees=None):
    """-> int64 array of node ids to cache on GPU `part`, cached order = slot order.
    owner: int array [N] (None: id % num_parts, pyfrontend.cpp:57).
    indices: the CSR's neighbour array.  Given, nodes are ranked by how often they
    occur in it -- their out-degree as a message SOURCE, which is what
    `graph.out_degree` is at memory_manager.py:54,61 and what decides how often
    sampling reaches a node.  Omitted, the row length is used (the same number
    on a symmetric graph such as the reference's datasets).  degrees: an explicit
    per-node degree array (then indptr may be None)."""
    if not 0.0 < fraction <= 1.0:
        raise ValueError("cache fraction must be in (0, 1]")  # memory_manager.py:28
    if degrees is not None:
        deg = np.asarray(degrees, dtype=np.int64)
        n = deg.size
    else:
        ip = np.asarray(indptr, dtype=np.int64)
        n = ip.size - 1
        deg = np.diff(ip) if indices is None else np.bincount(np.asarray(indices), minlength=n)
    own = (np.arange(n) % num_parts) if owner is None else np.asarray(owner)
    n_cached = int(fraction * n)                      # memory_manager.py:42
    mine = np.nonzero(own == part)[0]
    if fraction <= 1.0 / num_parts:                   # memory_manager.py:52-56 (reference: <= .25, P = 4)
        order = np.argsort(-deg[mine], kind="stable")
        return mine[order[:n_cached]].astype(np.int64)
    rest = np.nonzero(own != part)[0]                 # memory_manager.py:58-66
    order = np.argsort(-deg[rest], kind="stable")
    extra = rest[order[:max(0, n_cached - mine.size)]]
    return np.concatenate([mine, extra]).astype(np.int64)


def hit_rate(cache_ids, ids, num_nodes):
    """Fraction of `ids` served from HBM."""
    m = np.zeros(num_nodes, dtype=bool)
    m[np.asarray(cache_ids, dtype=np.int64)] = True
    ids = np.asarray(ids, dtype=np.int64)
    return float(m[ids].mean()) if ids.size else 1.0
