"""Randomised parity sweep (seeded, deterministic): random small graphs with hubs,
isolated nodes and duplicate-heavy neighbourhoods x random layer counts, fanouts
(incl. -1 and 1), part counts up to GS_MAX_PARTS, owner maps, batch sizes and both
sampling rules -- every field of every (layer, part) against the oracle, on the
single-GPU slicer and on the partitioned slicer (P ranks emulated on one device)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200"))

from oracle import oracle as O  # noqa: E402

pytestmark = pytest.mark.gpu


def _random_graph(rng):
    n = int(rng.integers(1, 400)) if rng.integers(0, 4) else int(rng.integers(400, 5000))
    kind = rng.integers(0, 4)
    if kind == 0:      # sparse, many isolated rows
        deg = rng.integers(0, 3, n)
    elif kind == 1:    # a few hubs that touch everything: dedup stress
        deg = rng.integers(0, 6, n)
        deg[rng.integers(0, n, max(1, n // 50))] = n
    elif kind == 2:    # dense
        deg = rng.integers(min(n, 8), min(n, 40) + 1, n)
    else:              # power-law-ish
        deg = np.minimum(n, (rng.pareto(1.5, n) * 3).astype(np.int64))
    rows = np.repeat(np.arange(n), deg)
    dst = rng.integers(0, n, rows.size)
    if rng.integers(0, 2):  # half of the graphs keep self loops (slicer.cpp:33 drops them as edges)
        keep = np.ones(rows.size, bool)
    else:
        keep = dst != rows
    key = np.unique(rows[keep] * n + dst[keep])
    r = key // n
    ip = np.zeros(n + 1, np.int64)
    ip[1:] = np.cumsum(np.bincount(r, minlength=n))
    return ip, (key - r * n).astype(np.int64)


def _random_config(rng, n):
    L = int(rng.integers(1, 6))
    fan = [int(rng.choice([-1, 1, 2, 3, 5, 10, 25])) for _ in range(L)]
    if sum(f < 0 for f in fan) > 2:   # keep the full-neighbour blow-up bounded
        fan = [f if i < 2 or f > 0 else 3 for i, f in enumerate(fan)]
    P = int(rng.integers(1, 9))
    batch = int(rng.integers(1, min(n, 64 if n < 400 else 700) + 1))
    return fan, P, batch


@pytest.mark.parametrize("case", range(64))
def test_single_gpu_slicer_random_case(case):
    import torch
    assert torch.cuda.is_available()
    from cslicer import engine as E
    from __graft_entry__ import compare_sample
    rng = np.random.default_rng(1000 + case)
    ip, ix = _random_graph(rng)
    n = ip.size - 1
    fan, P, batch = _random_config(rng, n)
    owner_kind = rng.integers(0, 3)
    own = None
    if owner_kind == 1:
        own = ((np.arange(n, dtype=np.int64) * P) // n).astype(np.int32)
    elif owner_kind == 2:
        own = rng.integers(0, P, n).astype(np.int32)
    replace = bool(rng.integers(0, 2))
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fan, P, max_batch=batch, num_slots=3, num_streams=2, seed=case, owner=own,
                  replace=replace, use_graph=bool(case % 2))
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX if replace else O.RNG_PHILOX_NOREPLACE,
                         seed=case, owner=own)
    for b in range(4):
        nb = batch if b < 3 else int(rng.integers(0, batch + 1))  # ragged / empty last batch
        seeds = rng.permutation(n)[:nb]
        v = sl.slice(seeds, b)
        compare_sample(v, orc.slice(seeds, b))
        v.release()
    sl.close()
    g.close()


@pytest.mark.parametrize("case", range(32))
def test_partitioned_slicer_random_case(case):
    import torch
    assert torch.cuda.is_available()
    from cslicer.partitioned import LocalGroup, owner_of, range_bounds
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_partitioned import _compare_rank
    rng = np.random.default_rng(5000 + case)
    ip, ix = _random_graph(rng)
    n = ip.size - 1
    fan, P, batch = _random_config(rng, n)
    P = min(P, n)
    mode = "mod" if rng.integers(0, 2) else "range"
    b = range_bounds(n, P)
    own = None if mode == "mod" else owner_of(np.arange(n), n, P, "range", b)
    replace = bool(rng.integers(0, 2))
    grp = LocalGroup(ip, ix, P, fan, owner_mode=mode, bounds=b, max_batch=batch, num_slots=2,
                     seed=case, replace=replace)
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX if replace else O.RNG_PHILOX_NOREPLACE,
                         seed=case, owner=own)
    for bi in range(3):
        nb = batch if bi < 2 else int(rng.integers(1, batch + 1))
        seeds = rng.permutation(n)[:nb]
        views = grp.slice(seeds, bi)
        want = orc.slice(seeds, bi)
        for q, v in enumerate(views):
            _compare_rank(v, want, q, owner=own)
            v.release()
    grp.close()


def test_limits_eight_layers_eight_parts_and_widest_fanout():
    """GS_MAX_LAYERS = 8 layers x GS_MAX_PARTS = 8 parts, and the widest accepted fanout
    (1023) on rows both shorter and longer than it, with and without replacement."""
    from cslicer import engine as E
    from cslicer import _native as N
    from __graft_entry__ import compare_sample
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=2, max_degree=3000)
    n = ip.size - 1
    assert int(np.diff(ip).max()) > 1023
    g = E.Graph(ip, ix)
    for fan, P, batch, replace in [([2] * 8, 8, 64, True), ([1023], 8, 256, True),
                                   ([1023, 2], 3, 128, False)]:
        sl = E.Slicer(g, fan, P, max_batch=batch, num_slots=2, seed=11, replace=replace)
        orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX if replace else O.RNG_PHILOX_NOREPLACE, seed=11)
        hubs = np.argsort(-np.diff(ip))[:batch // 2]          # the longest rows are in the batch
        rest = np.setdiff1d(np.random.default_rng(3).permutation(n)[:batch], hubs)[:batch - hubs.size]
        seeds = np.concatenate([hubs, rest])
        v = sl.slice(seeds, 0)
        compare_sample(v, orc.slice(seeds, 0))
        v.release()
        sl.close()
    for bad in ([1024], [0], [-2], [2] * 9):
        with pytest.raises((N.SlicerError, ValueError, IndexError)):
            E.Slicer(g, bad, 4, max_batch=8)
    with pytest.raises(N.SlicerError):
        E.Slicer(g, [2], 9, max_batch=8)
    g.close()
