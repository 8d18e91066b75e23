"""Host-side cache policy (python/utils/memory_manager.py:40-74):
hand-computed expectations of the rule, and golden vectors produced by the
reference's own MemoryManager (tests/golden/make_cache_policy_golden.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200"))

from cslicer.feature_cache import degree_ordered_cache, hit_rate  # noqa: E402


def _graph(deg):
    return np.concatenate([[0], np.cumsum(deg)]).astype(np.int64)


def test_small_fraction_takes_top_degree_nodes_of_own_partition():
    deg = np.array([5, 1, 9, 3, 7, 2, 8, 4, 6, 0, 11, 10])  # N = 12, P = 4: part 2 owns 2, 6, 10
    ip = _graph(deg)
    got = degree_ordered_cache(ip, 2 / 12, part=2, num_parts=4)
    assert got.tolist() == [10, 2]           # degrees 11, 9 (then 8)
    got = degree_ordered_cache(ip, 0.25, part=2, num_parts=4)
    assert got.tolist() == [10, 2, 6]        # exactly the partition, by degree
    owner = np.array([0, 0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3])
    got = degree_ordered_cache(ip, 0.25, owner=owner, part=1, num_parts=4)
    assert got.tolist() == [4, 3, 5]


def test_large_fraction_is_partition_plus_top_degree_of_the_rest():
    deg = np.array([5, 1, 9, 3, 7, 2, 8, 4, 6, 0, 11, 10])
    ip = _graph(deg)
    got = degree_ordered_cache(ip, 0.5, part=0, num_parts=4)  # 6 rows: own {0,4,8} + 3 best others
    assert got.tolist() == [0, 4, 8, 10, 11, 2]
    full = degree_ordered_cache(ip, 1.0, part=3, num_parts=4)
    assert sorted(full.tolist()) == list(range(12))


def test_ties_go_to_the_smaller_id_and_fraction_is_validated():
    ip = _graph(np.full(8, 3))
    assert degree_ordered_cache(ip, 0.25, part=1, num_parts=2).tolist() == [1, 3]
    for bad in (0.0, -0.1, 1.5):
        with pytest.raises(ValueError):
            degree_ordered_cache(ip, bad)
    assert hit_rate([1, 3], [1, 1, 2, 3], 8) == 0.75
    assert hit_rate([], [], 8) == 1.0


def test_source_degree_ranking_uses_the_neighbour_array():
    ip = _graph(np.array([2, 2, 2, 0]))
    ix = np.array([3, 3, 3, 0, 3, 1])          # node 3 is the most frequent source, then 0 / 1
    got = degree_ordered_cache(ip, 0.5, part=0, num_parts=1, indices=ix)
    assert got.tolist() == [3, 0]


def test_policy_matches_the_reference_memory_manager_golden():
    """tests/golden/cache_policy.npz = MemoryManager.local_to_global_id[g] computed by the
    UNMODIFIED reference module (distinct degrees, so sort ties cannot matter): same
    nodes in the same slot order for fractions below, at and above 1/4; and what its
    refresh_cache appends per batch is exactly the set this design reads from host."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "cache_policy.npz"))
    deg, part = z["out_degree"], z["partition_map"]
    for frac in (0.05, 0.25, 0.5, 1.0):
        for g in range(4):
            got = degree_ordered_cache(None, frac, owner=part, part=g, num_parts=4, degrees=deg)
            assert np.array_equal(got, z[f"cached_{frac}_{g}"]), (frac, g)
    for frac in (0.05, 0.25):
        last = z[f"last_layer_{frac}"]
        for g in range(4):
            cached = z[f"cached_{frac}_{g}"]
            mine = last[part[last] == g]
            missing = mine[~np.isin(mine, cached)]          # served from host memory by the gather kernel
            if frac >= 0.25:  # memory_manager.py:76-78: no refresh at >= 1/4, even where the partition is larger than the cache
                assert np.array_equal(z[f"refreshed_{frac}_{g}"], cached)
                continue
            assert np.array_equal(z[f"refreshed_{frac}_{g}"], np.concatenate([cached, missing]))
            assert abs(hit_rate(cached, mine, deg.size) - (1.0 - missing.size / max(1, mine.size))) < 1e-12
