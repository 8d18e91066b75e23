"""CPU tests of the checker itself: the oracle (CPU restatement) against the
golden vectors produced by the unmodified reference, against the compiled
reference when it is present, and known-answer tests for its generators."""
import json
import os

import numpy as np
import pytest

import refstream
from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")
META = json.load(open(os.path.join(GOLD, "golden.json")))


def load_case(name):
    z = np.load(os.path.join(GOLD, f"{name}.npz"))
    m = META[name]
    return z["indptr"], z["indices"], O.cut_batches(z["perm"], m["minibatch"]), m


@pytest.mark.parametrize("name", sorted(META))
def test_oracle_matches_reference_golden(name):
    ip, ix, batches, m = load_case(name)
    orc = O.OracleSlicer(ip, ix, [10, 10, 10], 4, O.RNG_MT19937)  # one engine, state carried
    got = [refstream.digest_oracle(orc.slice(b, i)) for i, b in enumerate(batches)]
    assert got == m["sha256"]


@pytest.mark.parametrize("name", sorted(META))
def test_oracle_draw_stream_mode_equals_mt19937(name):
    """Feeding the mt19937 stream as explicit draws (the replay mode the CUDA
    path uses) is the same computation."""
    ip, ix, batches, m = load_case(name)
    a = O.OracleSlicer(ip, ix, [10, 10, 10], 4, O.RNG_MT19937)
    b = O.OracleSlicer(ip, ix, [10, 10, 10], 4, O.RNG_DRAWS, draws=O.mt19937_stream(400000))
    for i, batch in enumerate(batches[:3]):
        assert refstream.digest_oracle(a.slice(batch, i)) == refstream.digest_oracle(b.slice(batch, i))
    assert a.draws_used == b.draws_used


def test_golden_permutation_is_reference_shuffle():
    z = np.load(os.path.join(GOLD, "mixed.npz"))
    assert np.array_equal(O.epoch_permutations(2000, 1)[0], z["perm"])


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (no /root/reference)")
def test_oracle_vs_compiled_reference_fresh_graph(tmp_path):
    from cslicer.dataset import write_dataset
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=99, num_nodes=1500, num_edges=20000)
    write_dataset(str(tmp_path), ip, ix)
    perm = O.ref_epoch_permutations(1500, 1)[0]
    assert np.array_equal(perm, O.epoch_permutations(1500, 1)[0])
    batches = O.cut_batches(perm, 200)
    ref = O.ref_slice(str(tmp_path), batches)
    orc = O.OracleSlicer(ip, ix, [10, 10, 10], 4, O.RNG_MT19937)
    for i, b in enumerate(batches):
        assert refstream.digest_reference(ref[i]) == refstream.digest_oracle(orc.slice(b, i))


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    kat = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
            [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for ctr, key, want in kat:
        assert list(O.philox4x32_10(ctr, key)) == want


def test_noreplace_pick_is_a_keyed_permutation():
    """ORC_RNG_PHILOX_NOREPLACE (python/utils/sampler.py:187): over
    j = 0..deg-1 the picks of a row are a permutation of its neighbours, so any
    `fanout` of them are distinct; different nodes / batches / layers permute
    differently; every neighbour is equally likely to be pick 0."""
    for deg in [1, 2, 3, 4, 5, 7, 8, 9, 16, 17, 31, 33, 100, 1000, 1025]:
        picks = [O.noreplace_pick(5, 2, 1, 40 + deg, j, deg) for j in range(deg)]
        assert sorted(picks) == list(range(deg))
    a = [O.noreplace_pick(5, 2, 1, 7, j, 50) for j in range(50)]
    assert a != [O.noreplace_pick(5, 2, 1, 8, j, 50) for j in range(50)]
    assert a != [O.noreplace_pick(5, 3, 1, 7, j, 50) for j in range(50)]
    assert a != [O.noreplace_pick(5, 2, 0, 7, j, 50) for j in range(50)]
    assert a != [O.noreplace_pick(6, 2, 1, 7, j, 50) for j in range(50)]
    # first pick uniform, (pick 0, pick 1) uniform over ordered pairs: Fisher-Yates
    # branch (deg <= 16), 12-round and 8-round Feistel branches
    for deg, n, lim1, lim2 in [(13, 26000, 40.0, 240.0), (20, 30000, 52.0, 500.0), (40, 40000, 82.0, 1800.0)]:
        p0 = np.array([O.noreplace_pick(1, 0, 0, v, 0, deg) for v in range(n)])
        p1 = np.array([O.noreplace_pick(1, 0, 0, v, 1, deg) for v in range(n)])
        first = np.bincount(p0, minlength=deg)
        pair = np.bincount(p0 * deg + p1, minlength=deg * deg).reshape(deg, deg)
        exp = n / deg
        assert ((first - exp) ** 2 / exp).sum() < lim1      # chi-square, deg-1 dof, p ~ 1e-4
        assert pair.trace() == 0
        off = pair[~np.eye(deg, dtype=bool)]
        expp = n / (deg * (deg - 1))
        assert ((off - expp) ** 2 / expp).sum() < lim2      # deg*(deg-1)-1 dof, p < 1e-4


def test_noreplace_oracle_samples_distinct_neighbours():
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    fan = [10, 5]
    orc = O.OracleSlicer(ip, ix, fan, 1, O.RNG_PHILOX_NOREPLACE, seed=4)
    seeds = np.random.default_rng(2).permutation(ip.size - 1)[:200]
    r = orc.slice(seeds, 0)
    d = r["layers"][0][0]
    # P = 1: out_nodes = the frontier, in_nodes[csr_indices] = self + sampled sources per row
    src = d["in_nodes"][d["csr_indices"]]
    for i, nd in enumerate(d["out_nodes"]):
        row = src[d["csr_indptr"][i]:d["csr_indptr"][i + 1]]
        nb = ix[ip[nd]:ip[nd + 1]]
        assert set(row.tolist()) <= set(nb.tolist())
        if nb.size >= fan[0]:  # sampled row: exactly `fanout` picks, all distinct (no duplicate edges in the graph)
            assert len(row) == fan[0] and len(set(row.tolist())) == fan[0]
        else:
            assert np.array_equal(row, nb)


def test_mt19937_matches_numpy_legacy_seeding():
    m = np.random.MT19937()
    m._legacy_seeding(5489)
    assert np.array_equal(m.random_raw(3000).astype(np.uint32), O.mt19937_stream(3000))
    assert O.mt19937_stream(10000)[9999] == 4123659995  # the C++ standard's check value


def test_oracle_structure_invariants():
    """Properties the consumer relies on (python/data/bipartite.py): every list
    indexes inside its target, clean CSR is consistent, edges are conserved."""
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    for P, fan in [(4, [10, 10, 10]), (8, [15, 10, 5]), (1, [5, 5]), (3, [-1, -1])]:
        orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=1)
        seeds = np.random.default_rng(1).permutation(2000)[:128]
        r = orc.slice(seeds, 0)
        assert np.array_equal(r["frontiers"][0], seeds)
        total = 0
        for l, layer in enumerate(r["layers"]):
            nxt = set(r["frontiers"][l + 1].tolist())
            assert set(r["frontiers"][l].tolist()) <= nxt
            for g, d in enumerate(layer):
                n_in, n_out = d["in_nodes"].size, d["out_nodes"].size
                assert len(set(d["in_nodes"].tolist())) == n_in
                assert d["csr_indptr"].size == n_out + 1 and d["csr_indptr"][-1] == d["csr_indices"].size
                assert (np.diff(d["csr_indptr"]) > 0).all()
                if d["csr_indices"].size:
                    assert d["csr_indices"].min() >= 0 and d["csr_indices"].max() < n_in
                for k in ("owned_out_nodes", "from_ids"):
                    assert ((d[k] >= 0) & (d[k] < n_out)).all()
                assert d["owned_out_nodes"].size + d["from_ids"].size == n_out
                for k in ("to_ids", "self_ids_out"):
                    assert ((d[k] >= -1) & (d[k] < n_out)).all()
                assert ((d["self_ids_in"] >= -1) & (d["self_ids_in"] < n_in)).all()
                assert d["self_ids_in"].size == d["self_ids_out"].size
                total += d["csr_indices"].size
        assert total == r["edges"]


def test_pool_run_counts_edges():
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    perm = O.epoch_permutations(2000, 1)[0]
    batches = O.cut_batches(perm, 500)
    sec, edges = O.pool_run(ip, ix, [10, 10], 4, batches, workers=2, rng_mode=O.RNG_PHILOX, seed=5)
    orc = O.OracleSlicer(ip, ix, [10, 10], 4, O.RNG_PHILOX, seed=5)
    want = sum(orc.slice(b, i)["edges"] for i, b in enumerate(batches))
    assert edges == want and sec > 0
