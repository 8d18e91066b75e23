"""SURVEY 8(f) N3 -- the split-parallel shuffle's oracle (oracle/shuffle_oracle.py), pinned
to the thread-for-thread statement of the reference's mergeKernel
(src/gnn/dist_sage.cu:193-199; merge order :120-135) and to the committed fixture
tests/golden/shuffle.npz (tests/golden/make_shuffle_golden.py).  CPU only."""
import os

import numpy as np

from oracle import oracle as O
from oracle import shuffle_oracle as S

GOLD = os.path.join(os.path.dirname(__file__), "golden", "shuffle.npz")


def _load():
    z = np.load(GOLD)
    P = int(z["P"])
    parts = [z[f"part_{g}"] for g in range(P)]
    fr = [[z[f"from_{g}_{q}"] for q in range(P)] for g in range(P)]
    to = [[z[f"to_{g}_{q}"] for q in range(P)] for g in range(P)]
    return z, P, parts, fr, to


def test_vectorised_oracle_equals_the_kernel_statement_and_the_golden():
    z, P, parts, fr, to = _load()
    slow = S.shuffle_forward(parts, fr, to, kernel=S.merge_kernel)   # one "block" per row, one "thread" per feature
    fast = S.shuffle_forward(parts, fr, to)
    bwd = S.shuffle_backward(parts, fr, to)
    for g in range(P):
        assert np.array_equal(slow[g], z[f"fwd_{g}"])      # tolerance 0: fp32 adds in the reference's order
        assert np.array_equal(fast[g], z[f"fwd_{g}"])
        assert np.array_equal(bwd[g], z[f"bwd_{g}"])


def test_fixture_is_rebuilt_bit_for_bit_by_its_script():
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "make_shuffle_golden", os.path.join(os.path.dirname(GOLD), "make_shuffle_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    z, again = np.load(GOLD), m.build()
    assert sorted(z.files) == sorted(again)
    for k in z.files:
        assert np.array_equal(z[k], again[k]), k


def test_merge_order_matters_and_is_ascending_source_rank():
    """fp32 addition is not associative: the oracle must add peers in ascending src order
    (dist_sage.cu:127) -- a different order gives different bits on a crafted row."""
    big, small = np.float32(1e8), np.float32(3.0)
    parts = [np.array([[0.0]], np.float32), np.array([[big]], np.float32), np.array([[-big]], np.float32),
             np.array([[small]], np.float32)]
    one = np.array([0], np.int64)
    none = np.zeros(0, np.int64)
    fr = [[none] * 4 for _ in range(4)]
    to = [[none] * 4 for _ in range(4)]
    for src in (1, 2, 3):
        fr[src][0], to[0][src] = one, one
    out = S.shuffle_forward(parts, fr, to)
    assert out[0][0, 0] == np.float32(np.float32(np.float32(0.0 + big) - big) + small) == small
    wrong = np.float32(np.float32(np.float32(0.0 + small) + big) - big)   # src order 3, 1, 2
    assert wrong != small


def test_backward_is_the_transpose_of_forward_on_consumed_rows():
    """<grad_out, forward(x)> = <backward(grad_out), x> when the gradient of rows a GPU
    does not own is zero (only owned rows are consumed: bipartite.py:69-73)."""
    z, P, parts, fr, to = _load()
    rng = np.random.default_rng(1)
    grads = [rng.standard_normal(p.shape).astype(np.float64) for p in parts]
    for g in range(P):                      # zero the gradient of every row that is sent away
        for q in range(P):
            if q != g:
                grads[g][fr[g][q]] = 0.0
    x64 = [p.astype(np.float64) for p in parts]
    fwd = [a.copy() for a in x64]
    for dest in range(P):
        for src in range(P):
            if src != dest:
                keep = to[dest][src] >= 0
                np.add.at(fwd[dest], to[dest][src][keep], x64[src][fr[src][dest]][keep])
    lhs = sum(float((gq * fq).sum()) for gq, fq in zip(grads, fwd))
    back = [a.copy() for a in grads]
    for dest in range(P):
        for src in range(P):
            if src != dest:
                rows = np.zeros((fr[src][dest].size, grads[src].shape[1]))
                keep = to[dest][src] >= 0
                rows[keep] = grads[dest][to[dest][src][keep]]
                back[src][fr[src][dest]] = rows
    rhs = sum(float((b * x).sum()) for b, x in zip(back, x64))
    assert abs(lhs - rhs) <= 1e-9 * max(1.0, abs(lhs))
    ob = S.shuffle_backward([g.astype(np.float32) for g in grads], fr, to)
    for g in range(P):
        assert np.allclose(ob[g], back[g], rtol=0, atol=1e-6)


def test_peer_lists_pair_up_on_a_sliced_sample():
    """On a real oracle sample the per-peer split is consistent with the reference's own
    lists: the union over peers of from_ids[g][q] is part g's from_ids, and every owner
    position in to_ids[q][g] points at the same node the sender's row belongs to."""
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=31)
    P = 4
    orc = O.OracleSlicer(ip, ix, [5, 5], P, O.RNG_PHILOX, seed=3)
    s = orc.slice(np.random.default_rng(5).permutation(ip.size - 1)[:200], 0)
    for l in range(2):
        fr, to = S.peer_lists(s, l, lambda v: v % P)
        for g in range(P):
            mine = np.sort(np.concatenate([fr[g][q] for q in range(P)]))
            assert np.array_equal(mine, np.sort(s["layers"][l][g]["from_ids"]))
            for q in range(P):
                if q == g:
                    continue
                snd = s["layers"][l][g]["out_nodes"][fr[g][q]]
                keep = to[q][g] >= 0
                assert np.array_equal(s["layers"][l][q]["out_nodes"][to[q][g][keep]], snd[keep])
                assert (snd % P == q).all()
