"""GPU parity tests proper: the CUDA path, called through the C ABI
(cslicer.engine -> libgsplit_slicer.so), against
  * the golden vectors produced by the unmodified reference (replay of its
    mt19937 draw stream: bit-exact on every field the reference fills),
  * the CPU oracle on the same seeded inputs (bit-exact incl. the repaired CSR)
    across layers / parts / fanouts / owner maps,
  * size-independent properties at BASELINE.json's full products shape.
Integer work: the bar is bit-exact everywhere."""
import json
import os

import numpy as np
import pytest

import refstream
from oracle import oracle as O

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
META = json.load(open(os.path.join(GOLD, "golden.json")))


def _engine():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from cslicer import engine
    return engine


def _compare(view, orc):
    from __graft_entry__ import compare_sample
    compare_sample(view, orc)


@pytest.mark.parametrize("name", sorted(META))
def test_replay_matches_reference_golden(name):
    """Same seed-derived edge choices on both sides -> bit-exact Bipartites
    (north_star).  The draw stream is std::mt19937(5489), state carried over
    batches exactly as one reference Slicer carries it."""
    E = _engine()
    z = np.load(os.path.join(GOLD, f"{name}.npz"))
    m = META[name]
    batches = O.cut_batches(z["perm"], m["minibatch"])
    g = E.Graph(z["indptr"], z["indices"])
    sl = E.Slicer(g, [10, 10, 10], 4, max_batch=m["minibatch"], num_slots=3, num_streams=3,
                  draws=O.mt19937_stream(400000))
    got = []
    for i, b in enumerate(batches):
        v = sl.slice(b, i)
        got.append(refstream.digest_view(v))
        v.release()
    assert got == m["sha256"]
    sl.close()
    g.close()


def test_config1_two_layers_is_prefix_of_reference_three_layers():
    """BASELINE config 1 (2 layers, fanout [10,10], 4 parts): the reference can
    only run 3 layers; its first two layers are the 2-layer answer."""
    E = _engine()
    z = np.load(os.path.join(GOLD, "mixed.npz"))
    batches = O.cut_batches(z["perm"], 256)
    g = E.Graph(z["indptr"], z["indices"])
    orc = O.OracleSlicer(z["indptr"], z["indices"], [10, 10, 10], 4, O.RNG_PHILOX, seed=9)
    sl = E.Slicer(g, [10, 10], 4, max_batch=256, num_slots=2, seed=9)
    for i, b in enumerate(batches[:3]):
        v = sl.slice(b, i)
        assert refstream.digest_view(v) == refstream.digest_oracle(orc.slice(b, i), layers=2)
        v.release()
    sl.close()
    g.close()


CASES = [
    # (fanouts, parts, owner, batch, graph)
    ([10, 10, 10], 4, "mod", 256, "tiny"),
    ([15, 10, 5], 8, "mod", 512, "small"),
    ([15, 10, 5], 1, "mod", 300, "small"),
    ([15, 10, 5], 2, "range", 300, "small"),
    ([25, 10], 4, "range", 1024, "small"),
    ([-1, -1], 4, "mod", 128, "small"),
    ([-1, 5, -1], 3, "random", 64, "tiny"),
    ([1], 8, "random", 1000, "tiny"),
    ([3, 3, 3, 3, 3], 5, "mod", 17, "tiny"),
]


@pytest.mark.parametrize("fanouts,P,owner,batch,shape", CASES)
def test_philox_matches_oracle(fanouts, P, owner, batch, shape):
    E = _engine()
    from cslicer.synth import owner_range, synth_graph
    ip, ix = synth_graph(shape, seed=21)
    n = ip.size - 1
    own = None
    if owner == "range":
        own = owner_range(n, P)
    elif owner == "random":
        own = np.random.default_rng(4).integers(0, P, n).astype(np.int32)
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fanouts, P, max_batch=batch, num_slots=4, num_streams=4, seed=77, owner=own)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX, seed=77, owner=own)
    rng = np.random.default_rng(8)
    pending = []
    for b in range(6):  # keep several minibatches in flight on different streams
        seeds = rng.permutation(n)[:batch if b != 5 else max(1, batch // 3)]  # ragged last batch
        pending.append((sl.submit(seeds, b), seeds, b))
        if len(pending) == 4:
            slot, s, bi = pending.pop(0)
            v = sl.wait(slot)
            _compare(v, orc.slice(s, bi))
            v.release()
    for slot, s, bi in pending:
        v = sl.wait(slot)
        _compare(v, orc.slice(s, bi))
        v.release()
    sl.close()
    g.close()


@pytest.mark.parametrize("cand_capacity,what", [(1 << 26, "32 tags: the sequence wraps every 8 samples"),
                                                (1 << 28, "8 tags < 4 (L + 1): untagged table, reset every sample"),
                                                (0, "default capacity")])
def test_tagged_dedup_table_wraps_and_matches_oracle(cand_capacity, what):
    """The dedup table keeps (tag, position) per node: later layers / samples use smaller
    tags, so nothing is reset between them and the table is refilled only when a slot runs
    out of tags.  40 samples on ONE slot cross several refills; duplicate seeds (an error
    that leaves claims behind) in the middle must not leak into the next sample."""
    E = _engine()
    from cslicer import _native as N
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=13)
    n = ip.size - 1
    fan = [10, 5, 5]
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fan, 4, max_batch=128, num_slots=1, num_streams=1, seed=3, cand_capacity=cand_capacity)
    orc = O.OracleSlicer(ip, ix, fan, 4, O.RNG_PHILOX, seed=3)
    rng = np.random.default_rng(17)
    for b in range(40):
        seeds = rng.permutation(n)[:128 if b % 7 else 33]
        if b in (11, 30):
            bad = seeds.copy()
            bad[5] = bad[50 % bad.size]
            with pytest.raises(N.SlicerError):
                sl.slice(bad, b)
        v = sl.slice(seeds, b)
        _compare(v, orc.slice(seeds, b))
        v.release()
    sl.close()
    g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("fanouts,P,shape,batch", [([15, 10, 5], 4, "small", 512), ([10, 10, 10], 8, "tiny", 256),
                                                   ([1, 2, 40], 2, "small", 100), ([-1, 7], 3, "tiny", 64)])
def test_without_replacement_matches_oracle(fanouts, P, shape, batch):
    """GS_RNG_PHILOX_NOREPLACE (python/utils/sampler.py:187 samples
    distinct neighbours): bit-exact against the oracle's restatement of the keyed
    permutation, and the picks of every sampled row are distinct."""
    E = _engine()
    from cslicer.synth import synth_graph
    ip, ix = synth_graph(shape, seed=21)
    n = ip.size - 1
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fanouts, P, max_batch=batch, num_slots=2, num_streams=2, seed=3, replace=False)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX_NOREPLACE, seed=3)
    rng = np.random.default_rng(8)
    for b in range(4):
        seeds = rng.permutation(n)[:batch]
        v = sl.slice(seeds, b)
        _compare(v, orc.slice(seeds, b))
        # distinct picks: per destination, the sources over all parts have no repeats
        for l in range(len(fanouts)):
            pairs = []
            for q in range(P):
                ind, idx = v.numpy(l, q, "csr_indptr"), v.numpy(l, q, "csr_indices")
                src = v.numpy(l, q, "in_nodes")[idx]
                dst = np.repeat(v.numpy(l, q, "out_nodes"), np.diff(ind))
                pairs.append(np.stack([dst, src], 1))
            pairs = np.concatenate(pairs)
            assert np.unique(pairs, axis=0).shape[0] == pairs.shape[0]
        v.release()
    sl.close()
    g.close()
    with pytest.raises(ValueError):
        E.Slicer(E.Graph(ip, ix), fanouts, P, draws=np.zeros(4, np.uint32), replace=False)


@pytest.mark.parametrize("group,slots,streams", [(2, 4, 2), (4, 4, 1), (3, 6, 2), (4, 8, 2)])
def test_groups_of_minibatches_per_launch_match_oracle(group, slots, streams):
    """gs_config.group_size: `group` consecutive slots share one launch sequence
    (blockIdx.y = member).  Full groups replay the CUDA graph, a partly filled group
    (a wait on a staged slot, flush) is launched eagerly with fewer members; ragged
    and empty batches inside a group; slots are refilled as they are released."""
    E = _engine()
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=21)
    n = ip.size - 1
    fan, P, batch = [15, 10, 5], 4, 384
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fan, P, max_batch=batch, num_slots=slots, num_streams=streams, seed=5, group_size=group)
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=5)
    rng = np.random.default_rng(8)
    sizes = [batch, batch // 3, 0, batch, 1, batch, 77, batch, batch, 5, batch]
    pending = []
    for b, nb in enumerate(sizes):
        seeds = rng.permutation(n)[:nb]
        if len(pending) == slots:
            slot, s_, bi = pending.pop(0)
            v = sl.wait(slot)
            _compare(v, orc.slice(s_, bi))
            v.release()
        pending.append((sl.submit(seeds, b), seeds, b))
    sl.flush()
    for slot, s_, bi in pending:  # the last group is partly filled: wait launches it
        v = sl.wait(slot)
        _compare(v, orc.slice(s_, bi))
        v.release()
    # a lone submit followed by wait: a group of one member out of `group`
    seeds = rng.permutation(n)[:batch]
    v = sl.slice(seeds, 99)
    _compare(v, orc.slice(seeds, 99))
    v.release()
    sl.close(); g.close()


def test_malformed_csr_is_rejected_at_create():
    """gs_graph_create checks the CSR on the device before narrowing the ids: a
    bad indices.bin must be an error, not an out-of-bounds atomic later."""
    E = _engine()
    from cslicer import _native as N
    ip = np.array([0, 2, 3, 3], np.int64)
    for bad_ix, what in (([1, 2, 3], b"indices[2]"), ([1, -1, 0], b"indices[1]"), ([1 << 32, 0, 1], b"indices[0]")):
        with pytest.raises(N.SlicerError) as e:
            E.Graph(ip, np.array(bad_ix, np.int64))
        assert e.value.code == N.ERR_INVALID and what in str(e.value).encode()
    with pytest.raises(N.SlicerError) as e:  # not monotone
        E.Graph(np.array([0, 3, 2, 3], np.int64), np.array([1, 2, 0], np.int64))
    assert e.value.code == N.ERR_INVALID and "indptr" in str(e.value)
    g = E.Graph(ip, np.array([1, 2, 0], np.int64))  # the well-formed version still loads
    g.close()


def test_graph_and_eager_paths_agree_and_device_views_match_host():
    E = _engine()
    import torch
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=5)
    g = E.Graph(ip, ix)
    a = E.Slicer(g, [15, 10, 5], 4, max_batch=512, num_slots=2, seed=3, use_graph=True)
    b = E.Slicer(g, [15, 10, 5], 4, max_batch=512, num_slots=2, seed=3, use_graph=False)
    seeds = np.random.default_rng(0).permutation(ip.size - 1)[:512]
    va, vb = a.slice(seeds, 0), b.slice(seeds, 0)
    assert refstream.digest_view(va) == refstream.digest_view(vb)
    for l in range(3):
        for p in range(4):
            for name in ("in_nodes", "csr_indices", "to_ids"):
                assert np.array_equal(va.tensor(l, p, name).cpu().numpy(), va.numpy(l, p, name))
    assert torch.equal(va.frontier_tensor(3).cpu(), torch.from_numpy(va.frontier_numpy(3).copy()))
    va.release(); vb.release()
    # host mirror: copy engine (default; size estimated from earlier samples, remainder topped
    # up in wait) against stores from the last kernel -- growing, shrinking and growing batches
    c = E.Slicer(g, [15, 10, 5], 4, max_batch=512, num_slots=3, seed=3, host_copy_engine=False)
    for bi, nb in enumerate([64, 512, 8, 300, 512, 1]):
        sb = np.random.default_rng(bi).permutation(ip.size - 1)[:nb]
        v1, v2 = a.slice(sb, bi), c.slice(sb, bi)
        assert refstream.digest_view(v1) == refstream.digest_view(v2)
        for name in ("out_nodes", "csr_indices", "self_ids_in"):
            assert np.array_equal(v1.tensor(2, 3, name).cpu().numpy(), v1.numpy(2, 3, name))
        v1.release(); v2.release()
    c.close()
    prof = b.profile(seeds, 0)
    assert len(prof) == b.launches_per_sample and all(ms >= 0 for _, ms in prof)
    a.close(); b.close(); g.close()


def test_edge_cases_empty_isolated_and_errors():
    E = _engine()
    from cslicer import _native as N
    # 6 nodes: 0->1,2 ; 1->0 ; 2 isolated-out ; 3 self-contained pair with 4 ; 5 isolated
    ip = np.array([0, 2, 3, 3, 4, 5, 5], np.int64)
    ix = np.array([1, 2, 0, 4, 3], np.int64)
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, [-1, -1], 2, max_batch=6, num_slots=2)
    orc = O.OracleSlicer(ip, ix, [-1, -1], 2, O.RNG_PHILOX)
    for seeds in ([5], [2, 5], [0, 1, 2, 3, 4, 5], [3], []):
        v = sl.slice(np.array(seeds, np.int64), 0)
        _compare(v, orc.slice(np.array(seeds, np.int64), 0))
        v.release()
    with pytest.raises(N.SlicerError) as e:  # out of range
        sl.slice(np.array([0, 6], np.int64), 0)
    assert e.value.code == N.ERR_SEED
    with pytest.raises(N.SlicerError) as e:  # duplicate seed
        sl.slice(np.array([1, 1], np.int64), 0)
    assert e.value.code == N.ERR_SEED
    with pytest.raises(N.SlicerError) as e:  # batch too large
        sl.slice(np.arange(6).repeat(2), 0)
    assert e.value.code == N.ERR_INVALID
    v = sl.slice(np.array([0, 3], np.int64), 1)  # slicer still healthy after errors
    _compare(v, orc.slice(np.array([0, 3], np.int64), 1))
    v.release()
    s1, s2 = sl.submit([0], 0), sl.submit([1], 1)
    with pytest.raises(N.SlicerError) as e:
        sl.submit([2], 2)
    assert e.value.code == N.ERR_BUSY
    sl.release(s1); sl.release(s2)
    # workspace overflow is reported, and the slot recovers
    ip2, ix2 = __import__("cslicer.synth", fromlist=["x"]).synth_graph("tiny", seed=1)
    g2 = E.Graph(ip2, ix2)
    small = E.Slicer(g2, [10, 10], 4, max_batch=64, num_slots=1, cand_capacity=150)
    with pytest.raises(N.SlicerError) as e:
        small.slice(np.arange(64), 0)
    assert e.value.code == N.ERR_OVERFLOW
    orc2 = O.OracleSlicer(ip2, ix2, [10, 10], 4, O.RNG_PHILOX)
    v = small.slice(np.arange(1), 0)  # <= 11 then <= 121 candidates: fits
    _compare(v, orc2.slice(np.arange(1), 0))
    v.release()
    small.close(); g2.close(); sl.close(); g.close()


def test_statistical_equivalence_of_native_sampling():
    """Philox picks vs the reference's mt19937 picks: both uniform over a row's
    neighbours with replacement (slicer.cpp:15-20).  Chi-square on a hub row."""
    E = _engine()
    deg = 40
    ip = np.array([0, deg] + [deg] * deg, np.int64)
    ix = np.arange(1, deg + 1, dtype=np.int64)
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, [10], 1, max_batch=1, num_slots=2, seed=123)
    counts = np.zeros(deg + 1)
    trials = 4000
    for b in range(trials):
        v = sl.slice([0], b)
        src = v.numpy(0, 0, "in_nodes")[v.numpy(0, 0, "csr_indices")]
        assert src.size == 10
        np.add.at(counts, src, 1)
        v.release()
    exp = trials * 10 / deg
    chi2 = ((counts[1:] - exp) ** 2 / exp).sum()
    assert counts[0] == 0 and chi2 < 80.0  # 39 dof: p(chi2 > 80) ~ 1e-4
    sl.close(); g.close()


def test_full_size_products_properties():
    """BASELINE config 2 at full size (2.4 M nodes / ~62 M edges, [15,10,5],
    batch 4096): properties that need no oracle run, plus an oracle check of one
    minibatch (the sequential oracle takes a few seconds here)."""
    E = _engine()
    import torch
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("products", seed=1234, device="cuda")
    n = ip.size - 1
    g = E.Graph(ip, ix)
    P = 4
    sl = E.Slicer(g, [15, 10, 5], P, max_batch=4096, num_slots=2, seed=1)
    seeds = np.random.default_rng(1).permutation(n)[:4096]
    v = sl.slice(seeds, 0)
    deg = np.diff(ip)
    tot = 0
    for l in range(3):
        fr, nx = v.frontier_numpy(l), v.frontier_numpy(l + 1)
        assert np.unique(nx).size == nx.size and set(fr.tolist()) <= set(nx.tolist())
        fan = [15, 10, 5][l]
        want_edges = np.minimum(deg[fr], np.where(deg[fr] >= fan, fan, deg[fr])).sum()
        got = 0
        union_in = []
        for p in range(P):
            ins, outs = v.numpy(l, p, "in_nodes"), v.numpy(l, p, "out_nodes")
            cp, ci = v.numpy(l, p, "csr_indptr"), v.numpy(l, p, "csr_indices")
            assert (ins % P == p).all() and np.unique(ins).size == ins.size
            assert cp[0] == 0 and cp[-1] == ci.size and (np.diff(cp) > 0).all()
            assert ci.min() >= 0 and ci.max() < ins.size
            own, frm = v.numpy(l, p, "owned_out_nodes"), v.numpy(l, p, "from_ids")
            assert (outs[own] % P == p).all() and (outs[frm] % P != p).all()
            assert own.size + frm.size == outs.size
            si, so = v.numpy(l, p, "self_ids_in"), v.numpy(l, p, "self_ids_out")
            assert si.size == (fr % P == p).sum() == so.size
            ok = so >= 0
            assert np.array_equal(outs[so[ok]], fr[fr % P == p][ok])
            got += ci.size
            union_in.append(ins)
        assert got == want_edges  # every sampled edge lands in exactly one part
        assert set(np.concatenate(union_in).tolist()) | set(fr.tolist()) == set(nx.tolist())
        tot += got
    assert tot == v.edges
    orc = O.OracleSlicer(ip, ix, [15, 10, 5], P, O.RNG_PHILOX, seed=1)
    _compare(v, orc.slice(seeds, 0))
    v.release()
    sl.close(); g.close()
    torch.cuda.empty_cache()


def _full_size_case(shape, fanouts, P, batch, owner="mod"):
    E = _engine()
    import torch
    from cslicer.synth import owner_range, synth_graph
    ip, ix = synth_graph(shape, seed=1234, device="cuda")
    n = ip.size - 1
    own = owner_range(n, P) if owner == "range" else None
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fanouts, P, max_batch=batch, num_slots=2, num_streams=2, seed=11, owner=own)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX, seed=11, owner=own)
    rng = np.random.default_rng(3)
    for b in range(2):
        seeds = rng.permutation(n)[:batch]
        v = sl.slice(seeds, b)
        _compare(v, orc.slice(seeds, b))
        # idempotence: the same batch index gives the same sample again, on another slot
        v2 = sl.slice(seeds, b)
        assert refstream.digest_view(v) == refstream.digest_view(v2)
        v.release(); v2.release()
    sl.close(); g.close()
    torch.cuda.empty_cache()


def test_full_size_reddit_shaped_dedup_stress():
    """BASELINE config 3: 233 K nodes / ~115 M edges (avg degree ~490), 2-layer
    fanout [25,10] -- every row draws, the frontier saturates towards all nodes."""
    _full_size_case("reddit", [25, 10], 4, 4096)


def test_full_size_full_neighbour_products():
    """BASELINE config 4's slicing half: fanout -1 on the products shape, 2 layers;
    bit-exact Bipartites (no rng on either side)."""
    _full_size_case("products", [-1, -1], 4, 1024, owner="range")


@pytest.mark.parametrize("dim,P,fanouts", [(602, 4, [-1, -1]), (128, 8, [15, 10, 5]), (7, 3, [5, 5])])
def test_fused_feature_gather_matches_numpy(dim, P, fanouts):
    """BASELINE config 4's gather half: rows of the last layer's in_nodes_g, fused
    into the sample's graph; fp32 copy, so equality is exact (tolerance 0)."""
    E = _engine()
    import torch
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=12)
    n = ip.size - 1
    feat = np.random.default_rng(4321).random((n, dim), dtype=np.float32)
    g = E.Graph(ip, ix)
    F = E.Features(feat)
    sl = E.Slicer(g, fanouts, P, max_batch=256, num_slots=2, seed=5, features=F)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX, seed=5)
    seeds = np.random.default_rng(2).permutation(n)[:256]
    v = sl.slice(seeds, 0)
    r = orc.slice(seeds, 0)
    _compare(v, r)
    tot = 0
    for p in range(P):
        want = feat[r["layers"][-1][p]["in_nodes"]]
        got = v.features(p).cpu().numpy()
        assert got.shape == want.shape and np.array_equal(got, want)
        tot += want.shape[0]
    assert v.feature_rows == tot
    # standalone gather on arbitrary ids, and a sharded store (owner = id % P: row id // P)
    ids = torch.from_numpy(np.random.default_rng(1).integers(0, n, 5000).astype(np.int32)).cuda()
    assert np.array_equal(F.gather(ids).cpu().numpy(), feat[ids.cpu().numpy()])
    shard = E.Features(feat[1::P], id_div=P)  # nodes with id % P == 1
    mine = ids[ids % P == 1]
    assert np.array_equal(shard.gather(mine).cpu().numpy(), feat[mine.cpu().numpy()])
    v.release()
    small = E.Slicer(g, fanouts, P, max_batch=256, num_slots=1, seed=5, features=F, feature_rows_capacity=8)
    from cslicer import _native as N
    with pytest.raises(N.SlicerError) as e:
        small.slice(seeds, 0)
    assert e.value.code == N.ERR_OVERFLOW
    small.close(); sl.close(); shard.close(); F.close(); g.close()


@pytest.mark.parametrize("fraction,dim", [(0.1, 602), (0.4, 128), (1.0, 7)])
def test_cached_feature_store_matches_numpy(fraction, dim):
    """table in host memory, degree-ordered HBM cache in front
    (python/utils/memory_manager.py:40-104).  A gather returns the same rows
    whether a row is served from HBM or from host memory: exact equality."""
    E = _engine()
    import torch
    from cslicer import _native as N
    from cslicer.feature_cache import degree_ordered_cache, hit_rate
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=12)
    n, P, fanouts = ip.size - 1, 4, [10, 5]
    feat = np.random.default_rng(4321).random((n, dim), dtype=np.float32)
    cache = degree_ordered_cache(ip, fraction, part=1, num_parts=P)
    assert cache.size == int(fraction * n) and np.unique(cache).size == cache.size
    F = E.Features(feat, cache_rows=cache)
    assert F.rows_in_hbm == cache.size
    ids = np.random.default_rng(1).integers(0, n, 20000).astype(np.int32)
    assert np.array_equal(F.gather(torch.from_numpy(ids).cuda()).cpu().numpy(), feat[ids])
    hr = hit_rate(cache, ids, n)
    assert (hr == 1.0) if fraction == 1.0 else (0.0 < hr < 1.0)
    # fused into the slicer: the block every part reads
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, fanouts, P, max_batch=256, num_slots=2, seed=5, features=F)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX, seed=5)
    for b in range(2):
        seeds = np.random.default_rng(b).permutation(n)[:256]
        v = sl.slice(seeds, b)
        r = orc.slice(seeds, b)
        _compare(v, r)
        for p in range(P):
            assert np.array_equal(v.features(p).cpu().numpy(), feat[r["layers"][-1][p]["in_nodes"]])
        v.release()
    sl.close()
    # nothing cached: every row comes from host memory
    F0 = E.Features(feat, cache_rows=np.zeros(0, np.int64))
    assert np.array_equal(F0.gather(torch.from_numpy(ids[:3000]).cuda()).cpu().numpy(), feat[ids[:3000]])
    F0.close()
    for bad in ([0, 0], [n], [-1]):
        with pytest.raises(N.SlicerError) as e:
            E.Features(feat, cache_rows=np.array(bad))
        assert e.value.code == N.ERR_INVALID
    with pytest.raises(ValueError):
        E.Features(feat.astype(np.float64), cache_rows=cache)
    F.close(); g.close()


def test_reference_facing_module_end_to_end(tmp_path):
    """cslicer/test_py.py's loop, verbatim in shape, against the golden hashes:
    the module's own epoch shuffle + replayed draws reproduce the reference's
    whole sample stream through the reference-named attributes only."""
    import cslicer as M
    z = np.load(os.path.join(GOLD, "hubs.npz"))
    m = META["hubs"]
    M.write_dataset(str(tmp_path / "hubs"), z["indptr"], z["indices"])
    csl = M.cslicer("hubs", 16, 32, 1, m["minibatch"], root=str(tmp_path),
                    draws=O.mt19937_stream(400000))
    ns = csl.getNoSamples()
    assert ns == len(m["sha256"])
    got = []
    for i in range(ns):
        s = csl.getSample()
        got.append(refstream.digest_module_sample(s))
    assert got == m["sha256"]
    with pytest.raises(RuntimeError):
        csl.getSample()
    csl.close()


def test_bipatites_that_outlive_their_sample_keep_the_loop_running(tmp_path):
    """A caller may keep one bipatite of every sample and drop the sample itself (the
    reference returns deep copies): the slot stays pinned by the view, and getSample()
    must make room by detaching the oldest view instead of failing with GS_ERR_BUSY."""
    import gc
    import cslicer as M
    z = np.load(os.path.join(GOLD, "hubs.npz"))
    m = META["hubs"]
    M.write_dataset(str(tmp_path / "hubs"), z["indptr"], z["indices"])
    csl = M.cslicer("hubs", 16, 32, 1, m["minibatch"], root=str(tmp_path), draws=O.mt19937_stream(400000))
    kept, digests = [], []
    for i in range(csl.getNoSamples()):
        s = csl.getSample()
        digests.append(refstream.digest_module_sample(s))
        kept.append(s.layers[0][0])  # keeps the SampleView alive
        del s
        gc.collect()
    assert digests == m["sha256"]
    assert all(isinstance(b.in_nodes, list) for b in kept)  # still readable (detached copies)
    csl.close()


def test_bench_byte_model_names_every_launch_of_the_chain():
    """bench.py's per-kernel roofline divides each launch's algorithmic bytes by its time:
    the byte model must know every launch gs_slicer_profile reports (a renamed launch once
    left the dominant kernel with 0 bytes)."""
    import sys
    E = _engine()
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=5)
    g = E.Graph(ip, ix)
    sl = E.Slicer(g, [15, 10, 5], 4, max_batch=256, num_slots=1, num_streams=1, seed=1, use_graph=False)
    seeds = np.random.default_rng(2).permutation(ip.size - 1)[:256]
    names = [k for k, _ in bench.kernel_profile(sl, [(0, seeds)], reps=1)]
    assert len(names) == sl.launches_per_sample == 13
    v = sl.slice(seeds, 0)
    model = bench.kernel_bytes(v, 4)
    assert sorted(model) == sorted(names), (sorted(model), sorted(names))
    assert all(b > 0 for b in model.values())
    v.release(); sl.close(); g.close()


def test_nvtx_ranges_can_be_switched_on():
    """GSPLIT_NVTX=1 (tracing): the C ABI wraps submit / wait in NVTX
    ranges; results and exit status are unchanged (run in a child: the switch is read once)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, GSPLIT_NVTX="1")
    r = subprocess.run([sys.executable, "-c", "import __graft_entry__ as g; g.smoke()"], cwd=root, env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "smoke ok" in r.stdout
