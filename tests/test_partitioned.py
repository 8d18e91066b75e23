"""Partitioned mode: rank g of P holds the CSR rows of the nodes it
owns and must produce part g of every layer bit-identically to the oracle (and
so to the single-GPU slicer), plus the replicated frontiers.

CPU part: host-side sharding helpers.  GPU part: P ranks emulated in one process
on one device (`LocalGroup`: same kernels, arenas and barriers as one process
per GPU), and -- when the box has >= 2 GPUs -- real one-process-per-GPU ranks
wired with CUDA IPC through torch.distributed."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))


def test_partition_csr_round_trip():
    from cslicer.partitioned import owner_of, partition_csr, range_bounds
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    n = ip.size - 1
    for mode in ("mod", "range"):
        for P in (1, 2, 3, 8):
            b = range_bounds(n, P)
            shards = partition_csr(ip, ix, P, mode, b)
            own = owner_of(np.arange(n), n, P, mode, b)
            assert sum(s[0].size - 1 for s in shards) == n
            assert sum(s[1].size for s in shards) == ix.size
            for v in list(range(0, n, max(1, n // 50))) + [n - 1]:
                g = int(own[v])
                li = v // P if mode == "mod" else v - b[g]
                lip, lix = shards[g]
                assert np.array_equal(lix[lip[li]:lip[li + 1]], ix[ip[v]:ip[v + 1]])


def test_renumber_by_owner_makes_contiguous_ranges():
    from cslicer.partitioned import renumber_by_owner
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    n = ip.size - 1
    P = 4
    own = np.random.default_rng(1).integers(0, P, n)
    nip, nix, b, new_of_old, old_of_new = renumber_by_owner(ip, ix, own, P)
    assert b[0] == 0 and b[-1] == n and nip[-1] == ix.size
    for q in range(P):
        assert np.all(own[old_of_new[b[q]:b[q + 1]]] == q)
        assert np.all(np.diff(old_of_new[b[q]:b[q + 1]]) > 0)      # relative order kept
    for v in range(0, n, 37):
        w = new_of_old[v]
        assert np.array_equal(nix[nip[w]:nip[w + 1]], new_of_old[ix[ip[v]:ip[v + 1]]])
    assert np.array_equal(old_of_new[new_of_old], np.arange(n))


def _compare_rank(view, orc, g, check_frontiers=True, owner=None):
    """Rank g's sample against part g of the oracle's; the slots [q != g] of
    from_ids / to_ids (empty in the reference, bipartite.h:19-21) carry the per-peer
    split, whose expectation follows from the oracle's node lists: the rows rank q
    owns that appear in out_nodes_g, in frontier order."""
    from cslicer import _native
    L, P = view.num_layers, view.num_parts
    if check_frontiers:
        for l in range(L + 1):
            a, b = view.frontier_numpy(l), orc["frontiers"][l]
            assert a.shape == b.shape and np.array_equal(a, b), f"rank {g}: frontier {l} differs"
    for l in range(L):
        for name in _native.FIELDS:
            a, b = view.numpy(l, g, name), orc["layers"][l][g][name]
            assert a.shape == b.shape and np.array_equal(a, b), \
                f"rank {g} layer {l} field {name} differs: {a[:8]} vs {b[:8]} (len {a.size}/{b.size})"
        fr = orc["frontiers"][l]
        own = (fr % P) if owner is None else np.asarray(owner)[fr]
        pos = [{int(v): i for i, v in enumerate(orc["layers"][l][q]["out_nodes"])} for q in range(P)]
        for q in range(P):
            if q == g:
                continue
            for nm in _native.FIELDS:
                if nm not in ("from_ids", "to_ids"):
                    assert view.field_len(l, q, nm) == 0
            want_from = [pos[g][int(v)] for v, o in zip(fr, own) if o == q and int(v) in pos[g]]
            want_to = [pos[g].get(int(v), -1) for v, o in zip(fr, own) if o == g and int(v) in pos[q]]
            assert view.numpy(l, q, "from_ids").tolist() == want_from, (l, g, q, "from")
            assert view.numpy(l, q, "to_ids").tolist() == want_to, (l, g, q, "to")
    assert view.edges == sum(orc["layers"][l][g]["csr_indices"].size for l in range(L))


def test_shard_generation_equals_rows_of_the_whole_graph():
    """synth_graph(rows=...) replays the generator and keeps one rank's rows: the
    shards of a contiguous-range owner map concatenate to the whole graph."""
    from cslicer.partitioned import partition_csr, range_bounds
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=3, chunk=50000)
    n, P = ip.size - 1, 3
    b = range_bounds(n, P)
    whole = partition_csr(ip, ix, P, "range", b)
    for g in range(P):
        lip, lix, nn, m = synth_graph("small", seed=3, chunk=50000, rows=(b[g], b[g + 1]))
        assert (nn, m) == (n, ix.size)
        assert np.array_equal(lip, whole[g][0]) and np.array_equal(lix, whole[g][1])


CASES = [
    # (fanouts, parts, owner, batch, graph)
    ([10, 10, 10], 4, "mod", 256, "tiny"),
    ([15, 10, 5], 8, "mod", 512, "small"),
    ([15, 10, 5], 2, "range", 300, "small"),
    ([25, 10], 4, "range", 1024, "small"),
    ([-1, -1], 4, "mod", 128, "small"),
    ([-1, 5, -1], 3, "range", 64, "tiny"),
    ([5, 5], 1, "mod", 100, "tiny"),
    ([3, 3, 3, 3, 3], 5, "mod", 17, "tiny"),
]


@pytest.mark.gpu
@pytest.mark.parametrize("fanouts,P,owner,batch,shape", CASES)
def test_local_group_matches_oracle(fanouts, P, owner, batch, shape):
    import torch
    assert torch.cuda.is_available()
    from cslicer.partitioned import LocalGroup, owner_of, range_bounds
    from cslicer.synth import synth_graph
    ip, ix = synth_graph(shape, seed=21)
    n = ip.size - 1
    b = range_bounds(n, P)
    own = None if owner == "mod" else owner_of(np.arange(n), n, P, "range", b)
    grp = LocalGroup(ip, ix, P, fanouts, owner_mode=owner, bounds=b, max_batch=batch, num_slots=2,
                     seed=77)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX, seed=77, owner=own)
    rng = np.random.default_rng(8)
    for bi in range(5):
        seeds = rng.permutation(n)[:batch if bi != 4 else max(1, batch // 3)]
        views = grp.slice(seeds, bi)
        want = orc.slice(seeds, bi)
        for g, v in enumerate(views):
            _compare_rank(v, want, g, owner=own)
            v.release()
    grp.close()


@pytest.mark.gpu
@pytest.mark.parametrize("fanouts,P,owner", [([15, 10, 5], 4, "mod"), ([3, 20], 3, "range")])
def test_local_group_without_replacement_matches_oracle(fanouts, P, owner):
    """replace=False: every rank draws the same keyed permutation
    for the rows it owns as the sequential oracle."""
    from cslicer.partitioned import LocalGroup, owner_of, range_bounds
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=21)
    n = ip.size - 1
    b = range_bounds(n, P)
    own = None if owner == "mod" else owner_of(np.arange(n), n, P, "range", b)
    grp = LocalGroup(ip, ix, P, fanouts, owner_mode=owner, bounds=b, max_batch=300, num_slots=2,
                     seed=5, replace=False)
    orc = O.OracleSlicer(ip, ix, fanouts, P, O.RNG_PHILOX_NOREPLACE, seed=5, owner=own)
    rng = np.random.default_rng(9)
    for bi in range(3):
        seeds = rng.permutation(n)[:300]
        views = grp.slice(seeds, bi)
        want = orc.slice(seeds, bi)
        for g, v in enumerate(views):
            _compare_rank(v, want, g, owner=own)
            v.release()
    grp.close()


@pytest.mark.gpu
def test_local_group_pipelined_slots_and_duplicate_seed():
    """Two minibatches in flight per rank (slots), then a bad batch: every rank
    reports the error, reset() brings the group back."""
    from cslicer import _native
    from cslicer.partitioned import LocalGroup
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=5)
    n = ip.size - 1
    grp = LocalGroup(ip, ix, 4, [10, 10], max_batch=128, num_slots=3, seed=1)
    orc = O.OracleSlicer(ip, ix, [10, 10], 4, O.RNG_PHILOX, seed=1)
    rng = np.random.default_rng(0)
    batches = [rng.permutation(n)[:128] for _ in range(3)]
    pend = [grp.submit(s, i) for i, s in enumerate(batches)]
    for i, slots in enumerate(pend):
        want = orc.slice(batches[i], i)
        for g, v in enumerate(grp.wait(slots)):
            _compare_rank(v, want, g)
            v.release()
    bad = batches[0].copy()
    bad[5] = bad[9]
    slots = grp.submit(bad, 7)
    for r, s in zip(grp.ranks, slots):
        with pytest.raises(_native.SlicerError):
            r.wait(s)
    for r in grp.ranks:
        r.reset()
    views = grp.slice(batches[1], 1)
    want = orc.slice(batches[1], 1)
    for g, v in enumerate(views):
        _compare_rank(v, want, g)
        v.release()
    grp.close()


@pytest.mark.gpu
def test_products_shape_partitioned_equals_replicated():
    """BASELINE config 2 at full size, 8 parts: every rank's part equals the
    single-GPU slicer's part (itself pinned to the oracle at small sizes)."""
    from cslicer import _native, engine
    from cslicer.partitioned import LocalGroup
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("products", seed=1234)
    n = ip.size - 1
    P, fan = 8, [15, 10, 5]
    g = engine.Graph(ip, ix)
    sl = engine.Slicer(g, fan, P, max_batch=4096, num_slots=1, num_streams=1, seed=3)
    grp = LocalGroup(ip, ix, P, fan, max_batch=4096, num_slots=1, seed=3)
    seeds = np.random.default_rng(1).permutation(n)[:4096]
    ref = sl.slice(seeds, 11)
    views = grp.slice(seeds, 11)
    for q, v in enumerate(views):
        for l in range(4):
            assert np.array_equal(v.frontier_numpy(l), ref.frontier_numpy(l))
        for l in range(3):
            for name in _native.FIELDS:
                assert np.array_equal(v.numpy(l, q, name), ref.numpy(l, q, name)), (q, l, name)
        v.release()
    ref.release()
    grp.close()
    sl.close()
    g.close()


@pytest.mark.gpu
def test_one_process_per_gpu_ipc():
    """Real ranks: torchrun-style spawn of 2 processes, arenas wired with CUDA
    IPC, parity of each rank's part against the oracle."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", "29541",
                        os.path.join(HERE, "pslicer_rank_main.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("owner,P,dim", [("mod", 4, 602), ("range", 3, 64)])
def test_partitioned_fused_feature_gather(owner, P, dim):
    """Config 4 in the partitioned layout: rank g holds only the feature rows of the
    nodes it owns; in_nodes of its part are owned by it, so the fused gather is
    local.  fp32 copy: equality is exact (tolerance 0)."""
    from cslicer.partitioned import LocalGroup, owner_of, range_bounds
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=12)
    n = ip.size - 1
    b = range_bounds(n, P)
    own = None if owner == "mod" else owner_of(np.arange(n), n, P, "range", b)
    feat = np.random.default_rng(4321).random((n, dim), dtype=np.float32)
    fan = [-1, 5]
    grp = LocalGroup(ip, ix, P, fan, owner_mode=owner, bounds=b, max_batch=256, num_slots=2, seed=5,
                     features=feat)
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=5, owner=own)
    rng = np.random.default_rng(2)
    for bi in range(3):
        seeds = rng.permutation(n)[:256]
        want = orc.slice(seeds, bi)
        for g, v in enumerate(grp.slice(seeds, bi)):
            _compare_rank(v, want, g, owner=own)
            rows = feat[want["layers"][-1][g]["in_nodes"]]
            got = v.features(g).cpu().numpy()
            assert got.shape == rows.shape and np.array_equal(got, rows)
            assert v.feature_rows == rows.shape[0]
            v.release()
    grp.close()


def _shuffle_reference(want, l, P, own_of, parts, backward):
    """numpy statement of the shuffle on the oracle's lists: node nd owned by q with
    a partial row on g != q (nd in out_nodes_g).  forward: owner row += partial rows
    in rank order (when the owner has a row at all); backward: contributor row =
    owner row, or zeros."""
    pos = [{int(v): i for i, v in enumerate(want["layers"][l][q]["out_nodes"])} for q in range(P)]
    out = [x.copy() for x in parts]
    for q in range(P):          # owner
        for g in range(P):      # contributor, rank order
            if g == q:
                continue
            for nd in want["frontiers"][l]:
                nd = int(nd)
                if own_of(nd) != q or nd not in pos[g]:
                    continue
                if not backward:
                    if nd in pos[q]:
                        out[q][pos[q][nd]] = out[q][pos[q][nd]] + parts[g][pos[g][nd]]
                else:
                    out[g][pos[g][nd]] = parts[q][pos[q][nd]] if nd in pos[q] else 0.0
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("P,dim", [(4, 64), (3, 50)])
def test_split_parallel_shuffle_forward_and_backward(P, dim):
    """Consumer of from_ids/to_ids: partial aggregates of the rows a
    rank does not own travel to the owner and are summed there in rank order; the
    backward pass is the transpose.  fp32 adds in a fixed order: tolerance 0."""
    import torch
    from cslicer.partitioned import LocalGroup
    from cslicer.synth import synth_graph
    from oracle import shuffle_oracle as SO
    ip, ix = synth_graph("tiny", seed=31)
    n = ip.size - 1
    fan = [5, 5]
    grp = LocalGroup(ip, ix, P, fan, max_batch=200, num_slots=2, seed=3, shuffle_rows_capacity=4096,
                     shuffle_max_dim=dim)
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=3)
    seeds = np.random.default_rng(5).permutation(n)[:200]
    views = grp.slice(seeds, 0)
    want = orc.slice(seeds, 0)
    rng = np.random.default_rng(9)
    for l in range(2):
        for backward in (False, True):
            parts = [rng.random((views[g].field_len(l, g, "out_nodes"), dim), dtype=np.float32) for g in range(P)]
            rows = [torch.from_numpy(x).cuda() for x in parts]
            grp.shuffle(views, l, rows, backward)
            ref = _shuffle_reference(want, l, P, lambda v: v % P, parts, backward)
            # the pinned oracle (oracle/shuffle_oracle.py: mergeKernel + merge order of the reference,
            # held to tests/golden/shuffle.npz by tests/test_shuffle_oracle.py)
            fr, to = SO.peer_lists(want, l, lambda v: v % P)
            pinned = SO.shuffle_backward(parts, fr, to) if backward else SO.shuffle_forward(parts, fr, to)
            for g in range(P):
                assert np.array_equal(rows[g].cpu().numpy(), pinned[g]), (l, backward, g)
                assert np.array_equal(ref[g], pinned[g]), (l, backward, g)
                # the per-peer lists the CUDA path emits in the slots the reference leaves empty
                for q in range(P):
                    if q != g:
                        assert np.array_equal(views[g].numpy(l, q, "from_ids"), fr[g][q]), (l, g, q)
                        assert np.array_equal(views[g].numpy(l, q, "to_ids"), to[g][q]), (l, g, q)
    for v in views:
        v.release()
    grp.close()


@pytest.mark.gpu
def test_missing_peer_times_out_with_err_peer_and_the_group_recovers():
    """Failure detection: a rank whose peer never arrives does not hang -- the device-side
    barrier gives up after timeout_ms, the sample fails with GS_ERR_PEER, later barriers
    of the broken group return at once, and gs_pslicer_reset brings both ranks back.
    (Two ranks emulated on one device and one stream, as in LocalGroup: the minibatch of
    rank 0 is enqueued alone, so its barriers can never be met.)"""
    import time
    import torch
    assert torch.cuda.is_available()
    from cslicer import _native as N
    from cslicer.partitioned import LocalGroup
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=5)
    n, P, fan = ip.size - 1, 2, [5, 5]
    grp = LocalGroup(ip, ix, P, fan, max_batch=64, num_slots=2, seed=3, timeout_ms=60)
    seeds = np.random.default_rng(1).permutation(n)[:64]
    r0 = grp.ranks[0]
    t0 = time.time()
    slot = r0.stage(seeds, 0)
    r0.enqueue(slot, 0, grp.num_segments)      # rank 1 never enqueues this minibatch
    r0.commit(slot)
    with pytest.raises(N.SlicerError) as e:
        r0.wait(slot)
    assert e.value.code == N.ERR_PEER
    assert time.time() - t0 < 5.0              # one timeout, not one per barrier
    for r in grp.ranks:
        r.reset()
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=3)
    for bi in range(2):
        views = grp.slice(seeds, bi)
        want = orc.slice(seeds, bi)
        for g, v in enumerate(views):
            _compare_rank(v, want, g)
            v.release()
    grp.close()


@pytest.mark.gpu
def test_malformed_shard_is_rejected_at_create():
    """gs_pslicer_create validates its CSR shard on the device like gs_graph_create: the
    kernels index the shard's rows and the peers' tables with these ids."""
    from cslicer import _native
    from cslicer.partitioned import PartitionedSlicer
    # rank 0 of 2 (owner = id % 2) of a 6-node graph owns nodes 0, 2, 4: three local rows
    ip = np.array([0, 2, 3, 4], np.int64)
    for bad_ix, what in (([1, 2, 6, 0], "indices_local[2]"), ([1, -1, 0, 3], "indices_local[1]")):
        with pytest.raises(_native.SlicerError) as e:
            PartitionedSlicer(ip, np.array(bad_ix, np.int64), 6, 0, 2, [2])
        assert e.value.code == _native.ERR_INVALID and what in str(e.value)
    with pytest.raises(_native.SlicerError) as e:  # not monotone
        PartitionedSlicer(np.array([0, 3, 2, 4], np.int64), np.array([1, 2, 3, 0], np.int64), 6, 0, 2, [2])
    assert e.value.code == _native.ERR_INVALID and "indptr_local" in str(e.value)
    ps = PartitionedSlicer(ip, np.array([1, 2, 5, 0], np.int64), 6, 0, 2, [2])  # the well-formed shard loads
    ps.close()
