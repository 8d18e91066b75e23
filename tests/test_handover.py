"""Split-parallel hand-over (cslicer/handover.py): every rank slices whole
minibatches, part g of each sample travels to rank g in one all-to-all.

CPU: the block format against its host restatement, and a world_size-2 gloo run
in which the parts come from the oracle.  GPU: `gs_slicer_pack_parts` (own kernel)
against the sample it packs, on one device and -- with >= 2 GPUs -- over NCCL."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
FAN = [15, 10, 5]


def _oracle_parts(orc_sample, g):
    return [{k: v for k, v in layer[g].items()} for layer in orc_sample["layers"]]


def test_block_format_round_trip_and_overflow():
    from cslicer import _native as N
    from cslicer.handover import HEADER_WORDS, PartBlock, pack_block_numpy, part_words
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    orc = O.OracleSlicer(ip, ix, FAN, 4, O.RNG_PHILOX, seed=9)
    seeds = np.random.default_rng(1).permutation(ip.size - 1)[:128]
    want = orc.slice(seeds, (5 << 32) | 7)
    for g in range(4):
        fl = _oracle_parts(want, g)
        need = part_words(fl)
        assert need >= HEADER_WORDS and need % 4 == 0
        b = PartBlock(pack_block_numpy(fl, g, 4, seeds.size, (5 << 32) | 7, need + 64))
        assert (b.part, b.num_layers, b.num_parts, b.n_seeds) == (g, 3, 4, 128)
        assert b.batch_index == (5 << 32) | 7 and b.words == need
        assert b.edges == sum(len(x["csr_indices"]) for x in fl)
        for l in range(3):
            for name in N.FIELDS:
                assert np.array_equal(b.field(l, name), fl[l][name]), (g, l, name)
        with pytest.raises(RuntimeError, match="overflow"):
            PartBlock(pack_block_numpy(fl, g, 4, seeds.size, 0, need - 4))
    with pytest.raises(RuntimeError, match="magic"):
        PartBlock(np.zeros(HEADER_WORDS, dtype=np.int32))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cslicer import _native as N
    from cslicer.handover import PartBlock, pack_block_numpy, part_words
    from cslicer.sharding import rank_batch_ids
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    n = ip.size - 1
    orc = O.OracleSlicer(ip, ix, FAN, world, O.RNG_PHILOX, seed=11)
    perm = np.random.default_rng(2).permutation(n)
    batches = [perm[i * 96:(i + 1) * 96] for i in range(2 * world)]
    ok = True
    for k in range(2):                                   # round k: rank r slices batch k*world + r
        b = int(rank_batch_ids(2 * world, rank, world)[k])
        mine = orc.slice(batches[b], b)
        words = 4 * ((max(part_words(_oracle_parts(mine, g)) for g in range(world)) * 3 // 2 + 3) // 4)
        w = torch.tensor([words])
        dist.all_reduce(w, op=dist.ReduceOp.MAX)         # equal splits: one block size for the round
        words = int(w[0])
        send = torch.from_numpy(np.stack([pack_block_numpy(_oracle_parts(mine, g), g, world, 96, b, words)
                                          for g in range(world)]))
        recv = torch.empty_like(send)
        dist.all_to_all_single(recv, send)
        for src in range(world):
            blk = PartBlock(recv[src].numpy())
            want = orc.slice(batches[k * world + src], k * world + src)
            ok &= blk.part == rank and blk.batch_index == k * world + src
            for l in range(len(FAN)):
                for name in N.FIELDS:
                    ok &= bool(np.array_equal(blk.field(l, name), want["layers"][l][rank][name]))
    out.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


def test_two_gloo_ranks_receive_their_part_of_every_minibatch():
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    got = dict(out.get() for _ in range(world))
    assert got == {0: True, 1: True}


@pytest.mark.gpu
@pytest.mark.parametrize("P,fan,batch", [(4, [15, 10, 5], 512), (8, [10, 10], 256), (3, [-1, 5], 128), (1, [5], 64)])
def test_pack_parts_matches_the_sample_and_the_oracle(P, fan, batch):
    import torch
    from cslicer import _native as N
    from cslicer.engine import Graph, Slicer
    from cslicer.handover import HEADER_WORDS, PartBlock, pack_parts
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=4)
    n = ip.size - 1
    g = Graph(ip, ix)
    sl = Slicer(g, fan, P, max_batch=batch, num_slots=2, num_streams=2, seed=31)
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=31)
    rng = np.random.default_rng(6)
    for bi in (0, 1, (3 << 32) | 9):
        seeds = rng.permutation(n)[:batch]
        slot = sl.submit(seeds, bi)
        words = 4 * (sl.slab_capacity_words() // 4 + HEADER_WORDS)
        dst = torch.empty((P, words), dtype=torch.int32, device="cuda:0")
        pack_parts(sl, slot, dst, torch.cuda.current_stream().cuda_stream)   # before wait: ordered on the slot's stream
        host = dst.cpu().numpy()
        v = sl.wait(slot)
        want = orc.slice(seeds, bi)
        for q in range(P):
            b = PartBlock(host[q])
            assert (b.part, b.num_layers, b.num_parts, b.n_seeds, b.batch_index) == (q, len(fan), P, batch, bi)
            for l in range(len(fan)):
                for name in N.FIELDS:
                    a = b.field(l, name)
                    assert np.array_equal(a, v.numpy(l, q, name)), (q, l, name)
                    assert np.array_equal(a, want["layers"][l][q][name]), (q, l, name)
            assert b.edges == sum(v.field_len(l, q, "csr_indices") for l in range(len(fan)))
        # a block that is too small says so instead of writing past its end
        small = torch.full((P, HEADER_WORDS + 8), -7, dtype=torch.int32, device="cuda:0")
        guard = small.clone()
        pack_parts(sl, slot, small, torch.cuda.current_stream().cuda_stream)
        sh = small.cpu().numpy()
        if max(PartBlock(host[q]).words for q in range(P)) > HEADER_WORDS + 8:
            q = int(np.argmax([PartBlock(host[q]).words for q in range(P)]))
            assert sh[q][1] == N.ERR_OVERFLOW
            assert np.array_equal(sh[q][HEADER_WORDS:], guard.cpu().numpy()[q][HEADER_WORDS:])
            with pytest.raises(RuntimeError, match="overflow"):
                PartBlock(sh[q])
        v.release()
    with pytest.raises(N.SlicerError):
        pack_parts(sl, 0, torch.empty((P, HEADER_WORDS + 2), dtype=torch.int32, device="cuda:0"))
    sl.close()
    g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("P,fan,batch", [(4, [15, 10, 5], 512), (8, [10, 10], 256), (3, [-1, 5], 128), (1, [5], 64)])
def test_direct_hand_over_blocks_match_the_sample_and_the_oracle(P, fan, batch):
    """gs_slicer_submit_direct on one GPU (all P blocks local): the blocks the slicing kernels
    write themselves hold exactly the lists of an ordinary sample and of the oracle, in the
    block format of gs_slicer_pack_parts; a block that is too small is an overflow status."""
    import ctypes as C
    import torch
    from cslicer import _native as N
    from cslicer.engine import Graph, Slicer
    from cslicer.handover import HEADER_WORDS, PartBlock
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=4)
    n = ip.size - 1
    g = Graph(ip, ix)
    sl = Slicer(g, fan, P, max_batch=batch, num_slots=2, num_streams=2, seed=31, host_output=False)
    orc = O.OracleSlicer(ip, ix, fan, P, O.RNG_PHILOX, seed=31)
    rng = np.random.default_rng(6)
    words = 4 * (sl.slab_capacity_words() // 4 + HEADER_WORDS)
    for bi in (0, 1, (3 << 32) | 9):
        seeds = rng.permutation(n)[:batch]
        dst = torch.full((P, words), -7, dtype=torch.int32, device="cuda:0")
        torch.cuda.synchronize()
        arr = (C.c_void_p * P)(*[dst[q].data_ptr() for q in range(P)])
        slot = sl.submit_direct(seeds, bi, arr, words)
        h = sl.wait_header(slot)
        assert h.status == 0
        sl._release(slot)
        host = dst.cpu().numpy()
        v = sl.slice(seeds, bi)  # the ordinary path on the same slicer, after a direct sample
        want = orc.slice(seeds, bi)
        for q in range(P):
            b = PartBlock(host[q])
            assert (b.part, b.num_layers, b.num_parts, b.n_seeds, b.batch_index) == (q, len(fan), P, batch, bi)
            for l in range(len(fan)):
                for name in N.FIELDS:
                    a = b.field(l, name)
                    assert np.array_equal(a, v.tensor(l, q, name).cpu().numpy()), (q, l, name)
                    assert np.array_equal(a, want["layers"][l][q][name]), (q, l, name)
            assert b.edges == sum(v.field_len(l, q, "csr_indices") for l in range(len(fan)))
            assert (host[q][b.words:] == -7).all(), "wrote past the words in use"
        v.release()
    # too small a block: overflow status in the sample header and in the block headers, nothing past the end
    small_w = HEADER_WORDS + 8
    small = torch.full((P, small_w + 64), -7, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    arr = (C.c_void_p * P)(*[small[q].data_ptr() for q in range(P)])
    slot = sl.submit_direct(seeds, 5, arr, small_w)
    with pytest.raises(N.SlicerError) as e:
        sl.wait_header(slot)
    assert e.value.code == N.ERR_OVERFLOW
    sl._release(slot)
    sh = small.cpu().numpy()
    assert (sh[:, small_w:] == -7).all() and (sh[:, 1] == N.ERR_OVERFLOW).all()
    v = sl.slice(seeds, 5)  # the slicer recovers
    v.release()
    sl.close()
    g.close()


@pytest.mark.gpu
def test_hand_over_between_real_ranks_nccl():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29547")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", "29547",
                        os.path.join(HERE, "handover_rank_main.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


@pytest.mark.parametrize("world,depth,slack", [(2, 1, 0), (2, 2, 3), (4, 3, 1), (8, 2, 0), (3, 4, 2)])
def test_ring_protocol_never_overwrites_an_unreleased_round(world, depth, slack):
    """Model of PeerScatter's device-side ordering, run against the real RingSchedule.
    Every rank has a comm stream (FIFO of release marks and barriers) and packs that
    wait for one of the rank's own barriers; a barrier completes when it is at the head
    of EVERY rank's comm stream and the rank's pack of that round is done.  Ranks post,
    consume and release in random interleavings; whenever a pack runs, the buffer it
    writes must have been released by every owner -- and the protocol must not stall."""
    from cslicer.handover import RingSchedule
    rng = np.random.default_rng(100 * world + 10 * depth + slack)
    rounds = 60
    rings = [RingSchedule(depth, slack) for _ in range(world)]
    nbuf = rings[0].nbuf
    comm = [[] for _ in range(world)]            # per rank: ("release", X) | ("barrier", X)
    packs = [dict() for _ in range(world)]       # per rank: round -> barrier index to wait for (or None)
    pack_done = [set() for _ in range(world)]
    barrier_done = set()
    released_on_device = [set() for _ in range(world)]   # release marks the comm stream has passed
    collected = [0] * world                      # rounds consumed so far (in order)
    holds = [dict() for _ in range(world)]       # per owner: buffer -> (round, source) payload check
    for _ in range(200000):
        progressed = False
        i = int(rng.integers(world))
        act = int(rng.integers(4))
        ring = rings[i]
        if act == 0 and ring.posted < rounds:    # host: post the next round if the discipline allows
            if ring.posted - depth < 0 or (ring.posted - depth) in ring.released:
                r, k, after = ring.next_post()
                assert k == r % nbuf
                packs[i][r] = after
                comm[i].append(("barrier", r))
                progressed = True
        elif act == 1:                           # device: run a pack whose barrier has completed
            ready = [r for r, after in packs[i].items()
                     if r not in pack_done[i] and (after is None or after in barrier_done)]
            if ready:                            # packs sit on different slot streams: any order
                r = int(rng.choice(ready))
                for g in range(world):           # the store into owner g's buffer r % nbuf
                    prev = r - nbuf
                    assert prev < 0 or prev in released_on_device[g], (i, g, r, prev)
                    holds[g][(r % nbuf, i)] = r
                pack_done[i].add(r)
                progressed = True
        elif act == 2:                           # device: comm streams advance
            while comm[i] and comm[i][0][0] == "release":
                released_on_device[i].add(comm[i].pop(0)[1])
                progressed = True
            heads = [c[0] if c else None for c in comm]
            if heads[i] and heads[i][0] == "barrier":
                r = heads[i][1]
                if all(h == ("barrier", r) for h in heads) and all(r in pack_done[g] for g in range(world)):
                    for c in comm:
                        c.pop(0)
                    barrier_done.add(r)
                    progressed = True
        else:                                    # host: consume + release the oldest delivered round
            r = collected[i]
            if r in barrier_done:
                for src in range(world):         # what the consumer reads is that round's payload
                    assert holds[i][(r % nbuf, src)] == r, (i, src, r, holds[i][(r % nbuf, src)])
                comm[i].append(("release", r))
                ring.release(r)
                collected[i] += 1
                progressed = True
        if all(c == rounds for c in collected):
            break
    assert all(c == rounds for c in collected), f"protocol stalled: {collected}"
    with pytest.raises(RuntimeError):            # discipline: `depth` un-released rounds at most
        ring = RingSchedule(depth, slack)
        for _ in range(depth + 1):
            ring.next_post()


@pytest.mark.parametrize("world,depth,slack", [(2, 1, 0), (2, 2, 3), (4, 3, 1), (8, 2, 0), (3, 4, 2)])
def test_flag_protocol_never_overwrites_an_unreleased_round(world, depth, slack):
    """Model of the flag barrier (barrier="flags", gs_flags_post) against the real RingSchedule.
    Flag words: arrived[owner][buffer][source] and released[source][owner], values only grow.
    A rank's writes for round r (direct / staged / packed alike) start once released[rank][q] >=
    r - nbuf + 1 for every owner q; when they are complete the rank stores r + 1 into
    arrived[q][r % nbuf][rank] of every owner; an owner consumes round r once its whole row holds
    r + 1 and releases in order by storing r + 1 into released[s][owner] of every source.  Random
    interleavings of hosts and devices: a write must never land in a buffer whose previous round
    the owner has not released, the consumer must read the round it expects, nothing stalls."""
    from cslicer.handover import RingSchedule
    rng = np.random.default_rng(7 + 100 * world + 10 * depth + slack)
    rounds = 60
    rings = [RingSchedule(depth, slack) for _ in range(world)]
    nbuf = rings[0].nbuf
    arrived = np.zeros((world, nbuf, world), np.int64)     # [owner][buffer][source]
    released = np.zeros((world, world), np.int64)          # [source][owner]
    writes = [dict() for _ in range(world)]                # per source: round -> "queued" | "done"
    signalled = [set() for _ in range(world)]
    owner_released = [0] * world                           # rounds the owner's host has released (in order)
    collected = [0] * world
    holds = [dict() for _ in range(world)]                 # per owner: (buffer, source) -> round
    idle = 0
    for _ in range(300000):
        i = int(rng.integers(world))
        act = int(rng.integers(4))
        ring = rings[i]
        progressed = False
        if act == 0 and ring.posted < rounds:              # host: submit / post the next round
            if ring.posted - depth < 0 or (ring.posted - depth) in ring.released:
                r, k, _after = ring.next_post()
                writes[i][r] = "queued"
                progressed = True
        elif act == 1:                                     # device: the writes of a round whose buffers are free
            ready = [r for r, st in writes[i].items() if st == "queued" and
                     (r < nbuf or all(released[i][q] >= r - nbuf + 1 for q in range(world)))]
            if ready:
                r = int(rng.choice(ready))                 # chains sit on different streams: any order
                for q in range(world):
                    assert r < nbuf or owner_released[q] >= r - nbuf + 1, (i, q, r)
                    holds[q][(r % nbuf, i)] = r
                writes[i][r] = "done"
                progressed = True
        elif act == 2:                                     # device: the round's flag kernel (signal after the writes)
            todo = sorted(r for r, st in writes[i].items() if st == "done" and r not in signalled[i])
            if todo:
                r = todo[0]                                # one comm stream per rank: in order
                if all(rr in signalled[i] for rr in range(r)):
                    for q in range(world):
                        arrived[q][r % nbuf][i] = r + 1
                    signalled[i].add(r)
                    progressed = True
        else:                                              # host + device: consume and release the oldest round
            r = collected[i]
            if r < rounds and all(arrived[i][r % nbuf][s] >= r + 1 for s in range(world)):
                for s in range(world):
                    assert holds[i][(r % nbuf, s)] == r, (i, s, r, holds[i][(r % nbuf, s)])
                owner_released[i] = r + 1
                for s in range(world):
                    released[s][i] = r + 1
                ring.release(r)
                collected[i] += 1
                progressed = True
        idle = 0 if progressed else idle + 1
        if all(c == rounds for c in collected):
            break
        assert idle < 20000, "the protocol stalled"
    assert all(c == rounds for c in collected)
