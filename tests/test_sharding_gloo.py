"""N > 1 host logic on CPU: world_size-2 gloo processes shard an epoch's batches
with no overlap and no gap, and the whole-job reduction (max time, summed
counters) is what bench.py reports."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cslicer.sharding import epoch_batches, rank_batch_ids, reduce_job


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    perm = np.random.default_rng(5).permutation(10_007).astype(np.int64)  # same on every rank
    mine = epoch_batches(perm, 512, rank, world)
    ids = torch.tensor([b for b, _ in mine], dtype=torch.int64)
    seeds = sum(len(s) for _, s in mine)
    gathered = [torch.zeros(16, dtype=torch.int64) for _ in range(world)]
    pad = torch.full((16,), -1, dtype=torch.int64)
    pad[:ids.numel()] = ids
    dist.all_gather(gathered, pad)
    ms, (tot_seeds, tot_batches) = reduce_job(dist, "cpu", 10.0 * (rank + 1), [seeds, len(mine)])
    if rank == 0:
        allb = sorted(int(x) for g in gathered for x in g.tolist() if x >= 0)
        out.put((allb, ms, tot_seeds, tot_batches))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_partition_the_epoch_and_reduce():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    allb, ms, tot_seeds, tot_batches = out.get()
    nb = (10_007 - 1) // 512 + 1
    assert allb == list(range(nb))          # disjoint, complete
    assert tot_batches == nb and tot_seeds == 10_007
    assert ms == 20.0                        # max over ranks


def test_rank_batch_ids_shapes():
    assert rank_batch_ids(10, 0, 4).tolist() == [0, 4, 8]
    assert rank_batch_ids(10, 3, 4).tolist() == [3, 7]
    assert rank_batch_ids(2, 3, 4).tolist() == []
    got = np.sort(np.concatenate([rank_batch_ids(37, r, 8) for r in range(8)]))
    assert got.tolist() == list(range(37))


def _pworker(rank, world, port, out):
    """Partitioned mode, host side only (no GPU): every rank shards the same CSR,
    keeps its own part, and the ranks exchange their 64-byte arena handles."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cslicer.partitioned import exchange_handles, owner_of, partition_csr
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=3)
    n = ip.size - 1
    lip, lix = partition_csr(ip, ix, world)[rank]
    handle = bytes([rank]) * 64                      # stands in for cudaIpcMemHandle_t
    got = exchange_handles(dist, handle)
    ok = got == [bytes([r]) * 64 for r in range(world)]
    # shard sizes add up to the graph and rows keep their neighbour lists
    sizes = torch.tensor([lip.size - 1, lix.size], dtype=torch.int64)
    dist.all_reduce(sizes)
    mine = np.flatnonzero(owner_of(np.arange(n), n, world) == rank)
    rows_ok = all(np.array_equal(lix[lip[k]:lip[k + 1]], ix[ip[v]:ip[v + 1]]) for k, v in enumerate(mine))
    if rank == 0:
        out.put((ok, rows_ok, sizes.tolist(), n, int(ix.size)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_shard_the_graph_and_exchange_handles():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_pworker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    ok, rows_ok, sizes, n, m = out.get()
    assert ok and rows_ok and sizes == [n, m]
