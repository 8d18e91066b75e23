"""One rank of the one-process-per-GPU partitioned slicer (launched by
tests/test_partitioned.py::test_one_process_per_gpu_ipc and usable by hand:
torchrun --nproc-per-node N tests/pslicer_rank_main.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "occ-gnn_b200"), ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist
    from cslicer.partitioned import (PartitionedPipeline, PartitionedSlicer, connect_process_group,
                                     partition_csr)
    from cslicer.synth import synth_graph
    from oracle import oracle as O
    from test_partitioned import _compare_rank, _shuffle_reference
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")  # carries 64-byte handles once; data moves over NVLink
    ip, ix = synth_graph("small", seed=21)
    n = ip.size - 1
    fan = [15, 10, 5]
    lip, lix = partition_csr(ip, ix, world)[rank]
    ps = PartitionedSlicer(lip, lix, n, rank, world, fan, device=local, max_batch=512, num_slots=2,
                           seed=77, num_edges_global=ix.size, shuffle_rows_capacity=1 << 16,
                           shuffle_max_dim=32, timeout_ms=1500)
    connect_process_group(ps, dist)
    orc = O.OracleSlicer(ip, ix, fan, world, O.RNG_PHILOX, seed=77)
    rng = np.random.default_rng(8)
    for bi in range(6):
        seeds = rng.permutation(n)[:512 if bi != 5 else 100]
        v = ps.slice(seeds, bi)
        _compare_rank(v, orc.slice(seeds, bi), rank)
        v.release()
    # split-parallel shuffle over NVLink: every rank can restate all ranks' rows from the seed
    seeds = rng.permutation(n)[:512]
    v = ps.slice(seeds, 50)
    want = orc.slice(seeds, 50)
    for l in range(len(fan)):
        for backward in (False, True):
            parts = [np.random.default_rng(1000 + 10 * l + g).random(
                (want["layers"][l][g]["out_nodes"].size, 32), dtype=np.float32) for g in range(world)]
            rows = torch.from_numpy(parts[rank]).cuda()
            ps.shuffle(v, l, rows, backward)
            ref = _shuffle_reference(want, l, world, lambda x: x % world, parts, backward)
            assert np.array_equal(rows.cpu().numpy(), ref[rank]), (l, backward)
    v.release()
    dist.barrier()
    # failure detection over real peers: the last rank sits one minibatch out; the others
    # time out with GS_ERR_PEER instead of hanging, then the whole group resets and goes on
    from cslicer import _native as N
    from cslicer.partitioned import reset_process_group
    seeds = rng.permutation(n)[:256]
    if rank != world - 1:
        try:
            ps.slice(seeds, 60)
            raise AssertionError("a minibatch without one of its ranks must fail")
        except N.SlicerError as e:
            assert e.code == N.ERR_PEER, e
    reset_process_group(ps, dist)
    for bi in (61, 62):
        seeds = rng.permutation(n)[:256]
        v = ps.slice(seeds, bi)
        _compare_rank(v, orc.slice(seeds, bi), rank)
        v.release()
    dist.barrier()
    ps.close()
    # several minibatches in flight per rank: independent lanes, barriers overlap
    pp = PartitionedPipeline(lip, lix, n, rank, world, fan, dist, lanes=3, device=local, max_batch=512,
                             num_slots=2, seed=77, num_edges_global=ix.size)
    batches = [rng.permutation(n)[:512] for _ in range(9)]
    tickets = [pp.submit(b, 100 + i) for i, b in enumerate(batches[:6])]
    for i, b in enumerate(batches):
        v = pp.wait(tickets[i])
        _compare_rank(v, orc.slice(b, 100 + i), rank)
        v.release()
        if i + 6 < len(batches):
            tickets.append(pp.submit(batches[i + 6], 100 + i + 6))
    dist.barrier()
    pp.close()
    print(f"rank {rank} ok", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
