"""One rank of the split-parallel hand-over over NCCL (launched by
tests/test_handover.py::test_hand_over_between_real_ranks_nccl and usable by hand:
torchrun --nproc-per-node N tests/handover_rank_main.py)."""
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
    from cslicer import _native as N
    from cslicer.engine import Graph, Slicer
    from cslicer.handover import HEADER_WORDS, PartScatter
    from cslicer.synth import synth_graph
    from oracle import oracle as O
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ip, ix = synth_graph("small", seed=21)
    n = ip.size - 1
    fan, batch, rounds = [15, 10, 5], 512, 5
    g = Graph(ip, ix, device=local)
    sl = Slicer(g, fan, world, max_batch=batch, num_slots=4, num_streams=4, seed=77, host_output=False)
    orc = O.OracleSlicer(ip, ix, fan, world, O.RNG_PHILOX, seed=77)   # the checker: any rank can restate any minibatch
    perm = np.random.default_rng(8).permutation(n)
    batches = [perm[(i * batch) % (n - batch):][:batch] for i in range(rounds * world)]
    words = sl.slab_capacity_words() // world + HEADER_WORDS
    for kind in ("nccl all-to-all", "peer stores", "peer stores, bulk-copy engine", "direct", "staged"):
        # 2 ranks would use plain stores: force the cp.async.bulk path over NVLink as well
        os.environ["GSPLIT_PACK_TMA"] = "1" if kind.endswith("engine") else "0"
        if kind == "direct":  # the slicing kernels write the owners' blocks themselves
            from cslicer.handover import DirectScatter
            ps = DirectScatter(sl, dist, block_words=words, depth=2)
        elif kind == "staged":  # ... local staging blocks, sent by the copy engines
            from cslicer.handover import StagedScatter
            ps = StagedScatter(sl, dist, block_words=words, depth=2)
        elif kind.startswith("peer stores"):
            from cslicer.handover import PeerScatter
            ps = PeerScatter(sl, dist, block_words=words, depth=2)
        else:
            ps = PartScatter(sl, dist, block_words=words, depth=2)
        tickets = []
        for k in range(rounds):                      # two rounds in flight
            b = k * world + rank
            slot = ps.submit(batches[b], b) if kind in ("direct", "staged") else sl.submit(batches[b], b)
            tickets.append(ps.post(slot))
            sl.release(slot)
            if k >= 1:
                _check(ps, tickets[k - 1], k - 1, world, rank, orc, batches, fan, N)
        _check(ps, tickets[-1], rounds - 1, world, rank, orc, batches, fan, N)
        if kind.startswith("peer stores") or kind in ("direct", "staged"):
            try:                                     # discipline: an un-released round blocks the buffer ring
                for k in range(4):
                    slot = ps.submit(batches[0], 0) if kind in ("direct", "staged") else sl.submit(batches[0], 0)
                    ps.post(slot)
                    sl.release(slot)
                raise AssertionError("post did not refuse to overwrite an un-released round")
            except RuntimeError:
                pass
            ps.close()
        dist.barrier()
    os.environ.pop("GSPLIT_PACK_TMA", None)
    print(f"rank {rank} ok", flush=True)
    sl.close()
    g.close()
    dist.destroy_process_group()


def _check(ps, ticket, k, world, rank, orc, batches, fan, N):
    blocks = ps.collect(ticket)
    assert len(blocks) == world
    for src, blk in enumerate(blocks):
        b = k * world + src
        assert blk.part == rank and blk.batch_index == b, (blk.part, blk.batch_index, b)
        want = orc.slice(batches[b], b)
        for l in range(len(fan)):
            for name in N.FIELDS:
                got = blk.field(l, name).cpu().numpy()
                assert np.array_equal(got, want["layers"][l][rank][name]), (k, src, l, name)
    if hasattr(ps, "release"):
        ps.release(ticket)


if __name__ == "__main__":
    main()
