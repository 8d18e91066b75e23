"""Where the hand-over loses against independent replicas (diagnostic, torchrun N >= 2):
  A direct hand-over as benchmarked
  B the same protocol, every block written to LOCAL memory (no NVLink payload)
  C like B without the per-round all-reduce (event only)
  D like A without the all-reduce (NVLink payload, no barrier: content not guaranteed -- timing only)
usage: torchrun --nproc-per-node 2 tools/diag_handover.py"""
import ctypes as C
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import bench
from cslicer.engine import Graph, Slicer
from cslicer.handover import DirectScatter, HEADER_WORDS

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ip, ix = bench.load_graph("products", f"cuda:{local}")
g = Graph(ip, ix, device=local)
n = ip.size - 1
sl = Slicer(g, bench.FANOUTS, world, max_batch=4096, num_slots=12, num_streams=12, seed=7, host_output=False, grid_ctas=2)
batches = bench.seed_batches(n, 64, rank, world)
words = (int(sl.slab_capacity_words()) // world + HEADER_WORDS + 3) & ~3
depth = 12


def run(ps, k, barrier=True):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    inflight = []
    e0.record()
    for i in range(k):
        if len(inflight) == depth:
            slot, tk = inflight.pop(0)
            sl.wait_header(slot); sl.release(slot)
            ps.collect(tk, parse=False); ps.release(tk)
        bi, seeds = batches[i % len(batches)]
        slot = ps.submit(seeds, bi)
        inflight.append((slot, ps.post(slot) if barrier else post_nobarrier(ps, slot)))
    for slot, tk in inflight:
        sl.wait_header(slot); sl.release(slot)
        ps.collect(tk, parse=False); ps.release(tk)
    e1.record(ps.stream); e1.synchronize(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def post_nobarrier(ps, slot):
    from cslicer import _native as N
    r, k = ps._pending.pop(slot)
    N.check(N.lib().gs_slicer_scatter_parts(ps.slicer._h, int(slot), ps._dst[k], ps.block_words, None,
                                            C.c_void_p(ps.stream.cuda_stream)))
    with torch.cuda.stream(ps.stream):
        ps.done[k].record(ps.stream)
    ps._posted += 1
    return r


class NoBarrier(DirectScatter):  # timing only: nothing tells a rank that a round has arrived
    def _round_barrier(self, r, k):
        with torch.cuda.stream(self.stream):
            self.done[k].record(self.stream)

    def _buffer_free_event(self, r, k, after):
        return 0

    def release(self, ticket):
        self.ring.release(ticket)


for name, cls, cls_kw, local_only in (("A direct, flag barrier", DirectScatter, dict(barrier="flags"), False),
                                      ("C direct, flags, LOCAL blocks", DirectScatter, dict(barrier="flags"), True),
                                      ("D direct, NO barrier", NoBarrier, {}, False),
                                      ("E LOCAL blocks, NO barrier", NoBarrier, {}, True)):
    ps = cls(sl, dist, words, depth=depth, **cls_kw)
    if local_only:
        for k in range(ps.nbuf):
            for q in range(world):
                ps._dst[k][q] = ps._base + 4 * (k * world + q) * ps.block_words
    run(ps, 64)
    dist.barrier(); torch.cuda.synchronize()
    rounds = 960
    ms = run(ps, rounds)
    t = torch.tensor([ms], device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{name:34s} {rounds * world / (t.item() * 1e-3):9.1f} minibatches/s (whole job)", flush=True)
    ps.close()
sl.close(); g.close()
dist.destroy_process_group()
