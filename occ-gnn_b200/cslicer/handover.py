"""Split-parallel hand-over: slice whole minibatches anywhere, deliver part g to GPU g.

The consumer of `sample.layers[l][g]` is GPU g (python/train.py:120-160 moves
bipartite g of every layer to device g; python/data/bipartite.py:20-58 builds its
tensors there).  The partitioned slicer (partitioned.py) produces part g ON rank g
but makes all ranks walk every minibatch together.  This is the throughput form of
the same contract on a machine where the topology fits every GPU (DESIGN.md
section 6.4): rank r slices its own minibatches b = r, r + R, ... with the
replicated slicer, `gs_slicer_pack_parts` (own kernel) packs the R parts of a sample
into R self-describing blocks, and ONE all-to-all over NVLink moves block g to rank
g.  After round k rank g holds part g of minibatches kR .. kR + R - 1, bit-identical
to what the single-GPU slicer (and the oracle) emit for those minibatches.

Block layout: include/gsplit_slicer.h (gs_slicer_pack_parts).  `pack_block_numpy`
restates it on the host for the CPU tests; `PartBlock` parses either.
"""
import ctypes as C

import numpy as np

from . import _native as N

MAGIC = 0x31505347
HEADER_WORDS = 16 + 2 * N.MAX_LAYERS * N.NUM_FIELDS


def _pad4(n):
    return (int(n) + 3) & ~3


def part_words(fields_by_layer):
    """int32 words a block needs for `fields_by_layer[l][name] -> array`."""
    return HEADER_WORDS + sum(_pad4(len(fl[name])) for fl in fields_by_layer for name in N.FIELDS)


def pack_block_numpy(fields_by_layer, part, num_parts, n_seeds, batch_index, block_words):
    """Host restatement of k_pack_parts (slicer_kernels.cu) for one part."""
    L = len(fields_by_layer)
    blk = np.zeros(block_words, dtype=np.int32)
    need = part_words(fields_by_layer)
    blk[0] = MAGIC
    blk[1] = 0 if need <= block_words else N.ERR_OVERFLOW
    blk[2], blk[3], blk[4], blk[5] = part, L, num_parts, n_seeds
    blk[6] = np.uint32(batch_index & 0xFFFFFFFF).astype(np.int32)
    blk[7] = np.uint32(batch_index >> 32).astype(np.int32)
    blk[8] = min(need, 0x7FFFFFFF)
    blk[9] = sum(len(fl["csr_indices"]) for fl in fields_by_layer)
    at = HEADER_WORDS
    for l, fl in enumerate(fields_by_layer):
        for f, name in enumerate(N.FIELDS):
            a = np.asarray(fl[name], dtype=np.int32)
            k = 16 + 2 * (l * N.NUM_FIELDS + f)
            blk[k], blk[k + 1] = at, a.size
            if need <= block_words:
                blk[at:at + a.size] = a
            at += _pad4(a.size)
    return blk


class PartBlock:
    """One received block: part `part` of minibatch `batch_index`, every list a
    zero-copy slice of the buffer it arrived in (numpy array or torch tensor)."""

    def __init__(self, buf, header=None):
        h = np.asarray(header if header is not None else buf[:HEADER_WORDS]).astype(np.int64)
        if int(h[0]) != MAGIC:
            raise RuntimeError("not a part block (bad magic)")
        self.status = int(h[1])
        if self.status == N.ERR_OVERFLOW:
            raise RuntimeError(f"part block overflow: the part needs {int(h[8])} words; raise block_words")
        if self.status != 0:
            raise RuntimeError(f"source rank raised status {self.status} for this minibatch")
        self.part, self.num_layers, self.num_parts, self.n_seeds = int(h[2]), int(h[3]), int(h[4]), int(h[5])
        self.batch_index = (int(h[6]) & 0xFFFFFFFF) | ((int(h[7]) & 0xFFFFFFFF) << 32)
        self.words, self.edges = int(h[8]), int(h[9])
        self._spans = [[(int(h[16 + 2 * (l * N.NUM_FIELDS + f)]), int(h[17 + 2 * (l * N.NUM_FIELDS + f)]))
                        for f in range(N.NUM_FIELDS)] for l in range(self.num_layers)]
        self._buf = buf

    def field(self, l, name):
        off, ln = self._spans[l][N.FIELDS.index(name)]
        return self._buf[off:off + ln]

    def field_len(self, l, name):
        return self._spans[l][N.FIELDS.index(name)][1]


def pack_parts(slicer, slot, dst, consumer_stream=None):
    """gs_slicer_pack_parts: dst = int32 CUDA tensor [num_parts, block_words] (torch is
    plumbing: it owns the buffer the collective sends)."""
    if dst.dim() != 2 or dst.shape[0] != slicer.num_parts or not dst.is_contiguous():
        raise ValueError("dst must be a contiguous [num_parts, block_words] int32 tensor")
    N.check(N.lib().gs_slicer_pack_parts(slicer._h, int(slot), C.c_void_p(dst.data_ptr()), int(dst.shape[1]),
                                         C.c_void_p(consumer_stream or 0)))


class PartScatter:
    """Collective hand-over for R = num_parts ranks, `depth` rounds in flight.

        k = ps.post(slot)        # pack the slot's sample + all-to-all (async; every rank calls it)
        blocks = ps.collect(k)   # [PartBlock of source rank 0, 1, ...]: part `rank` of R minibatches

    The exchange runs on its own stream, so the slicer keeps working on the next
    minibatches while blocks are in flight.  The blocks of a round stay valid until
    `depth` further rounds have been posted (their buffers are then reused)."""

    def __init__(self, slicer, dist, block_words, depth=2, group=None):
        import torch
        self.slicer, self.dist, self.group = slicer, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if slicer.num_parts != self.world:
            raise ValueError(f"hand-over needs num_parts == world size ({slicer.num_parts} != {self.world})")
        self.block_words = _pad4(block_words)
        if self.block_words < HEADER_WORDS:
            raise ValueError("block_words smaller than a block header")
        dev = torch.device("cuda", slicer.graph.device)
        self.depth = int(depth)
        self.send = [torch.empty((self.world, self.block_words), dtype=torch.int32, device=dev)
                     for _ in range(self.depth)]
        self.recv = [torch.empty((self.world, self.block_words), dtype=torch.int32, device=dev)
                     for _ in range(self.depth)]
        self.stream = torch.cuda.Stream(device=dev)
        self.done = [torch.cuda.Event() for _ in range(self.depth)]
        self._posted = 0
        self._torch = torch

    def post(self, slot):
        torch = self._torch
        k = self._posted % self.depth
        if self._posted >= self.depth:
            self.done[k].synchronize()  # the round that used these buffers has been delivered
        pack_parts(self.slicer, slot, self.send[k], self.stream.cuda_stream)
        with torch.cuda.stream(self.stream):
            self.dist.all_to_all_single(self.recv[k], self.send[k], group=self.group)
            self.done[k].record(self.stream)
        self._posted += 1
        return self._posted - 1

    def collect(self, ticket, parse=True):
        k = ticket % self.depth
        self.done[k].synchronize()
        if not parse:
            return self.recv[k]
        hdr = self.recv[k][:, :HEADER_WORDS].cpu().numpy()
        return [PartBlock(self.recv[k][r], hdr[r]) for r in range(self.world)]

    def nvlink_bytes_per_round(self):
        """bytes this rank sends to peers per round (the local block does not travel)"""
        return 4 * self.block_words * (self.world - 1)


class _DevWords:
    """__cuda_array_interface__ carrier for an int32 device buffer this module owns."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": "<i4",
                                         "data": (int(ptr), False), "version": 2, "strides": None}


class RingSchedule:
    """Host-side bookkeeping of PeerScatter's receive ring (no GPU involved, so the
    CPU tests can run the protocol against a model of streams and barriers).

    Round X lands in buffer X % nbuf, nbuf = depth + 1 + slack.  `next_post` -> (X,
    buffer, index of the barrier the pack has to wait for, or None while the ring is
    still fresh); it refuses to go on while round X - depth is un-released, which is
    what makes barrier X - 1 - slack carry the release of round X - nbuf."""

    def __init__(self, depth, slack):
        if depth < 1 or slack < 0:
            raise ValueError("depth >= 1 and slack >= 0")
        self.depth, self.slack, self.nbuf = int(depth), int(slack), int(depth) + 1 + int(slack)
        self.posted = 0
        self.released = set()

    def next_post(self):
        r = self.posted
        if r - self.depth >= 0 and (r - self.depth) not in self.released:
            raise RuntimeError(f"round {r - self.depth} must be released before round {r} is posted "
                               f"(depth={self.depth})")
        self.released.discard(r - self.depth - 1)
        after = r - 1 - self.slack
        self.posted += 1
        return r, r % self.nbuf, (after if r >= self.nbuf else None)

    def release(self, ticket):
        if not (0 <= ticket < self.posted):
            raise ValueError("unknown round")
        self.released.add(ticket)


class PeerScatter:
    """The hand-over with the exchange fused into the packing kernel: every rank's
    receive buffers are mapped into every other rank (CUDA IPC, handles all-gathered
    once through torch.distributed), and `gs_slicer_scatter_parts` writes block g of
    a sample straight into rank g's buffer -- the parts cross NVLink inside the
    kernel itself (16-byte stores, or cp.async.bulk pieces from 4 parts up); no send
    buffer, no library collective on the payload.  What is left for the library is a 4-byte all-reduce per round, used as
    the barrier that tells a rank all of its blocks have landed.

        k = ps.post(slot)        # scatter the slot's sample + barrier (async; every rank calls it)
        blocks = ps.collect(k)   # [PartBlock of source rank 0, 1, ...]
        ps.release(k)            # the blocks of round k may be overwritten
        ps.close()               # collective: unmap the peers' rings, free this rank's

    `depth` rounds may be un-released when `post` is called.  Ordering, all on the
    device: round X lands in buffer X % nbuf, nbuf = depth + 1 + slack; a rank's pack
    for round X waits (event wait on the slot's stream) for its own barrier
    X - 1 - slack, whose completion means every rank has queued that barrier --
    behind the releases it had made by then, i.e. of every round
    <= X - 1 - slack - depth = X - nbuf, the round this buffer held.  With slack > 0
    that barrier is several rounds old, so packs do not queue up behind barriers."""

    def __init__(self, slicer, dist, block_words, depth=2, group=None, slack=3, barrier="flags", timeout_ms=5000):
        """barrier="flags" (default): arrival and release travel as flag words stored into the peers'
        memory (gs_flags_post: one warp per round, no rank waits for a rank it does not need);
        "allreduce": a 4-byte library all-reduce per round (lock-step; kept for A/B)."""
        import torch
        if barrier not in ("flags", "allreduce"):
            raise ValueError("barrier must be 'flags' or 'allreduce'")
        self.barrier, self.timeout_ms = barrier, int(timeout_ms)
        self.slicer, self.dist, self.group = slicer, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if slicer.num_parts != self.world:
            raise ValueError(f"hand-over needs num_parts == world size ({slicer.num_parts} != {self.world})")
        self.block_words = _pad4(block_words)
        if self.block_words < HEADER_WORDS:
            raise ValueError("block_words smaller than a block header")
        self.device = slicer.graph.device
        dev = torch.device("cuda", self.device)
        self.ring = RingSchedule(depth, slack)
        self.depth, self.slack, self.nbuf = self.ring.depth, self.ring.slack, self.ring.nbuf
        # arena: the ring of blocks, then the flag words: arrived[nbuf][world] (written by the source
        # ranks), released[world] (written by the owner ranks)
        self._ring_words = self.nbuf * self.world * self.block_words
        self._flag_words = _pad4(self.nbuf * self.world + self.world)
        nbytes = 4 * (self._ring_words + self._flag_words)
        base, handle = C.c_void_p(), (C.c_char * 64)()
        rc = N.lib().gs_ipc_alloc(self.device, nbytes, C.byref(base), handle)
        self._base = base.value if rc == N.OK else None
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw) if rc == N.OK else b"", group=group)
        if any(len(h) != 64 for h in handles):  # collective verdict: every rank raises, none is left waiting
            if self._base:
                N.lib().gs_ipc_free(self.device, C.c_void_p(self._base))
                self._base = None
            raise RuntimeError(f"receive ring of {nbytes} bytes could not be allocated / exported on rank(s) "
                               f"{[q for q, h in enumerate(handles) if len(h) != 64]}")
        self._peer, err = [], None
        for q in range(self.world):
            if q == self.rank:
                self._peer.append(self._base)
                continue
            p = C.c_void_p()
            try:
                N.check(N.lib().gs_ipc_open(self.device, handles[q], C.byref(p)))
            except N.SlicerError as e:  # keep going: the verdict below is collective
                err = err or e
            self._peer.append(p.value)
        # every rank learns whether ALL mappings succeeded, so that a failure raises on every
        # rank instead of leaving the others inside the next collective
        ok = torch.tensor([0 if err else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok[0]) == 0:
            for q, p in enumerate(self._peer):
                if q != self.rank and p:
                    N.lib().gs_ipc_close(self.device, C.c_void_p(p))
            dist.barrier(group=group)
            N.lib().gs_ipc_free(self.device, C.c_void_p(self._base))
            self._base = None
            raise RuntimeError(f"peer mapping of the receive rings failed on some rank ({err or 'a peer'}): "
                               "use PartScatter (NCCL all-to-all) on this machine")
        # my block of round-buffer k lands in [k][rank] of every peer's arena
        self._dst = []
        for k in range(self.nbuf):
            arr = (C.c_void_p * self.world)()
            for q in range(self.world):
                arr[q] = self._peer[q] + 4 * (k * self.world + self.rank) * self.block_words
            self._dst.append(arr)
        self.recv = torch.as_tensor(_DevWords(self._base, (self.nbuf, self.world, self.block_words)), device=dev)
        flags = torch.as_tensor(_DevWords(self._base + 4 * self._ring_words, (self._flag_words,)), device=dev)
        flags.zero_()
        W, nb = self.world, self.nbuf
        arrived = lambda q, k, src: self._peer[q] + 4 * (self._ring_words + k * W + src)
        released = lambda q, owner: self._peer[q] + 4 * (self._ring_words + nb * W + owner)
        # round in buffer k: tell every owner q "the blocks of source `rank` are complete"; watch my own row
        self._sig_arrived = [(C.c_void_p * W)(*[arrived(q, k, self.rank) for q in range(W)]) for k in range(nb)]
        self._wait_arrived = [(C.c_void_p * W)(*[arrived(self.rank, k, src) for src in range(W)]) for k in range(nb)]
        # release: tell every source s "owner `rank` is done with your block"; watch what the owners told me
        self._sig_released = (C.c_void_p * W)(*[released(s, self.rank) for s in range(W)])
        self._wait_released = (C.c_void_p * W)(*[released(self.rank, q) for q in range(W)])
        self._status = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._rel_stream = torch.cuda.Stream(device=dev)
        self._rel_done = [torch.cuda.Event() for _ in range(nb)]
        self._next_release = 0
        self.stream = torch.cuda.Stream(device=dev)
        self.done = [torch.cuda.Event() for _ in range(self.nbuf)]
        self._flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self._torch = torch
        torch.cuda.synchronize(dev)
        dist.barrier(group=group)  # every arena is mapped everywhere before the first store

    def _flags(self, sig, n_sig, sig_value, wait, n_wait, wait_value, stream):
        N.check(N.lib().gs_flags_post(self.device, sig, n_sig, sig_value & 0xFFFFFFFF, wait, n_wait,
                                      wait_value & 0xFFFFFFFF, self.timeout_ms,
                                      C.c_void_p(self._status.data_ptr()), C.c_void_p(stream)))

    def _buffer_free_event(self, r, k, after):
        """-> handle of the event after which buffer k may be written for round r (0: at once)."""
        if self.barrier == "allreduce":  # the barrier behind which every owner released round r - nbuf
            return self.done[after % self.nbuf].cuda_event if after is not None else 0
        if r < self.nbuf:
            return 0
        # every owner has released round r - nbuf (it stored r - nbuf + 1 into my `released` row)
        self._flags(None, 0, 0, self._wait_released, self.world, r - self.nbuf + 1, self._rel_stream.cuda_stream)
        self._rel_done[k].record(self._rel_stream)
        return self._rel_done[k].cuda_event

    def _round_barrier(self, r, k):
        """On self.stream, behind the writes of this rank's blocks: tell the owners, wait for all sources."""
        torch = self._torch
        with torch.cuda.stream(self.stream):
            if self.barrier == "allreduce":
                self.dist.all_reduce(self._flag, group=self.group)  # 4 bytes; the payload went by stores
            else:
                self._flags(self._sig_arrived[k], self.world, r + 1, self._wait_arrived[k], self.world, r + 1,
                            self.stream.cuda_stream)
            self.done[k].record(self.stream)

    def post(self, slot):
        r, k, after = self.ring.next_post()
        ev = self._buffer_free_event(r, k, after)
        N.check(N.lib().gs_slicer_scatter_parts(self.slicer._h, int(slot), self._dst[k], self.block_words,
                                                C.c_void_p(ev), C.c_void_p(self.stream.cuda_stream)))
        self._round_barrier(r, k)
        return r

    def collect(self, ticket, parse=True):
        k = ticket % self.nbuf
        self.done[k].synchronize()
        if int(self._status[0]) != 0:
            raise RuntimeError(f"hand-over: a peer's flag did not arrive within {self.timeout_ms} ms (round {ticket})")
        if not parse:
            return self.recv[k]
        hdr = self.recv[k][:, :HEADER_WORDS].cpu().numpy()
        return [PartBlock(self.recv[k][r], hdr[r]) for r in range(self.world)]

    def release(self, ticket):
        """The consumer is done with round `ticket` (work queued so far on torch's
        current stream included): its buffer may be overwritten."""
        torch = self._torch
        cur = torch.cuda.current_stream(self.recv.device)
        if self.barrier == "allreduce":
            ev = torch.cuda.Event()
            ev.record(cur)
            self.stream.wait_event(ev)  # this rank's next barrier carries the release to the writers
        else:
            if ticket != self._next_release:
                raise RuntimeError(f"rounds are released in order with barrier='flags' (expected {self._next_release})")
            self._next_release += 1
            # behind the consumer's work: store ticket + 1 into every source rank's `released` row
            self._flags(self._sig_released, self.world, ticket + 1, None, 0, 0, cur.cuda_stream)
        self.ring.release(ticket)

    def nvlink_bytes_per_round(self):
        return 4 * self.block_words * (self.world - 1)

    def close(self):
        if getattr(self, "_base", None):
            self._torch.cuda.synchronize(self.recv.device)
            self.dist.barrier(group=self.group)  # nobody is still storing into a buffer about to go
            for q, p in enumerate(self._peer):
                if q != self.rank:
                    N.lib().gs_ipc_close(self.device, C.c_void_p(p))
            self.recv = None
            self.dist.barrier(group=self.group)
            N.lib().gs_ipc_free(self.device, C.c_void_p(self._base))
            self._base = None


class DirectScatter(PeerScatter):
    """The hand-over without a packing pass: `submit` hands the slicer the P blocks of the
    next ring buffer and the sample's OWN kernels write every list of part g into rank g's
    block as they produce it (gs_slicer_submit_direct) -- a list crosses NVLink once and is
    never written to, or read back from, the slicing GPU's memory.

        slot = ds.submit(seeds, b)   # instead of slicer.submit
        slicer.wait_header(slot)     # optional: status / sizes
        k = ds.post(slot)            # ordering + barrier only (no kernel)
        blocks = ds.collect(k); ...; ds.release(k); slicer._release(slot)

    The target buffer of round X is fixed at submit time and its kernels wait for this
    rank's barrier X - 1 - slack, so that barrier must have been posted already: at most
    `slack` rounds may be submitted ahead of the last post (checked)."""

    def __init__(self, slicer, dist, block_words, depth=2, group=None, slack=3, **kw):
        super().__init__(slicer, dist, block_words, depth=depth, group=group, slack=slack, **kw)
        self._pending = {}
        self._posted = 0

    def submit(self, seeds, batch_index=0):
        if self.barrier == "allreduce" and self.ring.posted - self._posted >= max(1, self.slack):
            raise RuntimeError(f"{self.ring.posted - self._posted} rounds submitted and not posted: "
                               f"at most slack={self.slack} may be in flight before post()")
        r, k, after = self.ring.next_post()
        ev = self._buffer_free_event(r, k, after)
        slot = self.slicer.submit_direct(seeds, batch_index, self._dst[k], self.block_words, ev)
        self._pending[slot] = (r, k)
        return slot

    def post(self, slot):
        r, k = self._pending.pop(slot)
        N.check(N.lib().gs_slicer_scatter_parts(self.slicer._h, int(slot), self._dst[k], self.block_words,
                                                None, C.c_void_p(self.stream.cuda_stream)))
        self._round_barrier(r, k)
        self._posted += 1
        return r


class StagedScatter(PeerScatter):
    """The hand-over through the copy engines: `submit` lets the slicing kernels write every part's
    lists ONCE, contiguously, into a LOCAL staging block (gs_slicer_submit_direct with local
    targets -- no slab write, no packing pass), and `post` -- called once the sample is complete,
    it waits for its header to learn the sizes -- sends block g to rank g with one copy-engine
    transfer per part (gs_slicer_push_parts): no SM stores cross NVLink, and the link runs at its
    bulk rate (measured on 8 B200: 307 GB/s sent per rank by copy engines against < 190 GB/s
    for stores issued by the slicing or packing kernels).  Barrier as in PeerScatter (flags).

        slot = ss.submit(seeds, b)        # several rounds ahead
        k = ss.post(slot)                 # blocks the host until THIS sample is sliced
        slicer._release(slot); blocks = ss.collect(k); ...; ss.release(k)"""

    deferred = True  # post() needs the finished sample: keep a few submits ahead of it

    def __init__(self, slicer, dist, block_words, depth=2, group=None, slack=3, **kw):
        super().__init__(slicer, dist, block_words, depth=depth, group=group, slack=slack, **kw)
        torch = self._torch
        dev = torch.device("cuda", self.device)
        self._stage = torch.empty((self.nbuf, self.world, self.block_words), dtype=torch.int32, device=dev)
        self._stage_dst = [(C.c_void_p * self.world)(*[self._stage[k][q].data_ptr() for q in range(self.world)])
                           for k in range(self.nbuf)]
        self._sent = [torch.cuda.Event() for _ in range(self.nbuf)]
        self._pending = {}

    def submit(self, seeds, batch_index=0):
        r, k, after = self.ring.next_post()
        # staging block k was last read by the transfers of round r - nbuf
        ev = self._sent[k].cuda_event if r >= self.nbuf else 0
        slot = self.slicer.submit_direct(seeds, batch_index, self._stage_dst[k], self.block_words, ev)
        self._pending[slot] = (r, k, after)
        return slot

    def post(self, slot):
        r, k, after = self._pending.pop(slot)
        self.slicer.wait_header(slot)  # raises on a failed sample; the header now holds the words in use
        ev = self._buffer_free_event(r, k, after)
        N.check(N.lib().gs_slicer_push_parts(self.slicer._h, int(slot), self._dst[k], C.c_void_p(ev),
                                             C.c_void_p(self.stream.cuda_stream)))
        self._sent[k].record(self.stream)  # behind every transfer of the round
        self._round_barrier(r, k)
        return r

