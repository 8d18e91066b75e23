"""Partitioned mode of the slicer: one rank per GPU, graph rows sharded by owner
(DESIGN.md section 6).  Rank g produces `sample.layers[l][g]` --
the bipartite GPU g of the reference's layout consumes (slicer.cpp:36-43) --
bit-identical to part g of the single-GPU slicer.

Three pieces:
  * `partition_csr` / `range_bounds`   -- host-side sharding of a CSR by owner;
  * `PartitionedSlicer`                -- one rank (gs_pslicer_* of the C ABI);
  * `connect_process_group`            -- one process per GPU: exchanges the CUDA
                                          IPC handles of the ranks' arenas through
                                          torch.distributed and freezes the ranks;
  * `LocalGroup`                       -- P ranks inside ONE process on one device,
                                          interleaved segment by segment on one
                                          stream (tests, single-GPU emulation).
All data movement between ranks happens inside the CUDA kernels (stores and
red.min into peer memory over NVLink); torch.distributed carries 64-byte
handles once, at set-up.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .engine import SampleView


def range_bounds(num_nodes, num_parts):
    """Contiguous owner ranges: part q owns [b[q], b[q+1])."""
    return [(num_nodes * q) // num_parts for q in range(num_parts + 1)]


def owner_of(ids, num_nodes, num_parts, owner_mode="mod", bounds=None):
    ids = np.asarray(ids)
    if owner_mode == "mod":
        return (ids % num_parts).astype(np.int32)
    b = np.asarray(bounds if bounds is not None else range_bounds(num_nodes, num_parts))
    return (np.searchsorted(b, ids, side="right") - 1).astype(np.int32)


def partition_csr(indptr, indices, num_parts, owner_mode="mod", bounds=None):
    """-> [(indptr_local, indices_local)] per rank.  Local row order is the
    owner's local index order (mod: id // P; range: id - b[rank]); neighbour ids
    stay global."""
    ip = np.asarray(indptr, dtype=np.int64)
    ix = np.asarray(indices, dtype=np.int64)
    n = ip.size - 1
    deg = np.diff(ip)
    out = []
    for g in range(num_parts):
        if owner_mode == "mod":
            rows = np.arange(g, n, num_parts, dtype=np.int64)
        else:
            b = bounds if bounds is not None else range_bounds(n, num_parts)
            rows = np.arange(b[g], b[g + 1], dtype=np.int64)
        d = deg[rows]
        lip = np.zeros(rows.size + 1, dtype=np.int64)
        np.cumsum(d, out=lip[1:])
        if owner_mode == "mod":
            # gather each owned row's slice
            idx = np.repeat(ip[rows] - lip[:-1], d) + np.arange(lip[-1], dtype=np.int64)
            lix = ix[idx]
        else:
            lix = ix[ip[rows[0]]:ip[rows[-1] + 1]] if rows.size else ix[:0]
        out.append((lip, np.ascontiguousarray(lix)))
    return out


def renumber_by_owner(indptr, indices, owner, num_parts):
    """Arbitrary owner map (e.g. METIS `partition_map_opt.bin`, metis.py:70-71) ->
    contiguous owner ranges, the form the partitioned slicer addresses arithmetically
    (SURVEY 8(f) N4).  Nodes are renumbered part after part, keeping their relative
    order inside a part.  -> (indptr', indices', bounds, new_of_old, old_of_new):
    row new_of_old[v] of the new CSR is row v of the old one with neighbour ids
    mapped; part q owns [bounds[q], bounds[q+1])."""
    ip = np.asarray(indptr, dtype=np.int64)
    ix = np.asarray(indices, dtype=np.int64)
    own = np.asarray(owner, dtype=np.int64)
    n = ip.size - 1
    if own.size != n or own.min(initial=0) < 0 or own.max(initial=0) >= num_parts:
        raise ValueError("owner map must have num_nodes entries in [0, num_parts)")
    old_of_new = np.argsort(own, kind="stable").astype(np.int64)
    new_of_old = np.empty(n, dtype=np.int64)
    new_of_old[old_of_new] = np.arange(n, dtype=np.int64)
    counts = np.bincount(own, minlength=num_parts)
    bounds = [0] + np.cumsum(counts).tolist()
    deg = np.diff(ip)[old_of_new]
    nip = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(deg, out=nip[1:])
    src = np.repeat(ip[old_of_new] - nip[:-1], deg) + np.arange(nip[-1], dtype=np.int64)
    nix = new_of_old[ix[src]]
    return nip, nix, bounds, new_of_old, old_of_new


class _Dev:
    def __init__(self, device):
        self.device = int(device)


class PartitionedSlicer:
    """One rank of the partitioned slicer (gs_pslicer)."""

    def __init__(self, indptr_local, indices_local, num_nodes, rank, num_parts, fanouts, *,
                 owner_mode="mod", bounds=None, device=0, max_batch=4096, num_slots=2, seed=0,
                 host_output=True, use_graph=True, stream=None, num_edges_global=0,
                 frontier_capacity=0, cand_capacity=0, slab_capacity=0, grid_ctas=0,
                 timeout_ms=5000, features_local=None, feature_rows_capacity=0,
                 shuffle_rows_capacity=0, shuffle_max_dim=0, replace=True):
        L = N.lib()
        ip = np.ascontiguousarray(indptr_local, dtype=np.int64)
        ix = np.ascontiguousarray(indices_local, dtype=np.int64)
        cfg = N.PConfig()
        L.gs_pconfig_default(C.byref(cfg))
        cfg.num_layers = len(fanouts)
        for i, f in enumerate(fanouts):
            cfg.fanouts[i] = int(f)
        cfg.num_parts, cfg.rank = int(num_parts), int(rank)
        cfg.owner_mode = N.OWNER_MOD if owner_mode == "mod" else N.OWNER_RANGE
        if owner_mode != "mod":
            b = bounds if bounds is not None else range_bounds(num_nodes, num_parts)
            for q in range(num_parts + 1):
                cfg.range_bounds[q] = int(b[q])
        cfg.seed = int(seed)
        cfg.rng_mode = N.RNG_PHILOX if replace else N.RNG_PHILOX_NOREPLACE
        cfg.max_batch, cfg.num_slots = int(max_batch), int(num_slots)
        cfg.host_output, cfg.use_graph = int(bool(host_output)), int(bool(use_graph))
        cfg.grid_ctas, cfg.timeout_ms = int(grid_ctas), int(timeout_ms)
        cfg.frontier_capacity, cfg.cand_capacity = int(frontier_capacity), int(cand_capacity)
        cfg.slab_capacity, cfg.num_edges_global = int(slab_capacity), int(num_edges_global)
        cfg.stream = stream
        feat = None
        if features_local is not None:  # [n_local, dim] fp32 rows of the owned nodes, local order
            feat = np.ascontiguousarray(features_local, dtype=np.float32)
            if feat.ndim != 2 or feat.shape[0] != ip.size - 1:
                raise ValueError("features_local must be [n_local, dim]")
            cfg.features_local_host, cfg.feature_dim = feat.ctypes.data, int(feat.shape[1])
            cfg.feature_rows_capacity = int(feature_rows_capacity)
        cfg.shuffle_rows_capacity, cfg.shuffle_max_dim = int(shuffle_rows_capacity), int(shuffle_max_dim)
        h = C.c_void_p()
        N.check(L.gs_pslicer_create(int(device), int(num_nodes), ip.size - 1, ix.size,
                                    ip.ctypes.data, ix.ctypes.data, C.byref(cfg), C.byref(h)))
        self._h = h
        self.graph = _Dev(device)  # SampleView reads .graph.device
        self.rank, self.num_parts, self.fanouts = int(rank), int(num_parts), list(fanouts)
        self.num_segments = L.gs_pslicer_num_segments(h)
        self.launches_per_sample = L.gs_pslicer_launches_per_sample(h)
        self.device_bytes = L.gs_pslicer_device_bytes(h)

    # ---- wiring
    def ipc_handle(self):
        buf = (C.c_ubyte * 64)()
        N.check(N.lib().gs_pslicer_ipc_handle(self._h, buf))
        return bytes(buf)

    def connect_ipc(self, peer_rank, handle):
        buf = (C.c_ubyte * 64).from_buffer_copy(handle)
        N.check(N.lib().gs_pslicer_connect_ipc(self._h, int(peer_rank), buf))

    def connect_local(self, peer_rank, peer):
        N.check(N.lib().gs_pslicer_connect_local(self._h, int(peer_rank), peer._h))

    def ready(self):
        N.check(N.lib().gs_pslicer_ready(self._h))

    # ---- one minibatch
    def submit(self, seeds, batch_index=0):
        s = np.ascontiguousarray(seeds, dtype=np.int64)
        slot = C.c_int32(-1)
        N.check(N.lib().gs_pslicer_submit(self._h, s.ctypes.data, s.size, int(batch_index), C.byref(slot)))
        return slot.value

    def stage(self, seeds, batch_index=0):
        s = np.ascontiguousarray(seeds, dtype=np.int64)
        slot = C.c_int32(-1)
        N.check(N.lib().gs_pslicer_stage(self._h, s.ctypes.data, s.size, int(batch_index), C.byref(slot)))
        return slot.value

    def enqueue(self, slot, seg_lo, seg_hi):
        N.check(N.lib().gs_pslicer_enqueue(self._h, slot, seg_lo, seg_hi))

    def commit(self, slot):
        N.check(N.lib().gs_pslicer_commit(self._h, slot))

    def wait(self, slot):
        hp = C.POINTER(N.SampleHeader)()
        try:
            N.check(N.lib().gs_pslicer_wait(self._h, slot, C.byref(hp)))
        except N.SlicerError:
            N.lib().gs_pslicer_release(self._h, slot)
            raise
        dev, host = C.c_void_p(), C.c_void_p()
        N.check(N.lib().gs_pslicer_slab(self._h, slot, C.byref(dev), C.byref(host)))
        return SampleView(self, slot, hp.contents, dev.value or 0, host.value or 0)

    def wait_header(self, slot):
        hp = C.POINTER(N.SampleHeader)()
        N.check(N.lib().gs_pslicer_wait(self._h, slot, C.byref(hp)))
        return hp.contents

    def _release(self, slot):
        if getattr(self, "_h", None):
            N.lib().gs_pslicer_release(self._h, slot)

    def _feature_slab(self, slot):
        ptr, dim, rows = C.c_void_p(), C.c_int32(), C.c_int32()
        N.check(N.lib().gs_pslicer_feature_slab(self._h, slot, C.byref(ptr), C.byref(dim), C.byref(rows)))
        return ptr.value, dim.value

    release = _release

    def slice(self, seeds, batch_index=0):
        return self.wait(self.submit(seeds, batch_index))

    def shuffle(self, view, layer, rows, backward=False, phase=-1):
        """Split-parallel shuffle of `rows` ([len(out_nodes), dim] fp32 CUDA tensor, the
        partial aggregates of layer `layer` of `view`), in place: forward adds the
        peers' partial rows into the rows this rank owns, backward hands the owners'
        rows back to the contributors.  Collective: every rank calls it for the same
        minibatch, in the same order.  Blocks until done when phase == -1."""
        import torch
        assert rows.is_cuda and rows.dtype == torch.float32 and rows.is_contiguous() and rows.dim() == 2
        if phase in (-1, 0):
            torch.cuda.current_stream(rows.device).synchronize()  # rows were produced on torch's stream
        N.check(N.lib().gs_pslicer_shuffle(self._h, view.slot, int(layer), rows.data_ptr(),
                                           int(rows.shape[1]), int(bool(backward)), int(phase)))
        if phase in (-1, 2):
            self.sync()
        return rows

    def sync(self):
        N.check(N.lib().gs_pslicer_sync(self._h))

    def reset(self):
        N.check(N.lib().gs_pslicer_reset(self._h))

    def close(self):
        if getattr(self, "_h", None):
            N.lib().gs_pslicer_destroy(self._h)
            self._h = None

    __del__ = close


def exchange_handles(dist, handle):
    """All-gather one 64-byte handle per rank through torch.distributed (any backend,
    gloo included: no GPU involved) -> list indexed by rank."""
    if len(handle) != 64:
        raise ValueError("a CUDA IPC memory handle is 64 bytes")
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, bytes(handle))
    if any(h is None or len(h) != 64 for h in out):
        raise RuntimeError("handle exchange failed")
    return out


def connect_process_group(ps, dist):
    """One process per GPU: all-gather the ranks' 64-byte CUDA IPC handles over
    torch.distributed (any backend), map every peer arena, freeze.  Collective:
    every rank of the group must call it."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if world != ps.num_parts or rank != ps.rank:
        raise ValueError("process group does not match the slicer's (rank, num_parts)")
    handles = exchange_handles(dist, ps.ipc_handle())
    for q in range(world):
        if q != rank:
            ps.connect_ipc(q, handles[q])
    ps.ready()
    dist.barrier()
    return ps


class PartitionedPipeline:
    """Several minibatches in flight per rank: `lanes` independent partitioned
    slicers (own arena, workspaces, stream and barrier flags each), minibatch b
    runs on lane b % lanes on every rank.  A lane's 3L+1 barriers then overlap
    with the kernels of the other lanes.  Collective constructor (one process
    per GPU): every rank builds the same lanes in the same order."""

    def __init__(self, indptr_local, indices_local, num_nodes, rank, num_parts, fanouts, dist, *,
                 lanes=4, **kw):
        self.lanes = [PartitionedSlicer(indptr_local, indices_local, num_nodes, rank, num_parts,
                                        fanouts, **kw) for _ in range(int(lanes))]
        for ps in self.lanes:
            connect_process_group(ps, dist)
        self.rank, self.num_parts = int(rank), int(num_parts)
        self._dist = dist
        self.launches_per_sample = self.lanes[0].launches_per_sample
        self.device_bytes = sum(ps.device_bytes for ps in self.lanes)
        self._next = 0

    def submit(self, seeds, batch_index=0):
        """-> ticket (lane, slot)"""
        k = self._next
        self._next = (k + 1) % len(self.lanes)
        return k, self.lanes[k].submit(seeds, batch_index)

    def wait(self, ticket):
        return self.lanes[ticket[0]].wait(ticket[1])

    def wait_header(self, ticket):
        return self.lanes[ticket[0]].wait_header(ticket[1])

    def release(self, ticket):
        self.lanes[ticket[0]].release(ticket[1])

    def slice(self, seeds, batch_index=0):
        return self.wait(self.submit(seeds, batch_index))

    def reset(self):
        """After any error (a GS_ERR_PEER timeout included): collective -- every rank
        calls it; nobody submits again before all ranks have reset (gs_pslicer_reset)."""
        for ps in self.lanes:
            ps.sync()
        self._dist.barrier()   # every rank has stopped writing into its peers' arenas
        for ps in self.lanes:
            ps.reset()
        self._dist.barrier()   # nobody signals before every arena is clean
        self._next = 0

    def close(self):
        for ps in self.lanes:
            ps.close()


def reset_process_group(ps, dist):
    """Collective recovery of one PartitionedSlicer per rank after an error:
    gs_pslicer_reset on every rank, then a host barrier before anyone submits."""
    ps.sync()
    dist.barrier()   # every rank has stopped writing into its peers' arenas
    ps.reset()
    dist.barrier()   # nobody signals before every arena is clean


class LocalGroup:
    """P ranks in one process on one device, sharing one stream: the segments of
    a minibatch are enqueued rank after rank, so every barrier wait finds its
    signals already ahead of it in the stream.  Same kernels, same arenas, same
    barriers as one process per GPU -- used by the tests and as a single-GPU
    emulation of the P-GPU layout."""

    def __init__(self, indptr, indices, num_parts, fanouts, *, owner_mode="mod", bounds=None,
                 device=0, max_batch=4096, num_slots=2, seed=0, host_output=True, **kw):
        import torch
        self.num_parts, self.device = int(num_parts), int(device)
        n = len(indptr) - 1
        self._stream = torch.cuda.Stream(device=device)
        shards = partition_csr(indptr, indices, num_parts, owner_mode, bounds)
        feats = kw.pop("features", None)  # [N, dim] -> every rank is given the rows it owns
        own = owner_of(np.arange(n), n, num_parts, owner_mode, bounds) if feats is not None else None
        self.ranks = [PartitionedSlicer(ip, ix, n, g, num_parts, fanouts, owner_mode=owner_mode,
                                        features_local=None if feats is None else feats[own == g],
                                        bounds=bounds, device=device, max_batch=max_batch,
                                        num_slots=num_slots, seed=seed, host_output=host_output,
                                        stream=self._stream.cuda_stream,
                                        num_edges_global=max(1, len(indices)), **kw)
                      for g, (ip, ix) in enumerate(shards)]
        for a in self.ranks:
            for b in self.ranks:
                if a is not b:
                    a.connect_local(b.rank, b)
        for a in self.ranks:
            a.ready()
        self.num_segments = self.ranks[0].num_segments

    def submit(self, seeds, batch_index=0):
        slots = [r.stage(seeds, batch_index) for r in self.ranks]
        for k in range(self.num_segments):
            for r, s in zip(self.ranks, slots):
                r.enqueue(s, k, k + 1)
        for r, s in zip(self.ranks, slots):
            r.commit(s)
        return slots

    def wait(self, slots):
        return [r.wait(s) for r, s in zip(self.ranks, slots)]

    def slice(self, seeds, batch_index=0):
        """-> [SampleView of rank 0, ..., rank P-1]"""
        return self.wait(self.submit(seeds, batch_index))

    def shuffle(self, views, layer, rows, backward=False):
        """rows[g]: rank g's [len(out_nodes_g), dim] tensor; all ranks, phase by phase."""
        for phase in (0, 1, 2):
            for r, v, x in zip(self.ranks, views, rows):
                r.shuffle(v, layer, x, backward, phase)
        return rows

    def close(self):
        for r in self.ranks:
            r.close()
