"""Array-level host API over the C ABI: Graph, Slicer, SampleView.

This is the layer the reference-facing classes in `cslicer/__init__.py` sit on
(and what bench.py times).  It adds nothing to the computation: it stages
seeds, submits, waits and wraps the slab spans as numpy (pinned host mirror) or
torch (device) views.
"""
import ctypes as C

import numpy as np

from . import _native as N


class _DevSpan:
    """Minimal __cuda_array_interface__ carrier so torch can wrap a slab span."""

    def __init__(self, ptr, n, owner):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<i4",
                                         "data": (int(ptr), False), "version": 2,
                                         "strides": None}
        self._owner = owner


class Graph:
    """Device-resident CSR (replaces the in-memory product of Dataset::read_graph,
    /root/reference/cslicer/dataset.cpp:18-36)."""

    def __init__(self, indptr, indices, device=0):
        ip = np.ascontiguousarray(indptr, dtype=np.int64)
        ix = np.ascontiguousarray(indices, dtype=np.int64)
        self.num_nodes, self.num_edges, self.device = int(ip.size - 1), int(ix.size), int(device)
        h = C.c_void_p()
        N.check(N.lib().gs_graph_create(self.device, self.num_nodes, self.num_edges,
                                        ip.ctypes.data, ix.ctypes.data, C.byref(h)))
        self._h = h

    @classmethod
    def from_directory(cls, directory, device=0, verify=True):
        """The reference's on-disk layout (meta.txt, indptr.bin, indices.bin) streamed straight
        into HBM by gs_graph_load: the host never holds the arrays; checksums, offsets and ids
        are verified on the device."""
        self = cls.__new__(cls)
        h = C.c_void_p()
        N.check(N.lib().gs_graph_load(int(device), str(directory).encode(), int(bool(verify)), C.byref(h)))
        self._h, self.device = h, int(device)
        self.num_nodes = int(N.lib().gs_graph_num_nodes(h))
        self.num_edges = int(N.lib().gs_graph_num_edges(h))
        return self

    def close(self):
        if getattr(self, "_h", None) and N is not None and N.lib is not None:
            N.lib().gs_graph_destroy(self._h)
            self._h = None

    __del__ = close


class Features:
    """Node features (gs_features).  rows: [num_rows, dim] fp32; node id lives at
    row id // id_div (id_div=1: full replica; =P: the shard of one part under
    owner = id % P).

    cache_rows=None: the whole table is copied to HBM.  cache_rows=<row indices>:
    the table stays in host memory (page-locked in place -- `rows` must then be a
    C-contiguous float32 array, which this object keeps alive) and only those rows
    are held in HBM; gathers serve every other row from host memory inside the
    same kernel (python/utils/memory_manager.py:40-104 without the refresh step)."""

    def __init__(self, rows, device=0, id_div=1, cache_rows=None):
        h = C.c_void_p()
        if cache_rows is None:
            r = np.ascontiguousarray(rows, dtype=np.float32)
        else:
            r = rows
            if not (isinstance(r, np.ndarray) and r.dtype == np.float32 and r.flags.c_contiguous):
                raise ValueError("a cached store maps `rows` in place: pass a C-contiguous float32 ndarray")
        if r.ndim != 2:
            raise ValueError("features must be [num_rows, dim]")
        self.num_rows, self.dim, self.device, self.id_div = int(r.shape[0]), int(r.shape[1]), int(device), int(id_div)
        if cache_rows is None:
            N.check(N.lib().gs_features_create(self.device, self.num_rows, self.dim, self.id_div,
                                               r.ctypes.data, C.byref(h)))
            self.rows_in_hbm = self.num_rows
        else:
            c = np.ascontiguousarray(cache_rows, dtype=np.int64)
            self._host_rows = r  # mapped in place by the library
            N.check(N.lib().gs_features_create_cached(self.device, self.num_rows, self.dim, self.id_div,
                                                      r.ctypes.data, c.ctypes.data, c.size, C.byref(h)))
            self.rows_in_hbm = int(c.size)
        self._h = h

    def gather(self, ids):
        """ids: int32 CUDA tensor -> [len(ids), dim] fp32 CUDA tensor (own kernel,
        enqueued on torch's current stream)."""
        import torch
        ids = ids.to(device=f"cuda:{self.device}", dtype=torch.int32).contiguous()
        out = torch.empty((ids.numel(), self.dim), dtype=torch.float32, device=ids.device)
        N.check(N.lib().gs_features_gather(self._h, ids.data_ptr(), ids.numel(), out.data_ptr(),
                                           torch.cuda.current_stream(ids.device).cuda_stream))
        return out

    def close(self):
        if getattr(self, "_h", None):
            N.lib().gs_features_destroy(self._h)
            self._h = None

    __del__ = close


class SampleView:
    """One completed sample in a slot: header + views into the slab."""

    def __init__(self, slicer, slot, hdr, dev_ptr, host_ptr):
        self._slicer, self.slot = slicer, slot
        self.num_layers, self.num_parts = hdr.num_layers, hdr.num_parts
        self.n_seeds, self.batch_index = hdr.n_seeds, hdr.batch_index
        self.edges, self.slab_used = int(hdr.edges), int(hdr.slab_used)
        self.draw_cursor = int(hdr.draw_cursor)
        L, P = self.num_layers, self.num_parts
        self.frontier_len = [hdr.frontier_len[l] for l in range(L + 1)]
        self.cand_len = [hdr.cand_len[l] for l in range(L)]
        self._frontier = [(hdr.frontier[l].off, hdr.frontier[l].len) for l in range(L + 1)]
        self._spans = [[[(hdr.field[l][g][f].off, hdr.field[l][g][f].len)
                         for f in range(N.NUM_FIELDS)] for g in range(P)] for l in range(L)]
        self._dev_ptr, self._host_ptr = dev_ptr, host_ptr
        self.feature_rows = int(hdr.feature_rows)
        self.feature_part_off = [int(hdr.feature_part_off[g]) for g in range(P + 1)]
        self._host = None
        self._released = False
        self._frontier_on_device = bool(hdr.reserved[2])
        self._fcache = {}
        self._fdev_ptr = slicer._frontier_base(slot) if self._frontier_on_device else dev_ptr
        if host_ptr:
            n = max(self.slab_used, 1)
            self._host = np.ctypeslib.as_array(
                C.cast(host_ptr, C.POINTER(C.c_int32)), shape=(n,))

    # ---- host views (zero-copy into the pinned mirror until detach/release)
    def _span(self, l, g, name):
        return self._spans[l][g][N.FIELDS.index(name)]

    def numpy(self, l, g, name):
        if self._host is None:
            raise RuntimeError("slicer was created with host_output=0: use .tensor()")
        off, ln = self._span(l, g, name)
        return self._host[off:off + ln]

    def frontier_numpy(self, l):
        """Frontier l as a host array.  Frontiers are by-products (not part of a
        reference Sample): the single-GPU slicer keeps them on the device only and
        this copies one out on demand."""
        off, ln = self._frontier[l]
        if self._frontier_on_device:
            if l not in self._fcache:
                if self._released:
                    raise RuntimeError("sample slot already released")
                out = np.empty(ln, dtype=np.int32)
                N.check(N.lib().gs_slicer_read_frontier(self._slicer._h, self.slot, int(l), out.ctypes.data))
                self._fcache[l] = out
            return self._fcache[l]
        if self._host is None:
            raise RuntimeError("slicer was created with host_output=0: use .frontier_tensor()")
        return self._host[off:off + ln]

    # ---- device views (torch is plumbing here: it only wraps the pointer)
    def _dev(self, off, ln, base=None):
        import torch
        if self._released:
            raise RuntimeError("sample slot already released")
        if ln == 0:
            return torch.empty(0, dtype=torch.int32, device=f"cuda:{self._slicer.graph.device}")
        return torch.as_tensor(_DevSpan((self._dev_ptr if base is None else base) + 4 * off, ln, self),
                               device=f"cuda:{self._slicer.graph.device}")

    def tensor(self, l, g, name):
        return self._dev(*self._span(l, g, name))

    def frontier_tensor(self, l):
        return self._dev(*self._frontier[l], base=self._fdev_ptr)

    def features(self, g):
        """Fused gather result for part g: rows of the last layer's in_nodes_g,
        [len(in_nodes_g), dim] fp32 device tensor (zero-copy view of the slot's feature slab)."""
        import torch
        if self._released:
            raise RuntimeError("sample slot already released")
        ptr, dim = self._slicer._feature_slab(self.slot)
        a, b = self.feature_part_off[g], self.feature_part_off[g + 1]
        dev = f"cuda:{self._slicer.graph.device}"
        if b == a:
            return torch.empty((0, dim), dtype=torch.float32, device=dev)

        class _F:
            pass
        f = _F()
        f.__cuda_array_interface__ = {"shape": (b - a, dim), "typestr": "<f4",
                                      "data": (ptr + 4 * a * dim, False), "version": 2,
                                      "strides": None}
        f._owner = self
        return torch.as_tensor(f, device=dev)

    def field_len(self, l, g, name):
        return self._span(l, g, name)[1]

    def detach(self):
        """Copy the host mirror out of the slot and give the slot back."""
        if not self._released:
            if self._host is not None:
                self._host = self._host.copy()
            if self._frontier_on_device:
                for l in range(self.num_layers + 1):
                    self.frontier_numpy(l)
            self.release()

    def release(self):
        if not self._released:
            self._released = True
            self._slicer._release(self.slot)

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass

    def algorithmic_bytes(self, owner_bytes=0, w=4):
        """Compulsory traffic of this sample per SURVEY 8(d): every input read
        once, every output written once (ids/positions `w` bytes, CSR row
        offsets of the graph 8 bytes each, own csr_indptr `w` bytes)."""
        L, P = self.num_layers, self.num_parts
        total = 0
        for l in range(L):
            F, Fn = self.frontier_len[l], self.frontier_len[l + 1]
            E = sum(self.field_len(l, g, "csr_indices") for g in range(P))
            total += F * (w + 16 + owner_bytes) + E * (w + owner_bytes) + Fn * w
            for g in range(P):
                n = {k: self.field_len(l, g, k) for k in N.FIELDS}
                total += (n["in_nodes"] + n["out_nodes"] + n["owned_out_nodes"] + n["from_ids"] +
                          n["to_ids"] + n["self_ids_in"] + n["self_ids_out"]) * w
                total += n["csr_indptr"] * w + n["csr_indices"] * w
        return total


class Slicer:
    """gs_slicer: per-slot workspaces + CUDA graphs over one Graph."""

    def __init__(self, graph, fanouts, num_parts=4, max_batch=4096, num_slots=4, num_streams=4,
                 seed=0, owner=None, draws=None, host_output=True, use_graph=True,
                 frontier_capacity=0, cand_capacity=0, slab_capacity=0, grid_ctas=0, l2_hints=True,
                 features=None, feature_rows_capacity=0, replace=True, host_copy_engine=True, group_size=1):
        """replace=False: the picks of a row are distinct (GS_RNG_PHILOX_NOREPLACE,
        the Python sibling's DGL sampling rule, python/utils/sampler.py:187).
        group_size: minibatches per launch sequence (must divide num_slots): submit()
        stages, the submit that fills a group launches it; wait() on a staged slot
        launches the partly filled group.  A replayed draw stream forces 1."""
        L = N.lib()
        self.graph = graph
        cfg = N.Config()
        L.gs_config_default(C.byref(cfg))
        cfg.num_layers = len(fanouts)
        for i, f in enumerate(fanouts):
            cfg.fanouts[i] = int(f)
        cfg.num_parts = int(num_parts)
        cfg.seed = int(seed)
        if not replace:
            if draws is not None:
                raise ValueError("replace=False cannot be combined with an explicit draw stream")
            cfg.rng_mode = N.RNG_PHILOX_NOREPLACE
        cfg.max_batch, cfg.num_slots, cfg.num_streams = int(max_batch), int(num_slots), int(num_streams)
        cfg.host_output, cfg.use_graph = int(bool(host_output)), int(bool(use_graph))
        cfg.frontier_capacity, cfg.cand_capacity = int(frontier_capacity), int(cand_capacity)
        cfg.slab_capacity, cfg.grid_ctas = int(slab_capacity), int(grid_ctas)
        cfg.l2_hints = int(bool(l2_hints))
        cfg.host_copy_engine = int(bool(host_copy_engine))
        cfg.group_size = 1 if draws is not None else int(group_size)
        self.group_size = cfg.group_size
        self.features = features
        if features is not None:
            cfg.features = features._h
            cfg.feature_rows_capacity = int(feature_rows_capacity)
        keep = []
        if owner is not None:
            o = np.ascontiguousarray(owner, dtype=np.int32)
            if o.size != graph.num_nodes:
                raise ValueError("owner map must have num_nodes entries")
            keep.append(o)
            cfg.owner_host = o.ctypes.data
        if draws is not None:
            d = np.ascontiguousarray(draws, dtype=np.uint32)
            keep.append(d)
            cfg.rng_mode, cfg.draws_host, cfg.n_draws = N.RNG_DRAWS, d.ctypes.data, d.size
        h = C.c_void_p()
        N.check(L.gs_slicer_create(graph._h, C.byref(cfg), C.byref(h)))
        self._h = h
        self.fanouts, self.num_parts, self.num_slots = list(fanouts), int(num_parts), int(num_slots)
        self.launches_per_sample = L.gs_slicer_launches_per_sample(h)
        self.slot_bytes = L.gs_slicer_slot_bytes(h)

    def slab_capacity_words(self):
        """int32 elements a slot's slab can hold: an upper bound for a hand-over block."""
        return int(N.lib().gs_slicer_slab_capacity(self._h))

    def submit(self, seeds, batch_index=0):
        s = np.ascontiguousarray(seeds, dtype=np.int64)
        slot = C.c_int32(-1)
        N.check(N.lib().gs_slicer_submit(self._h, s.ctypes.data, s.size, int(batch_index),
                                         C.byref(slot)))
        return slot.value

    def submit_direct(self, seeds, batch_index, dst_blocks, block_words, after_event=0):
        """gs_slicer_submit_direct: the sample's kernels write part g's lists straight into
        dst_blocks[g] (a ctypes array of num_parts device pointers, local or peer memory)."""
        s = np.ascontiguousarray(seeds, dtype=np.int64)
        slot = C.c_int32(-1)
        N.check(N.lib().gs_slicer_submit_direct(self._h, s.ctypes.data, s.size, int(batch_index), dst_blocks,
                                                int(block_words), C.c_void_p(after_event), C.byref(slot)))
        return slot.value

    def flush(self):
        """Launch the group being filled with the minibatches staged so far."""
        N.check(N.lib().gs_slicer_flush(self._h))

    def wait(self, slot):
        hp = C.POINTER(N.SampleHeader)()
        try:
            N.check(N.lib().gs_slicer_wait(self._h, slot, C.byref(hp)))
        except N.SlicerError:
            N.lib().gs_slicer_release(self._h, slot)
            raise
        dev, host = C.c_void_p(), C.c_void_p()
        N.check(N.lib().gs_slicer_slab(self._h, slot, C.byref(dev), C.byref(host)))
        return SampleView(self, slot, hp.contents, dev.value or 0, host.value or 0)

    def wait_header(self, slot):
        """Wait without building views (bench inner loop): -> (edges, slab_used)."""
        hp = C.POINTER(N.SampleHeader)()
        N.check(N.lib().gs_slicer_wait(self._h, slot, C.byref(hp)))
        return hp.contents

    def _release(self, slot):
        if getattr(self, "_h", None):
            N.lib().gs_slicer_release(self._h, slot)

    def _feature_slab(self, slot):
        ptr, dim = C.c_void_p(), C.c_int32()
        N.check(N.lib().gs_slicer_feature_slab(self._h, slot, C.byref(ptr), C.byref(dim)))
        return ptr.value, dim.value

    def _frontier_base(self, slot):
        ptr = C.c_void_p()
        N.check(N.lib().gs_slicer_frontiers(self._h, slot, C.byref(ptr)))
        return ptr.value or 0

    release = _release

    def slice(self, seeds, batch_index=0):
        return self.wait(self.submit(seeds, batch_index))

    def profile(self, seeds, batch_index=0):
        """-> list of (kernel name, ms) for one eager run with events between launches."""
        s = np.ascontiguousarray(seeds, dtype=np.int64)
        cap = 4 + 4 * N.MAX_LAYERS
        ms = (C.c_float * cap)()
        names = (C.c_char_p * cap)()
        n = C.c_int32(0)
        N.check(N.lib().gs_slicer_profile(self._h, s.ctypes.data, s.size, int(batch_index), ms,
                                          names, cap, C.byref(n)))
        return [(names[i].decode(), float(ms[i])) for i in range(n.value)]

    def close(self):
        if getattr(self, "_h", None):
            N.lib().gs_slicer_destroy(self._h)
            self._h = None

    __del__ = close


def reference_shuffle(nodes, reset):
    """In-place WorkerPool.cpp:40 epoch shuffle (libstdc++ random_shuffle + rand())."""
    assert nodes.dtype == np.int64 and nodes.flags.c_contiguous
    N.check(N.lib().gs_reference_shuffle(nodes.ctypes.data, nodes.size, int(bool(reset))))
    return nodes
