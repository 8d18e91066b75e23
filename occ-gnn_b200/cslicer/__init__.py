"""`cslicer` -- drop-in for the reference's pybind module of the same name.

Mirrors `PYBIND11_MODULE(cslicer, m)` (/root/reference/cslicer/pyfrontend.cpp:116-147):

    from cslicer import cslicer
    csl = cslicer(graphname, queue_size, no_worker_threads, number_of_epochs, minibatch_size)
    for i in range(csl.getNoSamples()):
        s = csl.getSample()          # s.layers[layer][gpu] -> bipatite
        b = s.layers[0][2]
        b.in_nodes, b.indptr, b.out_nodes, b.owned_out_nodes, b.indices,
        b.from_ids, b.to_ids, b.self_ids_in, b.self_ids_out, b.gpu_id

All slicing runs in hand-written sm_100a CUDA kernels behind the C ABI of
include/gsplit_slicer.h (libgsplit_slicer.so); this package only stages batches
and wraps results.  There is no CPU fallback.
"""
import collections
import os
import weakref

import numpy as np

from . import _native
from .dataset import Dataset, write_dataset  # noqa: F401
from .engine import Features, Graph, SampleView, Slicer, reference_shuffle  # noqa: F401

__all__ = ["cslicer", "sample", "bipatite", "test_list", "test_pyfront", "Graph", "Slicer", "Features",
           "Dataset", "write_dataset"]

# pyfrontend.cpp:24 hard-codes this root; overridable here by env or keyword.
DEFAULT_ROOT = "/data/sandeep/"
REF_LAYERS, REF_PARTS, REF_FANOUT = 3, 4, 10  # slicer.cpp:75, bipartite.h:19, slicer.cpp:10


class bipatite:  # sic: the name the reference registers (pyfrontend.cpp:128)
    """One layer's block on one GPU (PyBipartite, pybipartite.h:10-33).

    Attribute reads return fresh Python lists, as pybind's def_readwrite on an
    STL member does; `.numpy(name)` / `.tensor(name)` give zero-copy int32 views
    (host mirror / device).  With compat=True (default) `indptr` and `indices`
    reproduce the reference's values (all ones / empty); the repaired
    CSR is always available as `csr_indptr` / `csr_indices`.
    """

    _LISTS = ("in_nodes", "out_nodes", "owned_out_nodes", "self_ids_in", "self_ids_out",
              "csr_indptr", "csr_indices")

    def __init__(self, view=None, layer=0, part=0, num_parts=REF_PARTS, compat=True):
        self.__dict__["_v"] = view
        self.__dict__["_l"] = layer
        self.__dict__["_g"] = part
        self.__dict__["_P"] = num_parts
        self.__dict__["_compat"] = compat
        self.__dict__["_set"] = {}
        self.__dict__["gpu_id"] = part

    def numpy(self, name):
        if self._v is None:
            return np.zeros(0, np.int32)
        return self._v.numpy(self._l, self._g, name)

    def tensor(self, name):
        return self._v.tensor(self._l, self._g, name)

    def __getattr__(self, name):
        st = self.__dict__["_set"]
        if name in st:
            return st[name]
        if name in bipatite._LISTS:
            return self.numpy(name).tolist()
        if name in ("from_ids", "to_ids"):
            # only slot [own gpu] is ever filled (slicer.cpp:41-42)
            out = [[] for _ in range(self._P)]
            out[self._g] = self.numpy(name).tolist()
            return out
        if name == "indptr":
            if self._compat:
                return [1] * len(self.numpy("out_nodes"))  # bipartite.h:61-65
            return self.numpy("csr_indptr").tolist()
        if name == "indices":
            if self._compat:
                return []  # never filled in the reference (bipartite.h:17)
            return self.numpy("csr_indices").tolist()
        raise AttributeError(name)

    def __setattr__(self, name, value):  # def_readwrite: attributes are assignable
        self.__dict__["_set"][name] = value


class sample:
    """PySample (pybipartite.h:35-44): .layers[layer][gpu] -> bipatite."""

    def __init__(self, view=None, num_layers=REF_LAYERS, num_parts=REF_PARTS, compat=True):
        self._view = view
        if view is not None:
            num_layers, num_parts = view.num_layers, view.num_parts
        self.layers = [[bipatite(view, l, g, num_parts, compat) for g in range(num_parts)]
                       for l in range(num_layers)]

    # extras beyond the reference surface
    @property
    def view(self):
        return self._view

    def input_nodes(self):
        """Global ids whose features the sample reads (last frontier)."""
        return self._view.frontier_numpy(self._view.num_layers)

    def release(self):
        if self._view is not None:
            self._view.release()


class cslicer:
    """CSlicer (pyfrontend.cpp:25-89): same five positional arguments.

    `queue_size` -> sample slots in flight, `no_worker_threads` -> concurrent CUDA
    streams (each slot is one pre-captured CUDA graph).  Keyword extensions:
    root, dataset, fanouts, num_parts, owner_map ('mod' | 'range' | 'metis' | array),
    seed, device, draws (explicit 32-bit draw stream: replay mode), compat,
    shuffle ('reference' | 'none'), host_output, capacities, replace (False = distinct
    picks per row, the Python sibling's DGL rule, python/utils/sampler.py:187).
    """

    def __init__(self, name, queue_size, no_worker_threads, number_of_epochs, minibatch_size, *,
                 root=None, dataset=None, fanouts=None, num_layers=None, num_parts=REF_PARTS,
                 owner_map="mod", seed=0, device=0, draws=None, compat=True, shuffle="reference",
                 host_output=True, use_graph=True, frontier_capacity=0, cand_capacity=0,
                 slab_capacity=0, replace=True):
        if dataset is None:
            root = root or os.environ.get("GSPLIT_DATA_ROOT", DEFAULT_ROOT)
            dataset = Dataset(os.path.join(root, name), load_node_data=(owner_map == "metis"))
        self.name, self.dataset = name, dataset
        if fanouts is None:
            fanouts = [REF_FANOUT] * (num_layers or REF_LAYERS)
        self.fanouts, self.num_parts, self.compat = list(fanouts), int(num_parts), compat
        self.number_of_epochs, self.minibatch_size = int(number_of_epochs), int(minibatch_size)
        self.num_nodes = dataset.num_nodes
        owner = None
        if isinstance(owner_map, str):
            if owner_map == "range":
                from .synth import owner_range
                owner = owner_range(self.num_nodes, self.num_parts)
            elif owner_map == "metis":
                owner = dataset.partition_map
                if owner is None:
                    raise ValueError("dataset has no partition_map_opt.bin")
            elif owner_map != "mod":  # pyfrontend.cpp:57
                raise ValueError("owner_map: 'mod', 'range', 'metis' or an int array")
        else:
            owner = owner_map
        self.graph = Graph(dataset.indptr, dataset.indices, device)
        slots = max(2, int(queue_size))
        self.slicer = Slicer(self.graph, self.fanouts, self.num_parts, self.minibatch_size,
                             num_slots=slots, num_streams=max(1, int(no_worker_threads)),
                             seed=seed, owner=owner, draws=draws, host_output=host_output,
                             use_graph=use_graph, frontier_capacity=frontier_capacity,
                             cand_capacity=cand_capacity, slab_capacity=slab_capacity,
                             replace=replace)
        self._shuffle = shuffle
        self._nodes = np.arange(self.num_nodes, dtype=np.int64)  # WorkerPool.cpp:13-16
        self._per_epoch = (self.num_nodes - 1) // self.minibatch_size + 1
        self._total = self._per_epoch * self.number_of_epochs
        self._submitted = self._returned = 0
        self._inflight = collections.deque()  # slots in submission order
        self._live = collections.deque()      # weakrefs to samples handed out
        self._fill()

    # WorkerPool.cpp:37-51
    def _next_batch(self):
        e, j = divmod(self._submitted, self._per_epoch)
        if j == 0 and self._shuffle == "reference":
            reference_shuffle(self._nodes, reset=(e == 0))
        lo = j * self.minibatch_size
        return self._nodes[lo:lo + self.minibatch_size]

    def _fill(self):
        while self._submitted < self._total and len(self._inflight) + self._live_count() < \
                self.slicer.num_slots:
            b = self._next_batch()
            self._inflight.append(self.slicer.submit(b, self._submitted))
            self._submitted += 1

    def _live_count(self):
        # _live holds weak references to the SampleViews (not the samples): a slot stays taken
        # for as long as ANY object -- the sample or a bipatite taken from it -- keeps its view alive
        while self._live and (self._live[0]() is None or self._live[0]()._released):
            self._live.popleft()
        return sum(1 for r in self._live if r() is not None and not r()._released)

    def getNoSamples(self):  # pyfrontend.cpp:80-83
        return self._total

    def getSample(self):  # pyfrontend.cpp:73-78
        if self._returned >= self._total:
            raise RuntimeError("cslicer: all getNoSamples() samples were already returned")
        if not self._inflight:
            # every slot is pinned by a sample the caller still holds: copy the
            # oldest one out of its slot (the reference deep-copies every sample)
            for r in list(self._live):
                v = r()
                if v is not None and not v._released:
                    v.detach()
                    break
            self._fill()
        slot = self._inflight.popleft()
        view = self.slicer.wait(slot)
        s = sample(view, compat=self.compat)
        self._live.append(weakref.ref(view))
        self._returned += 1
        if self._live_count() >= self.slicer.num_slots and self._returned < self._total:
            for r in list(self._live):
                o = r()
                if o is not None and o is not view and not o._released:
                    o.detach()
                    break
        self._fill()
        return s

    def close(self):
        for slot in self._inflight:
            self.slicer.release(slot)
        self._inflight.clear()
        self.slicer.close()
        self.graph.close()


def test_list(l):  # pyfrontend.cpp:91-105
    l.append(10)
    return [1, 2, 3, 4]


def test_pyfront():  # pyfrontend.cpp:107-110: an empty 3 x 4 sample
    return sample(None)
