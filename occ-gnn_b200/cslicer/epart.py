"""Host-side mirror of Sampler.edge_partitioning (reference python/utils/sampler.py:64-134)
over the CUDA edge partitioner (csrc/epart.cu, gs_epart_* of include/gsplit_slicer.h):
same arguments, same return structure -- per layer a list of 4 dicts
{"src", "dest", "owned_nds", "shuffle_nds", "device"} -- with numpy int32 arrays
where the reference has int64 tensors."""
import ctypes as C

import numpy as np

from . import _native as N

EP_SRC, EP_DST, EP_OWNED, EP_SHUFFLE, EP_NUM_EXTRA = range(5)


class EdgePartitioner:
    def __init__(self, workload_assignment, node_gpu_mask=None, num_parts=4, max_edges=1 << 22, device=0):
        """workload_assignment: owner of every node (sampler.py:13); node_gpu_mask: bool [N, P],
        the memory manager's cache residency (sampler.py:93) -- needed for the last layer."""
        own = np.ascontiguousarray(workload_assignment, dtype=np.int32)
        self.num_nodes, self.num_parts = own.size, int(num_parts)
        res = None
        if node_gpu_mask is not None:
            res = np.ascontiguousarray(node_gpu_mask, dtype=np.uint8)
            if res.shape != (own.size, self.num_parts):
                raise ValueError("node_gpu_mask must be [num_nodes, num_parts]")
        h = C.c_void_p()
        N.check(N.lib().gs_epart_create(int(device), own.size, self.num_parts, own.ctypes.data,
                                        res.ctypes.data if res is not None else None, int(max_edges), C.byref(h)))
        self._h = h

    def _get(self, part, what, remote=0):
        host, n = C.c_void_p(), C.c_int64()
        N.check(N.lib().gs_epart_get(self._h, part, what, remote, C.byref(host), None, C.byref(n)))
        if what == EP_NUM_EXTRA:
            return n.value
        if n.value == 0:
            return np.empty(0, np.int32)
        return np.ctypeslib.as_array(C.cast(host, C.POINTER(C.c_int32)), shape=(n.value,)).copy()

    def partition_layer(self, src_ids, dest_ids, last_layer):
        s = np.ascontiguousarray(src_ids, dtype=np.int64)
        d = np.ascontiguousarray(dest_ids, dtype=np.int64)
        if s.shape != d.shape:
            raise ValueError("src_ids and dest_ids differ in length")
        N.check(N.lib().gs_epart_run(self._h, s.ctypes.data, d.ctypes.data, s.size, int(bool(last_layer))))
        P = self.num_parts
        parts = []
        for g in range(P):
            # sampler.py:114: the last layer fills shuffle_nds only for a part with extra edges
            fill = (not last_layer) or self._get(g, EP_NUM_EXTRA) > 0
            shuffle = {r: self._get(g, EP_SHUFFLE, r) for r in range(P) if r != g} if fill else {}
            parts.append({"src": self._get(g, EP_SRC), "dest": self._get(g, EP_DST),
                          "owned_nds": self._get(g, EP_OWNED), "shuffle_nds": shuffle, "device": g})
        return parts

    def edge_partitioning(self, blocks, layers=None):
        """blocks: [(src_ids, dest_ids)] per layer, as Sampler.sample returns them."""
        L = len(blocks)
        return [self.partition_layer(s, d, l == L - 1) for l, (s, d) in enumerate(blocks)]

    def close(self):
        if getattr(self, "_h", None):
            N.lib().gs_epart_destroy(self._h)
            self._h = None

    __del__ = close
