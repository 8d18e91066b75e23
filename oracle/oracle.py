"""oracle/oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-ends for
  * liborc.so              the CPU restatement (cslicer_oracle.c), parameterised;
  * _ref/libcslicer_ref.so the unmodified reference driven through ref_shim.cpp
                           (only the reference's own configuration: 3 layers,
                           fanout 10, 4 parts, owner = id % 4).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  Nothing here reads /root/reference at
run time; `build()` compiles from it when it is present.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liborc.so")
REF_LIB = os.path.join(HERE, "_ref", "libcslicer_ref.so")

MAX_LAYERS, MAX_PARTS = 8, 16
RNG_MT19937, RNG_PHILOX, RNG_DRAWS, RNG_PHILOX_NOREPLACE = 0, 1, 2, 3
FIELDS = ["in_nodes", "out_nodes", "owned_out_nodes", "from_ids", "to_ids",
          "self_ids_in", "self_ids_out", "indptr_compat", "csr_indptr", "csr_indices"]


def build(ref=True):
    """Compile the checker (and the reference, when its sources are present)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


class _Cfg(C.Structure):
    _fields_ = [("num_layers", C.c_int32), ("fanouts", C.c_int32 * MAX_LAYERS),
                ("num_parts", C.c_int32), ("rng_mode", C.c_int32), ("seed", C.c_uint64),
                ("owner", C.c_void_p), ("draws", C.c_void_p), ("n_draws", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build(ref=False)
        L = C.CDLL(LIB)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(_Cfg)]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_slice.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint64]
        L.orc_field_len.restype = C.c_int64
        L.orc_field_len.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_field_ptr.restype = C.POINTER(C.c_int64)
        L.orc_field_ptr.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_frontier_len.restype = C.c_int64
        L.orc_frontier_len.argtypes = [C.c_void_p, C.c_int]
        L.orc_frontier_ptr.restype = C.POINTER(C.c_int64)
        L.orc_frontier_ptr.argtypes = [C.c_void_p, C.c_int]
        L.orc_edges_last.restype = C.c_int64
        L.orc_edges_last.argtypes = [C.c_void_p]
        L.orc_draws_used.restype = C.c_uint64
        L.orc_draws_used.argtypes = [C.c_void_p]
        L.orc_epoch_permutations.argtypes = [C.c_int64, C.c_int, C.c_void_p]
        L.orc_pool_run.restype = C.c_double
        L.orc_pool_run.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(_Cfg),
                                   C.c_int, C.c_void_p, C.c_void_p, C.c_int64,
                                   C.POINTER(C.c_int64)]
        L.orc_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_philox_pick.restype = C.c_uint32
        L.orc_philox_pick.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int64, C.c_int]
        L.orc_noreplace_pick.restype = C.c_uint32
        L.orc_noreplace_pick.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int64, C.c_uint32, C.c_uint32]
        L.orc_mt19937_stream.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p]
        _lib = L
    return _lib


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _make_cfg(fanouts, num_parts, rng_mode, seed, owner, draws, keep):
    cfg = _Cfg()
    cfg.num_layers = len(fanouts)
    for i, f in enumerate(fanouts):
        cfg.fanouts[i] = int(f)
    cfg.num_parts = int(num_parts)
    cfg.rng_mode = int(rng_mode)
    cfg.seed = int(seed)
    if owner is not None:
        o = np.ascontiguousarray(owner, dtype=np.int32)
        keep.append(o)
        cfg.owner = o.ctypes.data
    if draws is not None:
        d = np.ascontiguousarray(draws, dtype=np.uint32)
        keep.append(d)
        cfg.draws = d.ctypes.data
        cfg.n_draws = d.size
    return cfg


class OracleSlicer:
    """One sequential restated slicer (its own masks and rng state)."""

    def __init__(self, indptr, indices, fanouts, num_parts=4, rng_mode=RNG_PHILOX,
                 seed=0, owner=None, draws=None):
        self._keep = []
        self.indptr, self.indices = _i64(indptr), _i64(indices)
        self.num_nodes = self.indptr.size - 1
        self.L, self.P = len(fanouts), int(num_parts)
        self._cfg = _make_cfg(fanouts, num_parts, rng_mode, seed, owner, draws, self._keep)
        self._h = lib().orc_create(self.num_nodes, self.indptr.ctypes.data,
                                   self.indices.ctypes.data, C.byref(self._cfg))
        if not self._h:
            raise ValueError("orc_create rejected the configuration")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    def slice(self, seeds, batch_index=0):
        """-> dict(layers=[[{field: np.int64 array}]*P]*L, frontiers=[...], edges=int)"""
        s = _i64(seeds)
        if lib().orc_slice(self._h, s.ctypes.data, s.size, int(batch_index)) != 0:
            raise ValueError("seed id out of range")
        L_ = lib()
        layers = []
        for l in range(self.L):
            parts = []
            for g in range(self.P):
                d = {}
                for fi, name in enumerate(FIELDS):
                    n = L_.orc_field_len(self._h, l, g, fi)
                    p = L_.orc_field_ptr(self._h, l, g, fi)
                    d[name] = np.ctypeslib.as_array(p, shape=(n,)).copy() if n else \
                        np.zeros(0, np.int64)
                d["gpu_id"] = g
                parts.append(d)
            layers.append(parts)
        fr = []
        for l in range(self.L + 1):
            n = L_.orc_frontier_len(self._h, l)
            p = L_.orc_frontier_ptr(self._h, l)
            fr.append(np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, np.int64))
        return dict(layers=layers, frontiers=fr, edges=int(L_.orc_edges_last(self._h)))

    @property
    def draws_used(self):
        return int(lib().orc_draws_used(self._h))


def epoch_permutations(num_nodes, epochs=1):
    out = np.empty((epochs, num_nodes), np.int64)
    lib().orc_epoch_permutations(num_nodes, epochs, out.ctypes.data)
    return out


def cut_batches(perm, minibatch):
    """WorkerPool.cpp:41-49: consecutive chunks, last one short."""
    return [perm[j:j + minibatch] for j in range(0, len(perm), minibatch)]


def _pack(batches):
    off = np.zeros(len(batches) + 1, np.int64)
    off[1:] = np.cumsum([len(b) for b in batches])
    flat = _i64(np.concatenate(batches)) if batches else np.zeros(0, np.int64)
    return flat, off


def pool_run(indptr, indices, fanouts, num_parts, batches, workers, rng_mode=RNG_MT19937,
             seed=0, owner=None):
    """Multithreaded restatement baseline -> (seconds, real edges)."""
    keep = []
    cfg = _make_cfg(fanouts, num_parts, rng_mode, seed, owner, None, keep)
    ip, ix = _i64(indptr), _i64(indices)
    flat, off = _pack(batches)
    e = C.c_int64(0)
    sec = lib().orc_pool_run(ip.size - 1, ip.ctypes.data, ix.ctypes.data, C.byref(cfg),
                             int(workers), flat.ctypes.data, off.ctypes.data, len(batches),
                             C.byref(e))
    return sec, int(e.value)


def philox4x32_10(ctr, key):
    c = np.asarray(ctr, np.uint32)
    k = np.asarray(key, np.uint32)
    o = np.zeros(4, np.uint32)
    lib().orc_philox4x32_10(c.ctypes.data, k.ctypes.data, o.ctypes.data)
    return o


def philox_pick(seed, batch_index, layer, node, slot):
    return int(lib().orc_philox_pick(seed, batch_index, layer, node, slot))


def noreplace_pick(seed, batch_index, layer, node, j, deg):
    return int(lib().orc_noreplace_pick(seed, batch_index, layer, node, j, deg))


def mt19937_stream(n, seed=5489):
    o = np.zeros(n, np.uint32)
    lib().orc_mt19937_stream(seed, n, o.ctypes.data)
    return o


# ------------------------------------------------------------- real reference
REF_FIELDS = ["in_nodes", "indptr", "out_nodes", "owned_out_nodes", "indices",
              "from_ids", "to_ids", "self_ids_in", "self_ids_out"]
_ref = None


def ref_available():
    return os.path.exists(REF_LIB)


def ref_lib():
    global _ref
    if _ref is None:
        R = C.CDLL(REF_LIB)
        R.ref_slice_batches.restype = C.c_int64
        R.ref_slice_batches.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_int64,
                                        C.POINTER(C.POINTER(C.c_int64))]
        R.ref_free.argtypes = [C.POINTER(C.c_int64)]
        R.ref_pool_time.restype = C.c_double
        R.ref_pool_time.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        R.ref_epoch_permutations.argtypes = [C.c_int64, C.c_int, C.c_void_p]
        _ref = R
    return _ref


def ref_slice(dataset_dir, batches):
    """Run the unmodified reference Slicer (one instance, rng state carried) on
    `batches` -> list of samples; sample[layer][part] = dict of np.int64 arrays
    (from_ids/to_ids: list of 4 arrays) + gpu_id."""
    batches = [b for b in batches if len(b)]
    flat, off = _pack(batches)
    out = C.POINTER(C.c_int64)()
    n = ref_lib().ref_slice_batches(dataset_dir.encode(), flat.ctypes.data, off.ctypes.data,
                                    len(batches), C.byref(out))
    buf = np.ctypeslib.as_array(out, shape=(max(n, 1),))[:n].copy()
    ref_lib().ref_free(out)
    pos = 0

    def take():
        nonlocal pos
        ln = int(buf[pos])
        a = buf[pos + 1:pos + 1 + ln]
        pos += 1 + ln
        return a

    samples = []
    for _ in batches:
        layers = []
        for _l in range(3):
            parts = []
            for _g in range(4):
                d = {}
                d["in_nodes"] = take()
                d["indptr"] = take()
                d["out_nodes"] = take()
                d["owned_out_nodes"] = take()
                d["indices"] = take()
                d["from_ids"] = [take() for _ in range(4)]
                d["to_ids"] = [take() for _ in range(4)]
                d["self_ids_in"] = take()
                d["self_ids_out"] = take()
                d["gpu_id"] = int(buf[pos])
                pos += 1
                parts.append(d)
            layers.append(parts)
        samples.append(layers)
    assert pos == n
    return samples


def ref_pool_time(dataset_dir, workers, batches):
    flat, off = _pack(batches)
    return ref_lib().ref_pool_time(dataset_dir.encode(), int(workers), flat.ctypes.data,
                                   off.ctypes.data, len(batches))


def ref_epoch_permutations(num_nodes, epochs=1):
    out = np.empty((epochs, num_nodes), np.int64)
    ref_lib().ref_epoch_permutations(num_nodes, epochs, out.ctypes.data)
    return out
