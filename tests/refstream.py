"""Canonical serialisation of one sample in the reference's own field order
(pybipartite.h:12-28), so the compiled reference, the CPU oracle and the CUDA
path can be compared through one SHA-256."""
import hashlib

import numpy as np

REF_ORDER = ["in_nodes", "indptr", "out_nodes", "owned_out_nodes", "indices", "from_ids",
             "to_ids", "self_ids_in", "self_ids_out"]


def _emit(chunks, a):
    a = np.asarray(a, dtype="<i8")
    chunks.append(np.array([a.size], dtype="<i8").tobytes())
    chunks.append(a.tobytes())


def _emit_part(chunks, get, g, P):
    """get(name) -> 1-D integer array for the reference-named field of part g."""
    _emit(chunks, get("in_nodes"))
    _emit(chunks, np.ones(len(get("out_nodes")), np.int64))  # bipartite.h:61-65
    _emit(chunks, get("out_nodes"))
    _emit(chunks, get("owned_out_nodes"))
    _emit(chunks, np.zeros(0, np.int64))                     # indices: never filled
    for name in ("from_ids", "to_ids"):
        for s in range(P):
            _emit(chunks, get(name) if s == g else np.zeros(0, np.int64))
    _emit(chunks, get("self_ids_in"))
    _emit(chunks, get("self_ids_out"))
    chunks.append(np.array([g], dtype="<i8").tobytes())


def digest_reference(sample):
    """sample: oracle.ref_slice() element (the real reference's output)."""
    chunks = []
    for layer in sample:
        P = len(layer)
        for g, d in enumerate(layer):
            _emit(chunks, d["in_nodes"])
            _emit(chunks, d["indptr"])
            _emit(chunks, d["out_nodes"])
            _emit(chunks, d["owned_out_nodes"])
            _emit(chunks, d["indices"])
            for name in ("from_ids", "to_ids"):
                for s in range(P):
                    _emit(chunks, d[name][s])
            _emit(chunks, d["self_ids_in"])
            _emit(chunks, d["self_ids_out"])
            chunks.append(np.array([d["gpu_id"]], dtype="<i8").tobytes())
    return hashlib.sha256(b"".join(chunks)).hexdigest()


def digest_oracle(res, layers=None):
    """res: OracleSlicer.slice() result."""
    chunks = []
    for layer in res["layers"][:layers]:
        P = len(layer)
        for g, d in enumerate(layer):
            _emit_part(chunks, lambda n, d=d: d[n], g, P)
    return hashlib.sha256(b"".join(chunks)).hexdigest()


def digest_view(view, layers=None):
    """view: cslicer.engine.SampleView (CUDA path, host mirror)."""
    chunks = []
    L = view.num_layers if layers is None else layers
    for l in range(L):
        for g in range(view.num_parts):
            _emit_part(chunks, lambda n, l=l, g=g: view.numpy(l, g, n), g, view.num_parts)
    return hashlib.sha256(b"".join(chunks)).hexdigest()


def digest_module_sample(s):
    """s: a `cslicer.sample` read through its reference-named attributes only."""
    chunks = []
    for layer in s.layers:
        P = len(layer)
        for b in layer:
            _emit(chunks, b.in_nodes)
            _emit(chunks, b.indptr)
            _emit(chunks, b.out_nodes)
            _emit(chunks, b.owned_out_nodes)
            _emit(chunks, b.indices)
            for lists in (b.from_ids, b.to_ids):
                assert len(lists) == P
                for x in lists:
                    _emit(chunks, x)
            _emit(chunks, b.self_ids_in)
            _emit(chunks, b.self_ids_out)
            chunks.append(np.array([b.gpu_id], dtype="<i8").tobytes())
    return hashlib.sha256(b"".join(chunks)).hexdigest()
