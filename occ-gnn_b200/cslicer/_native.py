"""ctypes binding of the C ABI in include/gsplit_slicer.h (libgsplit_slicer.so).

The library is the product: it is built in-tree by `__graft_entry__.build()`
(nvcc, sm_100a) and there is NO fallback -- a missing library raises
ImportError, a missing GPU raises RuntimeError from the first call that needs it.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# GSPLIT_SLICER_LIB: another build of the same library (A/B of compile-time variants)
LIB_PATH = os.environ.get("GSPLIT_SLICER_LIB") or os.path.join(HERE, "libgsplit_slicer.so")

MAX_LAYERS, MAX_PARTS, ABI_VERSION = 8, 8, 2
RNG_PHILOX, RNG_DRAWS, RNG_PHILOX_NOREPLACE = 0, 1, 2
OK, ERR_INVALID, ERR_CUDA, ERR_BUSY, ERR_OVERFLOW, ERR_SEED, ERR_STATE, ERR_PEER = range(8)
OWNER_MOD, OWNER_RANGE = 0, 1
FIELDS = ["in_nodes", "out_nodes", "owned_out_nodes", "from_ids", "to_ids", "self_ids_in",
          "self_ids_out", "csr_indptr", "csr_indices"]
NUM_FIELDS = len(FIELDS)

# every symbol include/gsplit_slicer.h declares (tests check the export table)
SYMBOLS = ["gs_last_error", "gs_abi_version", "gs_config_default", "gs_graph_create",
           "gs_graph_load", "gs_graph_destroy", "gs_graph_num_nodes", "gs_graph_num_edges", "gs_slicer_create",
           "gs_slicer_destroy", "gs_slicer_submit", "gs_slicer_flush", "gs_slicer_wait", "gs_slicer_slab",
           "gs_slicer_release", "gs_slicer_profile", "gs_slicer_launches_per_sample",
           "gs_slicer_slot_bytes", "gs_slicer_slab_capacity", "gs_reference_shuffle", "gs_bench_sector_probe",
           "gs_unique_first_noalloc", "gs_epart_create", "gs_epart_destroy", "gs_epart_run", "gs_epart_get",
           "gs_features_create",
           "gs_features_destroy", "gs_features_gather", "gs_features_create_cached", "gs_features_info", "gs_slicer_feature_slab",
           "gs_slicer_frontiers", "gs_slicer_read_frontier", "gs_slicer_pack_parts", "gs_slicer_scatter_parts", "gs_slicer_submit_direct", "gs_slicer_push_parts", "gs_flags_post",
           "gs_ipc_alloc", "gs_ipc_open", "gs_ipc_close", "gs_ipc_free",
           # partitioned mode (one rank per GPU)
           "gs_pconfig_default", "gs_pslicer_create", "gs_pslicer_destroy", "gs_pslicer_ipc_handle",
           "gs_pslicer_connect_ipc", "gs_pslicer_connect_local", "gs_pslicer_ready",
           "gs_pslicer_num_segments", "gs_pslicer_submit", "gs_pslicer_stage",
           "gs_pslicer_enqueue", "gs_pslicer_commit", "gs_pslicer_wait", "gs_pslicer_slab",
           "gs_pslicer_release", "gs_pslicer_feature_slab", "gs_pslicer_shuffle", "gs_pslicer_sync", "gs_pslicer_reset", "gs_pslicer_launches_per_sample",
           "gs_pslicer_device_bytes"]


class Span(C.Structure):
    _fields_ = [("off", C.c_int32), ("len", C.c_int32)]


class SampleHeader(C.Structure):
    _fields_ = [("status", C.c_int32), ("num_layers", C.c_int32), ("num_parts", C.c_int32),
                ("n_seeds", C.c_int32), ("batch_index", C.c_uint64), ("draw_cursor", C.c_uint64),
                ("edges", C.c_int64), ("slab_used", C.c_int64),
                ("frontier_len", C.c_int32 * (MAX_LAYERS + 1)),
                ("cand_len", C.c_int32 * MAX_LAYERS),
                ("frontier", Span * (MAX_LAYERS + 1)),
                ("field", ((Span * NUM_FIELDS) * MAX_PARTS) * MAX_LAYERS),
                ("draw_base", C.c_uint64 * MAX_LAYERS),
                ("late_status", C.c_int32), ("feature_rows", C.c_int32),
                ("feature_part_off", C.c_int32 * (MAX_PARTS + 1)),
                ("reserved", C.c_int32 * 3)]


class Config(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("num_layers", C.c_int32),
                ("fanouts", C.c_int32 * MAX_LAYERS), ("num_parts", C.c_int32),
                ("rng_mode", C.c_int32), ("seed", C.c_uint64), ("owner_host", C.c_void_p),
                ("draws_host", C.c_void_p), ("n_draws", C.c_uint64), ("max_batch", C.c_int32),
                ("num_slots", C.c_int32), ("num_streams", C.c_int32), ("host_output", C.c_int32),
                ("frontier_capacity", C.c_int64), ("cand_capacity", C.c_int64),
                ("slab_capacity", C.c_int64), ("grid_ctas", C.c_int32), ("use_graph", C.c_int32),
                ("l2_hints", C.c_int32), ("host_copy_engine", C.c_int32),
                ("features", C.c_void_p), ("feature_rows_capacity", C.c_int64),
                ("group_size", C.c_int32), ("reserved0", C.c_int32)]


class PConfig(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("num_layers", C.c_int32),
                ("fanouts", C.c_int32 * MAX_LAYERS), ("num_parts", C.c_int32),
                ("rank", C.c_int32), ("owner_mode", C.c_int32),
                ("range_bounds", C.c_int32 * (MAX_PARTS + 1)), ("seed", C.c_uint64),
                ("max_batch", C.c_int32), ("num_slots", C.c_int32), ("host_output", C.c_int32),
                ("grid_ctas", C.c_int32), ("timeout_ms", C.c_int32), ("use_graph", C.c_int32),
                ("frontier_capacity", C.c_int64), ("cand_capacity", C.c_int64),
                ("slab_capacity", C.c_int64), ("num_edges_global", C.c_int64),
                ("stream", C.c_void_p), ("features_local_host", C.c_void_p),
                ("feature_dim", C.c_int32), ("rng_mode", C.c_int32),
                ("feature_rows_capacity", C.c_int64), ("shuffle_rows_capacity", C.c_int64),
                ("shuffle_max_dim", C.c_int32), ("reserved1", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (nvcc, sm_100a). There is no CPU or PyTorch fallback for the slicer.")
    L = C.CDLL(LIB_PATH)
    L.gs_last_error.restype = C.c_char_p
    L.gs_abi_version.restype = C.c_int
    L.gs_config_default.argtypes = [C.POINTER(Config)]
    L.gs_graph_create.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                  C.POINTER(C.c_void_p)]
    L.gs_graph_load.argtypes = [C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
    L.gs_graph_destroy.argtypes = [C.c_void_p]
    L.gs_graph_num_nodes.restype = C.c_int64
    L.gs_graph_num_nodes.argtypes = [C.c_void_p]
    L.gs_graph_num_edges.restype = C.c_int64
    L.gs_graph_num_edges.argtypes = [C.c_void_p]
    L.gs_slicer_create.argtypes = [C.c_void_p, C.POINTER(Config), C.POINTER(C.c_void_p)]
    L.gs_slicer_destroy.argtypes = [C.c_void_p]
    L.gs_slicer_submit.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64,
                                   C.POINTER(C.c_int32)]
    L.gs_slicer_flush.argtypes = [C.c_void_p]
    L.gs_slicer_wait.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.POINTER(SampleHeader))]
    L.gs_slicer_slab.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p),
                                 C.POINTER(C.c_void_p)]
    L.gs_slicer_release.argtypes = [C.c_void_p, C.c_int32]
    L.gs_slicer_frontiers.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]
    L.gs_slicer_read_frontier.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    L.gs_slicer_profile.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64,
                                    C.POINTER(C.c_float), C.POINTER(C.c_char_p), C.c_int32,
                                    C.POINTER(C.c_int32)]
    L.gs_slicer_launches_per_sample.argtypes = [C.c_void_p]
    L.gs_slicer_slot_bytes.restype = C.c_int64
    L.gs_slicer_slot_bytes.argtypes = [C.c_void_p]
    L.gs_slicer_slab_capacity.restype = C.c_int64
    L.gs_slicer_slab_capacity.argtypes = [C.c_void_p]
    L.gs_slicer_pack_parts.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]
    L.gs_slicer_scatter_parts.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.c_int64,
                                          C.c_void_p, C.c_void_p]
    L.gs_ipc_alloc.argtypes = [C.c_int, C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]
    L.gs_ipc_open.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
    L.gs_ipc_close.argtypes = [C.c_int, C.c_void_p]
    L.gs_ipc_free.argtypes = [C.c_int, C.c_void_p]
    L.gs_reference_shuffle.argtypes = [C.c_void_p, C.c_int64, C.c_int]
    L.gs_bench_sector_probe.argtypes = [C.c_int, C.c_int64, C.c_int64, C.POINTER(C.c_double)]
    L.gs_slicer_submit_direct.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64, C.c_void_p, C.c_int64,
                                          C.c_void_p, C.POINTER(C.c_int32)]
    L.gs_slicer_push_parts.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gs_flags_post.argtypes = [C.c_int, C.c_void_p, C.c_int32, C.c_uint32, C.c_void_p, C.c_int32, C.c_uint32,
                                C.c_uint32, C.c_void_p, C.c_void_p]
    L.gs_epart_create.argtypes = [C.c_int, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.POINTER(C.c_void_p)]
    L.gs_epart_destroy.argtypes = [C.c_void_p]
    L.gs_epart_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32]
    L.gs_epart_get.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p),
                               C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    L.gs_unique_first_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                          C.POINTER(C.c_int64)]
    L.gs_features_create.argtypes = [C.c_int, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                     C.POINTER(C.c_void_p)]
    L.gs_features_destroy.argtypes = [C.c_void_p]
    L.gs_features_create_cached.argtypes = [C.c_int, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                            C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
    L.gs_features_info.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                                   C.POINTER(C.c_int64)]
    L.gs_features_gather.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.gs_slicer_feature_slab.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p),
                                         C.POINTER(C.c_int32)]
    L.gs_pconfig_default.argtypes = [C.POINTER(PConfig)]
    L.gs_pslicer_create.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                    C.POINTER(PConfig), C.POINTER(C.c_void_p)]
    L.gs_pslicer_destroy.argtypes = [C.c_void_p]
    L.gs_pslicer_ipc_handle.argtypes = [C.c_void_p, C.c_void_p]
    L.gs_pslicer_connect_ipc.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.gs_pslicer_connect_local.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.gs_pslicer_ready.argtypes = [C.c_void_p]
    L.gs_pslicer_num_segments.argtypes = [C.c_void_p]
    L.gs_pslicer_submit.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64, C.POINTER(C.c_int32)]
    L.gs_pslicer_stage.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64, C.POINTER(C.c_int32)]
    L.gs_pslicer_enqueue.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    L.gs_pslicer_commit.argtypes = [C.c_void_p, C.c_int32]
    L.gs_pslicer_wait.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.POINTER(SampleHeader))]
    L.gs_pslicer_slab.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    L.gs_pslicer_release.argtypes = [C.c_void_p, C.c_int32]
    L.gs_pslicer_feature_slab.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p),
                                          C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.gs_pslicer_shuffle.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                                     C.c_int32]
    L.gs_pslicer_sync.argtypes = [C.c_void_p]
    L.gs_pslicer_reset.argtypes = [C.c_void_p]
    L.gs_pslicer_launches_per_sample.argtypes = [C.c_void_p]
    L.gs_pslicer_device_bytes.restype = C.c_int64
    L.gs_pslicer_device_bytes.argtypes = [C.c_void_p]
    if L.gs_abi_version() != ABI_VERSION:
        raise ImportError("libgsplit_slicer.so ABI version mismatch; rebuild")
    _lib = L
    return L


class SlicerError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def check(rc):
    if rc != OK:
        raise SlicerError(rc, lib().gs_last_error().decode() or f"gs error {rc}")
