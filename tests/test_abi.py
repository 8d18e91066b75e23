"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports
every symbol include/gsplit_slicer.h declares, ctypes struct layouts agree with
the header, host-only entry points work, and nothing falls back to a CPU path."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import cslicer
from cslicer import _native as N
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gsplit_slicer.h")


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = N.lib()
    names = header_functions()
    assert len(names) >= 17
    for n in names:
        assert hasattr(L, n), f"{n} declared in the header but not exported"
    assert sorted(N.SYMBOLS) == names
    assert L.gs_abi_version() == N.ABI_VERSION


def test_struct_layouts_match_the_header(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "gsplit_slicer.h"\n'
                   'int main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(gs_sample_header),'
                   'sizeof(gs_config), offsetof(gs_sample_header, field),'
                   'offsetof(gs_config, max_batch), offsetof(gs_sample_header, draw_base),'
                   'sizeof(gs_pconfig), offsetof(gs_pconfig, stream), offsetof(gs_pconfig, feature_rows_capacity));'
                   'return 0;}')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert got == [C.sizeof(N.SampleHeader), C.sizeof(N.Config), N.SampleHeader.field.offset,
                   N.Config.max_batch.offset, N.SampleHeader.draw_base.offset,
                   C.sizeof(N.PConfig), N.PConfig.stream.offset, N.PConfig.feature_rows_capacity.offset]


def test_config_default_is_the_reference_configuration():
    cfg = N.Config()
    N.lib().gs_config_default(C.byref(cfg))
    assert cfg.num_layers == 3 and cfg.num_parts == 4 and list(cfg.fanouts)[:3] == [10, 10, 10]


def test_reference_shuffle_matches_libstdcxx_loop():
    want = O.epoch_permutations(5000, 2)  # first: both sides drive the process-global rand()
    nodes = np.arange(5000, dtype=np.int64)
    cslicer.reference_shuffle(nodes, reset=True)
    assert np.array_equal(nodes, want[0])
    cslicer.reference_shuffle(nodes, reset=False)
    assert np.array_equal(nodes, want[1])


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ip = np.array([0, 1, 2], np.int64)
    ix = np.array([1, 0], np.int64)
    with pytest.raises(N.SlicerError) as e:
        cslicer.Graph(ip, ix)
    assert e.value.code == N.ERR_CUDA and "no CPU fallback" in str(e.value)


def test_argument_validation_is_reported_not_crashed():
    h = C.c_void_p()
    ip = np.array([0, 1, 3], np.int64)  # indptr[N] != num_edges
    ix = np.array([1, 0], np.int64)
    rc = N.lib().gs_graph_create(0, 2, 2, ip.ctypes.data, ix.ctypes.data, C.byref(h))
    assert rc == N.ERR_INVALID and b"indptr" in N.lib().gs_last_error()
    assert N.lib().gs_slicer_create(None, None, C.byref(h)) == N.ERR_INVALID


def test_dataset_round_trip_and_checksums(tmp_path):
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("tiny", seed=11)
    meta = cslicer.write_dataset(str(tmp_path / "g"), ip, ix)
    assert meta["csum_offsets"] == int(ip.sum()) and meta["csum_edges"] == int(ix.sum())
    ds = cslicer.Dataset(str(tmp_path / "g"))
    assert ds.num_nodes == ip.size - 1 and np.array_equal(ds.indptr, ip) and np.array_equal(ds.indices, ix)
    assert ds.partition_map is not None and ds.fsize == 1
    ix2 = ix.copy()
    ix2[0] += 1
    ix2.astype("<i8").tofile(str(tmp_path / "g" / "indices.bin"))
    with pytest.raises(ValueError):
        cslicer.Dataset(str(tmp_path / "g"))


def test_synthetic_graph_is_clean_csr():
    from cslicer.synth import synth_graph
    ip, ix = synth_graph("small", seed=2)
    n = ip.size - 1
    assert n == 20000 and abs(ix.size - 300000) < 15000
    rows = np.repeat(np.arange(n), np.diff(ip))
    assert (ix != rows).all() and ix.min() >= 0 and ix.max() < n
    key = rows * n + ix
    assert (np.diff(key) > 0).all()  # rows sorted, no duplicate edges
    ip2, ix2 = synth_graph("small", seed=2)
    assert np.array_equal(ip, ip2) and np.array_equal(ix, ix2)


def test_module_surface_of_the_reference():
    # pyfrontend.cpp:116-147
    for name in ("cslicer", "sample", "bipatite", "test_list", "test_pyfront"):
        assert hasattr(cslicer, name)
    l = [1, 2, 3]
    assert cslicer.test_list(l) == [1, 2, 3, 4] and l[-1] == 10
    s = cslicer.test_pyfront()
    assert len(s.layers) == 3 and all(len(x) == 4 for x in s.layers)
    b = s.layers[1][2]
    assert b.gpu_id == 2 and b.in_nodes == [] and b.from_ids == [[], [], [], []]
    b.in_nodes = [5]  # def_readwrite
    assert b.in_nodes == [5]


def test_round2_entry_points_validate_their_arguments_without_a_device():
    """The entry points added in round 2 reject bad arguments before they touch a GPU, and none
    of them computes on the CPU when there is no device."""
    import torch
    L = N.lib()
    h = C.c_void_p()
    own = np.zeros(10, np.int32)
    # edge partitioner
    assert L.gs_epart_create(0, 10, 9, own.ctypes.data, None, 100, C.byref(h)) == N.ERR_INVALID      # parts > 8
    assert L.gs_epart_create(0, 0, 4, own.ctypes.data, None, 100, C.byref(h)) == N.ERR_INVALID       # no nodes
    bad = own.copy(); bad[3] = 4
    assert L.gs_epart_create(0, 10, 4, bad.ctypes.data, None, 100, C.byref(h)) == N.ERR_INVALID      # owner outside [0, P)
    assert b"owner[3]" in L.gs_last_error()
    assert L.gs_epart_run(None, None, None, 0, 0) == N.ERR_INVALID
    # flag barrier
    assert L.gs_flags_post(0, None, 9, 1, None, 0, 0, 0, None, None) == N.ERR_INVALID                # > GS_MAX_PARTS words
    assert L.gs_flags_post(0, None, 0, 0, None, 2, 1, 0, None, None) == N.ERR_INVALID                # wait without words / status
    # direct / staged hand-over
    slot = C.c_int32(-1)
    assert L.gs_slicer_submit_direct(None, None, 0, 0, None, 0, None, C.byref(slot)) == N.ERR_INVALID
    assert L.gs_slicer_push_parts(None, 0, None, None, None) == N.ERR_INVALID
    if not torch.cuda.is_available():
        assert L.gs_epart_create(0, 10, 4, own.ctypes.data, None, 100, C.byref(h)) == N.ERR_CUDA
        assert b"no CPU fallback" in L.gs_last_error()
        cnt = C.c_int64(-1)
        one = np.zeros(4, np.int32)
        assert L.gs_unique_first_noalloc(0, one.ctypes.data, 4, 16, one.ctypes.data, C.byref(cnt)) == N.ERR_CUDA


def test_edge_partitioner_mirror_checks_shapes_before_the_library():
    from cslicer.epart import EdgePartitioner
    with pytest.raises(ValueError):
        EdgePartitioner(np.zeros(10, np.int32), np.ones((10, 3), bool), num_parts=4)
