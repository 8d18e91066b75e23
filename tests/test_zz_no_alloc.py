"""gs_unique_first_noalloc(): the allocation-free dedup entry point (static device
scratch, include/gsplit_slicer.h) against the existing entry points.  Collected
last on purpose.  The existing path it is held to: one seed whose neighbour list
is `ids`, one layer with fan-out -1 and one part -> frontier 1 of
gs_slicer_submit() is the seed followed by the distinct ids in order of first
occurrence (slicer.cpp:22-43).  Integer work: bit-exact."""
import ctypes as C

import numpy as np
import pytest

from cslicer import _native as N

MAX_NODES, MAX_IDS = 1 << 20, 1 << 24


def _unique_first(ids):
    _, first = np.unique(ids, return_index=True)
    return ids[np.sort(first)]


def _via_slicer(engine, ids, num_nodes):
    """The existing entry points: node 0 is the only seed, its CSR row is `ids` (all in [1, num_nodes))."""
    indptr = np.full(num_nodes + 1, ids.size, np.int64)
    indptr[0] = 0
    g = engine.Graph(indptr, ids.astype(np.int64))
    s = engine.Slicer(g, [-1], num_parts=1, max_batch=4, num_slots=1, num_streams=1)
    v = s.slice(np.array([0], np.int64))
    f1 = v.frontier_numpy(1).copy()
    v.release()
    s.close()
    g.close()
    assert f1[0] == 0
    return f1[1:]


def _no_alloc(torch, ids, num_nodes):
    d_ids = torch.from_numpy(ids.astype(np.int32)).cuda()
    d_out = torch.full((max(1, ids.size),), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    cnt = C.c_int64(-1)
    N.check(N.lib().gs_unique_first_noalloc(0, d_ids.data_ptr(), ids.size, num_nodes, d_out.data_ptr(),
                                            C.byref(cnt)))
    out = d_out.cpu().numpy()
    assert (out[cnt.value:] == -7).all(), "wrote past the reported count"
    return out[:cnt.value]


@pytest.mark.gpu
def test_no_alloc_dedup_matches_the_slicer_on_two_calls_in_a_row():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from cslicer import engine
    rng = np.random.default_rng(20261105)
    # second call: fewer nodes and other ids, so anything the first call left in the
    # shared static array would show; third: the first input again
    a = rng.integers(1, 20000, size=50000)
    b = rng.integers(1, 3000, size=70001)
    for ids, nn in ((a, 20000), (b, 3000), (a, 20000)):
        got = _no_alloc(torch, ids, nn)
        assert np.array_equal(got, _unique_first(ids).astype(np.int32))
        assert np.array_equal(got, _via_slicer(engine, ids, nn)), "differs from gs_slicer_submit's frontier"
    # ragged ends: one id, all equal, already distinct
    for ids in (np.array([5]), np.full(1000, 9), rng.permutation(4096)):
        assert np.array_equal(_no_alloc(torch, ids, 4096), _unique_first(ids).astype(np.int32))
    # the largest table the static scratch holds
    big = rng.integers(0, MAX_NODES, size=1 << 21)
    assert np.array_equal(_no_alloc(torch, big, MAX_NODES), _unique_first(big).astype(np.int32))
    # an id outside [0, num_nodes) is reported, and the next call is clean
    bad = a.copy()
    bad[1234] = 20000
    with pytest.raises(N.SlicerError) as e:
        _no_alloc(torch, bad, 20000)
    assert e.value.code == N.ERR_INVALID and "ids[1234]" in str(e.value)
    assert np.array_equal(_no_alloc(torch, a, 20000), _unique_first(a).astype(np.int32))


def test_no_alloc_rejects_sizes_beyond_the_static_scratch():
    cnt = C.c_int64(-1)
    L = N.lib()
    one = np.zeros(4, np.int32)  # never dereferenced: the size check comes first
    assert L.gs_unique_first_noalloc(0, one.ctypes.data, 4, MAX_NODES + 1, one.ctypes.data, C.byref(cnt)) \
        == N.ERR_INVALID
    assert b"GS_NOALLOC_MAX_NODES" in L.gs_last_error()
    assert L.gs_unique_first_noalloc(0, one.ctypes.data, MAX_IDS + 1, 16, one.ctypes.data, C.byref(cnt)) \
        == N.ERR_INVALID
    assert L.gs_unique_first_noalloc(0, None, 4, 16, one.ctypes.data, C.byref(cnt)) == N.ERR_INVALID
    assert L.gs_unique_first_noalloc(0, None, 0, 16, None, C.byref(cnt)) == N.OK and cnt.value == 0
