"""SURVEY 8(f) row N1 on the GPU: the CUDA edge partitioner (csrc/epart.cu through the C ABI)
against the golden vectors of the unmodified Sampler.edge_partitioning and, on larger seeded
inputs, against the CPU restatement.  Integer work: bit-exact, order included."""
import numpy as np
import pytest

from cslicer import _native as N
from oracle import sampler_oracle as SO
from test_epart_oracle import GOLD, golden_case, same_partitioning

pytestmark = pytest.mark.gpu


def _ep(owner, mask, max_edges=1 << 20):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from cslicer.epart import EdgePartitioner
    return EdgePartitioner(owner, mask, 4, max_edges=max_edges)


@pytest.mark.parametrize("name", [str(c) for c in GOLD["cases"]])
def test_cuda_partitioner_matches_the_reference_golden(name):
    owner, mask, blocks, want = golden_case(name)
    ep = _ep(owner, mask)
    same_partitioning(ep.edge_partitioning(blocks), want)
    same_partitioning(ep.edge_partitioning(blocks), want)  # the workspace (marks) is clean after a run
    ep.close()


@pytest.mark.parametrize("seed,n,e,extra", [(1, 200000, 700000, 0.25), (2, 50000, 900000, 0.0), (3, 1000, 5, 0.5),
                                            (4, 300000, 1 << 20, 1.0)])
def test_cuda_partitioner_matches_the_restatement(seed, n, e, extra):
    rng = np.random.default_rng(seed)
    owner = rng.integers(0, 4, n)
    mask = rng.random((n, 4)) < extra
    mask[np.arange(n), owner] = True
    blocks = [(rng.integers(0, n, e), rng.integers(0, n, e)) for _ in range(2)]
    ep = _ep(owner, mask, max_edges=1 << 20)
    same_partitioning(ep.edge_partitioning(blocks), SO.edge_partitioning(blocks, owner, mask))
    ep.close()


def test_empty_layer_and_bad_ids():
    owner = np.arange(100) % 4
    mask = np.ones((100, 4), bool)
    ep = _ep(owner, mask, max_edges=64)
    z = np.empty(0, np.int64)
    parts = ep.edge_partitioning([(z, z)])
    assert all(p["src"].size == 0 and p["owned_nds"].size == 0 and p["shuffle_nds"] == {} for p in parts[0])
    with pytest.raises(N.SlicerError) as err:
        ep.partition_layer(np.array([1, 2, 100]), np.array([3, 4, 5]), False)
    assert err.value.code == N.ERR_INVALID and "edge 2" in str(err.value)
    good = ep.partition_layer(np.array([1, 2, 7]), np.array([3, 4, 5]), False)
    same_partitioning([good], SO.edge_partitioning([(np.array([1, 2, 7]), np.array([3, 4, 5])), (z, z)], owner, mask)[:1])
    with pytest.raises(N.SlicerError):
        ep.partition_layer(np.arange(65), np.arange(65), False)  # beyond max_edges
    ep.close()
