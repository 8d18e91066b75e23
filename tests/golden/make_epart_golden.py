"""Golden vectors for SURVEY 8(f) row N1: run the UNMODIFIED Sampler.edge_partitioning of the
reference (python/utils/sampler.py:64-134) in the build container and commit its inputs and
outputs.  dgl is not installed here and edge_partitioning does not use it: the two imports at
the top of sampler.py (dgl.sampling, data.bipartite) are satisfied with empty stand-in modules,
and `self` is a stand-in object with the two attributes the method reads
(workload_assignment, memory_manager.node_gpu_mask).
  python tests/golden/make_epart_golden.py      # writes tests/golden/epart.npz
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/python"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_sampler():
    for name in ("dgl", "dgl.sampling", "data", "data.bipartite"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["dgl.sampling"].sample_neighbors = None
    sys.modules["data.bipartite"].BipartiteGraph = None
    sys.path.insert(0, REF)
    from utils import sampler  # noqa: E402  (the reference's own file)
    return sampler.Sampler


def make_case(seed, n_nodes, batch, fanouts, cache_extra):
    """Blocks shaped like Sampler.sample's (sampler.py:176-207): per layer the sampled edges
    plus one self loop per frontier node; the next frontier is the sorted union."""
    rng = np.random.default_rng(seed)
    owner = rng.integers(0, 4, n_nodes)
    mask = rng.random((n_nodes, 4)) < cache_extra
    mask[np.arange(n_nodes), owner] = True  # a GPU always holds its own partition (memory_manager.py)
    last = np.unique(rng.choice(n_nodes, batch, replace=False))
    blocks = []
    for fan in fanouts:
        seeds = np.repeat(last, rng.integers(0, fan + 1, last.size))
        nbrs = rng.integers(0, n_nodes, seeds.size)
        src = np.concatenate([last, nbrs])   # "self_loop_dests" of sampler.py:201: neighbours + self
        dst = np.concatenate([last, seeds])
        blocks.append((src.astype(np.int64), dst.astype(np.int64)))
        last = np.unique(src)
    return owner.astype(np.int64), mask, blocks


def main():
    Sampler = load_reference_sampler()
    out = {}
    cases = [("a", 11, 3000, 64, [5, 4, 3], 0.3), ("b", 12, 500, 40, [3, 3], 0.0), ("c", 13, 2000, 100, [6], 0.6)]
    out["cases"] = np.array([c[0] for c in cases])
    for name, seed, n, batch, fanouts, extra in cases:
        owner, mask, blocks = make_case(seed, n, batch, fanouts, extra)
        stub = types.SimpleNamespace(workload_assignment=torch.from_numpy(owner),
                                     memory_manager=types.SimpleNamespace(node_gpu_mask=torch.from_numpy(mask)))
        tb = [(torch.from_numpy(s), torch.from_numpy(d)) for s, d in blocks]
        parts = Sampler.edge_partitioning(stub, tb, None)
        out[f"{name}.owner"] = owner
        out[f"{name}.mask"] = mask
        out[f"{name}.layers"] = np.int64(len(blocks))
        for l, (s, d) in enumerate(blocks):
            out[f"{name}.{l}.src"], out[f"{name}.{l}.dst"] = s, d
            for pe in parts[l]:
                g = pe["device"]
                out[f"{name}.{l}.{g}.src"] = pe["src"].numpy()
                out[f"{name}.{l}.{g}.dest"] = pe["dest"].numpy()
                out[f"{name}.{l}.{g}.owned_nds"] = pe["owned_nds"].numpy()
                out[f"{name}.{l}.{g}.shuffle_keys"] = np.array(sorted(pe["shuffle_nds"].keys()), np.int64)
                for r, v in pe["shuffle_nds"].items():
                    out[f"{name}.{l}.{g}.shuffle_nds.{r}"] = v.numpy()
    np.savez_compressed(os.path.join(HERE, "epart.npz"), **out)
    print("wrote epart.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
