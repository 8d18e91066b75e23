"""Generates tests/golden/*.npz + golden.json by RUNNING THE UNMODIFIED
REFERENCE (oracle/_ref/libcslicer_ref.so, built from /root/reference/cslicer by
oracle/Makefile).  Run in the build container only:  python tests/golden/make_golden.py

Each case stores the graph itself (so nothing depends on an RNG library version),
the batches the reference was fed (its own epoch permutation, WorkerPool.cpp:40-49)
and one SHA-256 per sample over every field in the reference's order
(tests/refstream.py).  Cases:
  lowdeg   500 nodes, max degree 9  -> the reference's rng-free all-neighbours branch
  mixed    2000 nodes, power-law degrees up to ~950 -> sampled branch, mt19937 replay
  hubs     300 nodes, every degree >= 10 -> every row draws
"""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "occ-gnn_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

from cslicer.dataset import write_dataset  # noqa: E402
from cslicer.synth import low_degree_graph, synth_graph  # noqa: E402
from oracle import oracle as O  # noqa: E402
import refstream  # noqa: E402


def dense_graph(n, lo, hi, seed):
    rng = np.random.default_rng(seed)
    rows, cols = [], []
    for i in range(n):
        d = int(rng.integers(lo, hi + 1))
        c = rng.choice(n - 1, size=d, replace=False)
        c = np.sort(np.where(c >= i, c + 1, c))
        rows.append(d)
        cols.append(c)
    indptr = np.zeros(n + 1, np.int64)
    indptr[1:] = np.cumsum(rows)
    return indptr, np.concatenate(cols).astype(np.int64)


def main():
    assert O.ref_available(), "build oracle/_ref first (make -C oracle ref)"
    cases = {
        "lowdeg": (low_degree_graph(500, 9, seed=7), 64),
        "mixed": (synth_graph("tiny", seed=1234), 256),
        "hubs": (dense_graph(300, 10, 40, seed=3), 50),
    }
    meta = {}
    for name, ((ip, ix), mb) in cases.items():
        n = ip.size - 1
        perm = O.ref_epoch_permutations(n, 1)[0]
        batches = O.cut_batches(perm, mb)
        d = tempfile.mkdtemp(prefix="golden_")
        write_dataset(d, ip, ix)
        ref = O.ref_slice(d, batches)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), indptr=ip, indices=ix, perm=perm)
        meta[name] = {"num_nodes": int(n), "num_edges": int(ix.size), "minibatch": mb,
                      "layers": 3, "fanout": 10, "parts": 4,
                      "max_degree": int(np.diff(ip).max()),
                      "sha256": [refstream.digest_reference(s) for s in ref]}
        print(name, n, ix.size, len(batches), "batches")
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
