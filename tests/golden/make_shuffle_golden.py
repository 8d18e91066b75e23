"""Fixture for the split-parallel shuffle: a fixed 4-GPU layer
(out_nodes, from_ids, to_ids per peer, partial rows) pushed through the thread-for-thread
statement of the reference's mergeKernel (src/gnn/dist_sage.cu:193-199, merge order of
:120-135).  The reference's CUDA engine cannot run in the build container (no GPU, and it
needs its own cmake build), so the golden is produced by `oracle.shuffle_oracle.merge_kernel`
-- the literal per-block / per-thread loop -- and every faster path (numpy-vectorised
oracle, the CUDA shuffle) is held to it.  Run from the repo root:
    python tests/golden/make_shuffle_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import shuffle_oracle as S  # noqa: E402


def build():
    rng = np.random.default_rng(20261105)
    P, dim, n_nodes = 4, 8, 60
    frontier = rng.permutation(n_nodes)[:40]
    own = lambda v: v % P
    # every frontier node gets a row on a random non-empty set of GPUs
    lay = [{"out_nodes": []} for _ in range(P)]
    for nd in frontier:
        gpus = [g for g in range(P) if rng.random() < 0.55] or [int(rng.integers(0, P))]
        for g in gpus:
            lay[g]["out_nodes"].append(int(nd))
    sample = {"layers": [[{"out_nodes": np.asarray(x["out_nodes"], np.int64)} for x in lay]],
              "frontiers": [frontier]}
    fr, to = S.peer_lists(sample, 0, own)
    parts = [rng.standard_normal((len(lay[g]["out_nodes"]), dim)).astype(np.float32) for g in range(P)]
    fwd = S.shuffle_forward(parts, fr, to, kernel=S.merge_kernel)
    bwd = S.shuffle_backward(parts, fr, to)
    out = {"frontier": frontier, "P": P, "dim": dim}
    for g in range(P):
        out[f"out_nodes_{g}"] = sample["layers"][0][g]["out_nodes"]
        out[f"part_{g}"] = parts[g]
        out[f"fwd_{g}"] = fwd[g]
        out[f"bwd_{g}"] = bwd[g]
        for q in range(P):
            out[f"from_{g}_{q}"] = fr[g][q]
            out[f"to_{g}_{q}"] = to[g][q]
    return out


if __name__ == "__main__":
    z = build()
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "shuffle.npz"), **z)
    print("wrote shuffle.npz:", {k: getattr(v, "shape", v) for k, v in z.items() if k.startswith("fwd")})
