"""TEST INFRASTRUCTURE -- CPU restatement (numpy) of Sampler.edge_partitioning, the Python
sibling's per-GPU edge routing (reference python/utils/sampler.py:64-134).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this file; the product is
occ-gnn_b200/csrc/epart.cu.

Pinned: tests/golden/epart.npz holds inputs and outputs of the UNMODIFIED reference method run
in the build container (tests/golden/make_epart_golden.py); tests/test_epart_oracle.py holds
this restatement to them entry for entry."""
import numpy as np


def _unique(a):
    return np.unique(a)  # sorted, like torch.unique (sampler.py:80,86,112,119)


def edge_partitioning(blocks, owner, mask, num_parts=4):
    """blocks: [(src_ids, dest_ids)] per layer (sampler.py:72,92); owner = workload_assignment;
    mask = memory_manager.node_gpu_mask [N, P].  -> per layer, per gpu: dict(src, dest,
    owned_nds, shuffle_nds {remote: ids}, device)."""
    owner = np.asarray(owner)
    out = []
    L = len(blocks)
    for layer in range(L - 1):                                    # sampler.py:70
        src, dst = (np.asarray(x) for x in blocks[layer])
        parts = []
        for g in range(num_parts):                                # :75
            sel = np.flatnonzero(owner[src] == g)                 # :76
            owned = _unique(dst[owner[dst] == g])                 # :80-81
            dsel = dst[sel]                                       # :82
            shuffle = {r: _unique(dsel[owner[dsel] == r]) for r in range(num_parts) if r != g}  # :83-87
            parts.append({"src": src[sel], "dest": dst[sel], "shuffle_nds": shuffle, "device": g, "owned_nds": owned})
        out.append(parts)
    src, dst = (np.asarray(x) for x in blocks[L - 1])             # :92
    emask = np.asarray(mask)[src, owner[dst]]                     # :94
    remote = np.flatnonzero(~emask)                               # :96
    natural = np.flatnonzero(emask)                               # :97
    parts = []
    for g in range(num_parts):                                    # :103
        nat = natural[owner[dst[natural]] == g]                   # :105
        ext = remote[owner[src[remote]] == g]                     # :106
        s_loc = np.concatenate([src[nat], src[ext]])              # :107
        d_loc = np.concatenate([dst[nat], dst[ext]])              # :110
        dn = _unique(dst)                                         # :112
        owned = dn[owner[dn] == g]                                # :113
        shuffle = {}
        if ext.size:                                              # :114
            for r in range(num_parts):
                if r != g:                                        # :116-117
                    shuffle[r] = _unique(dst[ext[owner[dst[ext]] == r]])  # :118-119
        parts.append({"src": s_loc, "dest": d_loc, "owned_nds": owned, "shuffle_nds": shuffle, "device": g})
    out.append(parts)
    return out
