"""CPU restatement of the reference's split-parallel shuffle.

TEST INFRASTRUCTURE -- a checker, never on the product path (only tests/ and
bench.py's CPU legs import oracle/).

What it follows, citations relative to /root/reference:

  * src/gnn/dist_sage.cu:193-199  `mergeKernel`: one CUDA block per received row x, one
    thread per feature y:  dest[indices[x] * dim + y] += src[x * dim + y].
  * src/gnn/dist_sage.cu:120-135  the merge loop: for dest in 0..P-1, for src in 0..3,
    src != dest: merge(temp[src][dest], temp[dest][dest], local_to_local_t) -- the rows a
    GPU received are added to its OWN partial rows peer after peer, in ascending src
    order (fp32 adds in that fixed order).
  * src/gnn/dist_sage.cu:100-110  the move before it: temp[src][dest] is a copy, on dest,
    of the rows src computed for destinations dest owns.
  * python/data/bipartite.py:88-99  `pull_for_remotes` / `push_from_remotes`: the Python
    sibling of the same step: rows are selected on the sender by position in its
    out_nodes (FROM side) and land on the owner by position in its out_nodes (TO side).
    (The sibling's push averages pairwise, `(local + remote) / 2`, and carries the comment
    "global_ids is wrong !!" -- bipartite.py:94-98; the CUDA engine's += is what the
    cslicer lists were designed for: to_ids = merge_indices, batch_slice_multi_gpu.py:183.)
  * python/layers/dist_sageconv.py:50-66  the order of the phases: local gather on every
    GPU, then every (src, dest) pair's pull + push, then the self concat.

In cslicer's lists (slicer.cpp:36-43): part g's `from_ids` are the positions in
out_nodes_g of rows g computed for OTHER owners; the owner q finds them through
`to_ids_q`.  The per-peer split of the two lists (bipartite.h:19-21 declares
from_ids[4] / to_ids[4]; the reference fills only the own slot) pairs entry k of
sender g's from_ids[q] with entry k of owner q's to_ids[g]; an owner without a row
of its own for that node holds -1 there (duplicate.cpp:35-39) and the partial row is
dropped -- what mergeKernel's assert(indices[x] < dest_dim1) would trip on.
"""
import numpy as np


def merge_kernel(src_rows, dest_rows, indices):
    """dist_sage.cu:193-199, thread for thread: block x, thread y."""
    n, dim = src_rows.shape
    assert indices.shape[0] == n                      # merge(): indices->s.dim1 == src->s.dim1 (:206)
    for x in range(n):                                # blockIdx.x
        assert indices[x] < dest_rows.shape[0]        # :197
        for y in range(dim):                          # threadIdx.x
            dest_rows[indices[x], y] = np.float32(dest_rows[indices[x], y] + src_rows[x, y])


def merge_rows(src_rows, dest_rows, indices):
    """The same update, vectorised (indices of one peer are distinct: a node has one row)."""
    if indices.size:
        dest_rows[indices] = dest_rows[indices] + src_rows


def shuffle_forward(parts, from_ids, to_ids, kernel=merge_rows):
    """parts[g]: [len(out_nodes_g), dim] fp32 partial aggregates of GPU g.
    from_ids[g][q], to_ids[q][g]: paired position lists (sender g, owner q).
    -> list of merged row blocks (dist_sage.cu:100-135: move, then merge per dest in
    ascending src order).  Entries with to_ids == -1 are dropped."""
    P = len(parts)
    out = [np.array(p, dtype=np.float32, copy=True) for p in parts]
    for dest in range(P):                             # :122
        for src in range(P):                          # :127, src != dest
            if src == dest:
                continue
            f, t = np.asarray(from_ids[src][dest]), np.asarray(to_ids[dest][src])
            assert f.shape == t.shape
            moved = np.asarray(parts[src], dtype=np.float32)[f]   # :100-110 copy of the sender's rows (pre-merge values)
            keep = t >= 0
            kernel(moved[keep], out[dest], t[keep])
    return out


def shuffle_backward(grads, from_ids, to_ids):
    """The transpose (autograd of pull + push with += on the owner): the gradient of an
    owner's row flows back unchanged to every contributor's partial row; a contributor
    whose partial row was dropped (to_ids == -1) receives zeros."""
    P = len(grads)
    out = [np.array(g, dtype=np.float32, copy=True) for g in grads]
    for dest in range(P):
        for src in range(P):
            if src == dest:
                continue
            f, t = np.asarray(from_ids[src][dest]), np.asarray(to_ids[dest][src])
            rows = np.zeros((f.size, grads[src].shape[1]), np.float32)
            keep = t >= 0
            rows[keep] = np.asarray(grads[dest], dtype=np.float32)[t[keep]]
            out[src][f] = rows
    return out


def peer_lists(sample, layer, owner_of):
    """Per-peer split of from_ids / to_ids of one oracle sample (the slots [q != g] the
    reference declares and leaves empty): from_ids[g][q] = positions in out_nodes_g of
    the rows owned by q, in frontier order; to_ids[q][g] = position of the same node in
    out_nodes_q, or -1."""
    lay = sample["layers"][layer]
    P = len(lay)
    pos = [{int(v): i for i, v in enumerate(lay[q]["out_nodes"])} for q in range(P)]
    fr = [[[] for _ in range(P)] for _ in range(P)]
    to = [[[] for _ in range(P)] for _ in range(P)]
    for nd in sample["frontiers"][layer]:
        nd = int(nd)
        q = int(owner_of(nd))
        for g in range(P):
            if g != q and nd in pos[g]:
                fr[g][q].append(pos[g][nd])
                to[q][g].append(pos[q].get(nd, -1))
    as_arr = lambda m: [[np.asarray(x, dtype=np.int64) for x in row] for row in m]
    return as_arr(fr), as_arr(to)
