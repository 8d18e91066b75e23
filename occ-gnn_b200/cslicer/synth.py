"""Seeded synthetic graphs of the shapes BASELINE.json names.

Degree sequence: discretised Pareto(alpha) scaled to the target mean and
capped at N-1; destinations uniform at random; rows sorted; no self loops and
no duplicate edges (as `convert_dgl_dataset.py:44-47` leaves them).  torch is
used as an array library only (CPU here, the GPU on the bench box for the
60 M - 1.6 B edge shapes); the result is a host CSR in int64.
"""
import numpy as np
import torch

SHAPES = {
    # name: (num_nodes, target_edges, alpha)
    "arxiv": (169_343, 1_166_243, 2.0),
    "products": (2_449_029, 61_859_140, 1.8),
    "reddit": (232_965, 114_848_857, 2.5),
    "papers100M": (111_059_956, 1_615_685_872, 1.8),
    "tiny": (2_000, 16_000, 2.0),
    "small": (20_000, 300_000, 1.8),
}


def power_law_degrees(num_nodes, mean_degree, alpha, gen, device, cap=None):
    cap = float(num_nodes - 1 if cap is None else cap)
    u = torch.rand(num_nodes, generator=gen, device=device, dtype=torch.float64)
    raw = (1.0 - u).clamp_min(1e-12).pow(-1.0 / alpha)
    scale = mean_degree / float(raw.clamp_max(cap).mean())
    for _ in range(8):  # fixed point for the cap's effect on the mean
        scale *= mean_degree / float((raw * scale).clamp_max(cap).mean())
    jitter = torch.rand(num_nodes, generator=gen, device=device, dtype=torch.float64)
    return (raw * scale).clamp_max(cap).add_(jitter).floor_().to(torch.int64)


def synth_graph(shape="tiny", seed=1234, device="cpu", num_nodes=None, num_edges=None,
                alpha=None, max_degree=None, chunk=1 << 27, rows=None):
    """-> (indptr int64[N+1], indices int64[M]) as numpy arrays on the host.

    rows=(lo, hi): keep only the rows lo <= r < hi of the SAME graph (the generator
    is replayed in full on `device`, only that shard reaches the host) ->
    (indptr_local int64[hi-lo+1], indices_local int64[m_local], num_nodes, num_edges_global);
    the shard a rank of the partitioned slicer holds under a contiguous-range owner map."""
    if isinstance(shape, str):
        n0, m0, a0 = SHAPES[shape]
    else:
        n0, m0, a0 = shape
    n = int(num_nodes or n0)
    m = int(num_edges or m0)
    alpha = float(alpha or a0)
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    deg = power_law_degrees(n, m / n, alpha, gen, dev, cap=max_degree)
    # Generate (row, dst) keys in row chunks to bound peak memory at the 1.6 B shape.
    row_end = torch.cumsum(deg, 0)
    out_deg = torch.zeros(n, dtype=torch.int64, device=dev)
    pieces = []
    shard = None if rows is None else (int(rows[0]), int(rows[1]))
    r0 = 0
    while r0 < n:
        base = int(row_end[r0 - 1]) if r0 else 0
        r1 = int(torch.searchsorted(row_end, torch.tensor(base + chunk, device=dev)))
        r1 = min(max(r1, r0 + 1), n)
        d = deg[r0:r1]
        rows = torch.repeat_interleave(torch.arange(r0, r1, device=dev), d)
        dst = torch.randint(0, n, (rows.numel(),), generator=gen, device=dev)
        keep = dst != rows
        key = (rows[keep] * n + dst[keep]).unique()  # sorted by (row, dst), duplicates dropped
        rows_u = torch.div(key, n, rounding_mode="floor")
        out_deg[r0:r1] = torch.bincount(rows_u - r0, minlength=r1 - r0)
        nbr = key - rows_u * n
        if shard is not None:
            nbr = nbr[(rows_u >= shard[0]) & (rows_u < shard[1])]
        if nbr.numel() or shard is None:
            pieces.append(nbr.cpu())
        del rows, dst, keep, key, rows_u, nbr
        r0 = r1
    if shard is not None:
        lo, hi = shard
        ip = torch.zeros(hi - lo + 1, dtype=torch.int64)
        ip[1:] = torch.cumsum(out_deg[lo:hi].cpu(), 0)
        ix = torch.cat(pieces) if pieces else torch.zeros(0, dtype=torch.int64)
        return ip.numpy(), ix.numpy(), n, int(out_deg.sum())
    indptr = torch.zeros(n + 1, dtype=torch.int64)
    indptr[1:] = torch.cumsum(out_deg.cpu(), 0)
    indices = torch.cat(pieces) if len(pieces) > 1 else pieces[0]
    return indptr.numpy(), indices.numpy()


def low_degree_graph(num_nodes, max_degree, seed=7):
    """Every row has < max_degree+1 neighbours: with max_degree <= 9 the
    reference takes its rng-free all-neighbours branch (slicer.cpp:11-14)."""
    rng = np.random.default_rng(seed)
    deg = rng.integers(0, max_degree + 1, size=num_nodes)
    rows = np.repeat(np.arange(num_nodes), deg)
    dst = rng.integers(0, num_nodes, size=rows.size)
    keep = dst != rows
    key = np.unique(rows[keep] * num_nodes + dst[keep])
    r = key // num_nodes
    indptr = np.zeros(num_nodes + 1, np.int64)
    indptr[1:] = np.cumsum(np.bincount(r, minlength=num_nodes))
    return indptr, (key - r * num_nodes).astype(np.int64)


def owner_range(num_nodes, num_parts):
    """Contiguous-range owner map id*P//N (the locality-friendly one)."""
    return ((np.arange(num_nodes, dtype=np.int64) * num_parts) // num_nodes).astype(np.int32)
