"""Multi-GPU layout of the slicing path: whole minibatches are the unit that
shards.  Rank r of R slices batches r, r+R, r+2R, ... of
the epoch's batch list; nothing is exchanged while slicing.  The reference has no
multi-process mode (one process, N threads popping one queue, WorkerPool.cpp:29-33);
this is the one-process-per-GPU form of the same "whoever is free takes the next
batch" model, made deterministic."""
import numpy as np


def rank_batch_ids(num_batches, rank, world):
    """Batch indices rank `rank` of `world` slices, in order."""
    if not (0 <= rank < world):
        raise ValueError("rank outside [0, world)")
    return np.arange(rank, num_batches, world, dtype=np.int64)


def epoch_batches(perm, minibatch, rank=0, world=1):
    """WorkerPool.cpp:41-49 batch cut of one epoch permutation, restricted to this
    rank: -> list of (global batch index, seed ids)."""
    nb = (len(perm) - 1) // minibatch + 1
    return [(int(b), perm[b * minibatch:(b + 1) * minibatch]) for b in rank_batch_ids(nb, rank, world)]


def reduce_job(dist, device, elapsed_ms, counters):
    """Whole-job timing: MAX of the ranks' device times, SUM of their counters.
    dist = torch.distributed (initialised) or None for a single process."""
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(elapsed_ms), [int(c) for c in counters]
    t = torch.tensor([float(elapsed_ms)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    c = torch.tensor([int(x) for x in counters], dtype=torch.int64, device=device)
    dist.all_reduce(c, op=dist.ReduceOp.SUM)
    return float(t[0]), [int(x) for x in c.tolist()]
