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


def bind_to_gpu_cpus(device):
    """Pin this process to the CPU cores next to GPU `device` (NVML's ideal-CPU
    mask) so the pinned sample mirrors it allocates afterwards -- first-touch
    pages -- land on the NUMA node the GPU's PCIe root hangs off.  With one
    process per GPU every rank writes 40-55 GB/s into host memory; without this
    all eight ranks share one node's memory controllers.  Returns the CPU list,
    or None when NVML / affinity is unavailable (never fatal)."""
    import os
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        # honour CUDA_VISIBLE_DEVICES: look the device up by PCI bus id, not by index
        bus = torch.cuda.get_device_properties(device).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(device), "pci_domain_id", 0)
        dev_id = getattr(torch.cuda.get_device_properties(device), "pci_device_id", 0)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{dom:08x}:{bus:02x}:{dev_id:02x}.0".encode())
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None
