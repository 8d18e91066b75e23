"""What the box's NVLink gives an all-to-all (diagnostic, torchrun): NCCL all_to_all_single of
equal blocks, and every rank storing blocks into its peers' IPC-mapped buffers with
cudaMemcpyAsync (copy engines) -- bytes SENT per rank per second."""
import ctypes as C
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from cslicer import _native as N

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = f"cuda:{local}"
for mb in (1, 4, 32):
    words = mb * (1 << 18)
    src = torch.ones(world * words, dtype=torch.int32, device=dev)
    dst = torch.empty_like(src)
    for _ in range(3):
        dist.all_to_all_single(dst, src)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        dist.all_to_all_single(dst, src)
    e1.record(); e1.synchronize()
    sent = 4 * words * (world - 1) * reps
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"NCCL all_to_all_single, {mb} MiB per peer: {sent / (t.item() * 1e-3) / 1e9:7.1f} GB/s sent per rank", flush=True)
# copy engines into IPC-mapped peer buffers
mb = 4
words = mb * (1 << 18)
base, handle = C.c_void_p(), (C.c_char * 64)()
N.check(N.lib().gs_ipc_alloc(local, 4 * world * words, C.byref(base), handle))
handles = [None] * world
dist.all_gather_object(handles, bytes(handle.raw))
peers = []
for q in range(world):
    if q == rank:
        peers.append(base.value)
    else:
        p = C.c_void_p()
        N.check(N.lib().gs_ipc_open(local, handles[q], C.byref(p)))
        peers.append(p.value)
src = torch.ones(world * words, dtype=torch.int32, device=dev)
rt = C.CDLL("libcudart.so.12")
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
streams = [torch.cuda.Stream() for _ in range(world)]
torch.cuda.synchronize(); dist.barrier()


def push(reps):
    for _ in range(reps):
        for q in range(world):
            if q == rank:
                continue
            rc = rt.cudaMemcpyAsync(peers[q] + 4 * rank * words, src.data_ptr() + 4 * q * words, 4 * words, 4, streams[q].cuda_stream)
            assert rc == 0, rc


push(3)
torch.cuda.synchronize(); dist.barrier()
import time
t0 = time.perf_counter()
push(20)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
t = torch.tensor([dt], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"copy engines -> peers' IPC buffers, {mb} MiB per peer: {4 * words * (world - 1) * 20 / t.item() / 1e9:7.1f} GB/s sent per rank", flush=True)
dist.barrier()
for q in range(world):
    if q != rank:
        N.lib().gs_ipc_close(local, C.c_void_p(peers[q]))
dist.barrier()
N.lib().gs_ipc_free(local, base)
dist.destroy_process_group()
