"""D2H copy-engine rate of one GPU into pinned host memory; run several copies of
this at the same START_AT (epoch seconds) on different GPUs to see what the HOST
sustains in aggregate.  Diagnostic only."""
import os, time
import torch
dev = int(os.environ.get("DEV", "0"))
torch.cuda.set_device(dev)
mb = int(os.environ.get("MB", "40"))
secs = float(os.environ.get("SECS", "3"))
n = mb << 20
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h = torch.empty(n, dtype=torch.uint8).pin_memory()
for _ in range(3):
    h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
start = float(os.environ.get("START_AT", "0"))
while time.time() < start:
    pass
t0 = time.perf_counter()
reps = 0
while time.perf_counter() - t0 < secs:
    for _ in range(20):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    reps += 20
dt = time.perf_counter() - t0
print(f"GPU {dev}: D2H {mb} MiB x {reps}: {n * reps / dt / 1e9:.1f} GB/s", flush=True)
