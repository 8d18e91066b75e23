import torch, time
for mb in [4, 40, 256]:
    n = mb * 1024 * 1024
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    for _ in range(3): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    t = time.perf_counter()
    reps = 20
    for _ in range(reps): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"D2H {mb} MiB: {n*reps/dt/1e9:.1f} GB/s")
# two streams concurrently
d1 = torch.empty(40<<20, dtype=torch.uint8, device="cuda"); d2 = torch.empty_like(d1)
h1 = torch.empty(40<<20, dtype=torch.uint8).pin_memory(); h2 = torch.empty(40<<20, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(20):
    with torch.cuda.stream(s1): h1.copy_(d1, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t
print(f"D2H 2 streams: {2*20*(40<<20)/dt/1e9:.1f} GB/s")
