"""Random-gather ceiling on this GPU: 4-byte and 8-byte random reads from tables
of several sizes (torch index_select as a well-tuned gather), reported as
elements/s and as 32-byte-sector GB/s.  Diagnostic only."""
import torch, time
dev = "cuda"
def run(table_elems, dtype, n=1 << 26):
    t = torch.zeros(table_elems, dtype=dtype, device=dev)
    idx = torch.randint(0, table_elems, (n,), device=dev, dtype=torch.int64)
    for _ in range(3): torch.index_select(t, 0, idx)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): out = torch.index_select(t, 0, idx)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    es = t.element_size()
    print(f"table {table_elems*es/1e6:8.1f} MB {str(dtype):12s}: {n/ms/1e6:7.2f} G elem/s, "
          f"sector traffic {n*32/ms/1e6:7.1f} GB/s, stream traffic (idx+out) {n*(8+es)/ms/1e6:6.1f} GB/s", flush=True)
for elems, dt in [(2_449_029, torch.int64), (2_449_029 * 8, torch.int64), (61_857_105, torch.int32),
                  (500_000_000, torch.int32), (2_000_000_000, torch.int32)]:
    run(elems, dt)
