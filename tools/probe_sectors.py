"""Random-sector rate of this GPU (gs_bench_sector_probe) for a few table footprints."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "occ-gnn_b200")); sys.path.insert(0, ROOT)
import bench
for mb in (4, 10, 40, 118, 247, 1000, 4000):
    r = bench.sector_probe(0, mb * 1000 * 1000, 1 << 26)
    print(f"table {mb:5d} MB: {r / 1e9:7.2f} G reads/s = {r * 32 / 1e12:5.2f} TB/s of 32-byte sectors", flush=True)
