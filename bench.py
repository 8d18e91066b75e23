#!/usr/bin/env python
"""bench.py -- sliced minibatches/s of the cslicer hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): ogbn-products-shaped synthetic graph
(2,449,029 nodes, ~61.9 M edges, seed 1234), 3 layers, fanout [15,10,5], batch
4096, split into `--parts` bipartites per layer (default 4, the reference's
value).  One step = `--mb-per-step` minibatches pushed through the slicer's slot
pipeline.  Whole minibatches are the sharding unit: at N > 1 every rank holds a
replica of the graph topology (0.3 GB of a 180 GB HBM) and slices its own
minibatches, no collective on the data path -> "scaling": "weak".

own arm      : CUDA slicer through the C ABI (cslicer.engine).
                 value = results left in HBM (header only to the host);
                 e2e   = host seeds in, every Bipartite list mirrored to pinned
                         host memory inside the timed region.
--impl reference : the reference algorithm on the host cores (multithreaded CPU
                 restatement, one slicer per thread as WorkerPool.cpp:29-33; the
                 compiled reference itself cannot express fanout [15,10,5]).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

# partitioned mode runs several lanes (streams) whose barrier kernels must be co-resident
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "occ-gnn_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

FANOUTS = [15, 10, 5]
BATCH = 4096
SHAPE = "products"
GRAPH_SEED = 1234
CACHE = os.path.join(ROOT, ".bench_cache")
HEADER_BYTES = 4888  # sizeof(gs_sample_header); checked against the library in own_arm


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# Libraries (NCCL's version banner, ...) write to fd 1; the contract is ONE JSON
# line on stdout, so fd 1 is pointed at stderr for the run and the line goes to
# the saved descriptor.
_REAL_STDOUT = None


def capture_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    if isinstance(obj, dict) and obj.get("n_gpus", 1) > 1 and "hand_over" in obj:
        # N > 1: the verdicts and every leg's value right behind the headline, so that they survive a
        # record that keeps only the head (or only the tail) of the line
        legs = {"hand_over": obj["hand_over"].get("value"), "hand_over_exchange": obj["hand_over"].get("headline_exchange"),
                "hand_over_parity": obj["hand_over"].get("parity"),
                "partitioned": (obj.get("partitioned") or {}).get("value"),
                "partitioned_parity": (obj.get("partitioned") or {}).get("parity"),
                "replicas": (obj.get("replicas") or {}).get("value")}
        head = ("metric", "value", "unit", "n_gpus", "parity")
        obj = {**{k: obj[k] for k in head if k in obj}, "legs": legs, **{k: v for k, v in obj.items() if k not in head},
               "legs_again": legs}
    line = json.dumps(obj) + "\n"
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        sys.stdout.write(line)
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line.encode())


def load_graph(shape, device):
    """Seeded synthetic CSR; cached on disk so the two arms share one graph."""
    from cslicer.synth import synth_graph
    os.makedirs(CACHE, exist_ok=True)
    f = os.path.join(CACHE, f"{shape}_{GRAPH_SEED}_{device.split(':')[0]}.npz")
    if os.path.exists(f):
        try:
            z = np.load(f)
            return z["indptr"], z["indices"]
        except Exception:
            pass
    t = time.time()
    ip, ix = synth_graph(shape, seed=GRAPH_SEED, device=device)
    log(f"[bench] generated {shape}: N={ip.size - 1} M={ix.size} in {time.time() - t:.1f}s on {device}")
    if int(os.environ.get("RANK", "0")) == 0 and ix.size < 200_000_000:
        try:
            tmp = f + f".tmp{os.getpid()}.npz"
            np.savez(tmp, indptr=ip, indices=ix)
            os.replace(tmp, f)
        except Exception as e:  # cache is best effort
            log("[bench] cache write failed:", e)
    return ip, ix


def graph_stats(shape, ip):
    """state alpha and the maximum degree of the generated graph."""
    from cslicer.synth import SHAPES
    deg = np.diff(ip)
    return {"degree_law": "discretised Pareto, destinations uniform", "alpha": SHAPES.get(shape, (0, 0, None))[2],
            "mean_degree": round(float(deg.mean()), 2), "max_degree": int(deg.max()),
            "rows_with_degree_0": int((deg == 0).sum())}


def host_description():
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return {"cpu_model": model, "nproc": os.cpu_count(), "usable_cores": host_cores()}


def load_graph_cpu_arm(shape):
    """The graph for a CPU-only arm: this process must not create a CUDA context
    (the driver classes an arm by the devices it touched).  The cached file is used
    when present; otherwise a CHILD process generates it (on the GPU when there is
    one: the 62 M-edge shape takes minutes on the host) and exits before any timing."""
    f = os.path.join(CACHE, f"{shape}_{GRAPH_SEED}_cuda.npz")
    fc = os.path.join(CACHE, f"{shape}_{GRAPH_SEED}_cpu.npz")
    for cand in (f, fc):
        if os.path.exists(cand):
            try:
                z = np.load(cand)
                return z["indptr"], z["indices"]
            except Exception:
                pass
    code = ("import sys; sys.argv=['bench.py']; import bench, torch; "
            f"bench.load_graph({shape!r}, 'cuda:0' if torch.cuda.is_available() else 'cpu')")
    subprocess.run([sys.executable, "-c", code], cwd=ROOT, check=True, stdout=subprocess.DEVNULL)
    for cand in (f, fc):
        if os.path.exists(cand):
            z = np.load(cand)
            return z["indptr"], z["indices"]
    raise RuntimeError("graph generation in the child process left no cache file")


def seed_batches(num_nodes, count, rank, world):
    """Reference behaviour (WorkerPool.cpp:40-49): a fixed permutation of all
    nodes cut into batches; rank r takes batches r, r+world, ..."""
    rng = np.random.default_rng(GRAPH_SEED + 1)
    perm = rng.permutation(num_nodes).astype(np.int64)
    per_epoch = num_nodes // BATCH
    out = []
    for i in range(count):
        b = (rank + i * world) % per_epoch
        out.append((b, perm[b * BATCH:(b + 1) * BATCH]))
    return out


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(gpu_index)], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f:
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


def run_pipeline(sl, batches, steps, mb_per_step, torch, timed=True):
    """Push steps*mb_per_step minibatches through the slot pipeline.
    -> (device ms between two CUDA events bracketing the region, wall s,
        total edges, total slab int32 elements)."""
    n = steps * mb_per_step
    depth = sl.num_slots
    inflight = []
    edges = slab = 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(n):
        if len(inflight) == depth:
            slot = inflight.pop(0)
            h = sl.wait_header(slot)
            edges += h.edges
            slab += h.slab_used
            sl.release(slot)
        bi, seeds = batches[i % len(batches)]
        inflight.append(sl.submit(seeds, bi))
    for slot in inflight:
        h = sl.wait_header(slot)
        edges += h.edges
        slab += h.slab_used
        sl.release(slot)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), time.perf_counter() - t0, edges, slab


def kernel_profile(sl, batches, reps=8):
    """Eager runs of ONE minibatch with a CUDA event between launches (outside the
    timed region) -> [(kernel@layer, mean ms)], launch order preserved."""
    acc, order = {}, []
    for r in range(reps):
        bi, seeds = batches[r % len(batches)]
        layer = -1
        for name, ms in sl.profile(seeds, bi):
            if name.startswith("rows"):
                layer += 1
            key = name if "finish" in name else f"{name}@{layer}"
            if r == 0:
                order.append(key)
            acc[key] = acc.get(key, 0.0) + ms
    return [(k, acc[k] / reps) for k in order]


def kernel_bytes(view, P):
    """Bytes each launch of one sample must move (its own inputs and outputs, w=4;
    DESIGN section 5) -- the per-kernel roofline numerator."""
    out = {}
    L = view.num_layers
    tail = {}
    for l in range(L):
        F, Fn, C = view.frontier_len[l], view.frontier_len[l + 1], view.cand_len[l]
        E = sum(view.field_len(l, g, "csr_indices") for g in range(P))
        n_in = sum(view.field_len(l, g, "in_nodes") for g in range(P))
        n_out = sum(view.field_len(l, g, "out_nodes") for g in range(P))
        lists = sum(view.field_len(l, g, k) for g in range(P)
                    for k in ("owned_out_nodes", "from_ids", "to_ids", "self_ids_in", "self_ids_out"))
        relabel = E * (4 + 4) + (E - n_in) * 4                            # csr_indices in/out + rank lookups
        rowside = F * (4 * 4 + 4) + n_out * 4 + (2 * n_out + lists) * 4     # nd, rowhas, ea, selfrank, relabel map; rowscan; lists out
        tail[l] = relabel + rowside
        rows = F * (4 + 16) + F * (16 + 4 + 4 + 4 + 4 + 4)                 # frontier + indptr pair in; rowinfo, off, selfrank, ea, rowhas, table mark out
        out[f"rows@{l}"] = rows
        expand = F * 20 + (C - F) * (4 + 4) + C * 5 + (C - F) * 4               # rowinfo + off, pick + atom.min per edge, stream out, rowhas
        # relabel and row side of layer l - 1 ride in this launch
        out[f"expand@{l}" if l == 0 else f"expand+relabel+rowside@{l}"] = expand + (tail[l - 1] if l else 0)
        out[f"cand_flags+rowcount@{l}"] = C * 5 + C * 4 + C * 1 + (E - n_in) * 4 + F * 8   # stream in, table / ea read, flags + moved words out; rowhas + ids
        out[f"cand_apply@{l}"] = C * 5 + (Fn + n_in + E) * 4 + F * 16              # stream in, lists + rowscan records out
    out["relabel+rowside+finish"] = tail[L - 1]
    return out


def random_accesses(view, P):
    """Random (data-dependent) 4-byte accesses one minibatch makes, by target:
    CSR (indptr pairs + neighbour picks), the slot's N-sized dedup table, and the
    C-sized position-indexed rank array."""
    csr = tab = pos = 0
    for l in range(view.num_layers):
        F, C = view.frontier_len[l], view.cand_len[l]
        E = sum(view.field_len(l, g, "csr_indices") for g in range(P))
        n_in = sum(view.field_len(l, g, "in_nodes") for g in range(P))
        csr += F + (C - F)          # indptr[nd], one neighbour pick per non-self candidate
        # (no per-row entry reset any more: the table is tagged)
        tab += (C - F)              # k_expand: atom.min
        tab += C                    # k_cand: read (table, or the row-indexed minimum)
        pos += E - n_in             # relabel: rank of the first occurrence, by position
    return csr, tab, pos


def sector_probe(device, table_bytes, n=1 << 26):
    """Random-sector ceiling of this GPU, measured by the library's own probe kernel
    (csrc/slicer_kernels.cu k_sector_probe: eight independent 4-byte reads in flight
    per thread, full occupancy) on a table of the given footprint -> reads/s."""
    import ctypes
    from cslicer import _native as N
    out = ctypes.c_double(0.0)
    N.check(N.lib().gs_bench_sector_probe(int(device), int(table_bytes), int(n), ctypes.byref(out)))
    return out.value


def sector_model(view, ip, fanouts):
    """SURVEY 8(d), last sentence of "Algorithmic bytes": the same compulsory traffic
    with every random gather charged one 32-byte sector -- the indptr pair of every
    frontier row (16 B -> 32 B) and every sampled pick (4 B -> 32 B; rows expanded in
    full are contiguous and stay at 4 B per neighbour).  -> (bytes, gathers)."""
    extra = gathers = 0
    deg = np.diff(ip)
    for l in range(view.num_layers):
        fr = view.frontier_numpy(l)
        d = deg[fr]
        fan = fanouts[l]
        picks = int(fan * (d >= fan).sum()) if fan >= 0 else 0
        extra += fr.size * (32 - 16) + picks * (32 - 4)
        gathers += fr.size + picks
    return view.algorithmic_bytes(owner_bytes=0, w=4) + extra, gathers


# One launch of k_expand_rows, layer 2 (the dominant launch: expansion of the third layer + relabel
# and row side of the second), default workload: 96.03 MB read + 7.07 MB written to DRAM (one warm
# minibatch, single slot) against 53.2 MB algorithmic -- the excess is 32-byte sectors moved for
# 4-byte neighbour picks and table atomics that miss L2.
NCU_TRAFFIC = {"workload": ("products", [15, 10, 5], 4096, 4), "kernel": "expand+relabel+rowside@2", "bytes": 103100416,
               "source": "profiles/r2_ncu_expand_rows_full_raw.csv (ncu --set full, tools/gpu_r2_final_n1.sh)"}


def own_arm(args):
    import torch
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    bound = None
    if world > 1:
        from cslicer.sharding import bind_to_gpu_cpus
        bound = bind_to_gpu_cpus(local)  # pinned sample mirrors land on the GPU's NUMA node
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    import ctypes
    from cslicer import _native
    from cslicer.engine import Graph, Slicer
    global HEADER_BYTES
    HEADER_BYTES = ctypes.sizeof(_native.SampleHeader)
    ip, ix = load_graph(args.shape, f"cuda:{local}")
    n = ip.size - 1
    g = Graph(ip, ix, device=local)
    P, mbs = args.parts, args.mb_per_step
    batches = seed_batches(n, max(64, mbs), rank, world)
    kw = dict(num_parts=P, max_batch=BATCH, num_slots=args.slots, num_streams=args.streams, seed=7,
              grid_ctas=args.grid_ctas, l2_hints=not args.no_l2_hints, group_size=args.group)
    dev = Slicer(g, FANOUTS, host_output=False, **kw)
    hst = Slicer(g, FANOUTS, host_output=True, host_copy_engine=not args.host_stores, **kw)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")

    # warm-up (>= 3 steps) for both modes
    W = max(3, args.warmup)
    run_pipeline(dev, batches, W, mbs, torch)
    run_pipeline(hst, batches, W, mbs, torch)
    # one sample for the byte model and sizes
    v = hst.slice(batches[0][1], batches[0][0])
    alg_bytes = v.algorithmic_bytes(owner_bytes=0, w=4)
    kbytes = kernel_bytes(v, P)
    sizes = dict(frontier=v.frontier_len, candidates=v.cand_len, edges=v.edges,
                 slab_bytes=v.slab_used * 4)
    ra_csr, ra_tab, ra_pos = random_accesses(v, P)
    sec_bytes, _ = sector_model(v, ip, FANOUTS)
    v.release()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local) if rank == 0 else None
    barrier()
    ms_dev, wall_dev, edges_dev, _ = run_pipeline(dev, batches, args.steps, mbs, torch)
    barrier()
    ms_e2e, wall_e2e, edges_e2e, slab_e2e = run_pipeline(hst, batches, args.steps, mbs, torch)
    barrier()
    clk = clocks.stop() if clocks else None
    from cslicer.sharding import reduce_job
    ms_dev, (edges_dev,) = reduce_job(dist, f"cuda:{local}", ms_dev, [edges_dev])
    ms_e2e, (edges_e2e, slab_e2e) = reduce_job(dist, f"cuda:{local}", ms_e2e, [edges_e2e, slab_e2e])
    total_mb = args.steps * mbs * world
    value = total_mb / (ms_dev * 1e-3)
    e2e = total_mb / (ms_e2e * 1e-3)

    out = None
    if rank == 0:
        dev_launches = dev.launches_per_sample
        prof = kernel_profile(dev, batches)
        chain_ms = sum(ms for _, ms in prof)
        top, top_ms = max(prof, key=lambda kv: kv[1])
        per_gpu_mb_s = value / world
        achieved = alg_bytes * per_gpu_mb_s / 1e9  # GB/s of compulsory traffic, pipelined steady state
        roof = {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": None, "peak_source": peak_src,
                "what": "algorithmic bytes of one minibatch (SURVEY 8(d), w=4) x minibatches/s per GPU, "
                        f"all {dev.launches_per_sample} kernels of the chain, slots pipelined on streams",
                "algorithmic_bytes_per_minibatch": int(alg_bytes),
                "chain_ms_eager_single_stream": round(chain_ms, 4),
                "chain_gbs_eager_single_stream": round(alg_bytes / (chain_ms * 1e-3) / 1e9, 2),
                "dominant_kernel": {"name": top, "avg_ms": round(top_ms, 4),
                                    "share_of_chain": round(top_ms / chain_ms, 3),
                                    "algorithmic_bytes": int(kbytes.get(top, 0)),
                                    "achieved_gbs": round(kbytes.get(top, 0) / (top_ms * 1e-3) / 1e9, 2),
                                    "frac": round(kbytes.get(top, 0) / (top_ms * 1e-3) / 1e9 / peak, 4)},
                "kernels_ms": {k: round(ms, 4) for k, ms in prof}}
        # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, one launch, from the
        # committed `ncu --set full` capture (valid for the default workload only)
        if (args.shape, FANOUTS, BATCH, P) == NCU_TRAFFIC["workload"] and top == NCU_TRAFFIC["kernel"]:
            roof["traffic"] = NCU_TRAFFIC["bytes"]
            roof["traffic_source"] = NCU_TRAFFIC["source"]
        # the sector-granular bound next to the algorithmic one, and the rate at which
        # this GPU can serve random sectors at all (own probe kernel, tables of the same footprint)
        roof["sector_bound"] = {"bytes_per_minibatch": int(sec_bytes), "achieved_gbs": round(sec_bytes * per_gpu_mb_s / 1e9, 2),
                                "frac": round(sec_bytes * per_gpu_mb_s / 1e9 / peak, 4),
                                "what": "algorithmic bytes with every random gather (indptr pair per row, every "
                                        "sampled pick) charged one 32-byte sector, x minibatches/s per GPU, / the same peak"}
        g_csr = sector_probe(local, ix.size * 4)
        g_tab = sector_probe(local, n * 4 * args.slots)
        floor_s = ra_csr / g_csr + (ra_tab + ra_pos) / g_tab
        roof["random_access"] = {
            "per_minibatch": {"csr_4B": int(ra_csr), "table_4B": int(ra_tab), "rank_by_position_4B": int(ra_pos)},
            "yardstick": {"kernel": "k_sector_probe (csrc/slicer_kernels.cu): 8 independent 4-byte reads in flight per "
                                    "thread, 8 CTAs of 256 threads per SM; one 32 B sector per read",
                          "csr_array_gsector_s": round(g_csr / 1e9, 2), "csr_array_gbs": round(g_csr * 32 / 1e9, 1),
                          "tables_all_slots_gsector_s": round(g_tab / 1e9, 2),
                          "tables_all_slots_gbs": round(g_tab * 32 / 1e9, 1)},
            "floor_us_per_minibatch": round(floor_s * 1e6, 1),
            "achieved_us_per_minibatch": round(1e6 / per_gpu_mb_s, 1),
            "frac_of_random_access_floor": round(floor_s * per_gpu_mb_s, 3),
            "note": "floor = this minibatch's random accesses at the probe's rate, nothing else counted"}
        out = {"metric": "sliced minibatches/s", "value": round(value, 2), "unit": "minibatches/s",
               "n_gpus": world, "steps": args.steps, "warmup": W,
               "ms_per_step": round(ms_dev / args.steps, 4), "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
               "config": {"workload": f"{args.shape}-shaped synthetic graph N={n} M={ix.size} seed {GRAPH_SEED}; "
                                      f"{len(FANOUTS)}-layer fanout {FANOUTS}, batch {BATCH}, {P} parts (owner=id%P)",
                          "graph": graph_stats(args.shape, ip), "host": host_description(),
                          "minibatches_per_step": mbs, "slots": args.slots, "streams": args.streams,
                          "minibatches_per_launch": args.group,
                          "sharding": "whole minibatches per rank, graph topology replicated",
                          "cpu_binding": (f"rank 0 on {len(bound)} cores next to its GPU" if bound else "none"),
                          "l2": f"inputs larger than L2: graph {(ip.size * 8 + ix.size * 4) / 1e9:.2f} GB + per-slot tables/workspaces "
                                f"{dev.slot_bytes * args.slots / 1e9:.2f} GB cycled every step",
                          "sizes_of_one_minibatch": sizes},
               "sampled_edges_per_s": round(edges_dev / (ms_dev * 1e-3), 1),
               "e2e": {"value": round(e2e, 2), "unit": "minibatches/s",
                       "h2d_bytes_per_step": mbs * (BATCH * 8 + 16),
                       "d2h_bytes_per_step": int(slab_e2e * 4 / (args.steps * world)) + mbs * HEADER_BYTES,
                       "ms_per_step": round(ms_e2e / args.steps, 4),
                       "sampled_edges_per_s": round(edges_e2e / (ms_e2e * 1e-3), 1)},
               "gpu_launches": dev.launches_per_sample * args.steps * mbs // args.group,
               "launches_per_minibatch": round(dev.launches_per_sample / args.group, 2),
               "wall_s": {"value_region": round(wall_dev, 4), "e2e_region": round(wall_e2e, 4)},
               "clocks": clk, "roofline": roof}
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(ip, ix, P, batches, args)
            try:
                out["feature_gather"] = feature_gather_line(ip, ix, g, args, torch, batches, peak)
            except Exception as e:
                out["feature_gather"] = {"unavailable": f"{type(e).__name__}: {e}"}
            try:  # the contiguous-range owner map next to id % P (owner read from a byte table)
                from cslicer.synth import owner_range
                rsl = Slicer(g, FANOUTS, host_output=False, owner=owner_range(n, P), **kw)
                run_pipeline(rsl, batches, 3, mbs, torch)
                ms_r, _, edges_r, _ = run_pipeline(rsl, batches, 3, mbs, torch)
                rsl.close()
                out["owner_range"] = {"value": round(3 * mbs / (ms_r * 1e-3), 2), "unit": "minibatches/s",
                                      "sampled_edges_per_s": round(edges_r / (ms_r * 1e-3), 1),
                                      "note": "same workload, owner = id*P//N from a uint8[N] table, samples left in HBM"}
            except Exception as e:
                out["owner_range"] = {"unavailable": f"{type(e).__name__}: {e}"}
            try:
                out["loader"] = loader_line(ip, ix, torch)
            except Exception as e:
                out["loader"] = {"unavailable": f"{type(e).__name__}: {e}"}
            try:
                out["edge_partitioning"] = edge_partitioning_line(n, P, sizes)
            except Exception as e:
                out["edge_partitioning"] = {"unavailable": f"{type(e).__name__}: {e}"}
            try:
                out["reference_config"] = reference_config(ip, ix, g, args, torch)
            except Exception as e:  # context only; never fail the bench line on it
                out["reference_config"] = {"unavailable": f"{type(e).__name__}: {e}"}
    dev.close()
    hst.close()
    if dist is not None and not args.no_hand_over:
        try:
            ho = hand_over_line(args, torch, dist, g, n, local, rank, world)
        except Exception as e:  # context only; never fail the bench line on it
            ho = {"unavailable": f"{type(e).__name__}: {e}"}
        if out is not None:
            # N > 1: the contract line is the layout north_star names -- part g of every minibatch ends
            # on GPU g.  Whole minibatches sliced where they fall and left there ("replicas") is kept
            # as an extra key: it says how fast N independent slicers are, not how the job scales.
            out["replicas"] = {"value": out["value"], "e2e": out["e2e"]["value"], "ms_per_step": out["ms_per_step"],
                               "sampled_edges_per_s": out["sampled_edges_per_s"],
                               "sharding": "whole minibatches per rank, graph topology replicated, parts stay "
                                           "where they were sliced (no exchange)"}
            if ho.get("value"):
                out["value"], out["ms_per_step"] = ho["value"], ho["ms_per_step"]
                out["config"]["minibatches_per_step"] = ho["minibatches_per_step"]
                how = {"staged": "written once into local part-contiguous blocks by the slicing kernels and sent to "
                                 "GPU g by the copy engines",
                       "direct": "written straight into GPU g's receive buffer by the slicing kernels",
                       "peer_stores": "delivered to GPU g by the packing kernel's own stores"}[ho["headline_exchange"]]
                out["config"]["sharding"] = ("hand-over: whole minibatches per rank on a replicated topology, "
                                             f"{world} parts, part g of every minibatch {how} over NVLink; barrier by "
                                             "flag words in the peers' memory (content-checked)")
                out["config"]["workload"] = out["config"]["workload"].replace(f"{P} parts", f"{world} parts")
                out["parity"] = ho["parity"]
                # per minibatch: the chain + the round's flag kernel + the release flag kernel (+ the packing kernel)
                extra = 3 if ho["headline_exchange"] == "peer_stores" else 2
                out["gpu_launches"] = int((dev_launches / args.group + extra) * args.steps * ho["minibatches_per_step"] / world)
                # the same minibatches as the replicas leg (rank r slices batches r, r + world, ...): edges per minibatch carry over
                out["sampled_edges_per_s"] = round(out["replicas"]["sampled_edges_per_s"] * ho["value"] / out["replicas"]["value"], 1) \
                    if out["replicas"]["value"] else None
                if ho["e2e"].get("value"):
                    out["e2e"] = ho["e2e"]
                # value-first ordering keeps these keys inside any truncated tail of the record
            out["hand_over"] = ho
    g.close()
    if dist is not None and not args.no_partitioned:
        try:
            pl = partitioned_line(args, torch, dist, ip, ix, local, rank, world)
        except Exception as e:  # context only; never fail the bench line on it
            pl = {"unavailable": f"{type(e).__name__}: {e}"}
        if out is not None:
            out["partitioned"] = pl
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        emit(out)


def list_digest(t, torch):
    """64-bit digest of an int32 device list (order-sensitive, wraps mod 2^64)."""
    if t.numel() == 0:
        return 0x9E3779B97F4A7C15
    x = t.to(torch.int64)
    i = torch.arange(1, x.numel() + 1, device=x.device, dtype=torch.int64)
    return int(((x + 0x632BE5AB) * (2 * i + 1)).sum().item()) & 0xFFFFFFFFFFFFFFFF


def part_digest(get_field, L, torch):
    """Digest of every list of one part: get_field(layer, name) -> int32 device tensor."""
    from cslicer import _native as N
    d = 0
    for l in range(L):
        for f, name in enumerate(N.FIELDS):
            d = (d * 0x100000001B3 + list_digest(get_field(l, name), torch) + 31 * l + f) & 0xFFFFFFFFFFFFFFFF
    return d


def hand_over_line(args, torch, dist, g, n, local, rank, world, steps=None, rounds=None):
    """Split-parallel hand-over at N > 1: every rank slices its own
    minibatches with `world` parts on the replicated topology and part g of every
    minibatch is delivered to rank g -- what the split-parallel training step
    consumes (python/train.py:120-160).  Two exchanges are measured: the packing
    kernel's own stores into the owners' buffers (PeerScatter: the headline) and an
    NCCL all-to-all of packed blocks.  Both end with a CONTENT check: every list of
    the parts a rank received is compared, on the device, with part `rank` of the
    same minibatch sliced by the rank's own single-GPU slicer.
    -> dict for the bench line (rank 0), None elsewhere."""
    from cslicer import _native as N
    from cslicer.engine import Slicer
    from cslicer.handover import HEADER_WORDS, DirectScatter, PartScatter, PeerScatter, StagedScatter
    steps = steps or args.steps
    rounds = rounds or max(16, args.mb_per_step // 4)
    kw = dict(num_parts=world, max_batch=BATCH, num_slots=args.slots, num_streams=args.streams, seed=7,
              grid_ctas=args.grid_ctas, group_size=args.group)
    sl = Slicer(g, FANOUTS, host_output=False, **kw)
    batches = seed_batches(n, 64, rank, world)  # rank r slices batches r, r + world, ...
    # one block size for the job: the largest part of a probe minibatch on any rank + 25 %
    slot = sl.submit(batches[0][1], batches[0][0])
    h = sl.wait_header(slot)
    need = max(HEADER_WORDS + sum((h.field[l][q][f].len + 3) & ~3 for l in range(len(FANOUTS))
                                  for f in range(N.NUM_FIELDS)) for q in range(world))
    slab_words = int(h.slab_used)
    sl.release(slot)
    t = torch.tensor([need], dtype=torch.int64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    need = int(t[0])
    block_words = (need * 5 // 4 + 3) & ~3
    depth = args.slots
    results = {}
    sent = 4 * block_words * (world - 1)
    all_kinds = ("staged", "staged_e2e", "direct", "direct_e2e", "peer_stores", "peer_stores_e2e", "nccl_all_to_all")
    only = [a for a in getattr(args, "hand_over_arms", "").split(",") if a]
    for kind in all_kinds:
        if only and kind.replace("_e2e", "") not in only:
            results[kind] = (float("nan"), "not run (--hand-over-arms)")
            continue
        try:  # constructor failures are collective (raised on every rank): the other arm still runs
            cls = PartScatter if kind == "nccl_all_to_all" else DirectScatter if kind.startswith("direct") else \
                StagedScatter if kind.startswith("staged") else PeerScatter
            ps = cls(sl, dist, block_words, depth=depth)
        except RuntimeError as e:
            log(f"[bench] hand-over arm {kind} unavailable: {e}")
            results[kind] = (float("nan"), "unavailable")
            continue
        mirror_words = min(block_words, (need * 21 // 20 + 3) & ~3) if kind.endswith("e2e") else 0
        try:
            results[kind] = _hand_over_run(ps, sl, batches, depth, steps, rounds if not mirror_words else max(8, rounds // 4),
                                           torch, dist, local, rank, world, n, mirror_words)
        except Exception as e:  # the flag waits time out, so a failing rank does not leave the others hanging
            log(f"[bench] hand-over arm {kind} failed on rank {rank}: {type(e).__name__}: {e}")
            results[kind] = (float("nan"), "failed")
        try:
            if hasattr(ps, "close"):
                ps.close()
        except Exception as e:
            log(f"[bench] hand-over arm {kind}: close failed: {type(e).__name__}: {e}")
        ps = None
    out = None
    if rank == 0:
        def rate(kind, r):  # NaN (arm unavailable or its content check failed) -> null: the line stays strict JSON
            ms = results[kind][0]
            return None if ms != ms else round(steps * r * world / (ms * 1e-3), 2)
        # headline: the faster of the two exchanges that need no library collective on the payload
        cands = [(results[kd][0], kd) for kd in ("staged", "direct", "peer_stores") if results[kd][0] == results[kd][0]]
        head = min(cands)[1] if cands else "peer_stores"
        ms = results[head][0]
        out = {"value": rate(head, rounds),
               "unit": "minibatches/s (whole job; rank g ends up holding part g of every minibatch)",
               "headline_exchange": head,
               "parity": results[head][1],
               "parity_check": "every list of every part received in the last round == part `rank` of the same "
                               "minibatch from the rank's own single-GPU slicer (torch.equal + 64-bit digests on the "
                               "device, all-reduce MIN of the verdict; a mismatch nulls the value)",
               "ms_per_step": None if ms != ms else round(ms / steps, 4), "minibatches_per_step": rounds * world,
               "e2e": {"value": rate(head + "_e2e", max(8, rounds // 4)), "unit": "minibatches/s",
                       "parity": results[head + "_e2e"][1],
                       "h2d_bytes_per_step": rounds * (BATCH * 8 + 16),
                       "d2h_bytes_per_step": int(max(8, rounds // 4) * world * need * 4),
                       "what": "host seeds in; every received part block copied to pinned host memory (one DMA "
                               "per block, size from the running estimate, header-checked) inside the timed region"},
               "barrier": "flag words stored into the peers' memory (gs_flags_post), no collective per round",
               "ranks": world, "parts": world, "rounds_in_flight": depth,
               "block_bytes": block_words * 4, "payload_bytes_per_minibatch": slab_words * 4,
               "nvlink_bytes_sent_per_rank_per_round": int(sent),
               "nvlink_gbs_per_rank": None if ms != ms else round(sent * steps * rounds / (ms * 1e-3) / 1e9, 2),
               "staged": {"value": rate("staged", rounds), "parity": results["staged"][1],
                          "e2e": rate("staged_e2e", max(8, rounds // 4)),
                          "exchange": "gs_slicer_submit_direct into LOCAL staging blocks (each part's lists written "
                                      "once, contiguously) + gs_slicer_push_parts: one copy-engine transfer per "
                                      "part to its owner; no SM store crosses NVLink"},
               "direct": {"value": rate("direct", rounds), "parity": results["direct"][1],
                          "e2e": rate("direct_e2e", max(8, rounds // 4)),
                          "exchange": "gs_slicer_submit_direct: the slicing kernels write every list of part g "
                                      "into rank g's receive buffer as they produce it (no slab write, no packing "
                                      "pass; each list crosses NVLink once)"},
               "peer_stores": {"value": rate("peer_stores", rounds), "parity": results["peer_stores"][1],
                               "e2e": rate("peer_stores_e2e", max(8, rounds // 4))},
               "exchange": "gs_slicer_scatter_parts: the packing kernel writes every block straight into its "
                           "owner rank's receive buffer (CUDA IPC over NVLink; cp.async.bulk pieces from 4 parts "
                           "up, 16-byte stores below); "
                           "overlapped with slicing",
               "same_with_nccl_all_to_all": {
                   "value": rate("nccl_all_to_all", rounds), "parity": results["nccl_all_to_all"][1],
                   "exchange": "gs_slicer_pack_parts into a send buffer + all_to_all_single of equal blocks"}}
    sl.close()
    return out


def _hand_over_run(ps, sl, batches, depth, steps, rounds, torch, dist, local, rank, world, n, mirror_words=0):
    """-> (device ms of steps * rounds rounds, max over ranks; parity verdict)."""
    from cslicer import _native as N
    from cslicer.handover import HEADER_WORDS, PartBlock
    from cslicer.sharding import reduce_job
    release = getattr(ps, "release", lambda tk: None)
    dev = f"cuda:{local}"
    host = cstream = None
    if mirror_words:
        host = [torch.empty((world, mirror_words), dtype=torch.int32).pin_memory() for _ in range(2)]
        cstream = torch.cuda.Stream(device=dev)
    cdone = [torch.cuda.Event(), torch.cuda.Event()]
    state = {"k": 0, "short": 0}

    def mirror(buf):  # the round's blocks -> pinned host memory, one DMA per block
        hb = host[state["k"] & 1]
        cdone[state["k"] & 1].synchronize()  # the copy that used this host buffer two rounds ago
        with torch.cuda.stream(cstream):
            cstream.wait_stream(torch.cuda.current_stream(dev))
            for src in range(world):
                hb[src].copy_(buf[src, :mirror_words], non_blocking=True)
            cdone[state["k"] & 1].record(cstream)
        state["k"] += 1
        return hb

    def run(k):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        inflight, last, last_host = [], None, None
        e0.record()
        pend, G = [], max(1, sl.group_size)
        if getattr(ps, "deferred", False):
            # post() needs the finished sample: submits run `ahead` rounds in front of it
            ahead, submitted = max(1, depth // 2), []

            def consume(slot, tk, keep=False):
                nonlocal last, last_host
                buf = ps.collect(tk, parse=False)
                if mirror_words:
                    hb = mirror(buf)
                    cstream.synchronize()
                    if keep:
                        last_host = hb
                if keep:
                    last = buf.clone()
                release(tk)

            for i in range(k):
                if len(submitted) == ahead:
                    slot = submitted.pop(0)
                    inflight.append((slot, ps.post(slot)))
                    sl.release(slot)
                if len(inflight) == depth - ahead:
                    consume(*inflight.pop(0))
                bi, seeds = batches[i % len(batches)]
                submitted.append(ps.submit(seeds, bi))
            for slot in submitted:
                inflight.append((slot, ps.post(slot)))
                sl.release(slot)
            for slot, tk in inflight:
                consume(slot, tk, keep=True)
            e1.record(ps.stream)
            e1.synchronize()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1), last, last_host
        for i in range(k):
            if len(inflight) + len(pend) == depth:
                slot, tk = inflight.pop(0)
                sl.wait_header(slot)
                sl.release(slot)
                buf = ps.collect(tk, parse=False)
                if mirror_words:
                    mirror(buf)
                    cstream.synchronize()  # the consumer holds the host copy before the round is released
                release(tk)
            bi, seeds = batches[i % len(batches)]
            pend.append(ps.submit(seeds, bi) if hasattr(ps, "submit") else sl.submit(seeds, bi))
            if len(pend) == G or i == k - 1:  # post once the group is launched (a post on a staged slot would launch it early)
                sl.flush()
                for slot in pend:
                    inflight.append((slot, ps.post(slot)))
                pend = []
        for slot, tk in inflight:
            sl.wait_header(slot)
            sl.release(slot)
            buf = ps.collect(tk, parse=False)
            if mirror_words:
                last_host = mirror(buf)
                cstream.synchronize()
            last = buf.clone()
            release(tk)
        e1.record(ps.stream)
        e1.synchronize()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), last, last_host
    run(2 * min(rounds, 32))
    dist.barrier()
    torch.cuda.synchronize()
    ms, last, last_host = run(steps * rounds)
    dist.barrier()
    ms, _ = reduce_job(dist, dev, ms, [0])
    # content check of what arrived last: part `rank` of the last round's minibatch of every source rank
    hdr = last[:, :HEADER_WORDS].cpu().numpy()
    k_last = (steps * rounds - 1) % len(batches)
    ok = 1
    L = len(FANOUTS)
    try:
        for src in range(world):
            b = PartBlock(last[src], hdr[src])
            bi, seeds = seed_batches(n, 64, src, world)[k_last]
            ok &= int(b.part == rank and b.batch_index == bi)
            v = sl.slice(seeds, bi)  # the same minibatch, sliced here: its part `rank` must be what arrived
            for l in range(L):
                for name in N.FIELDS:
                    a, w = b.field(l, name), v.tensor(l, rank, name)
                    ok &= int(a.numel() == w.numel() and bool(torch.equal(a, w)))
            ok &= int(part_digest(lambda l, nm: b.field(l, nm), L, torch) ==
                      part_digest(lambda l, nm: v.tensor(l, rank, nm), L, torch))
            v.release()
            if mirror_words:  # the host copy holds the same words (up to the words in use)
                used = int(hdr[src][8])
                ok &= int(used <= mirror_words and
                          bool(torch.equal(last_host[src][:used], last[src][:used].cpu())))
    except Exception as e:
        log(f"[bench] hand-over content check raised: {type(e).__name__}: {e}")
        ok = 0
    # collective verdict: a rank must not leave the sequence of collectives on its own
    t = torch.tensor([ok], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    good = int(t[0]) == 1
    return (ms if good else float("nan")), ("digest-equal" if good else "MISMATCH")


def partitioned_line(args, torch, dist, ip, ix, local, rank, world, steps=3, mbs=32, shard=None):
    """SURVEY 8(e) layout at N > 1: graph rows sharded by owner (rank g holds
    N/P of the CSR and of the dedup table), all ranks cooperate on EVERY
    minibatch and rank g produces part g; edges and new frontier nodes move by
    the kernels' own stores / red.min into peer memory over NVLink.  Capacity
    mode (papers100M-scale graphs, sharded features); the replicated mode above
    is the throughput mode.  -> dict for the bench line (rank 0), None elsewhere."""
    from cslicer.partitioned import PartitionedPipeline, partition_csr
    from cslicer.sharding import reduce_job
    okw = {}
    if shard is None:
        n, m_global = ip.size - 1, ix.size
        lip, lix = partition_csr(ip, ix, world)[rank]
    else:  # this rank's rows only (contiguous-range owner map); the full graph is nowhere on the host
        lip, lix, n, m_global, bounds = shard
        okw = dict(owner_mode="range", bounds=bounds)
    lanes, depth = args.lanes, 2 * args.lanes
    ps = PartitionedPipeline(lip, lix, n, rank, world, FANOUTS, dist, lanes=lanes, device=local,
                             max_batch=BATCH, num_slots=2, seed=7, host_output=False,
                             num_edges_global=m_global, grid_ctas=args.grid_ctas, timeout_ms=2000, **okw)
    batches = seed_batches(n, 64, 0, 1)  # every rank is given the same batches

    def run(k):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        inflight, edges = [], 0
        e0.record()
        for i in range(k):
            if len(inflight) == depth:
                tk = inflight.pop(0)
                edges += ps.wait_header(tk).edges
                ps.release(tk)
            bi, seeds = batches[i % len(batches)]
            inflight.append(ps.submit(seeds, bi))
        for tk in inflight:
            edges += ps.wait_header(tk).edges
            ps.release(tk)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), edges
    run(3 * mbs)
    dist.barrier()
    torch.cuda.synchronize()
    ms, edges = run(steps * mbs)
    dist.barrier()
    ms, (edges,) = reduce_job(dist, f"cuda:{local}", ms, [edges])
    v = ps.slice(batches[0][1], batches[0][0])
    fr = list(v.frontier_len)
    # content check: this rank's part against part `rank` of the same minibatch from the
    # single-GPU slicer (needs the whole graph on this rank: not available in shard mode)
    parity = "unchecked (the whole graph exists on no rank in shard mode; same kernels as the checked runs)"
    if shard is None:
        from cslicer import _native as N
        from cslicer.engine import Graph, Slicer
        ok = 1
        try:
            g1 = Graph(ip, ix, device=local)
            s1 = Slicer(g1, FANOUTS, num_parts=world, max_batch=BATCH, num_slots=1, num_streams=1, seed=7,
                        host_output=False)
            w = s1.slice(batches[0][1], batches[0][0])
            for l in range(len(FANOUTS)):
                for name in N.FIELDS:
                    a, b = v.tensor(l, rank, name), w.tensor(l, rank, name)
                    ok &= int(a.numel() == b.numel() and bool(torch.equal(a, b)))
            ok &= int(part_digest(lambda l, nm: v.tensor(l, rank, nm), len(FANOUTS), torch) ==
                      part_digest(lambda l, nm: w.tensor(l, rank, nm), len(FANOUTS), torch))
            w.release()
            s1.close()
            g1.close()
        except Exception as e:
            log(f"[bench] partitioned content check raised: {type(e).__name__}: {e}")
            ok = 0
        tk = torch.tensor([ok], dtype=torch.int32, device=f"cuda:{local}")
        dist.all_reduce(tk, op=dist.ReduceOp.MIN)
        parity = "digest-equal" if int(tk[0]) == 1 else "MISMATCH"
    v.release()
    # every rank must hold the same replicated frontier sizes (a cheap cross-rank invariant)
    t = torch.tensor(fr, dtype=torch.int64, device=f"cuda:{local}")
    tmax, tmin = t.clone(), t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    assert torch.equal(tmax, tmin), "ranks disagree on the frontier sizes"
    out = None
    if rank == 0:
        per_mb = ms / (steps * mbs)
        e_mb = edges / (steps * mbs)
        xbytes = e_mb * (world - 1) / world * 9 + sum(fr[1:]) * 8 * (world - 1) + BATCH * 4
        out = {"value": round(1e3 / per_mb, 2) if parity != "MISMATCH" else None,
               "unit": "minibatches/s (whole job, every rank works on every minibatch)", "parity": parity,
               "ms_per_minibatch": round(per_mb, 4), "sampled_edges_per_s": round(edges / (ms * 1e-3), 1),
               "ranks": world, "parts": world, "lanes_per_rank": lanes, "launches_per_minibatch_per_rank": ps.launches_per_sample,
               "barriers_per_minibatch": 3 * len(FANOUTS) + 1,
               "owner_map": "contiguous ranges" if shard is not None else "id % ranks",
               "frontier_sizes": [int(x) for x in fr],
               "device_bytes_per_rank": int(ps.device_bytes),
               "graph_shard_bytes_per_rank": int(lip.size * 8 + lix.size * 4),
               "nvlink_bytes_per_minibatch_est": int(xbytes),
               "nvlink_gbs_est": round(xbytes / (per_mb * 1e-3) / 1e9, 2),
               "note": "exchange = 9 B per remote edge (id, flag, red.min) + 8 B per new frontier node per "
                       "peer; NVLink 5 peak 900 GB/s per direction per GPU: the exchange is latency-, not "
                       "bandwidth-bound"}
    dist.barrier()
    ps.close()
    try:  # the split-parallel shuffle of partial rows, forward + backward, every layer
        sh = shuffle_line(torch, dist, lip, lix, n, m_global, local, rank, world, batches, okw)
    except Exception as e:
        log(f"[bench] shuffle leg raised: {type(e).__name__}: {e}")
        sh = {"unavailable": f"{type(e).__name__}: {e}"}
    if out is not None:
        out["shuffle"] = sh
    return out


def shuffle_line(torch, dist, lip, lix, n, m_global, local, rank, world, batches, okw, dim=128, reps=5):
    """gs_pslicer_shuffle (replaces pull_for_remotes / push_from_remotes + mergeKernel,
    python/data/bipartite.py:88-99, src/gnn/dist_sage.cu:193-220): partial rows of [|out_nodes|, dim]
    fp32 pushed into the owners' arenas over NVLink and merged there; backward hands them back.
    Timed per layer with the host blocked until done (max over ranks); bytes = rows this rank sends x
    dim x 4.  Checked against a restatement on the device: forward then backward of all-ones rows."""
    from cslicer.partitioned import PartitionedSlicer, connect_process_group
    import time as _t
    ps = PartitionedSlicer(lip, lix, n, rank, world, FANOUTS, device=local, max_batch=BATCH, num_slots=2, seed=7,
                           host_output=False, num_edges_global=m_global, shuffle_rows_capacity=1 << 19,
                           shuffle_max_dim=dim, timeout_ms=5000, **okw)
    connect_process_group(ps, dist)
    v = ps.slice(batches[0][1], batches[0][0])
    L = len(FANOUTS)
    sent = t_f = t_b = 0.0
    per_layer = []
    ok = 1
    for l in range(L):
        n_out = v.field_len(l, rank, "out_nodes")
        n_from = v.field_len(l, rank, "from_ids")
        rows = torch.ones((n_out, dim), dtype=torch.float32, device=f"cuda:{local}")
        ps.shuffle(v, l, rows, False)  # warm
        tf = tb = 0.0
        for _ in range(reps):
            rows.fill_(1.0)
            torch.cuda.synchronize(); dist.barrier()
            t0 = _t.perf_counter(); ps.shuffle(v, l, rows, False); tf += _t.perf_counter() - t0
            if _ == 0:
                # owned rows hold 1 + (number of peers that contributed); all entries of a row agree
                own = rows[v.tensor(l, rank, "owned_out_nodes").long()]
                ok &= int(bool((own >= 1).all()) and bool((own == own[:, :1]).all()) and bool((own == own.round()).all()))
            torch.cuda.synchronize(); dist.barrier()
            t0 = _t.perf_counter(); ps.shuffle(v, l, rows, True); tb += _t.perf_counter() - t0
        tt = torch.tensor([tf / reps, tb / reps], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        b = n_from * dim * 4
        per_layer.append({"rows_sent": int(n_from), "forward_ms": round(float(tt[0]) * 1e3, 4),
                          "backward_ms": round(float(tt[1]) * 1e3, 4)})
        sent += b; t_f += float(tt[0]); t_b += float(tt[1])
    v.release()
    tk = torch.tensor([ok], dtype=torch.int32, device=f"cuda:{local}")
    dist.all_reduce(tk, op=dist.ReduceOp.MIN)
    dist.barrier()
    ps.close()
    return {"dim": dim, "layers": per_layer, "bytes_sent_per_rank_forward": int(sent),
            "forward_gbs_per_rank": round(sent / t_f / 1e9, 2), "backward_gbs_per_rank": round(sent / t_b / 1e9, 2),
            "nvlink_peak_gbs_per_direction": 900, "frac_of_nvlink_forward": round(sent / t_f / 1e9 / 900, 4),
            "check": "merge-consistent" if int(tk[0]) == 1 else "INCONSISTENT",
            "what": "one minibatch of this workload, every layer's partial rows (fp32, dim 128): push into the owners' "
                    "arenas + merge (forward), hand back (backward); host-blocking calls, max over ranks; "
                    "bit-exactness against the restated mergeKernel is in tests/test_partitioned.py"}


def partitioned_shards_arm(args):
    """`--partitioned-shards` (N > 1 only): the partitioned slicer on a graph that is
    generated shard by shard -- every rank replays the seeded generator on its GPU
    and keeps only the rows it owns, so a papers100M-shaped graph (BASELINE config 5,
    8 parts) never exists whole on the host.  Prints one JSON line (evidence for
    SURVEY 8(e); not the contract line)."""
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world > 1, "--partitioned-shards needs torchrun with N > 1"
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from cslicer.partitioned import range_bounds
    from cslicer.synth import SHAPES, synth_graph
    n = SHAPES[args.shape][0]
    bounds = range_bounds(n, world)
    t = time.time()
    lip, lix, n, m = synth_graph(args.shape, seed=GRAPH_SEED, device=f"cuda:{local}",
                                 rows=(int(bounds[rank]), int(bounds[rank + 1])))
    torch.cuda.empty_cache()
    log(f"[bench] rank {rank}: rows {bounds[rank]}..{bounds[rank + 1]} of {args.shape} "
        f"(N={n} M={m}; local edges {lix.size}) in {time.time() - t:.1f}s")
    out = partitioned_line(args, torch, dist, None, None, local, rank, world,
                           shard=(lip, lix, n, m, bounds))
    if rank == 0:
        out = {"metric": "sliced minibatches/s (partitioned mode)", "n_gpus": world, "data": "synthetic",
               "config": {"workload": f"{args.shape}-shaped synthetic graph N={n} M={m} seed {GRAPH_SEED}; "
                                      f"{len(FANOUTS)}-layer fanout {FANOUTS}, batch {BATCH}, {world} parts = ranks"},
               **out}
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_port_rate(ip, ix, P, batches, workers, target_s):
    """Multithreaded CPU restatement (oracle/, the checker) on a bounded sample:
    one probe batch sizes the sample to about target_s seconds of work."""
    from oracle import oracle as O
    seeds = [b for _, b in batches]
    t1, _ = O.pool_run(ip, ix, FANOUTS, P, seeds[:1], workers=1, rng_mode=O.RNG_PHILOX, seed=7)
    nb = int(max(workers, min(4096, target_s / max(t1, 1e-3) * workers * 0.7)))
    nb = max(workers, (nb // workers) * workers)
    sample = [seeds[i % len(seeds)] for i in range(nb)]
    sec, edges = O.pool_run(ip, ix, FANOUTS, P, sample, workers=workers, rng_mode=O.RNG_PHILOX, seed=7)
    return nb / sec, edges / sec, nb, sec


def cpu_baseline(ip, ix, P, batches, args):
    cores = host_cores()
    rate, erate, nb, sec = cpu_port_rate(ip, ix, P, batches, cores, args.cpu_seconds)
    return {"value": round(rate, 3), "unit": "minibatches/s", "cores": cores, "kind": "port",
            "sample": f"{nb} minibatches of the same workload in {sec:.1f} s, {cores} threads "
                      "(one restated slicer per thread, WorkerPool.cpp:29-33)",
            "sampled_edges_per_s": round(erate, 1)}


def reference_config(ip, ix, g, args, torch):
    """The one configuration the UNMODIFIED reference can run (3 layers x fanout
    10 with replacement, 4 parts, owner = id % 4): the compiled reference
    (oracle/_ref, its own Slicer::run threads + ConQueues) on the host cores
    beside the CUDA path on the same graph and batches."""
    from cslicer.dataset import write_dataset
    from cslicer.engine import Slicer
    from oracle import oracle as O
    if not O.ref_available():
        return {"unavailable": "oracle/_ref/libcslicer_ref.so not built (no /root/reference at build time)"}
    n = ip.size - 1
    d = os.path.join(CACHE, f"{args.shape}_{GRAPH_SEED}_refdata")
    if not os.path.exists(os.path.join(d, "meta.txt")):
        write_dataset(d, ip, ix)
    cores = host_cores()
    batches = seed_batches(n, 8 * cores, 0, 1)
    seeds = [b for _, b in batches]
    t1 = O.ref_pool_time(d, 1, seeds[:4]) / 4
    nb = int(max(cores, min(len(seeds), 8.0 / max(t1, 1e-3) * cores * 0.6)))
    sec = O.ref_pool_time(d, cores, seeds[:nb])
    sl = Slicer(g, [10, 10, 10], num_parts=4, max_batch=BATCH, num_slots=args.slots,
                num_streams=args.streams, seed=7, grid_ctas=args.grid_ctas, host_output=False)
    run_pipeline(sl, batches, 3, args.mb_per_step, torch)
    ms, _, edges, _ = run_pipeline(sl, batches, 5, args.mb_per_step, torch)
    sl.close()
    hs = Slicer(g, [10, 10, 10], num_parts=4, max_batch=BATCH, num_slots=args.slots,
                num_streams=args.streams, seed=7, grid_ctas=args.grid_ctas, host_output=True)
    run_pipeline(hs, batches, 2, args.mb_per_step, torch)
    ms2, _, _, _ = run_pipeline(hs, batches, 3, args.mb_per_step, torch)
    hs.close()
    return {"config": "3-layer fanout [10,10,10] with replacement, batch 4096, 4 parts, owner=id%4 "
                      "(the only configuration the reference supports)",
            "reference_cpu": {"kind": "reference", "value": round(nb / sec, 3), "unit": "minibatches/s",
                              "cores": cores, "value_1_worker": round(1.0 / t1, 3),
                              "sample": f"{nb} minibatches, {cores} reference Slicer threads; stdout silenced"},
            "cuda_value": round(5 * args.mb_per_step / (ms * 1e-3), 2),
            "cuda_e2e": round(3 * args.mb_per_step / (ms2 * 1e-3), 2),
            "cuda_sampled_edges_per_s": round(edges / (ms * 1e-3), 1)}


def loader_line(ip, ix, torch):
    """the workload's graph written in the reference's on-disk layout, then
    loaded (a) by gs_graph_load -- files streamed through pinned staging buffers into HBM, checked
    and narrowed on the device -- and (b) the reference's way: whole files into host arrays
    (Dataset, dataset.cpp:18-36, with its checksums) and uploaded.  Page cache warm in both."""
    import shutil
    import cslicer
    from cslicer.engine import Graph
    d = tempfile.mkdtemp(prefix="gsplit_ds_")
    try:
        cslicer.write_dataset(d, ip, ix, features=np.zeros((ip.size - 1, 1), np.float32))
        nbytes = ip.size * 8 + ix.size * 8
        Graph.from_directory(d).close()
        t0 = time.perf_counter()
        g = Graph.from_directory(d)
        torch.cuda.synchronize()
        t_dev = time.perf_counter() - t0
        same = (g.num_nodes, g.num_edges) == (ip.size - 1, ix.size)
        g.close()
        t0 = time.perf_counter()
        ds = cslicer.Dataset(d, load_node_data=False)
        g = Graph(ds.indptr, ds.indices)
        torch.cuda.synchronize()
        t_host = time.perf_counter() - t0
        g.close()
    finally:
        shutil.rmtree(d, ignore_errors=True)
    return {"file_bytes": int(nbytes), "gs_graph_load_s": round(t_dev, 3), "gs_graph_load_gbs": round(nbytes / t_dev / 1e9, 2),
            "host_arrays_then_upload_s": round(t_host, 3), "host_arrays_then_upload_gbs": round(nbytes / t_host / 1e9, 2),
            "sizes_agree": bool(same),
            "what": "indptr.bin + indices.bin (int64) -> validated int32 CSR in HBM, checksums verified, page cache warm"}


def edge_partitioning_line(n, P, sizes, reps=3):
    """Sampler.edge_partitioning (python/utils/sampler.py:64-134) on blocks of
    this workload's size -- the CUDA partitioner through its host-buffer C ABI (copies inside the
    timed region) next to the numpy restatement of the reference method, same inputs."""
    from cslicer.epart import EdgePartitioner
    from oracle import sampler_oracle as SO
    rng = np.random.default_rng(7)
    owner = (np.arange(n) % P).astype(np.int32)
    mask = rng.random((n, P)) < 0.25
    mask[np.arange(n), owner] = True
    blocks = [(rng.integers(0, n, c), rng.integers(0, n, c)) for c in sizes["candidates"]]
    edges = sum(b[0].size for b in blocks)
    ep = EdgePartitioner(owner, mask, P, max_edges=max(b[0].size for b in blocks))
    ep.edge_partitioning(blocks)
    t0 = time.perf_counter()
    for _ in range(reps):
        got = ep.edge_partitioning(blocks)
    t_gpu = (time.perf_counter() - t0) / reps
    ep.close()
    t0 = time.perf_counter()
    want = SO.edge_partitioning(blocks, owner, mask, P)
    t_cpu = time.perf_counter() - t0
    same = all(np.array_equal(a[k], b[k]) for gl, wl in zip(got, want) for a, b in zip(gl, wl) for k in ("src", "dest", "owned_nds"))
    return {"edges_per_call": edges, "layers": len(blocks), "cuda_ms": round(t_gpu * 1e3, 3),
            "cuda_edges_per_s": round(edges / t_gpu, 1), "cpu_restatement_ms": round(t_cpu * 1e3, 1),
            "cpu_edges_per_s": round(edges / t_cpu, 1), "parity": "equal" if same else "DIFFERENT",
            "what": "host int64 edge arrays in, per-GPU lists back in pinned host memory (H2D + D2H inside the time); "
                    "25 % of the nodes cached per GPU on the last layer"}


def _time_gather(F, ids, torch, reps=5):
    """The gather kernel alone on the ids of one sample's input block (the fused
    gather shares its launch with the last relabel, so it is timed standalone)."""
    F.gather(ids)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        F.gather(ids)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def feature_gather_line(ip, ix, g, args, torch, batches, peak):
    """BASELINE config 4's other half: the fused input-feature gather (602-dim
    fp32, seed 4321) -- rows of the last layer's in_nodes_g, part after part,
    appended to every sample's CUDA graph.  HBM-bound: whole 2408-byte rows."""
    from cslicer.engine import Features, Slicer
    n, dim = ip.size - 1, 602
    gen = torch.Generator(device="cuda")
    gen.manual_seed(4321)
    rows = torch.rand((n, dim), generator=gen, device="cuda", dtype=torch.float32).cpu().numpy()
    F = Features(rows)
    sl = Slicer(g, FANOUTS, num_parts=args.parts, max_batch=BATCH, num_slots=3, num_streams=3, seed=7,
                grid_ctas=args.grid_ctas, host_output=False, features=F)
    run_pipeline(sl, batches, 1, 12, torch)
    ms, _, edges, _ = run_pipeline(sl, batches, 4, 12, torch)
    v = sl.slice(batches[0][1], batches[0][0])
    nrows = v.feature_rows
    ids = torch.cat([v.tensor(v.num_layers - 1, q, "in_nodes") for q in range(args.parts)]).clone()
    v.release()
    k_ms = _time_gather(F, ids, torch)
    nbytes = 2 * nrows * dim * 4
    sl.close()
    F.close()
    cached = cached_gather(ip, ix, g, args, rows, batches)
    del rows
    return {"dim": dim, "cached_store": cached, "rows_per_minibatch": int(nrows), "bytes_per_minibatch": int(nbytes),
            "kernel_ms": round(k_ms, 4), "achieved_gbs": round(nbytes / (k_ms * 1e-3) / 1e9, 1),
            "frac_of_hbm_peak": round(nbytes / (k_ms * 1e-3) / 1e9 / peak, 4),
            "minibatches_per_s_with_gather": round(48 / (ms * 1e-3), 2),
            "note": "algorithmic bytes = one read + one write of every gathered row; slicing + gather "
                    "pipelined on 3 slots"}


def cached_gather(ip, ix, g, args, rows, batches, fraction=0.25):
    """the same fused gather with the table left in host memory and
    the `fraction` highest-degree nodes cached in HBM (memory_manager.py:40-74); a
    miss is read from pinned host memory by the gather warp itself (PCIe-bound)."""
    from cslicer.engine import Features, Slicer
    from cslicer.feature_cache import degree_ordered_cache, hit_rate
    try:
        cache = degree_ordered_cache(ip, fraction, part=0, num_parts=1, indices=ix)
        F = Features(rows, cache_rows=cache)
        sl = Slicer(g, FANOUTS, num_parts=args.parts, max_batch=BATCH, num_slots=2, num_streams=2, seed=7,
                    grid_ctas=args.grid_ctas, host_output=False, features=F)
        v = sl.slice(batches[0][1], batches[0][0])
        ids = np.concatenate([v.tensor(v.num_layers - 1, q, "in_nodes").cpu().numpy()
                              for q in range(args.parts)])
        v.release()
        import torch
        k_ms = _time_gather(F, torch.from_numpy(ids).cuda(), torch, reps=3)
        hr = hit_rate(cache, ids, ip.size - 1)
        miss_bytes = (1.0 - hr) * ids.size * rows.shape[1] * 4
        sl.close()
        F.close()
        return {"cache_fraction": fraction, "policy": "highest source degree first (synthetic sources are uniform: weak skew)", "rows_in_hbm": int(cache.size),
                "hit_rate": round(hr, 4), "kernel_ms": round(k_ms, 3),
                "host_read_gbs": round(miss_bytes / (k_ms * 1e-3) / 1e9, 1),
                "note": "misses are zero-copy reads of pinned host rows inside the gather kernel"}
    except Exception as e:  # context only
        return {"unavailable": f"{type(e).__name__}: {e}"}


def reference_arm(args):
    """--impl reference: the reference algorithm on the host cores, same config,
    metric and unit.  The compiled reference (oracle/_ref) is fixed at 3 layers x
    fanout 10 x 4 parts, so the exact config is timed with the multithreaded CPU
    restatement ("port"); rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ip, ix = load_graph_cpu_arm(args.shape)
    n = ip.size - 1
    cores = host_cores()
    batches = seed_batches(n, 256, 0, 1)
    from oracle import oracle as O
    seeds = [b for _, b in batches]
    t1, _ = O.pool_run(ip, ix, FANOUTS, args.parts, seeds[:1], workers=1, rng_mode=O.RNG_PHILOX, seed=7)
    # size a step so (steps + warmup) steps end within ~2 minutes
    total_steps = args.steps + max(1, args.warmup)
    per_step = int(max(cores, min(4 * cores, 120.0 / total_steps / max(t1, 1e-3) * cores * 0.7)))
    per_step = max(cores, (per_step // cores) * cores)
    k = 0

    def step():
        nonlocal k
        sample = [seeds[(k + i) % len(seeds)] for i in range(per_step)]
        k += per_step
        return O.pool_run(ip, ix, FANOUTS, args.parts, sample, workers=cores,
                          rng_mode=O.RNG_PHILOX, seed=7)
    for _ in range(max(1, args.warmup)):
        step()
    sec = edges = 0.0
    for _ in range(args.steps):
        s_, e_ = step()
        sec += s_
        edges += e_
    value = args.steps * per_step / sec
    out = {"impl": "reference", "metric": "sliced minibatches/s", "value": round(value, 3),
           "unit": "minibatches/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
           "steps": args.steps, "warmup": max(1, args.warmup),
           "ms_per_step": round(sec / args.steps * 1e3, 2), "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
           "config": {"workload": f"{args.shape}-shaped synthetic graph N={n} M={ix.size} seed {GRAPH_SEED}; "
                                  f"{len(FANOUTS)}-layer fanout {FANOUTS}, batch {BATCH}, {args.parts} parts (owner=id%P)",
                      "minibatches_per_step": per_step},
           "sampled_edges_per_s": round(edges / sec, 1),
           "cpu_baseline": {"value": round(value, 3), "unit": "minibatches/s", "cores": cores,
                            "kind": "port",
                            "sample": f"{per_step} minibatches per step, {cores} threads, one restated "
                                      "slicer per thread (WorkerPool.cpp:29-33)"},
           "e2e": {"value": round(value, 3), "unit": "minibatches/s", "h2d_bytes_per_step": 0,
                   "d2h_bytes_per_step": 0}}
    emit(out)


def main():
    global BATCH, FANOUTS
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--shape", default=SHAPE)
    ap.add_argument("--parts", type=int, default=4)
    ap.add_argument("--batch", type=int, default=BATCH, help="seeds per minibatch (config 5: 8192)")
    ap.add_argument("--fanouts", default=",".join(map(str, FANOUTS)))
    ap.add_argument("--mb-per-step", type=int, default=640,
                    help="minibatches per step (default: 20 steps are > 1 s of device time at 10 K minibatches/s)")
    ap.add_argument("--slots", type=int, default=12)
    ap.add_argument("--streams", type=int, default=12)
    ap.add_argument("--group", type=int, default=1,
                    help="minibatches per launch sequence (gs_config.group_size); must divide --slots")
    ap.add_argument("--grid-ctas", type=int, default=2, help="CTAs per SM of a full-machine launch")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-l2-hints", action="store_true", help="A/B: plain loads/stores/reds on the dedup table")
    ap.add_argument("--host-stores", action="store_true",
                    help="e2e: mirror samples with stores from the last kernel instead of the copy engine")
    ap.add_argument("--lanes", type=int, default=8, help="partitioned mode: minibatches in flight per rank")
    ap.add_argument("--partitioned-shards", action="store_true",
                    help="N > 1: only the partitioned slicer, graph generated shard by shard (config 5 at scale)")
    ap.add_argument("--hand-over-arms", default="", help="N > 1: comma list of staged,direct,peer_stores,nccl_all_to_all (default: all)")
    ap.add_argument("--no-hand-over", action="store_true",
                    help="N > 1: skip the split-parallel hand-over (slice anywhere, part g -> rank g) measurement")
    ap.add_argument("--no-partitioned", action="store_true",
                    help="N > 1: skip the extra partitioned-mode (graph sharded by owner) measurement")
    args = ap.parse_args()
    BATCH = args.batch
    FANOUTS = [int(x) for x in args.fanouts.split(",")]
    capture_stdout()
    if args.impl == "reference":
        reference_arm(args)
    elif args.partitioned_shards:
        partitioned_shards_arm(args)
    else:
        own_arm(args)


if __name__ == "__main__":
    main()
