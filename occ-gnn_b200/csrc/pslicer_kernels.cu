// pslicer_kernels.cu -- sm_100a kernels of the partitioned slicer (one rank per
// GPU).  Same algorithm as slicer_kernels.cu -- candidate positions
// of the reference's sequential scan (slicer.cpp:25-64), minimum position =
// first occurrence (util/duplicate.cpp:14-26), prefix scans over positions =
// list order -- with the candidate stream of a layer spread over the ranks by
// owner of the edge source.  Exchange is done by the kernels themselves with
// stores and red.min into peer memory over NVLink (cudaIpc-mapped arenas); the
// only synchronisation is a flag barrier between phases.  No host round trip:
// a minibatch is one static launch sequence on one stream.
#include "pslicer_internal.cuh"
#include "slicer_device.cuh"

namespace {

// id / P and id % P without integer division on the hot path: shift / mask when
// P is a power of two, else multiply-high by ceil(2^32/P) (quotient exact or one
// too large for any 31-bit id) and one correction step.
__device__ __forceinline__ void ps_divmod(const PsWs &ws, int32_t v, int32_t &q, int32_t &r) {
  if (ws.pmask >= 0) {
    q = v >> ws.pshift;
    r = v & ws.pmask;
    return;
  }
  q = (int32_t)__umulhi((uint32_t)v, ws.pmagic);
  r = v - q * ws.P;
  if (r < 0) {
    r += ws.P;
    q--;
  }
}
__device__ __forceinline__ int ps_owner(const PsWs &ws, int32_t v) {
  if (ws.owner_mode == PS_OWNER_MOD) {
    int32_t q, r;
    ps_divmod(ws, v, q, r);
    return r;
  }
  int q = 0;
  while (q + 1 < ws.P && v >= ws.bounds[q + 1]) q++;
  return q;
}
// index of node v in its owner's local arrays (CSR shard rows, dedup table)
__device__ __forceinline__ int32_t ps_local(const PsWs &ws, int32_t v, int q) {
  if (ws.owner_mode == PS_OWNER_MOD) {
    int32_t d, r;
    ps_divmod(ws, v, d, r);
    return d;
  }
  return v - ws.bounds[q];
}
// candidates row `nd` (owned by this rank) contributes under fanout `fan`:
// 1 + (deg < fan or fan < 0 ? deg : fan)   (slicer.cpp:6-22)
__device__ __forceinline__ uint32_t ps_row_count(const PsWs &ws, int32_t nd, int fan) {
  const int32_t li = ps_local(ws, nd, ws.rank);
  const int32_t deg = (int32_t)(__ldg(ws.indptr + li + 1) - __ldg(ws.indptr + li));
  return 1u + (uint32_t)((fan >= 0 && deg >= fan) ? fan : deg);
}

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// ---------------------------------------------------------------- barrier
// signal: everything this rank wrote into peer arenas before this kernel is
// made visible (fence.sys), then flags[rank] = epoch in every peer's arena.
// wait: spin until every peer's flag reached this rank's epoch.  A status
// raised on any rank travels with the barrier so all ranks fail together, and
// a peer that never arrives turns into an error after timeout_ms, never a hang.
__global__ void pk_signal(PsWs ws) {
  __shared__ uint32_t s_e;
  if (threadIdx.x == 0) {
    s_e = *ws.epoch + 1u;
    *ws.epoch = s_e;
  }
  __syncthreads();
  const int q = threadIdx.x;
  if (q < ws.P) {
    const int32_t st = ws.hdr->status ? ws.hdr->status : ws.hdr->late_status;
    if (st != 0) *(volatile int32_t *)ws.peer[q].err = st;
    __threadfence_system();
    st_release_sys(ws.peer[q].flags + ws.rank, s_e);
  }
}
__global__ void pk_wait(PsWs ws) {
  const int q = threadIdx.x;
  const uint32_t e = *ws.epoch;
  bool timed_out = false;
  // epoch[1] != 0: an earlier barrier of this rank timed out; the group is broken
  // until gs_pslicer_reset(), so do not spend timeout_ms on every later barrier
  const bool broken = ws.epoch[1] != 0u;
  if (q < ws.P && !broken) {
    const unsigned long long t0 = globaltimer_ns();
    const unsigned long long lim = (unsigned long long)ws.timeout_ms * 1000000ull;
    while ((int32_t)(ld_acquire_sys(ws.me.flags + q) - e) < 0) {
      if (globaltimer_ns() - t0 > lim) {
        timed_out = true;
        break;
      }
      __nanosleep(200);
    }
  }
  const int any_to = __syncthreads_or(timed_out ? 1 : 0);
  if (threadIdx.x == 0) {
    const int32_t pe = *(volatile int32_t *)ws.me.err;
    if (any_to) ws.epoch[1] = 1u;
    if (ws.hdr->status == 0) {
      if (any_to || broken) ws.hdr->status = (int32_t)GS_ERR_PEER;
      else if (pe != 0) ws.hdr->status = pe;
    }
  }
}

// ----------------------------------------------------------------- pk_begin
// One CTA.  Header reset, seeds (replicated: every rank is given the same
// batch) narrowed into the slab as frontier 0; the candidate count of every
// seed this rank owns is stored into every rank's cnt0[].
__global__ void __launch_bounds__(1024) pk_begin(PsWs ws) {
  gs_sample_header *h = ws.hdr;
  const int n = ws.input->n_seeds;
  __shared__ int s_bad;
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  gs_span *spans = &h->field[0][0][0];
  for (int i = threadIdx.x; i < GS_MAX_LAYERS * GS_MAX_PARTS * GS_NUM_FIELDS; i += blockDim.x) {
    spans[i].off = 0;
    spans[i].len = 0;
  }
  for (int i = threadIdx.x; i <= GS_MAX_LAYERS; i += blockDim.x) {
    h->frontier_len[i] = 0;
    h->frontier[i].off = 0;
    h->frontier[i].len = 0;
    if (i < GS_MAX_LAYERS) {
      h->cand_len[i] = 0;
      h->draw_base[i] = 0;
    }
  }
  bool bad = (n < 0 || n > ws.Fcap || (int64_t)n > ws.slab_cap);
  if (!bad)
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int64_t v = ws.seeds_in[i];
      if (v < 0 || v >= (int64_t)ws.num_nodes) {
        s_bad = 1;
      } else {
        ws.slab[i] = (int32_t)v;
        if (ps_owner(ws, (int32_t)v) == ws.rank) {
          const uint32_t c = ps_row_count(ws, (int32_t)v, ws.fanouts[0]);
          for (int q = 0; q < ws.P; q++) ws.peer[q].cnt0[i] = c;
        }
      }
    }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    bad = bad || s_bad;
    h->status = bad ? (int32_t)GS_ERR_SEED : 0;
    h->num_layers = ws.L;
    h->num_parts = ws.P;
    h->n_seeds = n;
    h->batch_index = ws.input->batch_index;
    h->draw_cursor = 0;
    h->edges = 0;
    h->late_status = 0;
    h->reserved[0] = ws.input->want_host;  // kept on the device: k_finish must not poll host memory from every thread
    h->feature_rows = 0;
    const int nn = bad ? 0 : n;
    h->slab_used = ((int64_t)nn + 3) & ~3ll;
    h->frontier_len[0] = nn;
    h->frontier[0].off = 0;
    h->frontier[0].len = nn;
  }
}

// Layer l reads its rows' candidate counts from cnt0 (l == 0, written by the
// owners through pk_begin) or cnt[l & 1] (written by pk_front_apply).
__device__ __forceinline__ const uint32_t *ps_counts(const PsWs &ws, int layer) {
  return layer == 0 ? ws.me.cnt0 : ws.cnt[layer & 1];
}

// ------------------------------------------------------------- pk_rows_count
// Per-CTA sums of the replicated candidate counts, last-CTA spine -> per-CTA
// prefixes and C_l.  Identical on every rank.
__global__ void __launch_bounds__(GS_TPB) pk_rows_count(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ unsigned long long s_w64[GS_NW + 1];
  const int F = h->frontier_len[layer];
  const uint32_t *cnt = ps_counts(ws, layer);
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  unsigned long long csum = 0ull;
  for (int t = t0; t < t1; t++)
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const int i = t * GS_TILE + it * GS_TPB + threadIdx.x;
      if (i < F) csum += (unsigned long long)__ldcg(cnt + i);
    }
  unsigned long long ctot;
  block_excl_scan64(csum, s_w64, ctot);
  if (threadIdx.x == 0) ws.ftile_sum[blockIdx.x] = ctot;
  if (!last_block(ws.tickets + 0)) return;
  const int nb = (int)gridDim.x;
  const int chunk = (nb + GS_TPB - 1) / GS_TPB;
  const int a = min(nb, (int)threadIdx.x * chunk), b = min(nb, a + chunk);
  unsigned long long cs = 0ull;
  for (int t = a; t < b; t++) cs += __ldcg(ws.ftile_sum + t);
  unsigned long long cpre = block_excl_scan64(cs, s_w64, ctot);
  const bool ok = ctot <= (unsigned long long)ws.Ccap;
  if (ok)
    for (int t = a; t < b; t++) {
      ws.ftile_pre[t] = (uint32_t)cpre;
      cpre += __ldcg(ws.ftile_sum + t);
    }
  if (threadIdx.x == 0) {
    if (!ok) {
      h->status = (int32_t)GS_ERR_OVERFLOW;
      h->cand_len[layer] = 0;
    } else {
      h->cand_len[layer] = (int32_t)ctot;
      ws.off[F] = (uint32_t)ctot;
    }
  }
}

// -------------------------------------------------------------- pk_rows_scan
// Candidate offset of every row.  Every rank plants every row's self entry in
// its own dense stream (rowscan is sampled there); for the rows it owns it
// also records the CSR extent and initialises the node's table entry.
__global__ void __launch_bounds__(GS_TPB) pk_rows_scan(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_w[GS_NW + 1];
  const int F = h->frontier_len[layer];
  const int32_t *front = ws.slab + h->frontier[layer].off;
  const uint32_t *cnt = ps_counts(ws, layer);
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  uint32_t base = ws.ftile_pre[blockIdx.x];
  for (int t = t0; t < t1; t++) {
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const int i = t * GS_TILE + it * GS_TPB + threadIdx.x;
      const uint32_t c = i < F ? __ldcg(cnt + i) : 0u;
      uint32_t ctot;
      const uint32_t ex = block_excl_scan(c, s_w, ctot);
      if (i < F) {
        const int32_t nd = front[i];
        const uint32_t o = base + ex;
        const bool own = ps_owner(ws, nd) == ws.rank;
        ws.off[i] = o;
        ws.me.cand[o] = i;
        ws.me.cflag[o] = (uint8_t)(PS_CF_SELF | (own ? PS_CF_OWN : 0u));
        if (own) {
          const int32_t li = ps_local(ws, nd, ws.rank);
          const int64_t s = __ldg(ws.indptr + li), e = __ldg(ws.indptr + li + 1);
          ws.rowstart[i] = s;
          ws.rowdeg[i] = (int32_t)(e - s);
          ws.me.table[li] = make_uint2(GS_INF, o);
          ws.rowrem[i] = 0;
        }
        for (uint32_t tb = (o + GS_TILE - 1) / GS_TILE; (uint64_t)tb * GS_TILE < (uint64_t)o + c; tb++)
          ws.tile_first[tb] = (uint32_t)i;
      }
      base += ctot;
    }
  }
}

// ----------------------------------------------------------------- pk_expand
// Candidate-parallel expansion of the rows this rank owns.  Every sampled edge
// (p, src) is delivered to owner(src): the source id and an EDGE flag at
// position p of that rank's dense stream, and red.min(p) into its table entry
// of src -- plain 4-byte / 1-byte stores and one reduction over NVLink per edge
// (local when this rank owns src).  A pick equal to the destination is a self
// edge (slicer.cpp:33): nothing to deliver.
__global__ void __launch_bounds__(GS_TPB) pk_expand(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_off[GS_TILE + 2];
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *__restrict__ front = ws.slab + h->frontier[layer].off;
  const int fan = ws.fanouts[layer];
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const uint32_t batch = (uint32_t)h->batch_index;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int i_lo = (int)ws.tile_first[t];
    const int i_hi = (t + 1 < ntiles) ? (int)ws.tile_first[t + 1] : F - 1;
    const int cnt = i_hi - i_lo + 1;
    for (int k = threadIdx.x; k <= cnt; k += GS_TPB) s_off[k] = ws.off[i_lo + k];
    __syncthreads();
    uint32_t pos[GS_IPT];
    int row[GS_IPT], slot[GS_IPT];
    int32_t nd1[GS_IPT];
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const uint32_t p = (uint32_t)t * GS_TILE + it * GS_TPB + threadIdx.x;
      pos[it] = p;
      slot[it] = 0;
      row[it] = 0;
      if (p < C) {
        int lo = 0, hi = cnt - 1;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (s_off[mid] <= p) lo = mid; else hi = mid - 1;
        }
        slot[it] = (int)(p - s_off[lo]);
        row[it] = i_lo + lo;
      }
    }
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      if (slot[it] > 0) {
        nd1[it] = front[row[it]];
        if (ps_owner(ws, nd1[it]) != ws.rank) slot[it] = 0;  // another rank expands this row
      }
    }
    const int32_t *src[GS_IPT];
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      if (slot[it] > 0) {
        const int64_t rs = ws.rowstart[row[it]];
        const int32_t deg = ws.rowdeg[row[it]];
        int32_t pick = slot[it] - 1;
        if (fan >= 0 && deg >= fan) {
          if (ws.rng_mode == (int)GS_RNG_PHILOX_NOREPLACE) {
            const uint4 key = philox_block(ws.seed_lo, ws.seed_hi, batch, layer, nd1[it], GS_NR_KEY_BLOCK);
            pick = (int32_t)nr_pick(key, (uint32_t)pick, (uint32_t)deg);
          } else {
            const uint32_t r = philox_pick(ws.seed_lo, ws.seed_hi, batch, layer, nd1[it], slot[it] - 1);
            pick = (int32_t)(r % (uint32_t)deg);
          }
        }
        src[it] = ws.indices + rs + pick;
      }
    }
    int32_t v[GS_IPT];
#pragma unroll
    for (int it = 0; it < GS_IPT; it++)
      if (slot[it] > 0) v[it] = __ldg(src[it]);
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      if (slot[it] > 0 && v[it] != nd1[it]) {
        const int q = ps_owner(ws, v[it]);
        const PsArena &pa = ws.peer[q];
        pa.cand[pos[it]] = v[it];
        pa.cflag[pos[it]] = (uint8_t)PS_CF_EDGE;
        atomicMin(&pa.table[ps_local(ws, v[it], q)].x, pos[it]);
        if (q != ws.rank && !(ws.rowrem[row[it]] & (1u << q))) atomicOr(&ws.rowrem[row[it]], 1u << q);
      }
    }
    __syncthreads();
  }
  __threadfence_system();
}

// ------------------------------------------------------------- pk_cand_flags
// Classify the candidates this rank holds (sources it owns + self entries of
// rows it owns); per-CTA counts of in_nodes firsts and edges, spine, slab
// layout of in_nodes / csr_indices of B[layer][rank].
__global__ void __launch_bounds__(GS_TPB) pk_cand_flags(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_cnt[2];
  __shared__ uint32_t s_tot[PS_NCTR];
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *front = ws.slab + h->frontier[layer].off;
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  for (int t = t0; t < t1; t++) {
    uint32_t fl[GS_IPT];
    int32_t cv[GS_IPT];
    uint2 te[GS_IPT];
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const uint32_t p = (uint32_t)t * GS_TILE + it * GS_TPB + threadIdx.x;
      fl[it] = 0u;
      cv[it] = 0;
      if (p < C) {
        fl[it] = __ldcg(ws.me.cflag + p);
        if (fl[it]) cv[it] = __ldcg(ws.me.cand + p);
      }
      if ((fl[it] & PS_CF_SELF) && !(fl[it] & PS_CF_OWN)) fl[it] |= 0x80u;  // nothing to look up
    }
#pragma unroll
    for (int it = 0; it < GS_IPT; it++)
      if ((fl[it] & PS_CF_SELF) && !(fl[it] & 0x80u)) cv[it] = front[cv[it]];
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      te[it] = make_uint2(0u, 0u);
      if (fl[it] && !(fl[it] & 0x80u)) te[it] = __ldcg(ws.me.table + ps_local(ws, cv[it], ws.rank));
    }
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const uint32_t p = (uint32_t)t * GS_TILE + it * GS_TPB + threadIdx.x;
      bool ffirst = false, ifirst = false, edge = false;
      const uint32_t f = fl[it];
      if (f && !(f & 0x80u)) {
        const uint2 e = te[it];
        if (f & PS_CF_SELF) {
          ffirst = e.x > p;
          if (e.y != p) h->late_status = (int32_t)GS_ERR_SEED;  // duplicate seed
        } else {
          edge = true;
          ifirst = e.x == p;
          ffirst = ifirst && e.y > p;
        }
        ws.me.cflag[p] = (uint8_t)((f & (PS_CF_SELF | PS_CF_OWN | PS_CF_EDGE)) |
                                   (ffirst ? PS_CF_FFIRST : 0u) | (ifirst ? PS_CF_IFIRST : 0u));
      }
      const uint32_t mi = __ballot_sync(FULL, ifirst);
      const uint32_t me = __ballot_sync(FULL, edge);
      if (lane_id() == 0) {
        if (mi) atomicAdd(&s_cnt[0], (uint32_t)__popc(mi));
        if (me) atomicAdd(&s_cnt[1], (uint32_t)__popc(me));
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 2) ws.ctr[(size_t)threadIdx.x * ws.grid_max + blockIdx.x] = s_cnt[threadIdx.x];
  if (!last_block(ws.tickets + 1)) return;
  spine_scan<PS_NCTR>(ws.ctr, ws.grid_max, (int)gridDim.x, 2, s_tot, nullptr);
  if (threadIdx.x == 0) {
    bool ok = true;
    if (h->late_status != 0) {
      h->status = h->late_status;
      ok = false;
    }
    int64_t used = h->slab_used;
    const int g = ws.rank;
    h->field[layer][g][GS_F_IN_NODES] = slab_alloc(used, ws.slab_cap, s_tot[0], ok);
    h->field[layer][g][GS_F_CSR_INDICES] = slab_alloc(used, ws.slab_cap, s_tot[1], ok);
    ws.rowscan[F] = s_tot[1];
    if (ok) {
      h->slab_used = used;
      h->edges += s_tot[1];
    } else if (h->status == 0) {
      h->status = (int32_t)GS_ERR_OVERFLOW;
    }
  }
}

// ------------------------------------------------------------- pk_cand_apply
// Ranks from the spine + in-tile ballots, in position order: in_nodes,
// csr_indices (first occurrences write their final value, the rest ~id for
// pk_relabel), the relabel map, rowscan at every row's self entry.  Every node
// found to be new to the frontier is published at its position to every rank,
// together with the candidate count its row will have in the next layer.  The
// dense flags are left zeroed for the next layer.
__global__ void __launch_bounds__(GS_TPB) pk_cand_apply(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_wcnt[GS_NW][2];
  __shared__ uint32_t s_base[GS_NW][2];
  __shared__ uint32_t s_run[2][2];
  const int g = ws.rank;
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *__restrict__ front = ws.slab + h->frontier[layer].off;
  int32_t *in_nodes = ws.slab + h->field[layer][g][GS_F_IN_NODES].off;
  int32_t *csr_ix = ws.slab + h->field[layer][g][GS_F_CSR_INDICES].off;
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const int fan_next = layer + 1 < ws.L ? ws.fanouts[layer + 1] : 0;
  const bool need_cnt = layer + 1 < ws.L;
  const int lane = lane_id(), w = warp_id();
  const uint32_t ltm = lt_mask();
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  if (threadIdx.x < 2) s_run[t0 & 1][threadIdx.x] = ws.ctr[(size_t)threadIdx.x * ws.grid_max + blockIdx.x];
  for (int t = t0; t < t1; t++) {
    __syncthreads();
    const uint32_t wbase = (uint32_t)t * GS_TILE + (uint32_t)w * (GS_IPT * 32);
    uint32_t f[GS_IPT];
    int32_t ci[GS_IPT], cv[GS_IPT];
#pragma unroll
    for (int r = 0; r < GS_IPT; r++) {
      const uint32_t p = wbase + r * 32 + lane;
      f[r] = 0u;
      ci[r] = 0;
      if (p < C) {
        f[r] = ws.me.cflag[p];
        if (f[r]) {
          ci[r] = ws.me.cand[p];
          ws.me.cflag[p] = 0;
        }
      }
    }
#pragma unroll
    for (int r = 0; r < GS_IPT; r++) cv[r] = (f[r] & PS_CF_SELF) ? front[ci[r]] : ci[r];
    {
      uint32_t ni = 0u, ne = 0u;
#pragma unroll
      for (int r = 0; r < GS_IPT; r++) {
        ni += __popc(__ballot_sync(FULL, (f[r] & PS_CF_IFIRST) != 0u));
        ne += __popc(__ballot_sync(FULL, (f[r] & PS_CF_EDGE) != 0u));
      }
      if (lane == 0) {
        s_wcnt[w][0] = ni;
        s_wcnt[w][1] = ne;
      }
    }
    __syncthreads();
    if (threadIdx.x < 2 * GS_NW) {
      const int cc = threadIdx.x / GS_NW, ww = threadIdx.x % GS_NW;
      uint32_t a = s_run[t & 1][cc];
      for (int k = 0; k < ww; k++) a += s_wcnt[k][cc];
      s_base[ww][cc] = a;
      if (ww == GS_NW - 1) s_run[(t + 1) & 1][cc] = a + s_wcnt[ww][cc];
    }
    __syncthreads();
    uint32_t ri = s_base[w][0], re = s_base[w][1];
#pragma unroll
    for (int r = 0; r < GS_IPT; r++) {
      const uint32_t p = wbase + r * 32 + lane;
      const bool isself = (f[r] & PS_CF_SELF) != 0u, edge = (f[r] & PS_CF_EDGE) != 0u;
      const bool ffirst = (f[r] & PS_CF_FFIRST) != 0u, ifirst = (f[r] & PS_CF_IFIRST) != 0u;
      const uint32_t mi = __ballot_sync(FULL, ifirst);
      const uint32_t me = __ballot_sync(FULL, edge);
      const uint32_t ex = re + __popc(me & ltm);
      if (edge) {
        int32_t ixv = ~cv[r];
        if (ifirst) {
          const uint32_t rk = ri + __popc(mi & ltm);
          in_nodes[rk] = cv[r];
          ws.me.table[ps_local(ws, cv[r], g)].x = rk;
          ixv = (int32_t)rk;
        }
        csr_ix[ex] = ixv;
      }
      if (isself) ws.rowscan[ci[r]] = ex;
      if (ffirst) {
        const uint32_t c = need_cnt ? ps_row_count(ws, cv[r], fan_next) : 1u;
        const uint2 e = make_uint2((uint32_t)cv[r], c);
        for (int q = 0; q < ws.P; q++) ws.peer[q].nf[p] = e;
      }
      ri += __popc(mi);
      re += __popc(me);
    }
  }
  __threadfence_system();
}

// ------------------------------------------------------------ pk_front_count
// Two scans share this kernel and its spine: (1) the dense new-frontier marks
// of all ranks -> F_{l+1}; (2) the row-side classes of B[layer][rank]
// (bipartite.h:33-66): out = row has an edge in my part; owned = out and I own
// the row; to = I own the row and it sampled a remote source; self = I own it.
// Per-peer counters, 16 bits per rank in two 64-bit words (ranks 0-3 / 4-7): a
// CTA-wide scan of one word ranks a round of GS_TPB rows for four peers at once.
struct Pk16 {
  unsigned long long w[2];
};
__device__ __forceinline__ void pk16_add(Pk16 &v, int q) { v.w[q >> 2] += 1ull << (16 * (q & 3)); }
__device__ __forceinline__ uint32_t pk16_get(const Pk16 &v, int q) { return (uint32_t)(v.w[q >> 2] >> (16 * (q & 3))) & 0xFFFFu; }

// Row classes of B[layer][rank] (bipartite.h:33-66).  fromq: the rank that owns the
// row when this rank holds a partial aggregate for it (-1 otherwise); tomask: ranks
// that hold partial aggregates for a row this rank owns.
__device__ __forceinline__ void ps_row_class(const PsWs &ws, const int32_t *front, int i, int F,
                                             bool &out, bool &own, bool &to, bool &self, int &fromq,
                                             uint32_t &tomask) {
  out = own = to = self = false;
  fromq = -1;
  tomask = 0u;
  if (i < F) {
    const int q0 = ps_owner(ws, front[i]);
    const bool mine = q0 == ws.rank;
    out = ws.rowscan[i + 1] != ws.rowscan[i];
    own = out && mine;
    self = mine;
    if (mine) tomask = ws.rowrem[i];
    to = tomask != 0u;
    if (out && !mine) fromq = q0;
  }
}

__global__ void __launch_bounds__(GS_TPB) pk_front_count(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_cnt[PS_NCTR];
  __shared__ uint32_t s_tot[PS_NCTR];
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *front = ws.slab + h->frontier[layer].off;
  if (threadIdx.x < PS_NCTR) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  int t0, t1;
  cta_tiles((int)((C + GS_TILE - 1) / GS_TILE), t0, t1);
  uint32_t nf = 0u;
  for (int t = t0; t < t1; t++)
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const uint32_t p = (uint32_t)t * GS_TILE + it * GS_TPB + threadIdx.x;
      if (p < C && __ldcg(&ws.me.nf[p].y) != 0u) nf++;
    }
  cta_tiles((F + GS_TILE - 1) / GS_TILE, t0, t1);
  uint32_t n_out = 0u, n_own = 0u, n_to = 0u, n_self = 0u;
  Pk16 pf = {{0ull, 0ull}}, pt = {{0ull, 0ull}};  // per-peer from / to rows
  for (int t = t0; t < t1; t++)
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      bool o, w, to, sf;
      int fq;
      uint32_t tm;
      ps_row_class(ws, front, t * GS_TILE + it * GS_TPB + threadIdx.x, F, o, w, to, sf, fq, tm);
      n_out += o; n_own += w; n_to += to; n_self += sf;
      if (fq >= 0) pk16_add(pf, fq);
      for (int q = 0; q < ws.P; q++)
        if ((tm >> q) & 1u) pk16_add(pt, q);
    }
  const uint32_t v[5] = {nf, n_out, n_own, n_to, n_self};
#pragma unroll
  for (int c = 0; c < 5; c++) {
    uint32_t x = v[c];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x += __shfl_xor_sync(FULL, x, d);
    if (lane_id() == 0 && x) atomicAdd(&s_cnt[c], x);
  }
  for (int q = 0; q < ws.P; q++) {  // per-thread fields stay far below 2^16; sum them unpacked
    uint32_t a = pk16_get(pf, q), b = pk16_get(pt, q);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      a += __shfl_xor_sync(FULL, a, d);
      b += __shfl_xor_sync(FULL, b, d);
    }
    if (lane_id() == 0) {
      if (a) atomicAdd(&s_cnt[5 + q], a);
      if (b) atomicAdd(&s_cnt[5 + ws.P + q], b);
    }
  }
  __syncthreads();
  const int NCT = 5 + 2 * ws.P;
  if (threadIdx.x < NCT) ws.ctr[(size_t)threadIdx.x * ws.grid_max + blockIdx.x] = s_cnt[threadIdx.x];
  if (!last_block(ws.tickets + 2)) return;
  spine_scan<PS_NCTR>(ws.ctr, ws.grid_max, (int)gridDim.x, NCT, s_tot, nullptr);
  if (threadIdx.x == 0) {
    bool ok = true;
    const uint32_t fnext = s_tot[0];
    if (layer + 1 < ws.L && fnext > (uint32_t)ws.Fcap) ok = false;
    int64_t used = h->slab_used;
    const gs_span fs = slab_alloc(used, ws.slab_cap, fnext, ok);
    const uint32_t no = s_tot[1], nw = s_tot[2], nt = s_tot[3], ns = s_tot[4];
    gs_span *fl = h->field[layer][ws.rank];
    const gs_span ip = slab_alloc(used, ws.slab_cap, no + 1u, ok);
    fl[GS_F_OUT_NODES] = slab_alloc(used, ws.slab_cap, no, ok);
    fl[GS_F_CSR_INDPTR] = ip;
    fl[GS_F_OWNED_OUT] = slab_alloc(used, ws.slab_cap, nw, ok);
    fl[GS_F_FROM_IDS] = slab_alloc(used, ws.slab_cap, no - nw, ok);
    fl[GS_F_TO_IDS] = slab_alloc(used, ws.slab_cap, nt, ok);
    fl[GS_F_SELF_IN] = slab_alloc(used, ws.slab_cap, ns, ok);
    fl[GS_F_SELF_OUT] = slab_alloc(used, ws.slab_cap, ns, ok);
    // per-peer split, in the slots [q != rank] the reference leaves empty (bipartite.h:19-21):
    // FROM_IDS[q] = my out positions of rows rank q owns; TO_IDS[q] = my out position (or -1)
    // of rows I own that rank q holds a partial aggregate for.  Entry k of one pairs with
    // entry k of the other on rank q.
    for (int q = 0; q < ws.P; q++)
      if (q != ws.rank) {
        h->field[layer][q][GS_F_FROM_IDS] = slab_alloc(used, ws.slab_cap, s_tot[5 + q], ok);
        h->field[layer][q][GS_F_TO_IDS] = slab_alloc(used, ws.slab_cap, s_tot[5 + ws.P + q], ok);
      }
    if (ok) {
      ws.slab[ip.off + no] = (int32_t)ws.rowscan[F];
      h->slab_used = used;
      h->frontier[layer + 1] = fs;
      h->frontier_len[layer + 1] = (int32_t)fnext;
    } else {
      h->status = (int32_t)GS_ERR_OVERFLOW;
    }
  }
}

// ------------------------------------------------------------ pk_front_apply
// (1) compaction of the dense new-frontier marks into frontier l+1 and its
// candidate counts (identical on every rank), zeroing the marks behind it;
// (2) every row-side list of B[layer][rank].  Ranks come from a CTA-wide scan
// per round of GS_TPB consecutive items, so list order is position order.
__global__ void __launch_bounds__(GS_TPB) pk_front_apply(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_w[GS_NW + 1];
  const int g = ws.rank;
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *__restrict__ front = ws.slab + h->frontier[layer].off;
  int32_t *fnext = ws.slab + h->frontier[layer + 1].off;
  uint32_t *cnext = ws.cnt[(layer + 1) & 1];
  int t0, t1;
  cta_tiles((int)((C + GS_TILE - 1) / GS_TILE), t0, t1);
  uint32_t base = ws.ctr[blockIdx.x];
  for (int t = t0; t < t1; t++)
    for (int it = 0; it < GS_IPT; it++) {
      const uint32_t p = (uint32_t)t * GS_TILE + it * GS_TPB + threadIdx.x;
      uint2 e = make_uint2(0u, 0u);
      if (p < C) e = __ldcg(ws.me.nf + p);
      uint32_t tot;
      const uint32_t ex = block_excl_scan(e.y != 0u ? 1u : 0u, s_w, tot);
      if (e.y != 0u) {
        fnext[base + ex] = (int32_t)e.x;
        cnext[base + ex] = e.y;
        ws.me.nf[p] = make_uint2(0u, 0u);
      }
      base += tot;
    }
  const gs_span *fl = h->field[layer][g];
  int32_t *out_nodes = ws.slab + fl[GS_F_OUT_NODES].off, *ip = ws.slab + fl[GS_F_CSR_INDPTR].off;
  int32_t *owned = ws.slab + fl[GS_F_OWNED_OUT].off, *from = ws.slab + fl[GS_F_FROM_IDS].off;
  int32_t *to_ids = ws.slab + fl[GS_F_TO_IDS].off, *self_in = ws.slab + fl[GS_F_SELF_IN].off;
  int32_t *self_out = ws.slab + fl[GS_F_SELF_OUT].off;
  cta_tiles((F + GS_TILE - 1) / GS_TILE, t0, t1);
  uint32_t b_out = ws.ctr[(size_t)1 * ws.grid_max + blockIdx.x], b_own = ws.ctr[(size_t)2 * ws.grid_max + blockIdx.x];
  uint32_t b_to = ws.ctr[(size_t)3 * ws.grid_max + blockIdx.x], b_self = ws.ctr[(size_t)4 * ws.grid_max + blockIdx.x];
  __shared__ unsigned long long s_w64[GS_NW + 1];
  __shared__ uint32_t s_bf[GS_MAX_PARTS], s_bt[GS_MAX_PARTS];  // running per-peer bases of this CTA
  __shared__ int32_t s_fo[GS_MAX_PARTS], s_to[GS_MAX_PARTS];   // slab offsets of the per-peer lists
  if (threadIdx.x < ws.P) {
    const int q = threadIdx.x;
    s_bf[q] = ws.ctr[(size_t)(5 + q) * ws.grid_max + blockIdx.x];
    s_bt[q] = ws.ctr[(size_t)(5 + ws.P + q) * ws.grid_max + blockIdx.x];
    s_fo[q] = h->field[layer][q][GS_F_FROM_IDS].off;
    s_to[q] = h->field[layer][q][GS_F_TO_IDS].off;
  }
  __syncthreads();
  for (int t = t0; t < t1; t++)
    for (int it = 0; it < GS_IPT; it++) {
      const int i = t * GS_TILE + it * GS_TPB + threadIdx.x;
      bool o, w, to, sf;
      int fq;
      uint32_t tm;
      ps_row_class(ws, front, i, F, o, w, to, sf, fq, tm);
      uint32_t t_out, t_own, t_to, t_self;
      const uint32_t op = b_out + block_excl_scan(o ? 1u : 0u, s_w, t_out);
      const uint32_t ow = b_own + block_excl_scan(w ? 1u : 0u, s_w, t_own);
      const uint32_t pt = b_to + block_excl_scan(to ? 1u : 0u, s_w, t_to);
      const uint32_t ps = b_self + block_excl_scan(sf ? 1u : 0u, s_w, t_self);
      // per-peer ranks: 16-bit fields, four peers per CTA-wide scan
      Pk16 vf = {{0ull, 0ull}}, vt = {{0ull, 0ull}}, ef, et, tf, tt;
      if (fq >= 0) pk16_add(vf, fq);
      for (int q = 0; q < ws.P; q++)
        if ((tm >> q) & 1u) pk16_add(vt, q);
      const int nwords = ws.P > 4 ? 2 : 1;
      ef.w[1] = et.w[1] = tf.w[1] = tt.w[1] = 0ull;
      for (int k = 0; k < nwords; k++) {
        ef.w[k] = block_excl_scan64(vf.w[k], s_w64, tf.w[k]);
        et.w[k] = block_excl_scan64(vt.w[k], s_w64, tt.w[k]);
      }
      if (i < F) {
        const int32_t nd = front[i];
        if (o) {
          out_nodes[op] = nd;
          ip[op] = (int32_t)ws.rowscan[i];
          if (w) owned[ow] = (int32_t)op;
          else from[op - ow] = (int32_t)op;
        }
        const int32_t op_own = w ? (int32_t)op : -1;
        if (fq >= 0) ws.slab[s_fo[fq] + s_bf[fq] + pk16_get(ef, fq)] = (int32_t)op;
        for (int q = 0; q < ws.P; q++)
          if ((tm >> q) & 1u) ws.slab[s_to[q] + s_bt[q] + pk16_get(et, q)] = op_own;
        if (to) to_ids[pt] = op_own;
        if (sf) {
          const uint32_t x = ws.me.table[ps_local(ws, nd, g)].x;
          self_in[ps] = x == GS_INF ? -1 : (int32_t)x;
          self_out[ps] = op_own;
        }
      }
      b_out += t_out; b_own += t_own; b_to += t_to; b_self += t_self;
      __syncthreads();  // everyone has read the per-peer bases of this round
      if (threadIdx.x < ws.P) {
        s_bf[threadIdx.x] += pk16_get(tf, threadIdx.x);
        s_bt[threadIdx.x] += pk16_get(tt, threadIdx.x);
      }
      __syncthreads();
    }
}

// ---------------------------------------------------------------- pk_relabel
__global__ void __launch_bounds__(GS_TPB) pk_relabel(PsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const gs_span sp = h->field[layer][ws.rank][GS_F_CSR_INDICES];
  int32_t *ix = ws.slab + sp.off;
  const int stride = gridDim.x * GS_TPB;
  for (int e0 = blockIdx.x * GS_TPB + threadIdx.x; e0 < sp.len; e0 += GS_IPT * stride) {
    int32_t v[GS_IPT];
    uint32_t r[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int e = e0 + k * stride;
      v[k] = e < sp.len ? ix[e] : 0;
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) r[k] = v[k] < 0 ? ws.me.table[ps_local(ws, ~v[k], ws.rank)].x : 0u;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int e = e0 + k * stride;
      if (v[k] < 0) ix[e] = (int32_t)r[k];
    }
  }
}

// ----------------------------------------------------------------- pk_gather
// Fused input-feature gather: in_nodes of this rank's part are nodes it owns, so
// their rows are in the local shard.  One warp per 2408-byte row: streaming.
__global__ void __launch_bounds__(GS_TPB) pk_gather(PsWs ws) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const gs_span sp = h->field[ws.L - 1][ws.rank][GS_F_IN_NODES];
  const int R = sp.len;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const bool fits = (int64_t)R <= ws.feat_cap_rows;
    h->feature_rows = fits ? R : 0;
    for (int q = 0; q <= ws.P; q++) h->feature_part_off[q] = (fits && q > ws.rank) ? R : 0;
    if (!fits) h->late_status = (int32_t)GS_ERR_OVERFLOW;
  }
  if ((int64_t)R > ws.feat_cap_rows) return;
  const int32_t *ids = ws.slab + sp.off;
  const float *feat = ws.feat;
  const int dim = ws.feat_dim;
  auto src_of_row = [&](int64_t r) -> const float * {
    return feat + (int64_t)ps_local(ws, __ldg(ids + r), ws.rank) * dim;
  };
  gather_block(dim, (int64_t)R, ws.feat_out, src_of_row, (int64_t)blockIdx.x * GS_NW + warp_id(),
               (int64_t)gridDim.x * GS_NW, lane_id());
}

// ---------------------------------------------------------------- pk_shuffle
// Split-parallel shuffle of one layer's partial aggregates (dist_sage.cu:120-160,
// 193-199; bipartite.py:88-99).  push: one warp per row, straight into the
// owner's staging arena over NVLink (slot [sender][k]).  merge: after the
// barrier, the receiver folds one peer's rows in; the host launches the peers in
// rank order, and inside a peer every destination row is distinct, so the sums
// are deterministic.  backward is the transpose with copy instead of add.
__global__ void __launch_bounds__(GS_TPB) pk_shuffle_push(PsWs ws, int layer, const float *rows, int dim,
                                                          int backward) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const int lane = lane_id();
  const int gw = blockIdx.x * GS_NW + warp_id(), nw = gridDim.x * GS_NW;
  for (int q = 0; q < ws.P; q++) {
    if (q == ws.rank) continue;
    const gs_span sp = h->field[layer][q][backward ? GS_F_TO_IDS : GS_F_FROM_IDS];
    const int32_t *list = ws.slab + sp.off;
    float *dst = ws.peer[q].stage + (size_t)ws.rank * ws.stage_rows * ws.stage_dim;
    for (int k = gw; k < sp.len; k += nw) {
      const int32_t r = list[k];
      float *d = dst + (size_t)k * dim;
      if (r >= 0) {
        gather_row(rows, dim, (int64_t)r, d, 0, lane);
      } else {  // backward: the owner has no row for this node (reference semantics: -1)
        for (int c = lane; c < dim; c += 32) d[c] = 0.0f;
      }
    }
  }
  __threadfence_system();
}

__global__ void __launch_bounds__(GS_TPB) pk_shuffle_merge(PsWs ws, int layer, float *rows, int dim,
                                                           int backward, int q) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const int lane = lane_id();
  const int gw = blockIdx.x * GS_NW + warp_id(), nw = gridDim.x * GS_NW;
  const gs_span sp = h->field[layer][q][backward ? GS_F_FROM_IDS : GS_F_TO_IDS];
  const int32_t *list = ws.slab + sp.off;
  const float *src = ws.me.stage + (size_t)q * ws.stage_rows * ws.stage_dim;
  for (int k = gw; k < sp.len; k += nw) {
    const int32_t r = list[k];
    if (r < 0) continue;
    const float *sr = src + (size_t)k * dim;
    float *d = rows + (size_t)r * dim;
    if (backward) {
      for (int c = lane; c < dim; c += 32) d[c] = __ldcg(sr + c);
    } else {
      for (int c = lane; c < dim; c += 32) d[c] += __ldcg(sr + c);
    }
  }
}

// ----------------------------------------------------------------- pk_finish
// Sparse reset of this rank's table (every node touched is in the last
// frontier; this rank resets the ones it owns), header and slab -> pinned host.
__global__ void __launch_bounds__(GS_TPB) pk_finish(PsWs ws) {
  gs_sample_header *h = ws.hdr;
  const int tid = blockIdx.x * GS_TPB + threadIdx.x, nth = gridDim.x * GS_TPB;
  if (h->status == 0) {
    const int32_t *fl = ws.slab + h->frontier[ws.L].off;
    const int FL = h->frontier_len[ws.L];
    for (int i = tid; i < FL; i += nth) {
      const int32_t nd = fl[i];
      if (ps_owner(ws, nd) == ws.rank) ws.me.table[ps_local(ws, nd, ws.rank)] = make_uint2(GS_INF, GS_INF);
    }
  }
  {
    const int nw = (int)(sizeof(gs_sample_header) / 4);
    const int32_t *src = (const int32_t *)h;
    int32_t *dst = (int32_t *)ws.hdr_host;
    for (int i = tid; i < nw; i += nth) dst[i] = src[i];
  }
  if (h->reserved[0] && ws.slab_host && h->status == 0) {
    const int64_t n4 = h->slab_used >> 2;
    const int4 *src = (const int4 *)ws.slab;
    int4 *dst = (int4 *)ws.slab_host;
    for (int64_t i = tid; i < n4; i += nth) dst[i] = src[i];
  }
}

}  // namespace

void ps_launch_begin(const PsWs &ws, cudaStream_t st) { pk_begin<<<1, 1024, 0, st>>>(ws); }
void ps_launch_signal(const PsWs &ws, cudaStream_t st) { pk_signal<<<1, 32, 0, st>>>(ws); }
void ps_launch_wait(const PsWs &ws, cudaStream_t st) { pk_wait<<<1, 32, 0, st>>>(ws); }
void ps_launch_rows_count(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_rows_count<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_rows_scan(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_rows_scan<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_expand(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_expand<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_cand_flags(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_cand_flags<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_cand_apply(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_cand_apply<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_front_count(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_front_count<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_front_apply(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_front_apply<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_relabel(const PsWs &ws, int layer, int grid, cudaStream_t st) {
  pk_relabel<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void ps_launch_gather(const PsWs &ws, int grid, cudaStream_t st) { pk_gather<<<grid, GS_TPB, 0, st>>>(ws); }
void ps_launch_shuffle_push(const PsWs &ws, int layer, float *rows, int dim, int backward, int grid,
                            cudaStream_t st) {
  pk_shuffle_push<<<grid, GS_TPB, 0, st>>>(ws, layer, rows, dim, backward);
}
void ps_launch_shuffle_merge(const PsWs &ws, int layer, float *rows, int dim, int backward, int peer,
                             int grid, cudaStream_t st) {
  pk_shuffle_merge<<<grid, GS_TPB, 0, st>>>(ws, layer, rows, dim, backward, peer);
}
void ps_launch_finish(const PsWs &ws, int grid, cudaStream_t st) { pk_finish<<<grid, GS_TPB, 0, st>>>(ws); }
