// slicer_kernels.cu -- hand-written sm_100a kernels of the minibatch slicer.
//
// One layer of the reference's sequential slice_layer (slicer.cpp:25-64) +
// BiPartite::reorder (bipartite.cpp:3-17) becomes FOUR data-parallel launches, each
// a single pass over its stream; a launch hosts several roles (CTA ranges):
//
//   k_rows         frontier rows: CSR extents, candidate offsets (flat decoupled look-back
//                  scan -- the only cross-CTA wait of the chain), the rows' rank among the
//                  rows of their owner, and the mark "node v is row i of the current
//                  frontier" in the dedup table;
//   k_expand_rows  (fan > 0; k_expand for fan = -1) CSR expansion + Philox picks; one
//                  returning atom.min per edge source finds out whether the source is new
//                  to the layer or already a frontier row, and that answer makes every
//                  list length of the layer exact BEFORE any list is written: the last CTA
//                  lays the whole layer out in the sample slab (or in the parts' hand-over
//                  blocks).  Extra CTAs finish the PREVIOUS layer's lists: relabel and
//                  row side;
//   k_cand_flags   verdict per candidate (first occurrence / frontier slot), class counts
//                  per CTA run, spine scan by the last CTA; extra CTAs count the row-side
//                  classes the same way;
//   k_cand_apply   ranks inside a run from the spine prefix: next frontier, in_nodes, edge
//                  scratch (-> csr_indices), relabel map, per-row edge counts;
//   k_tail         relabel and row side of the last layer, fused feature gather, the
//                  slot's housekeeping for the next sample, header to the host.
//
// The sequential order the reference's lists are in (first occurrence in a scan
// over frontier rows and, inside a row, over [self, neighbours...]) is carried by
// a 32-bit *candidate position*: row i owns positions [off[i], off[i+1]), slot 0
// being the self entry.  "First occurrence" = minimum position (atom.min into a
// direct-address table), "list order" = ascending position (prefix scans over the
// candidate stream), so every output list is bit-identical to the sequential
// algorithm without any sort.
//
// All sizes live in device memory (gs_sample_header); tiles are handed out by
// tickets, so a whole minibatch -- or a group of them, blockIdx.y -- is one
// static CUDA graph of 4 L + 1 launches with no host round trip.  Integer / byte
// work, HBM- and latency-bound: no tensor cores (nothing here is a contraction).
#include <type_traits>

#include "slicer_device.cuh"

// Resident CTAs per SM the tiled kernels are compiled for (register caps 65536 /
// (GS_TPB * n)): the kernels wait on dependent random accesses, so what counts is
// how many warps of all the launches in flight fit an SM.  ptxas -v at these
// caps: k_rows 64 registers, the others 40 (a few dozen bytes of spills each);
// A/B of the caps: profiles/r2_variants_table.txt.
#ifndef GS_MINB_ROWS
#define GS_MINB_ROWS 4
#endif
#ifndef GS_MINB_EXPAND
#define GS_MINB_EXPAND 6
#endif
#ifndef GS_MINB_CAND
#define GS_MINB_CAND 6
#endif
#define GS_BOUNDS __launch_bounds__(GS_TPB)

namespace {

// --------------------------------------------------- L2 residency control
// The per-slot dedup tables are the only randomly re-read data; the CSR gathers
// are touched once.  With createpolicy hints the tables are kept in B200's 126 MB
// L2 (evict_last) while the picks pass through as evict_first.  ws.l2_hints = 0
// falls back to plain accesses (same results, for A/B measurements).
__device__ __forceinline__ uint64_t l2_policy_keep() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_stream() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
struct Tbl {  // table access with an optional keep-in-L2 policy
  uint32_t *t;
  uint64_t pol;
  bool hint;
  __device__ __forceinline__ Tbl(const GsWs &ws) : t(ws.table), pol(0), hint(ws.l2_hints != 0) {
    if (hint) pol = l2_policy_keep();
  }
  __device__ __forceinline__ uint32_t ld(int32_t v) const {
    if (!hint) return __ldcg(t + v);
    uint32_t r;
    asm volatile("ld.global.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(r) : "l"(t + v), "l"(pol));
    return r;
  }
  __device__ __forceinline__ void st(int32_t v, uint32_t x) const {
    if (!hint) { t[v] = x; return; }
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" :: "l"(t + v), "r"(x), "l"(pol) : "memory");
  }
  // returning minimum: the old value tells the caller what the node was before
  __device__ __forceinline__ uint32_t atom_min(int32_t v, uint32_t p) const {
    if (!hint) return atomicMin(t + v, p);
    uint32_t r;
    asm volatile("atom.global.min.L2::cache_hint.u32 %0, [%1], %2, %3;"
                 : "=r"(r) : "l"(t + v), "r"(p), "l"(pol) : "memory");
    return r;
  }
};
__device__ __forceinline__ int32_t ld_stream_i32(const int32_t *a, bool hint, uint64_t pol) {
  if (!hint) return __ldg(a);
  int32_t r;
  asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(r) : "l"(a), "l"(pol));
  return r;
}
// Final list entries are written once and not read again by the chain: streaming stores
// (evict-first) keep them from pushing the chain's intermediates out of L2.  -DGS_OUT_STREAM=0 for A/B.
__device__ __forceinline__ void st_out(int32_t *p, int32_t v) {
#if GS_OUT_STREAM
  __stcs(p, v);
#else
  *p = v;
#endif
}
__device__ __forceinline__ uint32_t ld_acquire(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(uint32_t *p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// owner = id % P (pyfrontend.cpp:57) without a hardware-unsupported modulo on the
// hot path: a mask when P is a power of two, else multiply-high by ceil(2^32/P)
// (quotient exact or one too large for any 31-bit id) and one correction.
__device__ __forceinline__ int owner_of(const GsWs &ws, int32_t v) {
  if (ws.owner) return (int)__ldg(ws.owner + v);
  if (ws.pmask >= 0) return (int)(v & ws.pmask);
  int32_t r = v - (int32_t)(__umulhi((uint32_t)v, ws.pmagic) * (uint32_t)ws.P);
  return r < 0 ? r + ws.P : r;
}

// Dedup-table entries are (tag | kind | payload), S = ws.pos_bits payload bits.
// A sample uses ONE tag; tags count DOWN from sample to sample, so whatever an
// earlier sample left behind is larger than anything this sample writes: it
// loses every atom.min and reads as "never seen".  Nothing is reset between
// layers or samples (DuplicateRemover::clear, duplicate.cpp:28-33, has no
// counterpart); k_tail refills the table when a slot runs out of tags.  Within
// a sample kind 0 (frontier row) < kind 1 (edge position): an atom.min of an
// edge position never displaces the row mark, and returns it.
struct Tag {
  uint32_t tg;    // this sample's tag, in place
  uint32_t kind;  // 1 << S
  uint32_t pmask; // payload mask
  int tsh;        // S + 1
  __device__ __forceinline__ Tag(const GsWs &ws, const GsState &st) {
    const int S = ws.pos_bits;
    tsh = S + 1;
    kind = 1u << S;
    pmask = kind - 1u;
    tg = ((1u << (31 - S)) - 1u - st.tag_idx) << tsh;
  }
  __device__ __forceinline__ bool mine(uint32_t e) const { return ((e ^ tg) >> tsh) == 0u; }
};

// bit q of a mask -> byte q of a packed counter word
// (multiply replicates the mask byte into all eight bytes, the AND keeps bit q in
// byte q, + 0x7F carries any kept bit into bit 7 of its byte)
__device__ __forceinline__ unsigned long long pk_spread(uint32_t mask) {
  const unsigned long long v = ((unsigned long long)(mask & 0xFFu) * 0x0101010101010101ull) & 0x8040201008040201ull;
  return ((v + 0x7F7F7F7F7F7F7F7Full) >> 7) & 0x0101010101010101ull;
}
// warp sum of a packed word whose bytes are <= 7 per lane: two redux instructions
__device__ __forceinline__ unsigned long long pk_redux(unsigned long long v) {
  const uint32_t lo = __reduce_add_sync(FULL, (uint32_t)v), hi = __reduce_add_sync(FULL, (uint32_t)(v >> 32));
  return ((unsigned long long)hi << 32) | lo;
}

__device__ __forceinline__ void stage_ws(GsWs &dst, const GsWs *__restrict__ src) {
  const uint32_t *s = reinterpret_cast<const uint32_t *>(src);
  uint32_t *d = reinterpret_cast<uint32_t *>(&dst);
  for (int i = threadIdx.x; i < (int)(sizeof(GsWs) / 4); i += blockDim.x) d[i] = __ldg(s + i);
  __syncthreads();
}

// ------------------------------------------------------------ flat look-back
// Single-pass prefix over UNITS (a tile of rows, a run of candidate tiles) handed
// out by a ticket.  A unit publishes its aggregate and then sums the aggregates of
// ALL its predecessors itself: thread j of the CTA reads unit base + j's counters
// (independent loads: one L2 round trip), a CTA-wide reduction per counter gives
// the exclusive prefix.  Nobody waits for anybody's PREFIX, only for aggregates,
// and every predecessor was drawn by a CTA that is already running -- so the wait
// is one status round trip whatever else shares the machine, and there is no chain
// through the units in flight (the classic decoupled look-back walks 32 units per
// round trip: with ~450 CTAs in flight that walk dominated the kernel).  Beyond
// GS_TPB units the sum restarts from the inclusive prefix that every GS_TPB-th
// unit publishes.  Status words carry the sample's scan epoch, so descriptors are
// never reset (k_tail zeroes them once per 2^30 uses).  Called by the whole CTA;
// s_mine[] must be visible (barrier) on entry, s_pre[] is on exit.
// Two halves, so that a caller can do the work that does not need the prefix between them
// (it then overlaps the wait for the predecessors).
template <typename T, int NC, int NCPAD>
__device__ __forceinline__ void lookback_publish(uint32_t *status, T *agg, int r, uint32_t ep, const T *s_mine,
                                                 T *s_pre) {
  const int tid = threadIdx.x;
  if (tid < NC) agg[(size_t)r * NCPAD + tid] = s_mine[tid];
  if (tid < NCPAD) s_pre[tid] = (T)0;
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    st_release(status + r, (ep << 2) | 1u);
  }
}
template <typename T, int NC, int NCPAD>
__device__ __forceinline__ void lookback_collect(uint32_t *status, T *agg, T *inc, int r, uint32_t ep,
                                                 const T *s_mine, T *s_pre) {
  const int tid = threadIdx.x;
  const int base = r & ~(GS_TPB - 1);
  {  // every predecessor of this chunk has published its aggregate (thread j watches unit base + j)
    const int j = base + tid;
    if (j < r)
      for (;;) {
        const uint32_t st = ld_acquire(status + j);
        if ((st >> 2) == ep && (st & 3u) != 0u) break;
        __nanosleep(64);
      }
    if (base > 0 && tid == GS_TPB - 1)  // ... and the unit before the chunk its inclusive prefix
      for (;;) {
        const uint32_t st = ld_acquire(status + base - 1);
        if ((st >> 2) == ep && (st & 3u) == 2u) break;
        __nanosleep(64);
      }
  }
  __syncthreads();
  // thread (ju, c) sums counter c over the units base + ju, base + ju + UPP, ...: independent
  // loads, one register of state; shared-memory atomics finish the sum
  constexpr int UPP = GS_TPB / NCPAD;
  const int c = tid % NCPAD, ju = tid / NCPAD;
  if (ju < UPP && c < NC) {
    T sum = (T)0;
    int j = base + ju;
    for (; j + 3 * UPP < r; j += 4 * UPP) {
      const T a0 = __ldcg(agg + (size_t)j * NCPAD + c), a1 = __ldcg(agg + (size_t)(j + UPP) * NCPAD + c),
              a2 = __ldcg(agg + (size_t)(j + 2 * UPP) * NCPAD + c), a3 = __ldcg(agg + (size_t)(j + 3 * UPP) * NCPAD + c);
      sum += a0 + a1 + a2 + a3;
    }
    for (; j < r; j += UPP) sum += __ldcg(agg + (size_t)j * NCPAD + c);
    if (base > 0 && ju == 0) sum += __ldcg(inc + (size_t)(base - 1) * NCPAD + c);
    if (sum != (T)0) atomicAdd(&s_pre[c], sum);
  }
  __syncthreads();
  if ((r & (GS_TPB - 1)) == GS_TPB - 1) {  // last unit of a chunk: publish the inclusive prefix
    if (tid < NC) inc[(size_t)r * NCPAD + tid] = s_pre[tid] + s_mine[tid];
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      st_release(status + r, (ep << 2) | 2u);
    }
  }
}

// ------------------------------------------------------------------- k_rows
// Sample header at the start of a sample (the CTA that draws tile 0 of layer 0
// runs this BEFORE it publishes its tile; every other CTA reports errors only
// after its look-back, which orders it behind that publication).
__device__ void init_header(const GsWs &ws, int n, int status) {
  gs_sample_header *h = ws.hdr;
  {
    GsTargets *tg = ws.tgt;
    const bool direct = ws.input->direct != 0;
    tg->direct = direct ? 1 : 0;
    tg->cap = direct ? ws.input->block_words : ws.slab_cap;
    for (int q = 0; q < GS_MAX_PARTS; q++) {
      tg->base[q] = direct && q < ws.P ? ws.input->part_dst[q] : ws.slab;
      tg->cursor[q] = GS_PART_HEADER_WORDS;
    }
  }
  h->status = status;
  h->num_layers = ws.L;
  h->num_parts = ws.P;
  h->n_seeds = n;
  h->batch_index = ws.input->batch_index;
  h->draw_cursor = *ws.draw_cursor;
  h->edges = 0;
  h->late_status = 0;
  h->feature_rows = 0;
  h->reserved[0] = ws.input->want_host;
  h->slab_used = 0;
  h->reserved[1] = (n + 3) & ~3;  // frontier slab in use (int32 elements)
  h->reserved[2] = 1;             // frontier spans index the frontier slab, not the sample slab
  for (int l = 0; l <= GS_MAX_LAYERS; l++) {
    h->frontier_len[l] = l == 0 ? n : 0;
    h->frontier[l].off = 0;
    h->frontier[l].len = l == 0 ? n : 0;
    if (l < GS_MAX_LAYERS) h->cand_len[l] = 0;
  }
}

// Per frontier row: CSR extent and number of candidates c = 1 + (deg < fan or
// fan < 0 ? deg : fan) (slicer.cpp:6-22), exclusive scan -> first candidate
// position; rank of the row among the rows of its owner (= its slot in
// self_ids_in / self_ids_out, slicer.cpp:33-34); the row's mark in the dedup
// table.  FIRST: the rows are the seeds, read as int64 from pinned host memory,
// checked and narrowed; claiming their table entries with atom.min doubles as
// the duplicate-seed check.
template <bool FIRST>
__device__ void rows_role(const GsWs &ws, int layer) {
  gs_sample_header *h = ws.hdr;
  __shared__ unsigned long long s_wc[GS_IPT * GS_NW][2 + GS_MAX_PARTS];   // totals of block (k, warp)
  __shared__ unsigned long long s_base[GS_IPT * GS_NW][2 + GS_MAX_PARTS]; // -> sums of the blocks before it
  __shared__ unsigned long long s_agg[GS_RS_NC], s_pre[GS_RS_NC];
  __shared__ int s_tile;
  if (!FIRST && h->status != 0) return;
  if (threadIdx.x < GS_RS_NC) s_agg[threadIdx.x] = 0ull;  // counters beyond 2 + P stay zero
  const int P = ws.P, NC = 2 + P;
  const int lane = lane_id(), w = warp_id();
  int F;
  const int32_t *front = nullptr;
  int first_status = 0;
  if (FIRST) {
    F = ws.input->n_seeds;
    if (F < 0) {
      first_status = GS_SKIPPED;
      F = 0;
    } else if (F > ws.Fcap || (int64_t)F > ws.fslab_cap) {
      first_status = (int)GS_ERR_SEED;
      F = 0;
    }
  } else {
    F = h->frontier_len[layer];
    front = ws.fslab + h->frontier[layer].off;
  }
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  uint32_t *cnt = ws.cnt + layer * GS_CNT_STRIDE;
  uint32_t *tick = ws.tickets + layer * GS_TICKETS;
  if (ntiles == 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      if (FIRST) init_header(ws, 0, first_status);
      h->cand_len[layer] = 0;
      ws.off[0] = 0u;
      h->draw_base[layer] = *ws.draw_cursor;
    }
    return;
  }
  const int fan = ws.fanouts[layer];
  const GsState stt = *ws.state;
  const uint32_t ep = stt.scan_epoch + 3u * (uint32_t)layer;
  // per-row arrays ping-pong by layer: rowside of layer l reads them in the launch in which
  // k_rows of layer l + 1 writes the next layer's
  uint32_t *rw_ea = ws.ea + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3), *rw_rowhas = ws.rowhas + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3),
           *rw_selfrank = ws.selfrank + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3);
  const Tag tag(ws, stt);
  const Tbl tbl(ws);
  const bool replay = ws.rng_mode == (int)GS_RNG_DRAWS;
  unsigned long long *rs_agg = ws.rs_val, *rs_inc = ws.rs_val + (size_t)ws.ftiles_cap * GS_RS_NC;
  bool bad = false;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(tick + 0, 1u);
  __syncthreads();
  int t = s_tile;
  while (t < ntiles) {
    if (FIRST && t == 0 && threadIdx.x == 0) init_header(ws, F, 0);
    // Rows are dealt lane-consecutive (row of k = tile + k * 256 + thread): every per-row load and
    // store of a warp then covers 32 consecutive rows = whole sectors.  The scan order is k-major:
    // block (k, warp) of 32 rows, blocks in the order k * GS_NW + warp.
    const int ib = t * GS_TILE + (int)threadIdx.x;
    int32_t nd[GS_IPT];
    int32_t deg[GS_IPT];
    int g0[GS_IPT];
    uint32_t c[GS_IPT];
    if (FIRST) {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        nd[k] = 0;
        if (ib + k * GS_TPB < F) {
          const int64_t v = ws.seeds_in[ib + k * GS_TPB];
          if (v < 0 || v >= (int64_t)ws.num_nodes) bad = true;
          else nd[k] = (int32_t)v;
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) nd[k] = ib + k * GS_TPB < F ? front[ib + k * GS_TPB] : 0;
    }
    {
      int64_t rs[GS_IPT], e_[GS_IPT];
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        rs[k] = 0; e_[k] = 0;
        if (ib + k * GS_TPB < F) {
          rs[k] = __ldg(ws.indptr + nd[k]);
          e_[k] = __ldg(ws.indptr + nd[k] + 1);
        }
      }
      // everything that does not need the prefix goes out right away
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        const int i = ib + k * GS_TPB;
        deg[k] = (int32_t)(e_[k] - rs[k]);
        g0[k] = 0;
        if (i < F) {
          g0[k] = owner_of(ws, nd[k]);
          ws.rowinfo[i] = make_int4((int)(uint32_t)rs[k],
                                    (int)((uint32_t)((uint64_t)rs[k] >> 32) | ((uint32_t)g0[k] << 29)), deg[k], nd[k]);
          rw_ea[i] = GS_INF;
          rw_rowhas[i] = 0u;
        }
      }
    }
    unsigned long long ex_c[GS_IPT];  // exclusive within the (k, warp) block
    uint32_t ex_s[GS_IPT], ex_g[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const bool in = ib + k * GS_TPB < F;
      const bool samp = in && fan >= 0 && deg[k] >= fan;
      c[k] = in ? 1u + (uint32_t)(samp ? fan : deg[k]) : 0u;
      const unsigned long long pk1 = in ? pk_one(g0[k]) : 0ull;
      const unsigned long long inc_c = warp_incl_scan64((unsigned long long)c[k]), inc_pk = pk_warp_incl(pk1);
      const uint32_t inc_s = warp_incl_scan(samp ? 1u : 0u);
      ex_c[k] = inc_c - c[k];
      ex_s[k] = inc_s - (samp ? 1u : 0u);
      ex_g[k] = pk_get(inc_pk - pk1, g0[k]);
      const unsigned long long tot_pk = __shfl_sync(FULL, inc_pk, 31);
      if (lane == 31) {
        s_wc[k * GS_NW + w][0] = inc_c;
        s_wc[k * GS_NW + w][1] = inc_s;
      }
      if (lane < P) s_wc[k * GS_NW + w][2 + lane] = pk_get(tot_pk, lane);
    }
    __syncthreads();
    if (threadIdx.x == 0) s_tile = (int)atomicAdd(tick + 0, 1u);  // next tile of this CTA (read after the barrier below)
    if (threadIdx.x < NC) {
      unsigned long long run = 0ull;
#pragma unroll 8
      for (int j = 0; j < GS_IPT * GS_NW; j++) {
        s_base[j][threadIdx.x] = run;
        run += s_wc[j][threadIdx.x];
      }
      s_agg[threadIdx.x] = run;
    }
    __syncthreads();
    lookback_publish<unsigned long long, 2 + GS_MAX_PARTS, GS_RS_NC>(ws.rs_status, rs_agg, t, ep, s_agg, s_pre);
    // the table marks go out while the predecessors' aggregates arrive
    bool dup = false;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int i = ib + k * GS_TPB;
      if (i < F) {
        if (FIRST) {
          ws.fslab[i] = nd[k];
          // a second claim of the same entry under this sample's tag is a duplicate seed
          const uint32_t old = atomicMin(ws.table + nd[k], tag.tg | (uint32_t)i);
          if (tag.mine(old)) dup = true;
        } else {
          tbl.st(nd[k], tag.tg | (uint32_t)i);
        }
      }
    }
    lookback_collect<unsigned long long, 2 + GS_MAX_PARTS, GS_RS_NC>(ws.rs_status, rs_agg, rs_inc, t, ep, s_agg, s_pre);
    if (FIRST && (bad || dup)) atomicCAS(&h->status, 0, (int)GS_ERR_SEED);  // after the look-back: behind init_header
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int i = ib + k * GS_TPB;
      if (i < F) {
        const unsigned long long *bs = s_base[k * GS_NW + w];
        const unsigned long long base_c = s_pre[0] + bs[0] + ex_c[k];
        ws.off[i] = (uint32_t)base_c;
        rw_selfrank[i] = (uint32_t)(s_pre[2 + g0[k]] + bs[2 + g0[k]]) + ex_g[k];
        if (replay) ws.samp_rank[i] = (uint32_t)(s_pre[1] + bs[1]) + ex_s[k];
        // candidate tiles that start inside this row (only the candidate-parallel k_expand of
        // fan = -1 layers looks rows up by position)
        const unsigned long long end = base_c + c[k];
        if (fan < 0)
          for (unsigned long long tb = (base_c + GS_TILE - 1) / GS_TILE;
               tb * GS_TILE < end && tb < (unsigned long long)ws.ctiles_cap; tb++)
            ws.tile_first[tb] = (uint32_t)i;
      }
    }
    if (t == ntiles - 1 && threadIdx.x == 0) {
      const unsigned long long ctot = s_pre[0] + s_agg[0];
      const unsigned long long stot = s_pre[1] + s_agg[1];
      if (ctot > (unsigned long long)ws.Ccap) {
        atomicCAS(&h->status, 0, (int)GS_ERR_OVERFLOW);
        h->cand_len[layer] = 0;
      } else {
        h->cand_len[layer] = (int32_t)ctot;
        ws.off[F] = (uint32_t)ctot;
        const uint64_t base = *ws.draw_cursor;
        h->draw_base[layer] = base;
        if (fan > 0) *ws.draw_cursor = base + (uint64_t)fan * (uint64_t)stot;
      }
      for (int q = 0; q < P; q++) cnt[GS_CNT_SELF(P, q)] = (uint32_t)(s_pre[2 + q] + s_agg[2 + q]);
    }
    __syncthreads();
    t = s_tile;
  }
}

// r % d for many r and one d (the picks of a row, slicer.cpp:16): the reciprocal is
// computed once per row, every remainder is then a multiply-high and at most three
// corrections.  inv <= floor(2^32 / d) by construction, so the quotient estimate
// never exceeds the true quotient and the loop makes the result exact.
struct Mod32 {
  uint32_t d, inv;
  __device__ __forceinline__ void set(uint32_t d_) {
    d = d_;
    const float f = __uint2float_ru(d_);
    uint32_t i0 = __float2uint_rz(__uint_as_float(__float_as_uint(__frcp_rz(f)) + 0x0ffffffeu));
    i0 += __umulhi(i0, (0u - d_) * i0);
    while ((unsigned long long)i0 * d_ > 0x100000000ull) i0--;  // never taken in practice; keeps the bound provable
    inv = d_ == 1u ? 0xFFFFFFFFu : i0;
  }
  __device__ __forceinline__ uint32_t rem(uint32_t r) const {
    uint32_t x = r - __umulhi(r, inv) * d;
    while (x >= d) x -= d;
    return x;
  }
};

// Packed per-part counters: one byte per part in a 32-bit word (P <= 4) or a 64-bit word.
template <typename PK> __device__ __forceinline__ PK pkt_one(int g) { return (PK)1 << (8 * g); }
template <typename PK> __device__ __forceinline__ uint32_t pkt_get(PK v, int g) { return (uint32_t)(v >> (8 * g)) & 0xFFu; }
__device__ __forceinline__ uint32_t pkt_redux(uint32_t v) { return __reduce_add_sync(FULL, v); }
__device__ __forceinline__ unsigned long long pkt_redux(unsigned long long v) { return pk_redux(v); }
__device__ __forceinline__ uint32_t pkt_incl(uint32_t v) { return warp_incl_scan(v); }
__device__ __forceinline__ unsigned long long pkt_incl(unsigned long long v) { return pk_warp_incl(v); }
__device__ __forceinline__ uint32_t pkt_spread(uint32_t mask, uint32_t) {  // bit q of mask -> byte q (q < 4)
  return ((mask & 0xFu) * 0x00204081u) & 0x01010101u;
}
__device__ __forceinline__ unsigned long long pkt_spread(uint32_t mask, unsigned long long) { return pk_spread(mask); }

// ----------------------------------------------------------------- k_expand
// Lays out every list of the layer.  Runs in the last CTA of k_expand: by then
// the per-layer totals are exact (see expand_role), so k_cand writes each list
// once, at its final place; a part's nine lists are contiguous.
__device__ void layer_alloc(const GsWs &ws, int layer, const uint32_t *cnt /* shared memory copy */) {
  gs_sample_header *h = ws.hdr;
  const int P = ws.P;
  const int F = h->frontier_len[layer];
  bool ok = true;
  const unsigned long long fnext = (unsigned long long)F + cnt[GS_CNT_NEW];
  if (layer + 1 < ws.L && fnext > (unsigned long long)ws.Fcap) ok = false;
  int64_t used = h->slab_used, fused = h->reserved[1];
  const gs_span fs = slab_alloc(fused, ws.fslab_cap, (uint32_t)fnext, ok);
  int64_t edges = 0;
  // direct hand-over: part q's lists go behind the header of block q (offsets relative to the
  // block); else every list shares the slot's slab (offsets into it)
  GsTargets *tg = ws.tgt;
  const bool direct = tg->direct != 0;
  const int64_t cap = tg->cap;
  for (int q = 0; q < P && ok; q++) {
    const uint32_t n_in = cnt[GS_CNT_IN(P, q)], n_e = cnt[GS_CNT_E(P, q)], n_out = cnt[GS_CNT_OUT(P, q)],
                   n_own = cnt[GS_CNT_OWN(P, q)], n_to = cnt[GS_CNT_TO(P, q)], n_self = cnt[GS_CNT_SELF(P, q)];
    gs_span *fl = h->field[layer][q];
    int64_t cur = direct ? tg->cursor[q] : used;
    fl[GS_F_IN_NODES] = slab_alloc(cur, cap, n_in, ok);
    fl[GS_F_OUT_NODES] = slab_alloc(cur, cap, n_out, ok);
    fl[GS_F_OWNED_OUT] = slab_alloc(cur, cap, n_own, ok);
    fl[GS_F_FROM_IDS] = slab_alloc(cur, cap, n_out - n_own, ok);
    fl[GS_F_TO_IDS] = slab_alloc(cur, cap, n_to, ok);
    fl[GS_F_SELF_IN] = slab_alloc(cur, cap, n_self, ok);
    fl[GS_F_SELF_OUT] = slab_alloc(cur, cap, n_self, ok);
    const gs_span ip = slab_alloc(cur, cap, n_out + 1u, ok);
    fl[GS_F_CSR_INDPTR] = ip;
    fl[GS_F_CSR_INDICES] = slab_alloc(cur, cap, n_e, ok);
    if (ok) tg->base[q][ip.off + n_out] = (int32_t)n_e;  // row pointer sentinel
    if (direct) {
      tg->cursor[q] = cur;
      used += cur;  // reported only
    } else {
      used = cur;
    }
    edges += n_e;
  }
  if (ok) {
    h->slab_used = direct ? 0 : used;
    h->reserved[1] = (int32_t)fused;
    h->frontier[layer + 1] = fs;
    h->frontier_len[layer + 1] = (int32_t)fnext;
    h->edges += edges;
  } else {
    atomicCAS(&h->status, 0, (int)GS_ERR_OVERFLOW);
  }
}

// A role's contiguous run of tiles: CTA bx of the nb CTAs that play the role in this launch.
__device__ __forceinline__ void role_tiles(int ntiles, int bx, int nb, int &t0, int &t1) {
  const int per = (ntiles + nb - 1) / nb;
  t0 = min(ntiles, bx * per);
  t1 = min(ntiles, t0 + per);
}
// last_block for a role played by n CTAs of the launch (own ticket)
__device__ __forceinline__ bool last_block_n(uint32_t *ticket, uint32_t n) {
  __shared__ uint32_t s_last_n;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t t = atomicAdd(ticket, 1u);
    s_last_n = (t == n - 1u) ? 1u : 0u;
    if (s_last_n) *ticket = 0u;
  }
  __syncthreads();
  if (s_last_n) __threadfence();
  return s_last_n != 0u;
}

// -DGS_DIAG=bits: knock-out builds for tools/build_variants.sh (WRONG results; they only show what a
// piece of work costs with all slots in flight).  1: row side writes nothing, 2: no relabel-map read.
// picks per work item of k_expand_rows (4 or 8): more picks in flight per dependent round trip
// (8: +2.4 % on the products workload, A/B on one B200 with tools/gpu_r2_call16.sh)
#ifndef GS_PB
#define GS_PB 8
#endif
#ifndef GS_OUT_STREAM
#define GS_OUT_STREAM 1
#endif
#ifndef GS_DIAG
#define GS_DIAG 0
#endif
// The candidate words ping-pong by layer: relabel / row side of layer l read them in the
// launch in which k_expand of layer l + 1 writes the next layer's stream.
__device__ __forceinline__ int32_t *cand_buf(const GsWs &ws, int layer) {
  return ws.cand + (size_t)(layer & 1) * (((size_t)ws.Ccap + 7) & ~(size_t)3);
}

// Candidate-parallel CSR row expansion (slicer.cpp:6-22).  A tile is GS_TILE
// consecutive candidate positions, so load is balanced whatever the degree
// skew; the rows a tile touches are found from tile_first[] and a binary
// search over their offsets staged in shared memory.  Slots >= 1 of a row are
// either all its neighbours in CSR order (coalesced reads) or `fan` picks with
// replacement.  Every real edge source v does ONE returning atom.min(position)
// on its table entry -- the parallel form of DuplicateRemover's first-occurrence
// rule (util/duplicate.cpp:14-26).  The value it returns is
//   * stale (another sample's tag): v is new to the frontier and this is the
//     first arrival for v        -> one more next-frontier node, one more in_node;
//   * a frontier-row mark (kind 0): v is row idx of the current frontier; its
//     minimum edge position is kept per ROW (ea[idx], an F-sized array) and the
//     candidate carries idx instead of v, so k_cand never needs v's self position
//     from an N-sized table;
//   * an edge position of this layer: nothing to count.
// One atomicOr per (thread, row) on rowhas[] finds the first edge of a row into
// each part: out_nodes / owned / to list lengths.  With the per-part edge counts
// these are ALL the list lengths of the layer, exact, before k_cand starts.
// RNG (gs_rng_mode) is a template parameter: the pick path of the other two
// modes costs instructions in a kernel bound by issue.
template <int RNG, typename PK>
__device__ void expand_role(const GsWs &ws, int layer, int nb_role) {
  gs_sample_header *h = ws.hdr;
  int32_t *const cand = cand_buf(ws, layer);
  if (h->status != 0) return;
  const Tbl tbl(ws);
  const uint64_t spol = ws.l2_hints ? l2_policy_stream() : 0ull;
  __shared__ uint32_t s_off[GS_TILE + 2];
  __shared__ uint32_t s_cnt[GS_CNT_STRIDE];
  const int P = ws.P;
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int fan = ws.fanouts[layer];
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const uint32_t batch = (uint32_t)h->batch_index;
  constexpr bool replay = RNG == (int)GS_RNG_DRAWS;
  constexpr bool norepl = RNG == (int)GS_RNG_PHILOX_NOREPLACE;
  const uint64_t dbase = h->draw_base[layer];
  const GsState stt = *ws.state;
  const Tag tag(ws, stt);
  const uint32_t tge = tag.tg | tag.kind;
  uint32_t *cnt = ws.cnt + layer * GS_CNT_STRIDE;
  const int lane = lane_id();
  // per-row arrays ping-pong by layer: rowside of layer l reads them in the launch in which
  // k_rows of layer l + 1 writes the next layer's
  uint32_t *rw_ea = ws.ea + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3), *rw_rowhas = ws.rowhas + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3);
  // fields of the shared-memory copy of ws used per candidate, once into registers
  const int32_t *__restrict__ indices = ws.indices;
  const uint8_t *__restrict__ owner_tbl = ws.owner;
  const int pmask = ws.pmask;
  const uint32_t pmagic = ws.pmagic, seed_lo = ws.seed_lo, seed_hi = ws.seed_hi;
  const uint64_t n_draws = ws.n_draws;
  const bool hints = ws.l2_hints != 0;
  auto own_of = [&](int32_t v) -> int {  // owner_of(ws, v) on the hoisted fields
    if (owner_tbl) return (int)__ldg(owner_tbl + v);
    if (pmask >= 0) return (int)(v & pmask);
    int32_t r = v - (int32_t)(__umulhi((uint32_t)v, pmagic) * (uint32_t)P);
    return r < 0 ? r + P : r;
  };
  if (threadIdx.x < GS_CNT_STRIDE) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  for (int t = blockIdx.x; t < ntiles; t += nb_role) {  // the expand CTAs come first in the launch
    const int i_lo = (int)ws.tile_first[t];
    const int i_hi = (t + 1 < ntiles) ? (int)ws.tile_first[t + 1] : F - 1;
    const int cntr = i_hi - i_lo + 1;  // <= GS_TILE + 1 rows, each has >= 1 candidate
    for (int k = threadIdx.x; k <= cntr; k += GS_TPB) s_off[k] = ws.off[i_lo + k];
    __syncthreads();
    // A thread owns GS_IPT CONSECUTIVE positions: one binary search, then a walk;
    // one Philox block serves up to four picks of a row; the stream goes out as
    // 16-byte / 4-byte vectors.
    const uint32_t p0 = (uint32_t)t * GS_TILE + GS_IPT * threadIdx.x;
    int row[GS_IPT], slot[GS_IPT];
    int32_t nd1[GS_IPT], deg[GS_IPT];
    int g0[GS_IPT];
    int64_t rs[GS_IPT];
    int lo = 0;
    if (p0 < C) {  // largest k < cntr with s_off[k] <= p0 (branch-free: s_off is ascending, s_off[0] <= p0)
#pragma unroll
      for (int step = GS_TILE; step > 0; step >>= 1)
        if (lo + step < cntr && s_off[lo + step] <= p0) lo += step;
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const uint32_t p = p0 + k;
      if (k > 0 && p < C && p >= s_off[lo + 1]) lo++;  // every row has >= 1 candidate: one step is enough
      row[k] = i_lo + lo;
      slot[k] = p < C ? (int)(p - s_off[lo]) : -1;  // 0 = self entry, -1 = past the end
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      if (k > 0 && row[k] == row[k - 1]) {
        nd1[k] = nd1[k - 1];
        rs[k] = rs[k - 1];
        deg[k] = deg[k - 1];
        g0[k] = g0[k - 1];
      } else {
        const int4 ri = ws.rowinfo[row[k]];
        rs[k] = (int64_t)(((uint64_t)((uint32_t)ri.y & 0x1FFFFFFFu) << 32) | (uint32_t)ri.x);
        g0[k] = (int)((uint32_t)ri.y >> 29);
        deg[k] = ri.z;
        nd1[k] = ri.w;
      }
    }
    const int32_t *src[GS_IPT];
    uint4 blk = make_uint4(0u, 0u, 0u, 0u);
    int blk_row = -1, blk_id = -1, mod_row = -1;
    Mod32 md;
    md.d = 1u; md.inv = 0u;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      src[k] = nullptr;
      if (slot[k] > 0) {
        int32_t pick = slot[k] - 1;
        if (fan >= 0 && deg[k] >= fan) {
          uint32_t r;
          if (replay) {
            const uint64_t di = dbase + (uint64_t)fan * ws.samp_rank[row[k]] + (uint32_t)pick;
            r = di < n_draws ? ws.draws[di] : 0u;
          } else {
            const int want = norepl ? GS_NR_KEY_BLOCK : (pick >> 2);
            if (blk_row != row[k] || blk_id != want) {
              blk_row = row[k];
              blk_id = want;
              blk = philox_block(seed_lo, seed_hi, batch, layer, nd1[k], blk_id);
            }
            r = norepl ? nr_pick(blk, (uint32_t)pick, (uint32_t)deg[k]) : philox_word(blk, pick & 3);
          }
          if (norepl) {
            pick = (int32_t)r;  // already a member of [0, deg)
          } else {
            if (mod_row != row[k]) {
              mod_row = row[k];
              md.set((uint32_t)deg[k]);
            }
            pick = (int32_t)md.rem(r);
          }
        }
        src[k] = indices + rs[k] + pick;
      }
    }
    int32_t v[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++)
      if (src[k]) v[k] = ld_stream_i32(src[k], hints, spol);
    // one returning atom.min per real edge source, all four in flight
    int g[GS_IPT];
    uint32_t old[GS_IPT];
    bool edge[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      edge[k] = src[k] != nullptr && v[k] != nd1[k];
      g[k] = 0;
      old[k] = 0u;
      if (edge[k]) {
        g[k] = own_of(v[k]);
        old[k] = tbl.atom_min(v[k], tge | (p0 + k));
      }
    }
    // first edge of a row into each part: one returning atomicOr per run of equal rows,
    // issued before the atom.min results are looked at (independent of them)
    uint32_t rbits[GS_IPT], was[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) rbits[k] = edge[k] ? 1u << g[k] : 0u;
#pragma unroll
    for (int k = 0; k + 1 < GS_IPT; k++)
      if (row[k + 1] == row[k]) {
        rbits[k + 1] |= rbits[k];
        rbits[k] = 0u;
      }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      was[k] = 0u;
      if (rbits[k]) was[k] = atomicOr(rw_rowhas + row[k], rbits[k]);
    }
    int32_t cvv[GS_IPT];
    uint32_t cf[GS_IPT];
    PK pk_in = 0, pk_e = 0, pk_out = 0, pk_own = 0, pk_to = 0;
    uint32_t n_new = 0u;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      if (slot[k] == 0) {
        cvv[k] = row[k];  // self entry: the row index; part = owner of the row's node
        cf[k] = GS_CK_SELF | ((uint32_t)g0[k] << GS_CF_PART_SHIFT);
      } else if (!edge[k]) {
        cvv[k] = 0;
        cf[k] = GS_CK_DEAD;
      } else {
        cvv[k] = v[k];
        cf[k] = GS_CK_EDGE | ((uint32_t)g[k] << GS_CF_PART_SHIFT);
        pk_e += pkt_one<PK>(g[k]);
        if (!tag.mine(old[k])) {  // first arrival for a node that is new to the frontier
          n_new++;
          pk_in += pkt_one<PK>(g[k]);
        } else if (!(old[k] & tag.kind)) {  // a row of the current frontier
          const uint32_t idx = old[k] & tag.pmask;
          if (atomicMin(rw_ea + idx, p0 + k) == GS_INF) pk_in += pkt_one<PK>(g[k]);
          cvv[k] = (int32_t)idx;
          cf[k] = GS_CK_FHIT | ((uint32_t)g[k] << GS_CF_PART_SHIFT);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++)
      if (rbits[k]) {
        const uint32_t nb = rbits[k] & ~was[k];
        if (nb) {
          const uint32_t own = 1u << g0[k];
          pk_out += pkt_spread(nb, (PK)0);
          if (nb & own) pk_own += pkt_one<PK>(g0[k]);
          if ((nb & ~own) && !(was[k] & ~own)) pk_to += pkt_one<PK>(g0[k]);
        }
      }
    if (p0 + GS_IPT <= C) {
      *reinterpret_cast<int4 *>(cand + p0) = make_int4(cvv[0], cvv[1], cvv[2], cvv[3]);
      *reinterpret_cast<uchar4 *>(ws.cflag + p0) =
          make_uchar4((unsigned char)cf[0], (unsigned char)cf[1], (unsigned char)cf[2], (unsigned char)cf[3]);
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (p0 + k < C) {
          cand[p0 + k] = cvv[k];
          ws.cflag[p0 + k] = (uint8_t)cf[k];
        }
    }
    // per-tile totals: a lane adds at most GS_IPT to a byte, a warp at most 128
    pk_in = pkt_redux(pk_in);
    pk_e = pkt_redux(pk_e);
    pk_out = pkt_redux(pk_out);
    pk_own = pkt_redux(pk_own);
    pk_to = pkt_redux(pk_to);
    n_new = __reduce_add_sync(FULL, n_new);
    if (lane < P) {
      uint32_t x;
      if ((x = pkt_get<PK>(pk_in, lane))) atomicAdd(&s_cnt[GS_CNT_IN(P, lane)], x);
      if ((x = pkt_get<PK>(pk_e, lane))) atomicAdd(&s_cnt[GS_CNT_E(P, lane)], x);
      if ((x = pkt_get<PK>(pk_out, lane))) atomicAdd(&s_cnt[GS_CNT_OUT(P, lane)], x);
      if ((x = pkt_get<PK>(pk_own, lane))) atomicAdd(&s_cnt[GS_CNT_OWN(P, lane)], x);
      if ((x = pkt_get<PK>(pk_to, lane))) atomicAdd(&s_cnt[GS_CNT_TO(P, lane)], x);
    }
    if (lane == 31 && n_new) atomicAdd(&s_cnt[GS_CNT_NEW], n_new);
    __syncthreads();
  }
  __syncthreads();
  if (threadIdx.x < 1 + 5 * P && s_cnt[threadIdx.x]) atomicAdd(cnt + threadIdx.x, s_cnt[threadIdx.x]);
  if (!last_block_n(ws.tickets + layer * GS_TICKETS + 1, (uint32_t)nb_role)) return;
  if (threadIdx.x < GS_CNT_STRIDE) s_cnt[threadIdx.x] = __ldcg(cnt + threadIdx.x);  // one round trip for all totals
  __syncthreads();
  if (threadIdx.x == 0) layer_alloc(ws, layer, s_cnt);
}

// Row-parallel expansion for sampled layers (fan > 0): a row has at most fan + 1
// candidates, so the work item is (row, block of four picks) -- exactly the unit
// one Philox block serves.  One thread per item: row record and offsets are read
// once, the Philox block and the reciprocal of the degree are computed once,
// the positions are off[row] + 1 + pick index -- no search, no staging, no
// redundant generator work (the candidate-parallel form below recomputes all of
// that per thread and is kept for fan = -1, where rows are unbounded).  Dedup,
// counting and the candidate stream are exactly those of expand_role.
template <int RNG, typename PK>
__device__ void expand_rows_role(const GsWs &ws, int layer, int nb_role) {
  gs_sample_header *h = ws.hdr;
  int32_t *const cand = cand_buf(ws, layer);
  __shared__ int32_t s_stage_c[GS_NW][32 * (GS_PB + 1)];
  __shared__ uint8_t s_stage_f[GS_NW][32 * (GS_PB + 1)];
  if (h->status != 0) return;
  const Tbl tbl(ws);
  const uint64_t spol = ws.l2_hints ? l2_policy_stream() : 0ull;
  __shared__ uint32_t s_cnt[GS_CNT_STRIDE];
  const int P = ws.P;
  const int F = h->frontier_len[layer];
  const int fan = ws.fanouts[layer];
  const uint32_t batch = (uint32_t)h->batch_index;
  constexpr bool replay = RNG == (int)GS_RNG_DRAWS;
  constexpr bool norepl = RNG == (int)GS_RNG_PHILOX_NOREPLACE;
  const uint64_t dbase = h->draw_base[layer];
  const GsState stt = *ws.state;
  const Tag tag(ws, stt);
  const uint32_t tge = tag.tg | tag.kind;
  uint32_t *cnt = ws.cnt + layer * GS_CNT_STRIDE;
  const int lane = lane_id();
  uint32_t *rw_ea = ws.ea + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3),
           *rw_rowhas = ws.rowhas + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3);
  const int32_t *__restrict__ indices = ws.indices;
  const uint8_t *__restrict__ owner_tbl = ws.owner;
  const int pmask = ws.pmask;
  const uint32_t pmagic = ws.pmagic, seed_lo = ws.seed_lo, seed_hi = ws.seed_hi;
  const uint64_t n_draws = ws.n_draws;
  const bool hints = ws.l2_hints != 0;
  auto own_of = [&](int32_t v) -> int {
    if (owner_tbl) return (int)__ldg(owner_tbl + v);
    if (pmask >= 0) return (int)(v & pmask);
    int32_t r = v - (int32_t)(__umulhi((uint32_t)v, pmagic) * (uint32_t)P);
    return r < 0 ? r + P : r;
  };
  const uint32_t nb = (uint32_t)(fan + GS_PB - 1) / GS_PB;            // items per row
  const uint32_t nbm = (uint32_t)((0x100000000ull + nb - 1) / nb);    // item / nb by multiply-high + one correction
  const uint32_t nitems = (uint32_t)F * nb;                           // F < 2^30 / ... : fits (F * (fan + 1) <= Ccap < 2^30)
  if (threadIdx.x < GS_CNT_STRIDE) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  for (uint32_t it0 = blockIdx.x * GS_TPB; it0 < nitems; it0 += (uint32_t)nb_role * GS_TPB) {
    const uint32_t item = it0 + threadIdx.x;
    PK pk_in = 0, pk_e = 0, pk_out = 0, pk_own = 0, pk_to = 0;
    PK pk_in2 = 0, pk_e2 = 0;  // picks 4..7 of an item: a byte of a warp total holds at most 32 x 4
    uint32_t n_new = 0u;
    // phase A: the item's row and the positions it owns
    uint32_t row = 0u, b = 0u, o = 0u;
    int4 ri = make_int4(0, 0, 0, 0);
    int nj = 0, j0 = 0;
    uint32_t lane_lo = GS_INF, lane_hi = 0u;
    if (item < nitems) {
      row = item;  // nb == 1: ceil(2^32 / nb) does not fit the magic word
      if (nb > 1u) {
        row = __umulhi(item, nbm);
        if (row * nb > item) row--;
      }
      b = item - row * nb;
      ri = ws.rowinfo[row];
      o = ws.off[row];
      const uint32_t c = ws.off[row + 1] - o;
      j0 = (int)((uint32_t)GS_PB * b);
      nj = min(GS_PB, (int)c - 1 - j0);  // picks of this item (<= 0: none; the row is shorter)
      if (nj > 0) {
        lane_lo = o + 1u + (uint32_t)j0;
        lane_hi = lane_lo + (uint32_t)nj;
      }
      if (b == 0u) {
        lane_lo = o;
        lane_hi = max(lane_hi, o + 1u);
      }
    }
    // The items of a warp own one contiguous range of positions (<= 32 * (GS_PB + 1)): staged in shared memory,
    // written out as whole sectors (per-lane stores of four words at a stride cost 4x the sectors).
    const uint32_t pmin = __reduce_min_sync(FULL, lane_lo), pmax = __reduce_max_sync(FULL, lane_hi);
    int32_t *st_c = s_stage_c[warp_id()];
    uint8_t *st_f = s_stage_f[warp_id()];
    if (item < nitems) {
      const int64_t rs = (int64_t)(((uint64_t)((uint32_t)ri.y & 0x1FFFFFFFu) << 32) | (uint32_t)ri.x);
      const int g0 = (int)((uint32_t)ri.y >> 29);
      const int32_t deg = ri.z, nd = ri.w;
      const bool samp = deg >= fan;
      if (b == 0u) {  // the row's self entry (slicer.cpp:7)
        st_c[o - pmin] = (int32_t)row;
        st_f[o - pmin] = (uint8_t)(GS_CK_SELF | ((uint32_t)g0 << GS_CF_PART_SHIFT));
      }
      if (nj > 0) {
        uint4 blk[GS_PB / 4];  // one Philox block serves four picks
        Mod32 md;
        md.d = 1u; md.inv = 0u;
#pragma unroll
        for (int q4 = 0; q4 < GS_PB / 4; q4++) blk[q4] = make_uint4(0u, 0u, 0u, 0u);
        if (samp) {
          if (!replay) {
#pragma unroll
            for (int q4 = 0; q4 < GS_PB / 4; q4++)
              if (q4 == 0 || (!norepl && nj > 4 * q4))
                blk[q4] = philox_block(seed_lo, seed_hi, batch, layer, nd,
                                       norepl ? GS_NR_KEY_BLOCK : (int)(b * (GS_PB / 4) + q4));
          }
          if (!norepl) md.set((uint32_t)deg);
        }
        int32_t v[GS_PB];
#pragma unroll
        for (int j = 0; j < GS_PB; j++) {
          v[j] = nd;
          if (j < nj) {
            int32_t pick = j0 + j;
            if (samp) {
              if (norepl) {
                pick = (int32_t)nr_pick(blk[0], (uint32_t)pick, (uint32_t)deg);
              } else {
                uint32_t r;
                if (replay) {
                  const uint64_t di = dbase + (uint64_t)fan * ws.samp_rank[row] + (uint32_t)pick;
                  r = di < n_draws ? ws.draws[di] : 0u;
                } else {
                  r = philox_word(blk[j >> 2], j & 3);
                }
                pick = (int32_t)md.rem(r);
              }
            }
            v[j] = ld_stream_i32(indices + rs + pick, hints, spol);
          }
        }
        int g[GS_PB];
        uint32_t old[GS_PB];
        bool edge[GS_PB];
        uint32_t rbits = 0u;
        const uint32_t p1 = o + 1u + (uint32_t)j0;
#pragma unroll
        for (int j = 0; j < GS_PB; j++) {
          edge[j] = j < nj && v[j] != nd;  // a sampled neighbour equal to the destination is a self edge (slicer.cpp:33)
          g[j] = 0;
          old[j] = 0u;
          if (edge[j]) {
            g[j] = own_of(v[j]);
            old[j] = tbl.atom_min(v[j], tge | (p1 + j));
            rbits |= 1u << g[j];
          }
        }
        uint32_t was = 0u;
        if (rbits) was = atomicOr(rw_rowhas + row, rbits);
#pragma unroll
        for (int j = 0; j < GS_PB; j++) {
          if (j < nj) {
            int32_t cv = 0;
            uint32_t cf = GS_CK_DEAD;
            if (edge[j]) {
              PK &a_e = (GS_PB > 4 && j >= 4) ? pk_e2 : pk_e, &a_in = (GS_PB > 4 && j >= 4) ? pk_in2 : pk_in;
              cv = v[j];
              cf = GS_CK_EDGE | ((uint32_t)g[j] << GS_CF_PART_SHIFT);
              a_e += pkt_one<PK>(g[j]);
              if (!tag.mine(old[j])) {  // first arrival for a node that is new to the frontier
                n_new++;
                a_in += pkt_one<PK>(g[j]);
              } else if (!(old[j] & tag.kind)) {  // a row of the current frontier
                const uint32_t idx = old[j] & tag.pmask;
                if (atomicMin(rw_ea + idx, p1 + j) == GS_INF) a_in += pkt_one<PK>(g[j]);
                cv = (int32_t)idx;
                cf = GS_CK_FHIT | ((uint32_t)g[j] << GS_CF_PART_SHIFT);
              }
            }
            st_c[p1 + j - pmin] = cv;
            st_f[p1 + j - pmin] = (uint8_t)cf;
          }
        }
        const uint32_t nbits = rbits & ~was;
        if (nbits) {  // first edge of the row into each of these parts
          const uint32_t own = 1u << g0;
          pk_out += pkt_spread(nbits, (PK)0);
          if (nbits & own) pk_own += pkt_one<PK>(g0);
          if ((nbits & ~own) && !(was & ~own)) pk_to += pkt_one<PK>(g0);
        }
      }
    }
    __syncwarp();
    if (pmax > pmin) {
      const uint32_t len = pmax - pmin;
      for (uint32_t j = (uint32_t)lane; j < len; j += 32u) {
        cand[pmin + j] = st_c[j];
        ws.cflag[pmin + j] = st_f[j];
      }
    }
    __syncwarp();
    pk_in = pkt_redux(pk_in);
    pk_e = pkt_redux(pk_e);
    pk_out = pkt_redux(pk_out);
    pk_own = pkt_redux(pk_own);
    pk_to = pkt_redux(pk_to);
    n_new = __reduce_add_sync(FULL, n_new);
    if (lane < P) {
      uint32_t x;
      if ((x = pkt_get<PK>(pk_in, lane))) atomicAdd(&s_cnt[GS_CNT_IN(P, lane)], x);
      if ((x = pkt_get<PK>(pk_e, lane))) atomicAdd(&s_cnt[GS_CNT_E(P, lane)], x);
      if ((x = pkt_get<PK>(pk_out, lane))) atomicAdd(&s_cnt[GS_CNT_OUT(P, lane)], x);
      if ((x = pkt_get<PK>(pk_own, lane))) atomicAdd(&s_cnt[GS_CNT_OWN(P, lane)], x);
      if ((x = pkt_get<PK>(pk_to, lane))) atomicAdd(&s_cnt[GS_CNT_TO(P, lane)], x);
    }
    if (lane == 31 && n_new) atomicAdd(&s_cnt[GS_CNT_NEW], n_new);
    if (GS_PB > 4) {
      pk_in2 = pkt_redux(pk_in2);
      pk_e2 = pkt_redux(pk_e2);
      if (lane < P) {
        uint32_t x;
        if ((x = pkt_get<PK>(pk_in2, lane))) atomicAdd(&s_cnt[GS_CNT_IN(P, lane)], x);
        if ((x = pkt_get<PK>(pk_e2, lane))) atomicAdd(&s_cnt[GS_CNT_E(P, lane)], x);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 1 + 5 * P && s_cnt[threadIdx.x]) atomicAdd(cnt + threadIdx.x, s_cnt[threadIdx.x]);
  if (!last_block_n(ws.tickets + layer * GS_TICKETS + 1, (uint32_t)nb_role)) return;
  if (threadIdx.x < GS_CNT_STRIDE) s_cnt[threadIdx.x] = __ldcg(cnt + threadIdx.x);  // one round trip for all totals
  __syncthreads();
  if (threadIdx.x == 0) layer_alloc(ws, layer, s_cnt);
}

// ------------------------------------------------------------------- k_cand
// Two lean passes over the candidate stream, both with every minimum in place:
//   edge of a NEW node v at p:   first occurrence iff table[v] holds p  -> in_nodes of its
//                                owner's part (bipartite.cpp:4) and the next frontier (slicer.cpp:45-49);
//   edge of frontier ROW idx:    first edge occurrence iff ea[idx] == p -> in_nodes; it is the node's
//                                frontier slot only if it precedes the row's self entry (p < off[idx]);
//   self entry of row i at s:    frontier slot iff no edge of the node came earlier (ea[i] > s).
// k_cand_flags does the random reads (table / ea / off), leaves the verdict in the
// flag byte (GS_CF_FIRST / GS_CF_FFIRST) and, for every edge that is NOT a first
// occurrence, the position of the first one in the candidate word; it counts the
// 1 + 2 P classes per CTA and its last CTA turns the counts into prefixes.
// k_cand_apply reads stream and prefixes only: ranks by packed warp scans, lists
// written at their final place.
// (Measured, profiles/README.md: the single-launch form of the two -- runs of tiles
// kept in shared memory between the sweeps, one look-back per run -- takes the same
// 73 us alone on layer 2 but needs 37 KB of shared memory and 80 registers per CTA
// and parks every CTA at a grid-wide wait, so that other minibatches' kernels find
// no SM slots: 4.1 K minibatches/s against the split form's rate.)
#define CS_NF 0
#define CS_IN(q) (1 + (q))
#define CS_E(q) (9 + (q))
#define CS_USED 17
#define GS_CF_FIRST 0x04u   // first occurrence as an edge source -> in_nodes of its part
#define GS_CF_FFIRST 0x08u  // first occurrence in the layer's stream -> next frontier

__device__ __forceinline__ void cand_load(const GsWs &ws, const int32_t *cand, uint32_t p0, uint32_t C,
                                          uint32_t (&fl)[GS_IPT], int32_t (&ci)[GS_IPT]) {
  if (p0 + GS_IPT <= C) {
    const int4 c4 = *reinterpret_cast<const int4 *>(cand + p0);
    const uchar4 f4 = *reinterpret_cast<const uchar4 *>(ws.cflag + p0);
    ci[0] = c4.x; ci[1] = c4.y; ci[2] = c4.z; ci[3] = c4.w;
    fl[0] = f4.x; fl[1] = f4.y; fl[2] = f4.z; fl[3] = f4.w;
  } else {
#pragma unroll
    for (int i = 0; i < GS_IPT; i++) {
      fl[i] = GS_CK_DEAD;
      ci[i] = 0;
      if (p0 + i < C) {
        fl[i] = ws.cflag[p0 + i];
        ci[i] = cand[p0 + i];
      }
    }
  }
}

template <typename PK>
__device__ void cand_flags_role(const GsWs &ws, int layer, int bx, int nb) {
  gs_sample_header *h = ws.hdr;
  int32_t *const cand = cand_buf(ws, layer);
  if (h->status != 0) return;
  __shared__ uint32_t s_cnt[GS_CS_NC];
  __shared__ uint32_t s_tot[GS_CS_NC];
  const Tbl tbl(ws);
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const GsState stt = *ws.state;
  const Tag tag(ws, stt);
  const uint32_t *rw_ea = ws.ea + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3);
  const int lane = lane_id();
  if (threadIdx.x < GS_CS_NC) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  uint32_t acc_i = 0u, acc_e = 0u, acc_f = 0u;  // lane q < 8: this warp's running counts of part q
  int t0, t1;
  role_tiles(ntiles, bx, nb, t0, t1);  // = cta_tiles of k_cand_apply, launched with nb CTAs
  for (int t = t0; t < t1; t++) {
    const uint32_t p0 = (uint32_t)t * GS_TILE + GS_IPT * threadIdx.x;
    uint32_t fl[GS_IPT];
    int32_t ci[GS_IPT];
    cand_load(ws, cand, p0, C, fl, ci);
    uint32_t x[GS_IPT], s[GS_IPT];
#pragma unroll
    for (int i = 0; i < GS_IPT; i++) {
      const uint32_t kind = fl[i] & GS_CK_MASK;
      x[i] = GS_INF; s[i] = 0u;
      if (kind == GS_CK_EDGE) {
        x[i] = tbl.ld(ci[i]) & tag.pmask;
      } else if (kind != GS_CK_DEAD) {
        x[i] = __ldcg(rw_ea + ci[i]);
        if (kind == GS_CK_FHIT) s[i] = __ldcg(ws.off + ci[i]);
      }
    }
    PK pi = 0, pe = 0;
    uint32_t nf = 0u;
    bool moved = false;
#pragma unroll
    for (int i = 0; i < GS_IPT; i++) {
      const uint32_t p = p0 + i, kind = fl[i] & GS_CK_MASK;
      const int g = (int)((fl[i] >> GS_CF_PART_SHIFT) & 7u);
      if (kind == GS_CK_SELF) {
        if (x[i] > p) { fl[i] |= GS_CF_FFIRST; nf++; }
      } else if (kind != GS_CK_DEAD) {
        pe += pkt_one<PK>(g);
        if (x[i] == p) {
          fl[i] |= GS_CF_FIRST;
          pi += pkt_one<PK>(g);
          if (kind == GS_CK_EDGE || p < s[i]) { fl[i] |= GS_CF_FFIRST; nf++; }
        } else {
          ci[i] = (int32_t)x[i];  // position of the source's first occurrence: all a later pass needs
          moved = true;
        }
      }
    }
    if (p0 + GS_IPT <= C) {
      if (moved) *reinterpret_cast<int4 *>(cand + p0) = make_int4(ci[0], ci[1], ci[2], ci[3]);
      *reinterpret_cast<uchar4 *>(ws.cflag + p0) =
          make_uchar4((unsigned char)fl[0], (unsigned char)fl[1], (unsigned char)fl[2], (unsigned char)fl[3]);
    } else {
#pragma unroll
      for (int i = 0; i < GS_IPT; i++)
        if (p0 + i < C) {
          cand[p0 + i] = ci[i];
          ws.cflag[p0 + i] = (uint8_t)fl[i];
        }
    }
    pi = pkt_redux(pi);
    pe = pkt_redux(pe);
    nf = __reduce_add_sync(FULL, nf);
    if (lane < (int)sizeof(PK)) {
      acc_i += pkt_get<PK>(pi, lane);
      acc_e += pkt_get<PK>(pe, lane);
    }
    acc_f += nf;
  }
  if (lane < 8) {
    if (acc_i) atomicAdd(&s_cnt[CS_IN(lane)], acc_i);
    if (acc_e) atomicAdd(&s_cnt[CS_E(lane)], acc_e);
  }
  if (lane == 8 && acc_f) atomicAdd(&s_cnt[CS_NF], acc_f);
  __syncthreads();
  if (threadIdx.x < CS_USED) ws.ccnt[(size_t)threadIdx.x * ws.cgrid_cap + bx] = s_cnt[threadIdx.x];
  if (!last_block_n(ws.tickets + layer * GS_TICKETS + 2, (uint32_t)nb)) return;
  spine_scan<GS_CS_NC>(ws.ccnt, ws.cgrid_cap, nb, CS_USED, s_tot, nullptr);
}

// Second pass: ranks and lists.  Same contiguous runs of tiles per CTA as
// cand_flags_role (cta_tiles), so a CTA starts from its spine prefix and carries
// the running class counts across its own tiles in shared memory.
template <typename PK>
__device__ void cand_apply_role(const GsWs &ws, int layer) {
  gs_sample_header *h = ws.hdr;
  int32_t *const cand = cand_buf(ws, layer);
  if (h->status != 0) return;
  __shared__ uint32_t s_wc[GS_NW][GS_CS_NC];    // per-warp class totals of the tile
  __shared__ uint32_t s_base[GS_NW][GS_CS_NC];  // CTA prefix + earlier tiles + earlier warps of the tile
  __shared__ uint32_t s_run[2][GS_CS_NC];       // running class counts at the start of a tile (ping-pong)
  __shared__ int32_t *s_in[GS_MAX_PARTS];
  __shared__ int32_t s_fo_ix[GS_MAX_PARTS];
  const int P = ws.P;
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const int32_t *__restrict__ front = ws.fslab + h->frontier[layer].off;
  int32_t *fnext = ws.fslab + h->frontier[layer + 1].off;
  const int lane = lane_id(), w = warp_id();
  const size_t rs_stride = sizeof(PK);  // rowscan: one record of 4 (P <= 4) or 8 words per row
  if (threadIdx.x < P) {
    s_in[threadIdx.x] = ws.tgt->base[threadIdx.x] + h->field[layer][threadIdx.x][GS_F_IN_NODES].off;
    // csr_indices go to the slot's edge scratch first (relabel_role writes the final list):
    // part q's edges start where the parts before it end
    int32_t at = 0;
    for (int q = 0; q < (int)threadIdx.x; q++) at += (h->field[layer][q][GS_F_CSR_INDICES].len + 3) & ~3;
    s_fo_ix[threadIdx.x] = at;
  }
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  if (threadIdx.x < GS_CS_NC)
    s_run[t0 & 1][threadIdx.x] =
        threadIdx.x < CS_USED ? ws.ccnt[(size_t)threadIdx.x * ws.cgrid_cap + blockIdx.x] : 0u;
  for (int t = t0; t < t1; t++) {
    __syncthreads();
    const uint32_t p0 = (uint32_t)t * GS_TILE + GS_IPT * threadIdx.x;
    const bool full = p0 + GS_IPT <= C;
    uint32_t fl[GS_IPT];
    int32_t ci[GS_IPT];
    cand_load(ws, cand, p0, C, fl, ci);
    PK pi = 0, pe = 0;
    uint32_t nf = 0u;
#pragma unroll
    for (int i = 0; i < GS_IPT; i++) {
      const uint32_t kind = fl[i] & GS_CK_MASK;
      const int g = (int)((fl[i] >> GS_CF_PART_SHIFT) & 7u);
      if (kind == GS_CK_EDGE || kind == GS_CK_FHIT) pe += pkt_one<PK>(g);
      if (fl[i] & GS_CF_FIRST) pi += pkt_one<PK>(g);
      nf += (fl[i] & GS_CF_FFIRST) ? 1u : 0u;
    }
    const PK inc_i = pkt_incl(pi), inc_e = pkt_incl(pe);
    const uint32_t inc_f = warp_incl_scan(nf);
    {
      const PK t_i = __shfl_sync(FULL, inc_i, 31), t_e = __shfl_sync(FULL, inc_e, 31);
      const uint32_t t_f = __shfl_sync(FULL, inc_f, 31);
      if (lane < 8) {
        const bool in = lane < (int)sizeof(PK);
        s_wc[w][CS_IN(lane)] = in ? pkt_get<PK>(t_i, lane) : 0u;
        s_wc[w][CS_E(lane)] = in ? pkt_get<PK>(t_e, lane) : 0u;
      }
      if (lane == 8) s_wc[w][CS_NF] = t_f;
    }
    __syncthreads();
    if (lane < CS_USED) {  // lane = class: this warp's base; warp 0 also advances the running counts
      uint32_t a = s_run[t & 1][lane];
      if (w == 0) {
        uint32_t all = 0u;
#pragma unroll
        for (int k = 0; k < GS_NW; k++) all += s_wc[k][lane];
        s_run[(t + 1) & 1][lane] = a + all;
      } else {
        for (int k = 0; k < w; k++) a += s_wc[k][lane];
      }
      s_base[w][lane] = a;
    }
    __syncwarp();
    const uint32_t *bw = s_base[w];
    uint32_t rf = bw[CS_NF] + inc_f - nf;
    PK ri = inc_i - pi, re = inc_e - pe;
    bool wrote = false;
#pragma unroll
    for (int i = 0; i < GS_IPT; i++) {
      const uint32_t kind = fl[i] & GS_CK_MASK;
      const int g = (int)((fl[i] >> GS_CF_PART_SHIFT) & 7u);
      if (kind == GS_CK_SELF) {
        if (fl[i] & GS_CF_FFIRST) fnext[rf++] = front[ci[i]];
        // edges of every part emitted before this row: its CSR row pointers, one record per row
        uint4 *rsp = reinterpret_cast<uint4 *>(ws.rowscan + (size_t)ci[i] * rs_stride);
        rsp[0] = make_uint4(bw[CS_E(0)] + pkt_get<PK>(re, 0), bw[CS_E(1)] + pkt_get<PK>(re, 1),
                            bw[CS_E(2)] + pkt_get<PK>(re, 2), bw[CS_E(3)] + pkt_get<PK>(re, 3));
        if (sizeof(PK) > 4)
          rsp[1] = make_uint4(bw[CS_E(4)] + pkt_get<PK>(re, 4), bw[CS_E(5)] + pkt_get<PK>(re, 5),
                              bw[CS_E(6)] + pkt_get<PK>(re, 6), bw[CS_E(7)] + pkt_get<PK>(re, 7));
      } else if (kind != GS_CK_DEAD) {
        const uint32_t ex = bw[CS_E(g)] + pkt_get<PK>(re, g);
        if (fl[i] & GS_CF_FIRST) {
          // the first edge occurrence of a source knows its rank: final csr_indices value,
          // in_nodes entry, relabel map entry; every other edge carries the position m of
          // that first occurrence and leaves ~m for relabel_role
          const uint32_t rk = bw[CS_IN(g)] + pkt_get<PK>(ri, g);
          const int32_t z = kind == GS_CK_FHIT ? front[ci[i]] : ci[i];  // a new node carries its id, a frontier row its index
          st_out(s_in[g] + rk, z);
          ws.ixs[s_fo_ix[g] + ex] = (int32_t)rk;
          if (fl[i] & GS_CF_FFIRST) fnext[rf++] = z;
          ci[i] = (int32_t)rk;
          wrote = true;
          ri += pkt_one<PK>(g);
        } else {
          ws.ixs[s_fo_ix[g] + ex] = ~ci[i];
        }
        re += pkt_one<PK>(g);
      }
    }
    if (wrote) {
      if (full) {
        *reinterpret_cast<int4 *>(cand + p0) = make_int4(ci[0], ci[1], ci[2], ci[3]);
      } else {
#pragma unroll
        for (int i = 0; i < GS_IPT; i++)
          if (p0 + i < C && (fl[i] & GS_CF_FIRST)) cand[p0 + i] = ci[i];
      }
    }
  }
}

// ------------------------------------------------------------------ rowside
// Row-side lists of layer `layer`, one dense pass over the frontier rows in the
// launch after k_cand.  Row i (node nd, owner g0) with has_g = "emitted >= 1 edge
// into part g" (rowhas, k_expand):  out_g = has_g (bipartite.h:59-62); owned_g =
// has_g && g == g0 (bipartite.h:56-58); from_g = has_g && g != g0, rank = rank(out_g)
// - rank(owned_g); to_g0 = any has_h, h != g0 (slicer.cpp:41); self_g0 = always
// (slicer.cpp:33-34), slot = selfrank (k_rows); csr_indptr_g = rowscan (k_cand);
// self_ids_in = in_nodes rank of the node's first edge occurrence (the relabel map
// at position ea[i]) or -1.  3 P counters, flat look-back per tile of rows.
#define FS_OUT(q) (q)
#define FS_OWN(q) (8 + (q))
#define FS_TO(q) (16 + (q))
#define FS_USED 24
// Count pass of the row side (rides in the k_cand_flags launch: rowhas is final once
// k_expand has run): the 3 P classes of every CTA run of rows, the last CTA turns
// them into exclusive prefixes.  Same runs as rowside_role (role_tiles, same nb).
template <typename PK>
__device__ void rowcount_role(const GsWs &ws, int layer, int bx, int nb) {
  const gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_cnt[GS_FS_NC];
  __shared__ uint32_t s_tot[GS_FS_NC];
  const int F = h->frontier_len[layer];
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  const int32_t *__restrict__ front = ws.fslab + h->frontier[layer].off;
  const uint32_t *rw_rowhas = ws.rowhas + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3);
  const int lane = lane_id();
  if (threadIdx.x < GS_FS_NC) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  uint32_t acc_o = 0u, acc_w = 0u, acc_t = 0u;  // lane q < 8: this warp's running counts of part q
  int t0, t1;
  role_tiles(ntiles, bx, nb, t0, t1);
  for (int t = t0; t < t1; t++) {
    const int i0 = t * GS_TILE + GS_IPT * (int)threadIdx.x;
    uint32_t has[GS_IPT] = {0u, 0u, 0u, 0u};
    int32_t nd[GS_IPT] = {0, 0, 0, 0};
    if (i0 + GS_IPT <= F) {
      const uint4 h4 = *reinterpret_cast<const uint4 *>(rw_rowhas + i0);
      const int4 n4 = *reinterpret_cast<const int4 *>(front + i0);
      has[0] = h4.x; has[1] = h4.y; has[2] = h4.z; has[3] = h4.w;
      nd[0] = n4.x; nd[1] = n4.y; nd[2] = n4.z; nd[3] = n4.w;
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (i0 + k < F) {
          has[k] = rw_rowhas[i0 + k];
          nd[k] = front[i0 + k];
        }
    }
    PK po = 0, pw = 0, pt = 0;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++)
      if (has[k]) {
        const int g0 = owner_of(ws, nd[k]);
        const uint32_t own = 1u << g0;
        po += pkt_spread(has[k], (PK)0);
        if (has[k] & own) pw += pkt_one<PK>(g0);
        if (has[k] & ~own) pt += pkt_one<PK>(g0);
      }
    po = pkt_redux(po);
    pw = pkt_redux(pw);
    pt = pkt_redux(pt);
    if (lane < (int)sizeof(PK)) {
      acc_o += pkt_get<PK>(po, lane);
      acc_w += pkt_get<PK>(pw, lane);
      acc_t += pkt_get<PK>(pt, lane);
    }
  }
  if (lane < 8) {
    if (acc_o) atomicAdd(&s_cnt[FS_OUT(lane)], acc_o);
    if (acc_w) atomicAdd(&s_cnt[FS_OWN(lane)], acc_w);
    if (acc_t) atomicAdd(&s_cnt[FS_TO(lane)], acc_t);
  }
  __syncthreads();
  if (threadIdx.x < FS_USED) ws.fcnt[(size_t)threadIdx.x * ws.cgrid_cap + bx] = s_cnt[threadIdx.x];
  if (!last_block_n(ws.tickets + layer * GS_TICKETS + 3, (uint32_t)nb)) return;
  spine_scan<GS_FS_NC>(ws.fcnt, ws.cgrid_cap, nb, FS_USED, s_tot, nullptr);
}

// Apply pass of the row side, in the launch after k_cand_apply (next to the next layer's
// k_rows, or in k_tail): CTA bx starts from rowcount_role's prefix of its run of rows and
// carries the running class counts across its tiles -- no wait on any other CTA.
template <typename PK>
__device__ void rowside_role(const GsWs &ws, int layer, int bx, int nb) {
  gs_sample_header *h = ws.hdr;
  int32_t *const cand = cand_buf(ws, layer);
  // Rows are dealt lane-consecutive (row = tile + k * 256 + thread): in every store below the 32 lanes
  // of a warp append to a list at CONSECUTIVE places, so a warp store touches a few sectors instead of
  // one per lane (four consecutive rows per thread cost 4x the sector writes; measured, profiles/README.md).
  __shared__ uint32_t s_wc[GS_IPT * GS_NW][GS_FS_NC];   // class counts of (k, warp) = 32 consecutive rows
  __shared__ uint32_t s_base[GS_IPT * GS_NW][GS_FS_NC]; // -> list places before them
  __shared__ uint32_t s_run[GS_FS_NC];
  __shared__ int32_t s_fo[GS_MAX_PARTS][GS_NUM_FIELDS];
  __shared__ int32_t *s_tb[GS_MAX_PARTS];  // list base of every part: the slab, or its hand-over block
  if (h->status != 0) return;
  const int P = ws.P;
  const int F = h->frontier_len[layer];
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  int t0, t1;
  role_tiles(ntiles, bx, nb, t0, t1);
  if (t0 >= t1) return;
  if (threadIdx.x < P) s_tb[threadIdx.x] = ws.tgt->base[threadIdx.x];
  const int32_t *__restrict__ front = ws.fslab + h->frontier[layer].off;
  const uint32_t *rw_ea = ws.ea + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3), *rw_rowhas = ws.rowhas + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3),
                 *rw_selfrank = ws.selfrank + (size_t)(layer & 1) * (((size_t)ws.Fcap + 3) & ~(size_t)3);  // ping-pong by layer (see rows_role)
  const int lane = lane_id(), w = warp_id();
  const size_t rs_stride = sizeof(PK);  // rowscan: one record of 4 (P <= 4) or 8 words per row
  if (threadIdx.x < P * GS_NUM_FIELDS)
    s_fo[threadIdx.x / GS_NUM_FIELDS][threadIdx.x % GS_NUM_FIELDS] =
        h->field[layer][threadIdx.x / GS_NUM_FIELDS][threadIdx.x % GS_NUM_FIELDS].off;
  if (threadIdx.x < GS_FS_NC)
    s_run[threadIdx.x] = threadIdx.x < FS_USED ? ws.fcnt[(size_t)threadIdx.x * ws.cgrid_cap + bx] : 0u;
  __syncthreads();
  for (int t = t0; t < t1; t++) {
    const int ib = t * GS_TILE + (int)threadIdx.x;  // row of k: ib + k * GS_TPB
    uint32_t has[GS_IPT], a[GS_IPT], sr[GS_IPT];
    int32_t nd[GS_IPT];
    int g0[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int i = ib + k * GS_TPB;
      has[k] = 0u; nd[k] = 0; a[k] = GS_INF; sr[k] = 0u;
      if (i < F) {
        has[k] = rw_rowhas[i];
        nd[k] = front[i];
        a[k] = rw_ea[i];
        sr[k] = rw_selfrank[i];
      }
    }
    // self_ids_in: the relabel map at the position of the node's first edge occurrence
    int32_t sin[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) sin[k] = (a[k] == GS_INF || (GS_DIAG & 2)) ? -1 : __ldcg(cand + a[k]);
    PK ro[GS_IPT], rw[GS_IPT], rt[GS_IPT];  // exclusive class counts within the (k, warp) block
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      PK po = 0, pw = 0, pt = 0;
      g0[k] = -1;
      if (ib + k * GS_TPB < F) {
        g0[k] = owner_of(ws, nd[k]);
        const uint32_t own = 1u << g0[k];
        po = pkt_spread(has[k], (PK)0);
        if (has[k] & own) pw = pkt_one<PK>(g0[k]);
        if (has[k] & ~own) pt = pkt_one<PK>(g0[k]);
      }
      const PK inc_o = pkt_incl(po), inc_w = pkt_incl(pw), inc_t = pkt_incl(pt);
      ro[k] = inc_o - po; rw[k] = inc_w - pw; rt[k] = inc_t - pt;
      const PK t_o = __shfl_sync(FULL, inc_o, 31), t_w = __shfl_sync(FULL, inc_w, 31), t_t = __shfl_sync(FULL, inc_t, 31);
      if (lane < 8) {
        const bool in = lane < (int)sizeof(PK);
        s_wc[k * GS_NW + w][FS_OUT(lane)] = in ? pkt_get<PK>(t_o, lane) : 0u;
        s_wc[k * GS_NW + w][FS_OWN(lane)] = in ? pkt_get<PK>(t_w, lane) : 0u;
        s_wc[k * GS_NW + w][FS_TO(lane)] = in ? pkt_get<PK>(t_t, lane) : 0u;
      }
    }
    __syncthreads();
    if (threadIdx.x < FS_USED) {
      uint32_t run = s_run[threadIdx.x];
#pragma unroll 8
      for (int j = 0; j < GS_IPT * GS_NW; j++) {
        s_base[j][threadIdx.x] = run;
        run += s_wc[j][threadIdx.x];
      }
      s_run[threadIdx.x] = run;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int i = ib + k * GS_TPB;
      const uint32_t *bs = s_base[k * GS_NW + w];
      if (i < F && !(GS_DIAG & 1)) {
        const int gq = g0[k];
        const uint32_t own = 1u << gq;
        int32_t op_own = -1;  // position of nd in its own part's out_nodes, or -1
        // the row's record (edges of every part emitted before it) in one or two 16-byte loads
        uint4 r0 = make_uint4(0u, 0u, 0u, 0u), r1 = r0;
        if (has[k] & 0x0Fu) r0 = *reinterpret_cast<const uint4 *>(ws.rowscan + (size_t)i * rs_stride);
        if (sizeof(PK) > 4 && (has[k] & 0xF0u)) r1 = *reinterpret_cast<const uint4 *>(ws.rowscan + (size_t)i * rs_stride + 4);
        for (uint32_t hb = has[k]; hb; hb &= hb - 1u) {
          const int q = __ffs((int)hb) - 1;
          const uint32_t op = bs[FS_OUT(q)] + pkt_get<PK>(ro[k], q);
          const uint32_t ow = bs[FS_OWN(q)] + pkt_get<PK>(rw[k], q);
          const uint4 rr = q < 4 ? r0 : r1;
          const uint32_t rsq = (q & 3) == 0 ? rr.x : (q & 3) == 1 ? rr.y : (q & 3) == 2 ? rr.z : rr.w;
          st_out(s_tb[q] + s_fo[q][GS_F_OUT_NODES] + op, nd[k]);
          st_out(s_tb[q] + s_fo[q][GS_F_CSR_INDPTR] + op, (int32_t)rsq);
          if (q == gq) {
            st_out(s_tb[q] + s_fo[q][GS_F_OWNED_OUT] + ow, (int32_t)op);
            op_own = (int32_t)op;
          } else {
            st_out(s_tb[q] + s_fo[q][GS_F_FROM_IDS] + (op - ow), (int32_t)op);
          }
        }
        if (has[k] & ~own) st_out(s_tb[gq] + s_fo[gq][GS_F_TO_IDS] + bs[FS_TO(gq)] + pkt_get<PK>(rt[k], gq), op_own);
        st_out(s_tb[gq] + s_fo[gq][GS_F_SELF_IN] + sr[k], sin[k]);
        st_out(s_tb[gq] + s_fo[gq][GS_F_SELF_OUT] + sr[k], op_own);
      }
    }
  }
}

// ---------------------------------------------------------------- relabel
// csr_indices: an edge that was not the first occurrence of its source holds ~m,
// m the position of the first one; cand[m] is that source's rank in in_nodes_g
// (DuplicateRemover::replace, duplicate.cpp:35-39 -- what bipartite.cpp:5 would do
// if `indices` were ever filled).  Runs in the launch after k_cand, next to the
// next layer's k_rows (or in k_tail).  bx / nb: this role's CTA index and count.
__device__ void relabel_role(const GsWs &ws, int layer, int bx, int nb) {
  const gs_sample_header *h = ws.hdr;
  int32_t *const cand = cand_buf(ws, layer);
  if (h->status != 0) return;
  const int P = ws.P;
  int32_t at = 0;
  for (int q = 0; q < P; q++) {
    const gs_span sp = h->field[layer][q][GS_F_CSR_INDICES];
    const int32_t *__restrict__ src = ws.ixs + at;  // k_cand_apply's scratch (same layout rule)
    int32_t *dst = ws.tgt->base[q] + sp.off;        // the final list: slab, or the part's hand-over block
    at += (sp.len + 3) & ~3;
    const int stride = nb * GS_TPB;
    for (int e0 = bx * GS_TPB + threadIdx.x; e0 < sp.len; e0 += GS_IPT * stride) {
      int32_t v[GS_IPT], r[GS_IPT];
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        const int e = e0 + k * stride;
        v[k] = e < sp.len ? __ldcs(src + e) : 0;  // >= 0: the rank itself (a first occurrence)
      }
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) r[k] = v[k] < 0 ? __ldcg(cand + (uint32_t)~v[k]) : v[k];
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        const int e = e0 + k * stride;
        if (e < sp.len) st_out(dst + e, r[k]);  // every entry, whole sectors
      }
    }
  }
}

// ------------------------------------------------------------ feature gather
// One warp per row: out[r, :] = rows[id / div, :].  602 fp32 = 2408 B rows are
// 8-byte but not 16-byte multiples, so the vector width is chosen from dim
// (float4 / float2 / float); rows are contiguous, so unlike the id work above this
// streams whole lines and is bounded by plain HBM bandwidth.  Fused into every
// sample's last launch: the rows of the last layer's in_nodes_g, part after part
// -- the block of input features GPU g reads for this minibatch
// (python/data/bipartite.py:61-67 gather, memory_manager.py:75-104 refresh).
// With a cached store (FeatSrc) rows outside the HBM cache are read in place from
// host memory by the same warp, so there is no per-batch refresh step and no
// global_to_local table to patch and undo (memory_manager.py:77-104).
__device__ void gather_role(const GsWs &ws, int bx, int nb) {
  const gs_sample_header *h = ws.hdr;
  __shared__ int32_t s_off[GS_MAX_PARTS], s_start[GS_MAX_PARTS + 1];
  const int P = ws.P, l = ws.L - 1;
  if (threadIdx.x == 0) {
    int32_t acc = 0;
    for (int q = 0; q < P; q++) {
      s_off[q] = h->field[l][q][GS_F_IN_NODES].off;
      s_start[q] = acc;
      acc += h->field[l][q][GS_F_IN_NODES].len;
    }
    s_start[P] = acc;
  }
  __syncthreads();
  const int R = s_start[P];
  if ((int64_t)R > ws.feat_cap_rows) return;  // k_tail's last CTA reports the overflow
  const FeatSrc fs{ws.feat, ws.feat_slot, ws.feat_host};
  const int dim = ws.feat_dim, div = ws.feat_div;
  const int32_t *slab = ws.slab;  // the fused gather is not combined with the direct hand-over
  auto src_of_row = [&](int64_t r) -> const float * {
    int q = 0;
    while (q + 1 < P && r >= s_start[q + 1]) q++;
    const int32_t id = slab[s_off[q] + ((int)r - s_start[q])];
    return fs.row((int64_t)(id / div), dim);
  };
  gather_block(dim, (int64_t)R, ws.feat_out, src_of_row, (int64_t)bx * GS_NW + warp_id(), (int64_t)nb * GS_NW,
               lane_id());
}

// ------------------------------------------------------------------ kernels
// WIDE: more than 4 parts -> packed per-part counters are 64-bit words (else 32-bit).
template <bool FIRST>
__global__ void __launch_bounds__(GS_TPB, GS_MINB_ROWS) k_rows(const GsWs *__restrict__ wsv, int layer) {
  __shared__ GsWs ws;
  stage_ws(ws, wsv + blockIdx.y);
  rows_role<FIRST>(ws, layer);
}

// Extra CTAs of an expand launch: relabel and row side of the PREVIOUS layer (they need
// k_cand_apply's output and nothing of this layer; none of their arrays is written here).
template <typename PK>
__device__ __forceinline__ void lists_roles(const GsWs &ws, int layer, int bx, int grid_relabel, int grid_rowside) {
  if (bx < grid_relabel) relabel_role(ws, layer, bx, grid_relabel);
  else rowside_role<PK>(ws, layer, bx - grid_relabel, grid_rowside);
}

template <int RNG, bool WIDE>
__global__ void __launch_bounds__(GS_TPB, GS_MINB_EXPAND) k_expand(const GsWs *__restrict__ wsv, int layer, int grid_expand,
                                                                    int grid_relabel) {
  typedef typename std::conditional<WIDE, unsigned long long, uint32_t>::type PK;
  __shared__ GsWs ws;
  stage_ws(ws, wsv + blockIdx.y);
  const int bx = (int)blockIdx.x;
  if (bx < grid_expand) expand_role<RNG, PK>(ws, layer, grid_expand);
  else lists_roles<PK>(ws, layer - 1, bx - grid_expand, grid_relabel, (int)gridDim.x - grid_expand - grid_relabel);
}

template <int RNG, bool WIDE>
__global__ void __launch_bounds__(GS_TPB, GS_MINB_EXPAND) k_expand_rows(const GsWs *__restrict__ wsv, int layer,
                                                                         int grid_expand, int grid_relabel) {
  typedef typename std::conditional<WIDE, unsigned long long, uint32_t>::type PK;
  __shared__ GsWs ws;
  stage_ws(ws, wsv + blockIdx.y);
  const int bx = (int)blockIdx.x;
  if (bx < grid_expand) expand_rows_role<RNG, PK>(ws, layer, grid_expand);
  else lists_roles<PK>(ws, layer - 1, bx - grid_expand, grid_relabel, (int)gridDim.x - grid_expand - grid_relabel);
}

template <bool WIDE>
__global__ void __launch_bounds__(GS_TPB, GS_MINB_CAND) k_cand_flags(const GsWs *__restrict__ wsv, int layer,
                                                                      int grid_cand) {
  typedef typename std::conditional<WIDE, unsigned long long, uint32_t>::type PK;
  __shared__ GsWs ws;
  stage_ws(ws, wsv + blockIdx.y);
  const int bx = (int)blockIdx.x;
  if (bx < grid_cand) cand_flags_role<PK>(ws, layer, bx, grid_cand);
  else rowcount_role<PK>(ws, layer, bx - grid_cand, (int)gridDim.x - grid_cand);
}

template <bool WIDE>
__global__ void __launch_bounds__(GS_TPB, GS_MINB_CAND) k_cand_apply(const GsWs *__restrict__ wsv, int layer) {
  typedef typename std::conditional<WIDE, unsigned long long, uint32_t>::type PK;
  __shared__ GsWs ws;
  stage_ws(ws, wsv + blockIdx.y);
  cand_apply_role<PK>(ws, layer);
}

// Last launch of a sample: relabel and row side of the last layer, the fused
// feature gather, and the slot's housekeeping -- every CTA takes its share of the
// rare dense refills (dedup table out of tags: 1 sample in 2^(31-S) - 1; scan
// descriptors: once per 2^30 scans), the last CTA to finish closes the header,
// mirrors it to pinned host memory and prepares the slot's state for the next sample.
template <bool WIDE>
__global__ void __launch_bounds__(GS_TPB, GS_MINB_CAND) k_tail(const GsWs *__restrict__ wsv, int grid_relabel, int grid_rowside, int grid_gather) {
  typedef typename std::conditional<WIDE, unsigned long long, uint32_t>::type PK;
  __shared__ GsWs ws;
  stage_ws(ws, wsv + blockIdx.y);
  gs_sample_header *h = ws.hdr;
  const int st = h->status;
  const int L = ws.L, P = ws.P;
  const int bx = (int)blockIdx.x;
  if (st == 0) {
    if (bx < grid_relabel) relabel_role(ws, L - 1, bx, grid_relabel);
    else if (bx < grid_relabel + grid_rowside) rowside_role<PK>(ws, L - 1, bx - grid_relabel, grid_rowside);
    else if (ws.feat) gather_role(ws, bx - grid_relabel - grid_rowside, grid_gather);
  }
  const GsState stt = *ws.state;
  const uint32_t ntags = 1u << (31 - ws.pos_bits);
  const bool refill = stt.tag_idx + 1u > ntags - 1u;  // the next sample would have no tag left
  const bool rescan = stt.scan_epoch + 6u * (uint32_t)L >= (1u << 30);
  const int64_t tid = (int64_t)bx * GS_TPB + threadIdx.x, nth = (int64_t)gridDim.x * GS_TPB;
  if (refill) {
    uint4 *t4 = (uint4 *)ws.table;
    const int64_t n4 = (int64_t)ws.num_nodes >> 2;
    const uint4 inf = make_uint4(GS_INF, GS_INF, GS_INF, GS_INF);
    for (int64_t i = tid; i < n4; i += nth) t4[i] = inf;
    if (tid < (ws.num_nodes & 3)) ws.table[(n4 << 2) + tid] = GS_INF;
  }
  if (rescan) {
    for (int64_t i = tid; i < ws.ftiles_cap; i += nth) ws.rs_status[i] = 0u;
  }
  if (!last_block(ws.tickets + L * GS_TICKETS)) return;
  for (int i = threadIdx.x; i < L * GS_CNT_STRIDE; i += GS_TPB) ws.cnt[i] = 0u;
  for (int i = threadIdx.x; i < L * GS_TICKETS; i += GS_TPB) ws.tickets[i] = 0u;
  if (threadIdx.x == 0) {
    if (ws.feat && st == 0) {
      int32_t acc = 0;
      for (int q = 0; q <= P; q++) {
        h->feature_part_off[q] = acc;
        if (q < P) acc += h->field[L - 1][q][GS_F_IN_NODES].len;
      }
      const bool fits = (int64_t)acc <= ws.feat_cap_rows;
      h->feature_rows = fits ? acc : 0;
      if (!fits) h->late_status = (int32_t)GS_ERR_OVERFLOW;
    }
    GsState nx;
    nx.tag_idx = refill ? 1u : stt.tag_idx + 1u;
    nx.scan_epoch = rescan ? 1u : stt.scan_epoch + 3u * (uint32_t)L;
    nx.pad[0] = nx.pad[1] = 0u;
    *ws.state = nx;
  }
  __syncthreads();
  if (ws.tgt->direct) {
    // direct hand-over: the header of every part's block (layout of gs_slicer_pack_parts), last,
    // so that a block whose header is in place is complete
    const int status = h->status ? h->status : h->late_status;
    const int ns = L * GS_NUM_FIELDS;
    // words in use of every block, for the host (gs_slicer_push_parts sizes its copies with them)
    if (threadIdx.x < P) h->feature_part_off[threadIdx.x] = (int32_t)ws.tgt->cursor[threadIdx.x];
    __syncthreads();
    for (int q = 0; q < P; q++) {
      int32_t *blk = ws.tgt->base[q];
      for (int k = threadIdx.x; k < GS_PART_HEADER_WORDS; k += GS_TPB) {
        int32_t v = 0;
        if (k == 0) v = (int32_t)GS_PART_MAGIC;
        else if (k == 1) v = status;
        else if (k == 2) v = q;
        else if (k == 3) v = L;
        else if (k == 4) v = P;
        else if (k == 5) v = h->n_seeds;
        else if (k == 6) v = (int32_t)(uint32_t)(h->batch_index & 0xFFFFFFFFull);
        else if (k == 7) v = (int32_t)(uint32_t)(h->batch_index >> 32);
        else if (k == 8) v = status == 0 ? (int32_t)ws.tgt->cursor[q] : GS_PART_HEADER_WORDS;
        else if (k == 9) {
          if (status == 0)
            for (int l = 0; l < L; l++) v += h->field[l][q][GS_F_CSR_INDICES].len;
        } else if (k >= 16 && k < 16 + 2 * ns && status == 0) {
          const int j = (k - 16) >> 1;
          const gs_span sp = h->field[j / GS_NUM_FIELDS][q][j % GS_NUM_FIELDS];
          v = ((k - 16) & 1) ? sp.len : sp.off;
        }
        blk[k] = v;
      }
    }
  }
  {
    const int nw = (int)(sizeof(gs_sample_header) / 4);
    const int32_t *src = (const int32_t *)h;
    int32_t *dst = (int32_t *)ws.hdr_host;
    for (int i = threadIdx.x; i < nw; i += GS_TPB) dst[i] = src[i];
  }
}

__global__ void GS_BOUNDS k_gather_ids(FeatSrc fs, int dim, int div,
                                                       const int32_t *ids, int64_t n, float *out) {
  auto src_of_row = [&](int64_t r) -> const float * { return fs.row((int64_t)(__ldg(ids + r) / div), dim); };
  gather_block(dim, n, out, src_of_row, (int64_t)blockIdx.x * GS_NW + warp_id(), (int64_t)gridDim.x * GS_NW,
               lane_id());
}

// ------------------------------------------------------------- k_pack_parts
// Split-parallel hand-over: blockIdx.y = part g.  Every list of part g (all
// layers, the nine fields) is copied into block g of `dst` behind a span table,
// 16-byte vectors, consecutive threads on consecutive addresses (slab spans and
// block offsets are both 16-byte aligned).  Block g is either a piece of a local
// send buffer (the P blocks then travel in one all-to-all) or rank g's receive
// buffer itself, mapped over CUDA IPC: then this kernel's stores ARE the
// exchange.  Streams 2 x 4 bytes per list element.
// TMA: a CTA stages 16 KB pieces of the lists in shared memory (coalesced 16-byte
// loads, four in flight per thread) and ONE thread hands each piece to the bulk-copy
// engine (cp.async.bulk shared -> global, double-buffered).  The NVLink round trip
// of a remote store is then the copy engine's business: no warp waits on it, and a
// grid of one CTA per SM is enough to keep 4.7 MB in flight, so the samples being
// sliced on the other streams keep their SM slots.
#define GS_PK_CHUNK 4096  // int32 words per piece
template <bool TMA>
__global__ void __launch_bounds__(GS_TPB) k_pack_parts(GsWs ws, GsPartDst dst, int64_t block_words) {
  const gs_sample_header *h = ws.hdr;
  __shared__ int32_t s_src[GS_MAX_LAYERS * GS_NUM_FIELDS], s_len[GS_MAX_LAYERS * GS_NUM_FIELDS],
      s_dst[GS_MAX_LAYERS * GS_NUM_FIELDS];
  __shared__ int64_t s_total;
  __shared__ int32_t s_edges;
  const int g = blockIdx.y, L = ws.L, ns = L * GS_NUM_FIELDS;
  int32_t *blk = dst.blk[g];  // may be peer memory: the stores below then travel over NVLink
  const int status = h->status ? h->status : h->late_status;
  if (threadIdx.x == 0) {
    int64_t at = GS_PART_HEADER_WORDS;
    int32_t edges = 0;
    for (int k = 0; k < ns; k++) {
      const gs_span sp = status == 0 ? h->field[k / GS_NUM_FIELDS][g][k % GS_NUM_FIELDS] : gs_span{0, 0};
      s_src[k] = sp.off;
      s_len[k] = sp.len;
      s_dst[k] = (int32_t)min(at, (int64_t)0x7FFFFFF0);
      at += ((int64_t)sp.len + 3) & ~3ll;
      if (k % GS_NUM_FIELDS == GS_F_CSR_INDICES) edges += sp.len;
    }
    s_total = at;
    s_edges = edges;
  }
  __syncthreads();
  const bool fits = s_total <= block_words;
  if (blockIdx.x == 0) {
    for (int k = threadIdx.x; k < GS_PART_HEADER_WORDS; k += GS_TPB) {
      int32_t v = 0;
      if (k == 0) v = (int32_t)GS_PART_MAGIC;
      else if (k == 1) v = status != 0 ? status : (fits ? 0 : (int32_t)GS_ERR_OVERFLOW);
      else if (k == 2) v = g;
      else if (k == 3) v = L;
      else if (k == 4) v = ws.P;
      else if (k == 5) v = h->n_seeds;
      else if (k == 6) v = (int32_t)(uint32_t)(h->batch_index & 0xFFFFFFFFull);
      else if (k == 7) v = (int32_t)(uint32_t)(h->batch_index >> 32);
      else if (k == 8) v = (int32_t)min(s_total, (int64_t)0x7FFFFFFF);
      else if (k == 9) v = s_edges;
      else if (k >= 16 && k < 16 + 2 * ns) {
        const int j = (k - 16) >> 1;
        v = ((k - 16) & 1) ? s_len[j] : s_dst[j];
      }
      blk[k] = v;
    }
  }
  if (!fits || status != 0) return;
  if (TMA) {
    __shared__ __align__(128) int32_t s_buf[TMA ? 2 : 1][TMA ? GS_PK_CHUNK : 4];
    const int G = (int)gridDim.x;
    int it = 0, cbase = 0;  // pieces are dealt round-robin over the part's CTAs
    for (int k = 0; k < ns; k++) {
      const int words = (s_len[k] + 3) & ~3;  // the padding words of a span belong to it in both layouts
      const int nch = (words + GS_PK_CHUNK - 1) / GS_PK_CHUNK;
      for (int j = ((int)blockIdx.x - cbase % G + G) % G; j < nch; j += G, it++) {
        const int w0 = j * GS_PK_CHUNK, nw = min(GS_PK_CHUNK, words - w0);
        int32_t *sb = s_buf[it & 1];
        // the bulk store that read this buffer two pieces ago must be done reading it
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncthreads();
        const int4 *src = reinterpret_cast<const int4 *>(ws.slab + s_src[k] + w0);
        for (int i = threadIdx.x; i < (nw >> 2); i += GS_TPB) reinterpret_cast<int4 *>(sb)[i] = src[i];
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
          const uint32_t sa = (uint32_t)__cvta_generic_to_shared(sb);
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                       :: "l"(blk + s_dst[k] + w0), "r"(sa), "r"(nw * 4) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      cbase += nch;
    }
    // the block is complete only when every bulk store has been performed
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    return;
  }
  const int tid = blockIdx.x * GS_TPB + threadIdx.x, nth = gridDim.x * GS_TPB;
  for (int k = 0; k < ns; k++) {
    const int n4 = (s_len[k] + 3) >> 2;  // the padding words of a span belong to it in both layouts
    const int4 *src = reinterpret_cast<const int4 *>(ws.slab + s_src[k]);
    int4 *out = reinterpret_cast<int4 *>(blk + s_dst[k]);
    // four independent 16-byte loads in flight per thread before the (possibly remote) stores
    int i = tid;
    for (; i + 3 * nth < n4; i += 4 * nth) {
      const int4 a = src[i], b = src[i + nth], c = src[i + 2 * nth], d = src[i + 3 * nth];
      out[i] = a;
      out[i + nth] = b;
      out[i + 2 * nth] = c;
      out[i + 3 * nth] = d;
    }
    for (; i < n4; i += nth) out[i] = src[i];
  }
}

__global__ void k_fill_u32(uint32_t *p, size_t n, uint32_t v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    p[i] = v;
}
__global__ void k_narrow_ids(const int64_t *src, int32_t *dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    dst[i] = (int32_t)src[i];
}

// CSR validation (gs_graph_create): the kernels index table[] / indptr[] with the
// ids they read from `indices`, so a bad file must be rejected here, not found as
// silent corruption later.  err[0]: 1 + first index i with indptr[i] > indptr[i+1]
// (or an end outside [0, M]); err[1]: 1 + first edge index with an id outside [0, N).
__global__ void k_check_indptr(const int64_t *indptr, int64_t n, int64_t m, unsigned long long *err) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t a = indptr[i], b = indptr[i + 1];
    if (a > b || a < 0 || b > m) atomicMin(err + 0, (unsigned long long)i + 1ull);
  }
}
__global__ void k_check_ids(const int64_t *ids, size_t n, int64_t base, int64_t num_nodes, unsigned long long *err) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int64_t v = ids[i];
    if (v < 0 || v >= num_nodes) atomicMin(err + 1, (unsigned long long)(base + (int64_t)i) + 1ull);
  }
}

// Wrapping 64-bit sum of an int64 array (the dataset's checksums, dataset.cpp:27,35): one
// shuffle reduction per warp, one atomic per CTA.
__global__ void __launch_bounds__(256) k_sum_i64(const int64_t *src, size_t n, unsigned long long *out) {
  unsigned long long acc = 0ull;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    acc += (unsigned long long)src[i];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(FULL, acc, d);
  __shared__ unsigned long long s_w[8];
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0ull;
    for (int k = 0; k < 8; k++) t += s_w[k];
    atomicAdd(out, t);
  }
}

// Random-sector yardstick (bench.py, roofline.random_access): every thread keeps
// eight independent 4-byte reads in flight, each from a pseudo-random word of a
// table of `elems` words -- one 32-byte DRAM sector per read once the table is
// larger than L2.  This is the access pattern of the neighbour picks and of the
// dedup-table reads; its rate is the ceiling the sector-granular bound is
// divided by.
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull;
  return x ^ (x >> 33);
}
__global__ void __launch_bounds__(256) k_sector_probe(const uint32_t *__restrict__ table, uint64_t elems, uint64_t n,
                                                      uint32_t seed, uint32_t *sink) {
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (uint64_t)gridDim.x * blockDim.x;
  uint32_t acc = 0u;
  for (uint64_t i = tid; i < n; i += nth * 8ull) {
    uint32_t v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint64_t j = i + (uint64_t)k * nth;
      const uint64_t idx = __umul64hi(mix64(j + ((uint64_t)seed << 32)), elems);
      v[k] = j < n ? __ldg(table + idx) : 0u;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) acc += v[k];
  }
  if (acc == 0x9E3779B9u) sink[0] = acc;  // keeps the loads alive; practically never taken
}

}  // namespace

// ---------------------------------------------------------------- launchers
void gs_kernels_prepare() {}
void gs_launch_rows(const GsWs *wsv, int members, int layer, int grid_rows, cudaStream_t st) {
  const dim3 g(grid_rows, members);
  if (layer == 0) k_rows<true><<<g, GS_TPB, 0, st>>>(wsv, 0);
  else k_rows<false><<<g, GS_TPB, 0, st>>>(wsv, layer);
}
template <bool WIDE>
static void launch_expand_t(const GsWs *wsv, dim3 g, int rng_mode, bool by_rows, int layer, int ge, int gr,
                            cudaStream_t st) {
  if (by_rows) {  // sampled layer: one thread per (row, block of four picks)
    if (rng_mode == (int)GS_RNG_DRAWS) k_expand_rows<(int)GS_RNG_DRAWS, WIDE><<<g, GS_TPB, 0, st>>>(wsv, layer, ge, gr);
    else if (rng_mode == (int)GS_RNG_PHILOX_NOREPLACE) k_expand_rows<(int)GS_RNG_PHILOX_NOREPLACE, WIDE><<<g, GS_TPB, 0, st>>>(wsv, layer, ge, gr);
    else k_expand_rows<(int)GS_RNG_PHILOX, WIDE><<<g, GS_TPB, 0, st>>>(wsv, layer, ge, gr);
    return;
  }
  if (rng_mode == (int)GS_RNG_DRAWS) k_expand<(int)GS_RNG_DRAWS, WIDE><<<g, GS_TPB, 0, st>>>(wsv, layer, ge, gr);
  else if (rng_mode == (int)GS_RNG_PHILOX_NOREPLACE) k_expand<(int)GS_RNG_PHILOX_NOREPLACE, WIDE><<<g, GS_TPB, 0, st>>>(wsv, layer, ge, gr);
  else k_expand<(int)GS_RNG_PHILOX, WIDE><<<g, GS_TPB, 0, st>>>(wsv, layer, ge, gr);
}
void gs_launch_expand(const GsWs *wsv, int members, bool wide, int rng_mode, bool by_rows, int layer, int grid,
                      int grid_relabel, int grid_rowside, cudaStream_t st) {
  // grid_relabel / grid_rowside: CTAs that finish the PREVIOUS layer's lists (0 for layer 0)
  const dim3 g(grid + grid_relabel + grid_rowside, members);
  if (wide) launch_expand_t<true>(wsv, g, rng_mode, by_rows, layer, grid, grid_relabel, st);
  else launch_expand_t<false>(wsv, g, rng_mode, by_rows, layer, grid, grid_relabel, st);
}
void gs_launch_cand(const GsWs *wsv, int members, bool wide, int pass, int layer, int grid, int grid_rowcount,
                    cudaStream_t st) {
  // pass 0: flags, 1: apply.  Both must be given the same `grid`: a CTA's run of tiles is the same in both.
  // grid_rowcount (pass 0): CTAs of the row-side count pass; the row side is later run with as many.
  const dim3 g(grid, members);
  if (pass == 0) {
    const dim3 g0(grid + grid_rowcount, members);
    if (wide) k_cand_flags<true><<<g0, GS_TPB, 0, st>>>(wsv, layer, grid);
    else k_cand_flags<false><<<g0, GS_TPB, 0, st>>>(wsv, layer, grid);
  } else {
    if (wide) k_cand_apply<true><<<g, GS_TPB, 0, st>>>(wsv, layer);
    else k_cand_apply<false><<<g, GS_TPB, 0, st>>>(wsv, layer);
  }
}
void gs_launch_tail(const GsWs *wsv, int members, bool wide, int grid_relabel, int grid_rowside, int grid_gather,
                    cudaStream_t st) {
  const dim3 g(grid_relabel + grid_rowside + grid_gather, members);
  if (wide) k_tail<true><<<g, GS_TPB, 0, st>>>(wsv, grid_relabel, grid_rowside, grid_gather);
  else k_tail<false><<<g, GS_TPB, 0, st>>>(wsv, grid_relabel, grid_rowside, grid_gather);
}
void gs_launch_gather_ids(const float *feat, const int32_t *slot_of, const float *host_rows, int dim,
                          int div, int64_t rows, const int32_t *ids, int64_t n, float *out, int grid,
                          cudaStream_t st) {
  (void)rows;
  k_gather_ids<<<grid, GS_TPB, 0, st>>>(FeatSrc{feat, slot_of, host_rows}, dim, div, ids, n, out);
}
void gs_launch_pack_parts(const GsWs &ws, const GsPartDst &dst, int64_t block_words, int grid_x, bool tma,
                          cudaStream_t st) {
  if (tma) k_pack_parts<true><<<dim3(grid_x, ws.P), GS_TPB, 0, st>>>(ws, dst, block_words);
  else k_pack_parts<false><<<dim3(grid_x, ws.P), GS_TPB, 0, st>>>(ws, dst, block_words);
}
void gs_launch_fill_u32(uint32_t *p, size_t n, uint32_t v, cudaStream_t st) {
  k_fill_u32<<<1184, 256, 0, st>>>(p, n, v);
}
void gs_launch_narrow_ids(const int64_t *src, int32_t *dst, size_t n, cudaStream_t st) {
  k_narrow_ids<<<1184, 256, 0, st>>>(src, dst, n);
}
void gs_launch_check_indptr(const int64_t *indptr, int64_t n, int64_t m, unsigned long long *err, cudaStream_t st) {
  k_check_indptr<<<1184, 256, 0, st>>>(indptr, n, m, err);
}
void gs_launch_check_ids(const int64_t *ids, size_t n, int64_t base, int64_t num_nodes, unsigned long long *err,
                         cudaStream_t st) {
  k_check_ids<<<1184, 256, 0, st>>>(ids, n, base, num_nodes, err);
}
void gs_launch_sum_i64(const int64_t *src, size_t n, unsigned long long *out, cudaStream_t st) {
  k_sum_i64<<<592, 256, 0, st>>>(src, n, out);
}
void gs_launch_sector_probe(const uint32_t *table, uint64_t elems, uint64_t n, uint32_t seed, uint32_t *sink,
                            int grid, cudaStream_t st) {
  k_sector_probe<<<grid, 256, 0, st>>>(table, elems, n, seed, sink);
}
