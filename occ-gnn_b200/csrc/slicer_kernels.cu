// slicer_kernels.cu -- hand-written sm_100a kernels of the minibatch slicer.
//
// One layer of the reference's sequential slice_layer (slicer.cpp:25-64) +
// BiPartite::reorder (bipartite.cpp:3-17) becomes eight data-parallel passes.
// The sequential order the reference's lists are in (first occurrence in a
// scan over frontier rows and, inside a row, over [self, neighbours...]) is
// carried by a 32-bit *candidate position*: row i owns positions
// [off[i], off[i+1]), slot 0 being the self entry.  "First occurrence" = minimum
// position (atomicMin into a direct-address table), "list order" = ascending
// position (prefix scans over the candidate stream), so every output list is
// bit-identical to the sequential algorithm without any sort.
//
// All sizes live in device memory (gs_sample_header); every kernel is a
// fixed-size grid striding over tiles, so a whole minibatch is one static CUDA
// graph with no host round trip.  Integer / byte work, HBM- and latency-bound:
// no tensor cores (nothing here is a contraction).
#include "slicer_device.cuh"

// Resident CTAs per SM the tiled kernels are compiled for (0: ptxas' choice,
// 46-55 registers -> 4-5 CTAs; n: register cap 65536 / (GS_TPB * n)).  Several
// minibatches run concurrently on their own streams, so what counts is how many
// warps of ALL of them fit an SM.  Measured (tools/build_variants.sh): 6 (40
// registers, few spills) = 0 within 0.3 %, 8 (32 registers, spills) -7 %.
#ifndef GS_MINB
#define GS_MINB 0
#endif
#if GS_MINB > 0
#define GS_BOUNDS __launch_bounds__(GS_TPB, GS_MINB)
#else
#define GS_BOUNDS __launch_bounds__(GS_TPB)
#endif

namespace {

// --------------------------------------------------- L2 residency control
// The per-slot dedup tables are the only randomly re-read data (4-5 touches per
// candidate); the CSR gathers and the candidate / slab streams are touched once.
// With createpolicy hints the tables are kept in B200's 126 MB L2 (evict_last)
// while the single-use traffic passes through as evict_first.  ws.l2_hints = 0
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
  __device__ __forceinline__ void min(int32_t v, uint32_t p) const {
    if (!hint) { atomicMin(t + v, p); return; }
    asm volatile("red.global.min.L2::cache_hint.u32 [%0], %1, %2;" :: "l"(t + v), "r"(p), "l"(pol) : "memory");
  }
};
__device__ __forceinline__ int32_t ld_stream_i32(const int32_t *a, bool hint, uint64_t pol) {
  if (!hint) return __ldg(a);
  int32_t r;
  asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(r) : "l"(a), "l"(pol));
  return r;
}

// rowscan is stored part-major ([q][Fcap+1]) so that consecutive rows are contiguous
__device__ __forceinline__ size_t rs_at(const GsWs &ws, int q, int i) { return (size_t)q * ((size_t)ws.Fcap + 1) + (size_t)i; }

// owner = id % P (pyfrontend.cpp:57) without a hardware-unsupported modulo on the
// hot path: a mask when P is a power of two, else multiply-high by ceil(2^32/P)
// (quotient exact or one too large for any 31-bit id) and one correction.
__device__ __forceinline__ int owner_of(const GsWs &ws, int32_t v) {
  if (ws.owner) return (int)__ldg(ws.owner + v);
  if (ws.pmask >= 0) return (int)(v & ws.pmask);
  int32_t r = v - (int32_t)(__umulhi((uint32_t)v, ws.pmagic) * (uint32_t)ws.P);
  return r < 0 ? r + ws.P : r;
}

// Tagged dedup table (ws.tag_shift = S > 0).  A sample draws L + 1 consecutive tag
// indices e (seed claim, then one per layer); the stored tag is ntags - 1 - e, so
// values written later are SMALLER: red.min overwrites anything left by earlier
// layers or samples, and a reader that finds another tag sees "no occurrence".
// Index 0 is never used (its tag is the all-ones fill pattern).
__device__ __forceinline__ uint32_t table_tag(const GsWs &ws, int use) {
  if (!ws.tag_shift) return 0u;
  const uint32_t e = __ldcg(ws.epoch + 1) + (uint32_t)use;
  return ((1u << (32 - ws.tag_shift)) - 1u - e) << ws.tag_shift;
}

// ------------------------------------------------------------------ k_begin
// One CTA.  Resets the sample header, narrows the seed ids (int64, pinned
// host memory read over PCIe) into the slab as frontier 0.
__global__ void __launch_bounds__(1024) k_begin(GsWs ws) {
  gs_sample_header *h = ws.hdr;
  const int n = ws.input->n_seeds;
  __shared__ int s_bad;
  __shared__ uint32_t s_tag;
  if (threadIdx.x == 0) {
    s_bad = 0;
    s_tag = 0u;
    if (ws.tag_shift) {
      const uint32_t ntags = 1u << (32 - ws.tag_shift);
      uint32_t e0 = ws.epoch[0];
      if (e0 + (uint32_t)ws.L + 1u > ntags) e0 = 1u;  // out of tags: the previous k_finish refilled the table
      ws.epoch[1] = e0;
      ws.epoch[0] = e0 + (uint32_t)ws.L + 1u;
      s_tag = (ntags - 1u - e0) << ws.tag_shift;
    }
  }
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
  bool bad = (n < 0 || n > ws.Fcap || (int64_t)n > ws.fslab_cap);
  if (!bad)
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      int64_t v = ws.seeds_in[i];
      if (v < 0 || v >= (int64_t)ws.num_nodes) {
        s_bad = 1;
      } else {
        ws.fslab[i] = (int32_t)v;
        // the table is all-INF between minibatches: a second claim of the same
        // entry is a duplicate seed (k_rows_scan resets the claimed entries)
        if (ws.tag_shift) {
          // every other entry carries a larger tag (or the fill pattern): the min always
          // takes the claim, and an old value with OUR tag is an earlier claim of this sample
          const uint32_t old = atomicMin(ws.table + v, s_tag | (uint32_t)i);
          if (((old ^ s_tag) >> ws.tag_shift) == 0u) s_bad = 1;
        } else if (atomicCAS(ws.table + v, GS_INF, (uint32_t)i) != GS_INF) {
          s_bad = 1;
        }
      }
    }
  __syncthreads();
  if (threadIdx.x == 0) {
    bad = bad || s_bad;
    h->status = bad ? (int32_t)GS_ERR_SEED : 0;
    h->num_layers = ws.L;
    h->num_parts = ws.P;
    h->n_seeds = n;
    h->batch_index = ws.input->batch_index;
    h->draw_cursor = *ws.draw_cursor;
    h->edges = 0;
    h->late_status = 0;
    h->reserved[0] = ws.input->want_host;  // kept on the device: k_finish must not poll host memory from every thread  // late error raised inside a tiled kernel
    const int nn = bad ? 0 : n;
    h->slab_used = 0;
    h->reserved[1] = (nn + 3) & ~3;  // frontier slab in use (int32 elements)
    h->reserved[2] = 1;              // frontier spans index the frontier slab, not the sample slab
    h->frontier_len[0] = nn;
    h->frontier[0].off = 0;
    h->frontier[0].len = nn;
  }
}

// ------------------------------------------------------------- k_rows_count
// Per frontier row: CSR extent and number of candidates c = 1 + (deg < fan or
// fan < 0 ? deg : fan)  (slicer.cpp:6-22).  Tile sums + last-CTA spine give the
// candidate offsets' per-CTA prefixes and C_l.
__global__ void GS_BOUNDS k_rows_count(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ unsigned long long s_w64[GS_NW + 1];
  __shared__ uint32_t s_w[GS_NW + 1];
  const int F = h->frontier_len[layer];
  const int32_t *front = ws.fslab + h->frontier[layer].off;
  const int fan = ws.fanouts[layer];
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  {
    unsigned long long csum = 0ull;
    uint32_t ssum = 0u;
    for (int t = t0; t < t1; t++) {
#pragma unroll
      for (int it = 0; it < GS_IPT; it++) {
        const int i = t * GS_TILE + it * GS_TPB + threadIdx.x;
        if (i < F) {
          const int32_t nd = front[i];
          const int64_t s = __ldg(ws.indptr + nd), e = __ldg(ws.indptr + nd + 1);
          const int32_t deg = (int32_t)(e - s);
          ws.rowstart[i] = s;
          ws.rowdeg[i] = deg;
          const bool samp = fan >= 0 && deg >= fan;
          csum += 1ull + (unsigned long long)(samp ? fan : deg);
          ssum += samp ? 1u : 0u;
        }
      }
    }
    unsigned long long ctot;
    uint32_t stot;
    block_excl_scan64(csum, s_w64, ctot);
    block_excl_scan(ssum, s_w, stot);
    if (threadIdx.x == 0) {
      ws.ftile_sum[blockIdx.x] = ctot;
      ws.ftile_samp[blockIdx.x] = stot;
    }
  }
  if (!last_block(ws.tickets + 0)) return;
  // spine: exclusive prefixes over the CTAs' sums
  const int nb = (int)gridDim.x;
  const int chunk = (nb + GS_TPB - 1) / GS_TPB;
  const int a = min(nb, (int)threadIdx.x * chunk), b = min(nb, a + chunk);
  unsigned long long cs = 0ull;
  uint32_t ss = 0u;
  for (int t = a; t < b; t++) {
    cs += __ldcg(ws.ftile_sum + t);
    ss += __ldcg(ws.ftile_samp + t);
  }
  unsigned long long ctot;
  uint32_t stot;
  unsigned long long cpre = block_excl_scan64(cs, s_w64, ctot);
  uint32_t spre = block_excl_scan(ss, s_w, stot);
  const bool ok = ctot <= (unsigned long long)ws.Ccap;
  if (ok)
    for (int t = a; t < b; t++) {
      ws.ftile_pre[t] = (uint32_t)cpre;
      ws.ftile_pre[ws.ftiles_max + t] = spre;
      cpre += __ldcg(ws.ftile_sum + t);
      spre += __ldcg(ws.ftile_samp + t);
    }
  if (threadIdx.x == 0) {
    if (!ok) {
      h->status = (int32_t)GS_ERR_OVERFLOW;
      h->cand_len[layer] = 0;
    } else {
      h->cand_len[layer] = (int32_t)ctot;
      ws.off[F] = (uint32_t)ctot;
      const uint64_t base = *ws.draw_cursor;
      h->draw_base[layer] = base;
      if (fan > 0) *ws.draw_cursor = base + (uint64_t)fan * (uint64_t)stot;
    }
  }
}

// -------------------------------------------------------------- k_rows_scan
// Candidate offset of every row (k_expand writes the stream, self entries
// included); resets the row node's table entry to "no edge-source
// occurrence yet" (entries touched in earlier layers all belong to frontier
// nodes), records which row each candidate tile starts in, and zeroes the
// frontier-first counters k_cand_flags accumulates into with atomics.
__global__ void GS_BOUNDS k_rows_scan(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const Tbl tbl(ws);
  __shared__ uint32_t s_w[GS_NW + 1];
  const int F = h->frontier_len[layer];
  const int32_t *front = ws.fslab + h->frontier[layer].off;
  const int fan = ws.fanouts[layer];
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  const bool replay = ws.rng_mode == (int)GS_RNG_DRAWS;
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  for (int k = blockIdx.x * GS_TPB + threadIdx.x; k < ws.ctiles_max; k += gridDim.x * GS_TPB)
    ws.ctile_cnt[k] = 0u;
  uint32_t base = ws.ftile_pre[blockIdx.x];  // running over this CTA's run of tiles
  uint32_t sbase = ws.ftile_pre[ws.ftiles_max + blockIdx.x];
  for (int t = t0; t < t1; t++) {
#pragma unroll
    for (int it = 0; it < GS_IPT; it++) {
      const int i = t * GS_TILE + it * GS_TPB + threadIdx.x;
      uint32_t c = 0u, sflag = 0u;
      int32_t nd = 0;
      if (i < F) {
        nd = front[i];
        const int32_t deg = ws.rowdeg[i];
        const bool samp = fan >= 0 && deg >= fan;
        c = 1u + (uint32_t)(samp ? fan : deg);
        sflag = samp ? 1u : 0u;
      }
      uint32_t ctot, stot = 0u;
      const uint32_t ex = block_excl_scan(c, s_w, ctot);
      uint32_t sex = 0u;
      if (replay) sex = block_excl_scan(sflag, s_w, stot);
      if (i < F) {
        const uint32_t o = base + ex;
        ws.off[i] = o;
        if (replay) ws.samp_rank[i] = sbase + sex;
        if (!ws.tag_shift) tbl.st(nd, GS_INF);  // tagged table: entries of earlier layers are stale by their tag
        for (uint32_t tb = (o + GS_TILE - 1) / GS_TILE; (uint64_t)tb * GS_TILE < (uint64_t)o + c; tb++)
          ws.tile_first[tb] = (uint32_t)i;
      }
      base += ctot;
      sbase += stot;
    }
  }
}

// ----------------------------------------------------------------- k_expand
// Candidate-parallel CSR row expansion (slicer.cpp:6-22).  A tile is GS_TILE
// consecutive candidate positions, so load is balanced whatever the degree
// skew; the rows a tile touches are found from tile_first[] and a binary
// search over their offsets staged in shared memory.  Slots >= 1 of a row are
// either all its neighbours in CSR order (coalesced reads) or `fan` picks with
// replacement.  Every real edge source does atomicMin(position) into the
// direct-address table -- the parallel form of DuplicateRemover's first-
// occurrence rule (util/duplicate.cpp:14-26).
// RNG (gs_rng_mode) is a template parameter: the pick path of the other two
// modes costs instructions in a kernel bound by issue.
template <int RNG>
__global__ void GS_BOUNDS k_expand(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const Tbl tbl(ws);
  const uint64_t spol = ws.l2_hints ? l2_policy_stream() : 0ull;
  __shared__ uint32_t s_off[GS_TILE + 2];
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *__restrict__ front = ws.fslab + h->frontier[layer].off;
  const int fan = ws.fanouts[layer];
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const uint32_t batch = (uint32_t)h->batch_index;
  constexpr bool replay = RNG == (int)GS_RNG_DRAWS;
  constexpr bool norepl = RNG == (int)GS_RNG_PHILOX_NOREPLACE;
  const uint64_t dbase = h->draw_base[layer];
  const uint32_t tg = table_tag(ws, 1 + layer);
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int i_lo = (int)ws.tile_first[t];
    const int i_hi = (t + 1 < ntiles) ? (int)ws.tile_first[t + 1] : F - 1;
    const int cnt = i_hi - i_lo + 1;  // <= GS_TILE + 1 rows, each has >= 1 candidate
    for (int k = threadIdx.x; k <= cnt; k += GS_TPB) s_off[k] = ws.off[i_lo + k];
    __syncthreads();
    // A thread owns GS_IPT CONSECUTIVE positions (the kernel is bound by
    // instruction issue): one binary search, then a walk; one Philox block serves
    // up to four picks of a row; the stream goes out as 16-byte / 4-byte vectors.
    const uint32_t p0 = (uint32_t)t * GS_TILE + GS_IPT * threadIdx.x;
    int row[GS_IPT], slot[GS_IPT];
    int32_t nd1[GS_IPT], deg[GS_IPT];
    int64_t rs[GS_IPT];
    int lo = 0;
    if (p0 < C) {
      int hi = cnt - 1;  // largest k with s_off[k] <= p0
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (s_off[mid] <= p0) lo = mid; else hi = mid - 1;
      }
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const uint32_t p = p0 + k;
      if (k > 0 && p < C && p >= s_off[lo + 1]) lo++;  // every row has >= 1 candidate: one step is enough
      row[k] = i_lo + lo;
      slot[k] = p < C ? (int)(p - s_off[lo]) : 0;  // 0 = self entry (or dead position)
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      if (k > 0 && row[k] == row[k - 1]) {
        nd1[k] = nd1[k - 1];
        rs[k] = rs[k - 1];
        deg[k] = deg[k - 1];
      } else {
        nd1[k] = front[row[k]];
        rs[k] = ws.rowstart[row[k]];
        deg[k] = ws.rowdeg[row[k]];
      }
    }
    const int32_t *src[GS_IPT];
    uint4 blk = make_uint4(0u, 0u, 0u, 0u);
    int blk_row = -1, blk_id = -1;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      src[k] = nullptr;
      if (slot[k] > 0) {
        int32_t pick = slot[k] - 1;
        if (fan >= 0 && deg[k] >= fan) {
          uint32_t r;
          if (replay) {
            const uint64_t di = dbase + (uint64_t)fan * ws.samp_rank[row[k]] + (uint32_t)pick;
            r = di < ws.n_draws ? ws.draws[di] : 0u;
          } else {
            const int want = norepl ? GS_NR_KEY_BLOCK : (pick >> 2);
            if (blk_row != row[k] || blk_id != want) {
              blk_row = row[k];
              blk_id = want;
              blk = philox_block(ws.seed_lo, ws.seed_hi, batch, layer, nd1[k], blk_id);
            }
            r = norepl ? nr_pick(blk, (uint32_t)pick, (uint32_t)deg[k]) : philox_word(blk, pick & 3);
          }
          if (norepl) {
            pick = (int32_t)r;  // already a member of [0, deg)
          } else {
            pick = (int32_t)(r % (uint32_t)deg[k]);
          }
        }
        src[k] = ws.indices + rs[k] + pick;
      }
    }
    int32_t v[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++)
      if (src[k]) v[k] = ld_stream_i32(src[k], ws.l2_hints != 0, spol);
    int32_t cvv[GS_IPT];
    uint32_t cf[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      cvv[k] = row[k];  // self entry: the row index
      cf[k] = GS_CF_SELF;
      if (src[k]) {
        cvv[k] = v[k];
        cf[k] = v[k] != nd1[k] ? 0u : GS_CF_LOOP;
        if (cf[k] == 0u) tbl.min(v[k], tg | (p0 + k));
      }
    }
    if (p0 + GS_IPT <= C) {
      *reinterpret_cast<int4 *>(ws.cand + p0) = make_int4(cvv[0], cvv[1], cvv[2], cvv[3]);
      *reinterpret_cast<uchar4 *>(ws.cflag + p0) =
          make_uchar4((unsigned char)cf[0], (unsigned char)cf[1], (unsigned char)cf[2], (unsigned char)cf[3]);
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (p0 + k < C) {
          ws.cand[p0 + k] = cvv[k];
          ws.cflag[p0 + k] = (uint8_t)cf[k];
        }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------- k_cand_flags
// With every minimum in place, classify each candidate.  An edge candidate is
// the first occurrence of its source as an edge source (-> in_nodes of the
// owner's bipartite, bipartite.cpp:4) when the table holds its own position;
// it then also claims the node's slot in the next frontier (slicer.cpp:45-49).
// A frontier row's self entry at s sees the node's minimum edge position a: the
// node's frontier slot is at min(a, s), so when s < a the self entry takes the
// slot and revokes the edge candidate's claim (kill[a] = 1, one less first in
// the CTA that owns position a).  Every other edge candidate replaces its node
// id by the position of the first occurrence: that is all k_cand_apply and
// k_relabel need to relabel it.  Per-CTA counts of the 2P+1 classes, then the
// last CTA scans them, fixes F_{l+1}, |in_nodes_g|, E_g and lays the lists out.
__global__ void GS_BOUNDS k_cand_flags(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const Tbl tbl(ws);
  __shared__ uint32_t s_cnt[GS_NC_MAX];
  __shared__ uint32_t s_tot[GS_NC_MAX];
  __shared__ uint32_t s_w[GS_NW + 1];
  const int P = ws.P, NC = 2 * P + 1;
  const int F = h->frontier_len[layer];
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *front = ws.fslab + h->frontier[layer].off;
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  const int per = (ntiles + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles per CTA (cta_tiles)
  // tile / per by multiply-high: exact while tile * per < 2^32 (ntiles <= 65536)
  const bool per_inv_ok = per > 1 && ntiles <= 65536;
  const uint32_t per_inv = per_inv_ok ? (uint32_t)((0x100000000ull + (uint32_t)per - 1) / (uint32_t)per) : 0u;
  // running class counts of this warp over the CTA's tiles: lane q < P keeps part q
  uint32_t acc_i = 0u, acc_e = 0u, acc_f = 0u;
  const uint32_t tg = table_tag(ws, 1 + layer);
  const int tsh = ws.tag_shift;
  if (threadIdx.x < GS_NC_MAX) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  for (int t = t0; t < t1; t++) {
    // A thread owns GS_IPT consecutive positions: vector loads / stores of the
    // stream, its GS_IPT random table reads in flight together, lane-local
    // packed counts and one warp reduction per tile.
    unsigned long long pk_i = 0ull, pk_e = 0ull;  // this lane's firsts / edges per part
    uint32_t n_ff = 0u;
    uint32_t fl[GS_IPT], te[GS_IPT];
    int32_t ci[GS_IPT], cv[GS_IPT];
    const uint32_t p0 = (uint32_t)t * GS_TILE + GS_IPT * threadIdx.x;
    const bool full = p0 + GS_IPT <= C;
    if (full) {
      const int4 c4 = *reinterpret_cast<const int4 *>(ws.cand + p0);
      const uchar4 f4 = *reinterpret_cast<const uchar4 *>(ws.cflag + p0);
      ci[0] = c4.x; ci[1] = c4.y; ci[2] = c4.z; ci[3] = c4.w;
      fl[0] = f4.x; fl[1] = f4.y; fl[2] = f4.z; fl[3] = f4.w;
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        fl[k] = GS_CF_LOOP;  // dead position
        ci[k] = 0;
        if (p0 + k < C) {
          fl[k] = ws.cflag[p0 + k];
          ci[k] = ws.cand[p0 + k];
        }
      }
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) cv[k] = (fl[k] & GS_CF_SELF) ? front[ci[k]] : ci[k];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      te[k] = 0u;
      if (!(fl[k] & GS_CF_LOOP)) te[k] = tbl.ld(cv[k]);
    }
    if (tsh) {  // this layer's tag -> the position; any other tag -> no occurrence
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) te[k] = ((te[k] ^ tg) >> tsh) == 0u ? (te[k] ^ tg) : GS_INF;
    }
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const uint32_t p = p0 + k;
      const uint32_t f = fl[k];
      if (f & GS_CF_SELF) {
        const uint32_t a = te[k];
        ws.rowa[ci[k]] = a;
        const bool ffirst = a > p;
        if (ffirst && a != GS_INF) {  // revoke the claim of the edge candidate at a
          ws.kill[a] = 1;
          const uint32_t ta = a / GS_TILE;
          const int cta = per == 1 ? (int)ta : (per_inv_ok ? (int)__umulhi(ta, per_inv) : (int)ta / per);
          atomicSub(&ws.ctile_cnt[min(cta, (int)gridDim.x - 1)], 1u);
        }
        n_ff += ffirst ? 1u : 0u;
        fl[k] = GS_CF_SELF | (ffirst ? GS_CF_FFIRST : 0u);
      } else if (!(f & GS_CF_LOOP)) {
        const bool ifirst = te[k] == p;  // also claims the frontier slot, unless revoked
        const int g = owner_of(ws, cv[k]);
        pk_e += pk_one(g);
        if (ifirst) {
          pk_i += pk_one(g);
          n_ff++;
        } else {
          ci[k] = (int32_t)te[k];
        }
        fl[k] = (ifirst ? (GS_CF_FFIRST | GS_CF_IFIRST) : 0u) | ((uint32_t)g << GS_CF_PART_SHIFT);
      }
    }
    if (full) {
      *reinterpret_cast<int4 *>(ws.cand + p0) = make_int4(ci[0], ci[1], ci[2], ci[3]);
      *reinterpret_cast<uchar4 *>(ws.cflag + p0) =
          make_uchar4((unsigned char)fl[0], (unsigned char)fl[1], (unsigned char)fl[2], (unsigned char)fl[3]);
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (p0 + k < C) {
          ws.cand[p0 + k] = ci[k];
          ws.cflag[p0 + k] = (uint8_t)fl[k];
        }
    }
    // one packed reduction per tile instead of 2P+1 ballots per round
    pk_i = pk_warp_sum(pk_i);
    pk_e = pk_warp_sum(pk_e);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) n_ff += __shfl_xor_sync(FULL, n_ff, d);
    // every lane holds the warp totals: lane q keeps the running count of part q
    acc_i += pk_get(pk_i, lane_id() & 7);
    acc_e += pk_get(pk_e, lane_id() & 7);
    acc_f += n_ff;
  }
  if (lane_id() < P) {
    if (acc_i) atomicAdd(&s_cnt[1 + lane_id()], acc_i);
    if (acc_e) atomicAdd(&s_cnt[1 + P + lane_id()], acc_e);
  }
  if (lane_id() == 0 && acc_f) atomicAdd(&s_cnt[0], acc_f);
  __syncthreads();
  if (threadIdx.x == 0) atomicAdd(&ws.ctile_cnt[blockIdx.x], s_cnt[0]);  // zeroed by k_rows_scan; peers subtract
  else if (threadIdx.x < NC) ws.ctile_cnt[(size_t)threadIdx.x * ws.ctiles_max + blockIdx.x] = s_cnt[threadIdx.x];
  if (!last_block(ws.tickets + 1)) return;
  spine_scan<GS_NC_MAX>(ws.ctile_cnt, ws.ctiles_max, (int)gridDim.x, NC, s_tot, s_w);
  if (threadIdx.x == 0) {
    bool ok = true;
    const uint32_t fnext = s_tot[0];
    if (layer + 1 < ws.L && fnext > (uint32_t)ws.Fcap) ok = false;
    int64_t used = h->slab_used, fused = h->reserved[1];
    gs_span fs = slab_alloc(fused, ws.fslab_cap, fnext, ok);
    int64_t edges = 0;
    for (int q = 0; q < P && ok; q++) {
      h->field[layer][q][GS_F_IN_NODES] = slab_alloc(used, ws.slab_cap, s_tot[1 + q], ok);
      h->field[layer][q][GS_F_CSR_INDICES] = slab_alloc(used, ws.slab_cap, s_tot[1 + P + q], ok);
      ws.rowscan[rs_at(ws, q, F)] = s_tot[1 + P + q];
      edges += s_tot[1 + P + q];
    }
    if (ok) {
      h->slab_used = used;
      h->reserved[1] = (int32_t)fused;
      h->frontier[layer + 1] = fs;
      h->frontier_len[layer + 1] = (int32_t)fnext;
      h->edges += edges;
    } else if (h->status == 0) {
      h->status = (int32_t)GS_ERR_OVERFLOW;
    }
  }
}

// ------------------------------------------------------------- k_cand_apply
// Second pass over the candidate stream: turn the class flags into ranks
// (tile prefix from the spine + in-tile ballot ranks, all in position order)
// and scatter:  next frontier, in_nodes_g, the edge sources of part g in CSR
// order (ids for now; k_relabel turns them into in_nodes positions), and for
// every row's self entry the number of part-g edges emitted before the row
// (rowscan), which is both the CSR row pointer and the row's edge count.
// First-occurrence winners record the node's rank in in_nodes_g at their own
// position (irank) -- the relabel map (DuplicateRemover::mask, duplicate.cpp:21)
// indexed by position of first occurrence instead of by node, so that it is a
// C-sized, L2-resident array and this kernel never touches the N-sized table.
__global__ void GS_BOUNDS k_cand_apply(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  // The kernels of this pipeline are bound by instruction issue, not by memory:
  // per-part ranks come from two packed warp scans (one byte per part) instead
  // of 2P ballot/popc pairs per round.
  __shared__ uint32_t s_wc[GS_NW][GS_NC_MAX];  // per-warp class totals of the tile
  __shared__ uint32_t s_base[GS_NW][GS_NC_MAX];  // CTA prefix + earlier tiles + earlier warps of the tile
  __shared__ uint32_t s_run[2][GS_NC_MAX];       // running class counts at the start of a tile (ping-pong)
  __shared__ int32_t s_in_off[GS_MAX_PARTS], s_ix_off[GS_MAX_PARTS];
  const int P = ws.P, NC = 2 * P + 1;
  const uint32_t C = (uint32_t)h->cand_len[layer];
  const int32_t *__restrict__ front = ws.fslab + h->frontier[layer].off;
  int32_t *fnext = ws.fslab + h->frontier[layer + 1].off;
  const int ntiles = (int)((C + GS_TILE - 1) / GS_TILE);
  if (threadIdx.x < P) {
    s_in_off[threadIdx.x] = h->field[layer][threadIdx.x][GS_F_IN_NODES].off;
    s_ix_off[threadIdx.x] = h->field[layer][threadIdx.x][GS_F_CSR_INDICES].off;
  }
  const int lane = lane_id(), w = warp_id();
  const uint32_t ltm = lt_mask();
  // A warp owns GS_IPT*32 consecutive positions of the tile (GS_IPT rounds of
  // 32), so in-warp order is position order and one cross-warp prefix per tile
  // is enough.
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  if (threadIdx.x < NC) s_run[t0 & 1][threadIdx.x] = ws.ctile_cnt[(size_t)threadIdx.x * ws.ctiles_max + blockIdx.x];
  for (int t = t0; t < t1; t++) {
    __syncthreads();
    // A lane owns GS_IPT consecutive positions, a warp 32 * GS_IPT: in-warp order
    // is position order, ONE packed scan per warp and tile gives every lane the
    // per-part counts before its first position; it then ranks its own items.
    const uint32_t p0 = (uint32_t)t * GS_TILE + GS_IPT * threadIdx.x;
    const bool full = p0 + GS_IPT <= C;
    uint32_t f[GS_IPT];
    int32_t ci[GS_IPT], cv[GS_IPT];
    if (full) {
      const int4 c4 = *reinterpret_cast<const int4 *>(ws.cand + p0);
      const uchar4 f4 = *reinterpret_cast<const uchar4 *>(ws.cflag + p0);
      ci[0] = c4.x; ci[1] = c4.y; ci[2] = c4.z; ci[3] = c4.w;
      f[0] = f4.x; f[1] = f4.y; f[2] = f4.z; f[3] = f4.w;
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        f[k] = GS_CF_LOOP;  // dead position: neither edge nor first
        ci[k] = 0;
        if (p0 + k < C) {
          f[k] = ws.cflag[p0 + k];
          ci[k] = ws.cand[p0 + k];
        }
      }
    }
    unsigned long long pi = 0ull, pe = 0ull;
    uint32_t nf = 0u;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      // an edge candidate's claim on the frontier slot may have been revoked by
      // the node's earlier self entry (k_cand_flags); leave kill[] zeroed
      if ((f[k] & (GS_CF_FFIRST | GS_CF_SELF)) == GS_CF_FFIRST && ws.kill[p0 + k]) {
        f[k] &= ~GS_CF_FFIRST;
        ws.kill[p0 + k] = 0;
      }
      cv[k] = (f[k] & GS_CF_SELF) ? front[ci[k]] : ci[k];
      const int g = (int)(f[k] >> GS_CF_PART_SHIFT);
      if (!(f[k] & (GS_CF_SELF | GS_CF_LOOP))) pe += pk_one(g);
      if (f[k] & GS_CF_IFIRST) pi += pk_one(g);
      nf += (f[k] & GS_CF_FFIRST) ? 1u : 0u;
    }
    // exclusive packed prefixes over the warp's lanes; lane 31 publishes the totals
    const unsigned long long inc_i = pk_warp_incl(pi), inc_e = pk_warp_incl(pe);
    const uint32_t inc_f = warp_incl_scan(nf);
    {  // lane = class: unpack this warp's totals (held by lane 31) once, one word per class
      const unsigned long long tot_i = __shfl_sync(FULL, inc_i, 31), tot_e = __shfl_sync(FULL, inc_e, 31);
      const uint32_t tot_f = __shfl_sync(FULL, inc_f, 31);
      if (lane < NC)
        s_wc[w][lane] = lane == 0 ? tot_f : (lane <= P ? pk_get(tot_i, lane - 1) : pk_get(tot_e, lane - 1 - P));
    }
    __syncthreads();
    if (lane < NC) {  // this warp's base = running counts + earlier warps; warp 0 also advances the running counts
      const int cc = lane;
      uint32_t a = s_run[t & 1][cc];
      if (w == 0) {
        uint32_t all = 0u;
#pragma unroll
        for (int k = 0; k < GS_NW; k++) all += s_wc[k][cc];
        s_run[(t + 1) & 1][cc] = a + all;
      } else {
        for (int k = 0; k < w; k++) a += s_wc[k][cc];
      }
      s_base[w][cc] = a;
    }
    __syncwarp();
    uint32_t rf = s_base[w][0] + inc_f - nf;
    unsigned long long ri = inc_i - pi, re = inc_e - pe;  // packed counts of this warp before the lane's item
    uint32_t myrank[GS_IPT];
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int g = (int)(f[k] >> GS_CF_PART_SHIFT);
      myrank[k] = 0u;
      if (f[k] & GS_CF_FFIRST) fnext[rf++] = cv[k];
      if (f[k] & GS_CF_SELF) {
        uint32_t *rsp = ws.rowscan + ci[k];
        for (int q = 0; q < P; q++, rsp += (size_t)ws.Fcap + 1) *rsp = s_base[w][1 + P + q] + pk_get(re, q);
      } else if (!(f[k] & GS_CF_LOOP)) {
        const uint32_t ex = s_base[w][1 + P + g] + pk_get(re, g);
        // An edge whose source occurs here for the first time knows its own
        // rank: it writes the final csr_indices value and records the rank at
        // its position (irank).  Every other edge carries the position m of
        // that first occurrence and leaves ~m (negative) for k_relabel.
        int32_t ixv = ~cv[k];
        if (f[k] & GS_CF_IFIRST) {
          const uint32_t rk = s_base[w][1 + g] + pk_get(ri, g);
          ws.slab[s_in_off[g] + rk] = cv[k];
          myrank[k] = rk;
          ixv = (int32_t)rk;
          ri += pk_one(g);
        }
        ws.slab[s_ix_off[g] + ex] = ixv;
        re += pk_one(g);
      }
    }
    if (full) {
      *reinterpret_cast<uint4 *>(ws.irank + p0) = make_uint4(myrank[0], myrank[1], myrank[2], myrank[3]);
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (p0 + k < C) ws.irank[p0 + k] = myrank[k];
    }
  }
}

// ------------------------------------------------------------ k_front_count
// Row-side classes.  Row i (node nd, owner g0) with has_g = "emitted >= 1 edge
// into part g":  out_g = has_g (bipartite.h:59-62); owned_g = has_g && g == g0
// (bipartite.h:56-58); to_g0 = any has_h, h != g0 (slicer.cpp:41);
// self_g0 = always (slicer.cpp:33-34).  from_g = out_g && g != g0 needs no
// counter: its rank is rank(out_g) - rank(owned_g).
// bit q of a mask -> byte q of a packed counter word
// (multiply replicates the mask byte into all eight bytes, the AND keeps bit q in
// byte q, + 0x7F carries any kept bit into bit 7 of its byte)
__device__ __forceinline__ unsigned long long pk_spread(uint32_t mask, int /*P*/) {
  const unsigned long long v = ((unsigned long long)(mask & 0xFFu) * 0x0101010101010101ull) & 0x8040201008040201ull;
  return ((v + 0x7F7F7F7F7F7F7F7Full) >> 7) & 0x0101010101010101ull;
}

__global__ void GS_BOUNDS k_front_count(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_cnt[GS_NF_MAX];
  __shared__ uint32_t s_tot[GS_NF_MAX];
  __shared__ uint32_t s_w[GS_NW + 1];
  const int P = ws.P, NF = 4 * P;
  const int F = h->frontier_len[layer];
  const int32_t *front = ws.fslab + h->frontier[layer].off;
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  if (threadIdx.x < GS_NF_MAX) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  uint32_t acc_out = 0u, acc_own = 0u, acc_to = 0u, acc_self = 0u;  // lane q < P: this warp's counts of part q
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  for (int t = t0; t < t1; t++) {
    // a thread owns GS_IPT consecutive rows: per part, GS_IPT + 1 consecutive
    // rowscan values give the rows' edge counts; classes are counted in packed
    // words (one byte per part) and reduced once per warp and tile
    const int i0 = t * GS_TILE + GS_IPT * (int)threadIdx.x;
    uint32_t has[GS_IPT] = {0u, 0u, 0u, 0u};
    if (i0 < F) {
      const int nk = min(GS_IPT, F - i0);
      for (int q = 0; q < P; q++) {
        const uint32_t *rs = ws.rowscan + rs_at(ws, q, i0);
        uint32_t prev = rs[0];
#pragma unroll
        for (int k = 0; k < GS_IPT; k++)
          if (k < nk) {
            const uint32_t nx = rs[k + 1];
            if (nx != prev) has[k] |= 1u << q;
            prev = nx;
          }
      }
    }
    unsigned long long p_out = 0ull, p_own = 0ull, p_to = 0ull, p_self = 0ull;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++)
      if (i0 + k < F) {
        const int g0 = owner_of(ws, front[i0 + k]);
        p_out += pk_spread(has[k], P);
        p_self += pk_one(g0);
        if ((has[k] >> g0) & 1u) p_own += pk_one(g0);
        if (has[k] & ~(1u << g0)) p_to += pk_one(g0);
      }
    if (i0 + GS_IPT <= F) {
      *reinterpret_cast<uchar4 *>(ws.rowhas + i0) =
          make_uchar4((unsigned char)has[0], (unsigned char)has[1], (unsigned char)has[2], (unsigned char)has[3]);
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (i0 + k < F) ws.rowhas[i0 + k] = (uint8_t)has[k];
    }
    p_out = pk_warp_sum(p_out);
    p_own = pk_warp_sum(p_own);
    p_to = pk_warp_sum(p_to);
    p_self = pk_warp_sum(p_self);
    // every lane holds the warp totals: lane q keeps the running counts of part q
    const int lq = lane_id() & 7;
    acc_out += pk_get(p_out, lq);
    acc_own += pk_get(p_own, lq);
    acc_to += pk_get(p_to, lq);
    acc_self += pk_get(p_self, lq);
  }
  if (lane_id() < P) {
    const int q = lane_id();
    if (acc_out) atomicAdd(&s_cnt[q], acc_out);
    if (acc_own) atomicAdd(&s_cnt[P + q], acc_own);
    if (acc_to) atomicAdd(&s_cnt[2 * P + q], acc_to);
    if (acc_self) atomicAdd(&s_cnt[3 * P + q], acc_self);
  }
  __syncthreads();
  if (threadIdx.x < NF) ws.ftile_cnt[(size_t)threadIdx.x * ws.ftiles_max + blockIdx.x] = s_cnt[threadIdx.x];
  if (!last_block(ws.tickets + 2)) return;
  spine_scan<GS_NF_MAX>(ws.ftile_cnt, ws.ftiles_max, (int)gridDim.x, NF, s_tot, s_w);
  if (threadIdx.x == 0) {
    bool ok = true;
    int64_t used = h->slab_used;
    for (int q = 0; q < P && ok; q++) {
      const uint32_t n_out = s_tot[q], n_own = s_tot[P + q], n_to = s_tot[2 * P + q],
                     n_self = s_tot[3 * P + q];
      gs_span *fl = h->field[layer][q];
      const gs_span ip = slab_alloc(used, ws.slab_cap, n_out + 1u, ok);
      fl[GS_F_OUT_NODES] = slab_alloc(used, ws.slab_cap, n_out, ok);
      fl[GS_F_CSR_INDPTR] = ip;
      fl[GS_F_OWNED_OUT] = slab_alloc(used, ws.slab_cap, n_own, ok);
      fl[GS_F_FROM_IDS] = slab_alloc(used, ws.slab_cap, n_out - n_own, ok);
      fl[GS_F_TO_IDS] = slab_alloc(used, ws.slab_cap, n_to, ok);
      fl[GS_F_SELF_IN] = slab_alloc(used, ws.slab_cap, n_self, ok);
      fl[GS_F_SELF_OUT] = slab_alloc(used, ws.slab_cap, n_self, ok);
      // row pointer sentinel = E_g (left in rowscan[F] by k_cand_flags' spine)
      if (ok) ws.slab[ip.off + n_out] = (int32_t)ws.rowscan[rs_at(ws, q, F)];
    }
    if (ok) h->slab_used = used;
    else h->status = (int32_t)GS_ERR_OVERFLOW;
  }
}

// ------------------------------------------------------------ k_front_apply
__global__ void GS_BOUNDS k_front_apply(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  __shared__ uint32_t s_wc[GS_NW][GS_NF_MAX];  // per-warp class totals of the tile (kind * P + part)
  __shared__ uint32_t s_base[GS_NW][GS_NF_MAX];
  __shared__ uint32_t s_run[2][GS_NF_MAX];
  __shared__ int32_t s_off[GS_MAX_PARTS][GS_NUM_FIELDS];
  const int P = ws.P, NF = 4 * P;
  const int F = h->frontier_len[layer];
  const int32_t *__restrict__ front = ws.fslab + h->frontier[layer].off;
  const int ntiles = (F + GS_TILE - 1) / GS_TILE;
  if (threadIdx.x < P * GS_NUM_FIELDS)
    s_off[threadIdx.x / GS_NUM_FIELDS][threadIdx.x % GS_NUM_FIELDS] =
        h->field[layer][threadIdx.x / GS_NUM_FIELDS][threadIdx.x % GS_NUM_FIELDS].off;
  const int lane = lane_id(), w = warp_id();
  int t0, t1;
  cta_tiles(ntiles, t0, t1);
  if (threadIdx.x < NF) s_run[t0 & 1][threadIdx.x] = ws.ftile_cnt[(size_t)threadIdx.x * ws.ftiles_max + blockIdx.x];
  for (int t = t0; t < t1; t++) {
    __syncthreads();
    // a lane owns GS_IPT consecutive rows (row order = list order); four packed
    // scans per warp and tile give it the per-part ranks before its first row
    const int i0 = t * GS_TILE + GS_IPT * (int)threadIdx.x;
    uint32_t has[GS_IPT] = {0u, 0u, 0u, 0u};
    int g0[GS_IPT];
    int32_t nd[GS_IPT] = {0, 0, 0, 0};
    if (i0 + GS_IPT <= F) {
      const uchar4 h4 = *reinterpret_cast<const uchar4 *>(ws.rowhas + i0);
      const int4 n4 = *reinterpret_cast<const int4 *>(front + i0);
      has[0] = h4.x; has[1] = h4.y; has[2] = h4.z; has[3] = h4.w;
      nd[0] = n4.x; nd[1] = n4.y; nd[2] = n4.z; nd[3] = n4.w;
    } else {
#pragma unroll
      for (int k = 0; k < GS_IPT; k++)
        if (i0 + k < F) {
          has[k] = ws.rowhas[i0 + k];
          nd[k] = front[i0 + k];
        }
    }
    unsigned long long p_out = 0ull, p_own = 0ull, p_to = 0ull, p_self = 0ull;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      g0[k] = -1;
      if (i0 + k < F) {
        g0[k] = owner_of(ws, nd[k]);
        p_out += pk_spread(has[k], P);
        p_self += pk_one(g0[k]);
        if ((has[k] >> g0[k]) & 1u) p_own += pk_one(g0[k]);
        if (has[k] & ~(1u << g0[k])) p_to += pk_one(g0[k]);
      }
    }
    const unsigned long long i_out = pk_warp_incl(p_out), i_own = pk_warp_incl(p_own),
                             i_to = pk_warp_incl(p_to), i_self = pk_warp_incl(p_self);
    {  // lanes 0..7 unpack the part counts of the four kinds from lane 31's totals: one word per class
      const unsigned long long t_out = __shfl_sync(FULL, i_out, 31), t_own = __shfl_sync(FULL, i_own, 31),
                               t_to = __shfl_sync(FULL, i_to, 31), t_self = __shfl_sync(FULL, i_self, 31);
      if (lane < P) {
        s_wc[w][lane] = pk_get(t_out, lane);
        s_wc[w][P + lane] = pk_get(t_own, lane);
        s_wc[w][2 * P + lane] = pk_get(t_to, lane);
        s_wc[w][3 * P + lane] = pk_get(t_self, lane);
      }
    }
    __syncthreads();
    if (lane < NF) {  // lane = class (kind * P + part): this warp's base, warp 0 also the next tile's start
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
    unsigned long long r_out = i_out - p_out, r_own = i_own - p_own, r_to = i_to - p_to, r_self = i_self - p_self;
#pragma unroll
    for (int k = 0; k < GS_IPT; k++) {
      const int i = i0 + k;
      if (i >= F) break;
      const int gq = g0[k];
      int32_t op_own = -1;  // position of nd in its own part's out_nodes, or -1
      for (int q = 0; q < P; q++)
        if ((has[k] >> q) & 1u) {
          const uint32_t op = s_base[w][q] + pk_get(r_out, q);
          const uint32_t ow = s_base[w][P + q] + pk_get(r_own, q);
          ws.slab[s_off[q][GS_F_OUT_NODES] + op] = nd[k];
          ws.slab[s_off[q][GS_F_CSR_INDPTR] + op] = (int32_t)ws.rowscan[rs_at(ws, q, i)];
          if (q == gq) {
            ws.slab[s_off[q][GS_F_OWNED_OUT] + ow] = (int32_t)op;
            op_own = (int32_t)op;
          } else {
            ws.slab[s_off[q][GS_F_FROM_IDS] + (op - ow)] = (int32_t)op;
          }
        }
      if (has[k] & ~(1u << gq)) {
        ws.slab[s_off[gq][GS_F_TO_IDS] + s_base[w][2 * P + gq] + pk_get(r_to, gq)] = op_own;
        r_to += pk_one(gq);
      }
      const uint32_t p_self_rank = s_base[w][3 * P + gq] + pk_get(r_self, gq);
      const uint32_t a = ws.rowa[i];  // position of the node's first edge occurrence, or INF: never an edge source
      ws.slab[s_off[gq][GS_F_SELF_IN] + p_self_rank] = a == GS_INF ? -1 : (int32_t)__ldcg(ws.irank + a);
      ws.slab[s_off[gq][GS_F_SELF_OUT] + p_self_rank] = op_own;
      r_out += pk_spread(has[k], P);
      r_self += pk_one(gq);
      if ((has[k] >> gq) & 1u) r_own += pk_one(gq);
    }
  }
}

// ---------------------------------------------------------------- k_relabel
// csr_indices: edge-source ids -> positions in in_nodes_g through the table
// (DuplicateRemover::replace, duplicate.cpp:35-39 -- what bipartite.cpp:5
// would do if `indices` were ever filled).
__global__ void GS_BOUNDS k_relabel(GsWs ws, int layer) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
  const int P = ws.P;
  for (int q = 0; q < P; q++) {
    const gs_span sp = h->field[layer][q][GS_F_CSR_INDICES];
    int32_t *ix = ws.slab + sp.off;
    const int stride = gridDim.x * GS_TPB;
    for (int e0 = blockIdx.x * GS_TPB + threadIdx.x; e0 < sp.len; e0 += GS_IPT * stride) {
      int32_t v[GS_IPT];
      uint32_t r[GS_IPT];
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        const int e = e0 + k * stride;
        v[k] = e < sp.len ? ix[e] : 0;  // >= 0: already final (written by the first occurrence)
      }
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) r[k] = v[k] < 0 ? __ldcg(ws.irank + (uint32_t)~v[k]) : 0u;
#pragma unroll
      for (int k = 0; k < GS_IPT; k++) {
        const int e = e0 + k * stride;
        if (v[k] < 0) ix[e] = (int32_t)r[k];
      }
    }
  }
}

// ----------------------------------------------------------------- k_finish
// Reset of the table (DuplicateRemover::clear, duplicate.cpp:28-33: only touched
// entries -- every node touched in any layer is in the last frontier; dense
// fill instead when the touched set covers most of the table's sectors),
// then the sample header and, when asked, the used part of the slab are
// mirrored to pinned host memory with 16-byte stores.
__global__ void GS_BOUNDS k_finish(GsWs ws) {
  gs_sample_header *h = ws.hdr;
  const Tbl tbl(ws);
  const int tid = blockIdx.x * GS_TPB + threadIdx.x, nth = gridDim.x * GS_TPB;
  if (h->status == 0 && ws.tag_shift) {
    // tagged table: nothing to reset, unless the NEXT sample would run out of tags
    // (k_begin applies the same test and restarts the tag sequence)
    if (__ldcg(ws.epoch) + (uint32_t)ws.L + 1u > (1u << (32 - ws.tag_shift))) {
      uint4 *t4 = (uint4 *)ws.table;
      const int64_t n4 = (int64_t)ws.num_nodes >> 2;
      const uint4 inf = make_uint4(GS_INF, GS_INF, GS_INF, GS_INF);
      for (int64_t i = tid; i < n4; i += nth) t4[i] = inf;
      if (tid < (ws.num_nodes & 3)) ws.table[(n4 << 2) + tid] = GS_INF;
    }
  } else if (h->status == 0) {
    const int L = ws.L;
    const int32_t *fl = ws.fslab + h->frontier[L].off;
    const int FL = h->frontier_len[L];
    if ((int64_t)FL * 8 > (int64_t)ws.num_nodes) {
      // most 32-byte sectors of the table hold a touched entry: a streaming
      // fill of the whole table moves fewer sectors than FL scattered stores
      uint4 *t4 = (uint4 *)ws.table;
      const int64_t n4 = (int64_t)ws.num_nodes >> 2;
      const uint4 inf = make_uint4(GS_INF, GS_INF, GS_INF, GS_INF);
      for (int64_t i = tid; i < n4; i += nth) t4[i] = inf;
      if (tid < (ws.num_nodes & 3)) ws.table[(n4 << 2) + tid] = GS_INF;
    } else {
      for (int i = tid; i < FL; i += nth) tbl.st(fl[i], GS_INF);
    }
  }
  {
    const int nw = (int)(sizeof(gs_sample_header) / 4);
    const int32_t *src = (const int32_t *)h;
    int32_t *dst = (int32_t *)ws.hdr_host;
    for (int i = tid; i < nw; i += nth) dst[i] = src[i];
  }
  if (h->reserved[0] && ws.slab_host && h->status == 0) {
    const int64_t n4 = h->slab_used >> 2;  // spans are 4-element aligned
    const int4 *src = (const int4 *)ws.slab;
    int4 *dst = (int4 *)ws.slab_host;
    for (int64_t i = tid; i < n4; i += nth) dst[i] = src[i];
  }
}

// ------------------------------------------------------------ feature gather
// One warp per row: out[r, :] = rows[id / div, :].  602 fp32 = 2408 B rows are
// 8-byte but not 16-byte multiples, so the vector width is chosen from dim
// (float4 / float2 / float); rows are contiguous, so unlike the id work above this
// kernel streams whole lines and is bounded by plain HBM bandwidth.
// Fused into every sample's graph: the rows of the last layer's in_nodes_g, part
// after part -- the block of input features GPU g reads for this minibatch
// (python/data/bipartite.py:61-67 gather, memory_manager.py:75-104 refresh).
// With a cached store (FeatSrc) rows outside the HBM cache are read in place from
// host memory by the same warp, so there is no per-batch refresh step and no
// global_to_local table to patch and undo (memory_manager.py:77-104).
__global__ void GS_BOUNDS k_gather_sample(GsWs ws) {
  gs_sample_header *h = ws.hdr;
  if (h->status != 0) return;
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
    if (blockIdx.x == 0) {
      const bool fits = (int64_t)acc <= ws.feat_cap_rows;
      h->feature_rows = fits ? acc : 0;
      for (int q = 0; q <= P; q++) h->feature_part_off[q] = q <= P ? s_start[q] : 0;
      if (!fits) h->late_status = (int32_t)GS_ERR_OVERFLOW;  // folded into status by k_finish
    }
  }
  __syncthreads();
  const int R = s_start[P];
  if ((int64_t)R > ws.feat_cap_rows) return;
  const FeatSrc fs{ws.feat, ws.feat_slot, ws.feat_host};
  const int dim = ws.feat_dim, div = ws.feat_div;
  const int32_t *slab = ws.slab;
  auto src_of_row = [&](int64_t r) -> const float * {
    int q = 0;
    while (q + 1 < P && r >= s_start[q + 1]) q++;
    const int32_t id = slab[s_off[q] + ((int)r - s_start[q])];
    return fs.row((int64_t)(id / div), dim);
  };
  gather_block(dim, (int64_t)R, ws.feat_out, src_of_row, (int64_t)blockIdx.x * GS_NW + warp_id(),
               (int64_t)gridDim.x * GS_NW, lane_id());
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

}  // namespace

// ---------------------------------------------------------------- launchers
void gs_launch_begin(const GsWs &ws, cudaStream_t st) { k_begin<<<1, 1024, 0, st>>>(ws); }
void gs_launch_rows_count(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_rows_count<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_rows_scan(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_rows_scan<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_expand(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  if (ws.rng_mode == (int)GS_RNG_DRAWS) k_expand<(int)GS_RNG_DRAWS><<<grid, GS_TPB, 0, st>>>(ws, layer);
  else if (ws.rng_mode == (int)GS_RNG_PHILOX_NOREPLACE) k_expand<(int)GS_RNG_PHILOX_NOREPLACE><<<grid, GS_TPB, 0, st>>>(ws, layer);
  else k_expand<(int)GS_RNG_PHILOX><<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_cand_flags(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_cand_flags<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_cand_apply(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_cand_apply<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_front_count(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_front_count<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_front_apply(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_front_apply<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_relabel(const GsWs &ws, int layer, int grid, cudaStream_t st) {
  k_relabel<<<grid, GS_TPB, 0, st>>>(ws, layer);
}
void gs_launch_finish(const GsWs &ws, int grid, cudaStream_t st) {
  k_finish<<<grid, GS_TPB, 0, st>>>(ws);
}
void gs_launch_gather_sample(const GsWs &ws, int grid, cudaStream_t st) {
  k_gather_sample<<<grid, GS_TPB, 0, st>>>(ws);
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
