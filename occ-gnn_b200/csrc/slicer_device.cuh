// slicer_device.cuh -- device helpers shared by the replicated slicer
// (slicer_kernels.cu) and the partitioned slicer (pslicer_kernels.cu).
#pragma once
#include "slicer_internal.cuh"

#define FULL 0xFFFFFFFFu

namespace {

// ------------------------------------------------------------------ helpers
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }
__device__ __forceinline__ uint32_t lt_mask() { return (1u << lane_id()) - 1u; }

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
  const int lane = lane_id();
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t n = __shfl_up_sync(FULL, v, d);
    if (lane >= d) v += n;
  }
  return v;
}
__device__ __forceinline__ unsigned long long warp_incl_scan64(unsigned long long v) {
  const int lane = lane_id();
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    unsigned long long n = __shfl_up_sync(FULL, v, d);
    if (lane >= d) v += n;
  }
  return v;
}

// Exclusive scan of one value per thread over the CTA.  s_w: GS_NW+1 words.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *s_w, uint32_t &total) {
  const int lane = lane_id(), w = warp_id();
  uint32_t inc = warp_incl_scan(v);
  if (lane == 31) s_w[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t x = lane < GS_NW ? s_w[lane] : 0u;
    uint32_t xi = warp_incl_scan(x);
    if (lane < GS_NW) s_w[lane] = xi - x;
    if (lane == GS_NW - 1) s_w[GS_NW] = xi;
  }
  __syncthreads();
  uint32_t res = s_w[w] + inc - v;
  total = s_w[GS_NW];
  __syncthreads();
  return res;
}
__device__ __forceinline__ unsigned long long block_excl_scan64(unsigned long long v,
                                                               unsigned long long *s_w,
                                                               unsigned long long &total) {
  const int lane = lane_id(), w = warp_id();
  unsigned long long inc = warp_incl_scan64(v);
  if (lane == 31) s_w[w] = inc;
  __syncthreads();
  if (w == 0) {
    unsigned long long x = lane < GS_NW ? s_w[lane] : 0ull;
    unsigned long long xi = warp_incl_scan64(x);
    if (lane < GS_NW) s_w[lane] = xi - x;
    if (lane == GS_NW - 1) s_w[GS_NW] = xi;
  }
  __syncthreads();
  unsigned long long res = s_w[w] + inc - v;
  total = s_w[GS_NW];
  __syncthreads();
  return res;
}

// Last-CTA election: every CTA calls this after its tile results are written.
// Returns true in exactly one CTA (the last to arrive), with all other CTAs'
// writes visible to it.  The ticket is re-armed for the next launch.
__device__ __forceinline__ bool last_block(uint32_t *ticket) {
  __shared__ uint32_t s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1) ? 1u : 0u;
    if (s_last) *ticket = 0u;
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0u;
}

// Tiles are dealt to CTAs in contiguous runs, so a CTA can carry running
// offsets across its own tiles and the cross-CTA spine only has gridDim.x
// entries per counter (not one per tile).  Count and apply kernels of a pair
// are launched with the same grid and therefore see the same runs.
__device__ __forceinline__ void cta_tiles(int ntiles, int &t0, int &t1) {
  const int per = (ntiles + (int)gridDim.x - 1) / (int)gridDim.x;
  t0 = min(ntiles, (int)blockIdx.x * per);
  t1 = min(ntiles, t0 + per);
}

// slab bump allocator (single thread).  16-byte aligned spans.  `used` is the
// caller's register copy of gs_sample_header::slab_used: a layout of ~30 spans
// must not pay a global round trip per span.
__device__ __forceinline__ gs_span slab_alloc(int64_t &used, int64_t cap, uint32_t len, bool &ok) {
  gs_span s;
  const int64_t next = used + (((int64_t)len + 3) & ~3ll);
  if (next > cap || next > 0x7FFFFFFFll) {
    ok = false;
    s.off = 0;
    s.len = 0;
    return s;
  }
  s.off = (int32_t)used;
  s.len = (int32_t)len;
  used = next;
  return s;
}

// Philox4x32-10 (Salmon et al. SC'11).  Draw for (seed, batch, layer, row's
// global node id, pick slot): counter-based, so any thread can produce any
// pick and the result does not depend on the schedule.
// One Philox block = four draws: picks 4*blk .. 4*blk+3 of (seed, batch, layer, node).
__device__ __forceinline__ uint4 philox_block(uint32_t k0, uint32_t k1, uint32_t batch, int layer,
                                              int32_t node, int blk) {
  uint32_t c0 = (uint32_t)node, c1 = 0u, c2 = batch, c3 = ((uint32_t)layer << 24) | (uint32_t)blk;
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ uint32_t philox_word(const uint4 &b, int w) {
  return w == 0 ? b.x : (w == 1 ? b.y : (w == 2 ? b.z : b.w));
}
__device__ __forceinline__ uint32_t philox_pick(uint32_t k0, uint32_t k1, uint32_t batch,
                                                int layer, int32_t node, int slot) {
  uint32_t c0 = (uint32_t)node, c1 = 0u, c2 = batch,
           c3 = ((uint32_t)layer << 24) | (uint32_t)(slot >> 2);
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  const int w = slot & 3;
  return w == 0 ? c0 : (w == 1 ? c1 : (w == 2 ? c2 : c3));
}

// GS_RNG_PHILOX_NOREPLACE: pick j of a row is perm(j), perm a keyed bijection of
// [0, deg) -- so the `fan` picks of a row are distinct (a uniform fan-subset,
// what dgl.sampling.sample_neighbors(replace=False) draws at
// python/utils/sampler.py:187) while every position still computes its own pick
// in O(1), without reading the row.  Keys: the Philox block with block id
// GS_NR_KEY_BLOCK of (seed, batch, layer, node).
//   deg <= 16: Fisher-Yates over sixteen 4-bit entries packed in one register,
//              run up to step j (exactly uniform up to 2^-28 draw bias);
//   deg  > 16: unbalanced Feistel network over ceil(log2 deg) bits (12 rounds on
//              5-bit domains, 8 above: pair/triple statistics measured uniform),
//              cycle-walked back into [0, deg) (domain < 2*deg: < 2 walks expected).
#define GS_NR_KEY_BLOCK 0x00FFFFFF
__device__ __forceinline__ uint32_t nr_mix(uint32_t r, uint32_t key) {
  uint32_t f = r * 0x9E3779B1u + key;
  f ^= f >> 16; f *= 0x85EBCA6Bu;
  f ^= f >> 13; f *= 0xC2B2AE35u;
  return f ^ (f >> 16);
}
__device__ __forceinline__ uint32_t nr_pick(const uint4 &key, uint32_t j, uint32_t deg) {
  if (deg <= 1u) return 0u;
  if (deg <= 16u) {
    unsigned long long perm = 0xFEDCBA9876543210ull;
    for (uint32_t i = 0; i <= j; i++) {
      const uint32_t t = i + __umulhi(nr_mix(i, philox_word(key, i & 3)), deg - i);
      const unsigned long long d = ((perm >> (4 * i)) ^ (perm >> (4 * t))) & 15ull;
      perm ^= (d << (4 * i)) ^ (d << (4 * t));
    }
    return (uint32_t)(perm >> (4 * j)) & 15u;
  }
  const int bits = 32 - __clz((int)(deg - 1u));
  const int rounds = bits <= 5 ? 12 : 8;
  uint32_t x = j;
  do {
    int wl = bits >> 1, wr = bits - wl;
    for (int r = 0; r < rounds; r++) {
      const uint32_t left = x >> wr, right = x & ((1u << wr) - 1u);
      const uint32_t rk = philox_word(key, r & 3) + (uint32_t)(r >> 2) * 0x632BE5ABu;
      const uint32_t f = nr_mix(right, rk) & ((1u << wl) - 1u);
      x = (right << wl) | (left ^ f);
      const int t = wl; wl = wr; wr = t;
    }
  } while (x >= deg);
  return x;
}

// Exclusive scan of `nc` counter arrays (cnt[c*stride + b], b < n = number of
// CTAs) by one CTA, in place; totals[c] receives each array's sum.  One warp
// per counter, eight independent coalesced loads in flight per lane, shuffles
// only -- the spine is on every layer's critical path, so no barrier and no
// dependent-load chain inside it.
template <int NCMAX>
__device__ void spine_scan(uint32_t *cnt, int stride, int n, int nc, uint32_t *totals,
                           uint32_t * /*unused*/) {
  const int w = warp_id(), lane = lane_id();
  for (int c = w; c < nc; c += GS_NW) {
    uint32_t *a = cnt + (size_t)c * stride;
    uint32_t carry = 0u;
    for (int base = 0; base < n; base += 32 * 8) {
      uint32_t v[8];
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const int idx = base + k * 32 + lane;
        v[k] = idx < n ? __ldcg(a + idx) : 0u;
      }
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const int idx = base + k * 32 + lane;
        const uint32_t inc = warp_incl_scan(v[k]);
        if (idx < n) a[idx] = carry + inc - v[k];
        carry += __shfl_sync(FULL, inc, 31);
      }
    }
    if (lane == 0) totals[c] = carry;
  }
  __syncthreads();
}

// Per-part counters packed one byte per part (P <= 8) in a 64-bit word: a lane
// that contributes an item of part g adds pk_one(g).  One packed scan / reduce
// over the warp replaces P ballot+popc pairs; a byte holds at most 32 * GS_IPT
// = 128 < 256 items.
__device__ __forceinline__ unsigned long long pk_one(int g) { return 1ull << (8 * g); }
__device__ __forceinline__ uint32_t pk_get(unsigned long long v, int g) { return (uint32_t)(v >> (8 * g)) & 0xFFu; }
__device__ __forceinline__ unsigned long long pk_warp_sum(unsigned long long v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
  return v;
}
__device__ __forceinline__ unsigned long long pk_warp_incl(unsigned long long v) {
  const int lane = lane_id();
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long n = __shfl_up_sync(FULL, v, d);
    if (lane >= d) v += n;
  }
  return v;
}

// One warp copies one feature row; vector width chosen from dim (float4 / float2 / float).
template <typename V>
__device__ __forceinline__ void copy_row(const float *__restrict__ src, float *__restrict__ dst,
                                         int nvec, int lane) {
  const V *s = reinterpret_cast<const V *>(src);
  V *d = reinterpret_cast<V *>(dst);
  int k = lane;
  for (; k + 96 < nvec; k += 128) {  // four independent loads in flight per lane
    const V a = __ldg(s + k), b = __ldg(s + k + 32), c = __ldg(s + k + 64), e = __ldg(s + k + 96);
    d[k] = a; d[k + 32] = b; d[k + 64] = c; d[k + 96] = e;
  }
  for (; k < nvec; k += 32) d[k] = __ldg(s + k);
}
// Feature store with an HBM cache in front of a host-resident table (SURVEY
// 8(f) N2; python/utils/memory_manager.py:40-104): slot_of[row] >= 0 -> the row
// sits in HBM at that slot, else it is read in place from pinned, mapped host
// memory over PCIe by the same warp.  slot_of == NULL: every row is in HBM.
struct FeatSrc {
  const float *rows;       // HBM rows (all of them, or the cached ones)
  const int32_t *slot_of;  // [num_rows] or NULL
  const float *host_rows;  // device alias of the host table (cached stores only)
  __device__ __forceinline__ const float *row(int64_t r, int dim) const {
    if (slot_of) {
      const int32_t s = __ldg(slot_of + r);
      return s >= 0 ? rows + (int64_t)s * dim : host_rows + r * dim;
    }
    return rows + r * dim;
  }
};
__device__ __forceinline__ void gather_row_from(const float *src, int dim, float *out, int64_t dst_row,
                                                int lane) {
  float *dst = out + dst_row * dim;
  if ((dim & 3) == 0) copy_row<float4>(src, dst, dim >> 2, lane);
  else if ((dim & 1) == 0) copy_row<float2>(src, dst, dim >> 1, lane);
  else copy_row<float>(src, dst, dim, lane);
}
// ---- gather of R feature rows into a dense [R][dim] block; src_of_row(r) -> the row's source.
// Vector width from dim (float4 / float2 / float).  Rows longer than one warp pass: one warp
// per row, four loads in flight per lane (copy_row).  Rows that fit one pass (<= 32 vectors,
// e.g. 128 x fp32 = 512 B): four rows per warp at a time, so a lane still has four independent
// loads in flight -- 4.15 -> 6.00 TB/s on 128-wide rows (0.95 of this GPU's plain copy).
// (Tried and dropped: a warp owning a contiguous piece of the OUTPUT so that every store of a
// 2408-byte-row gather is a full aligned sector -- 5.73 vs 5.78 TB/s, no gain: the 602-wide
// case is bound on the read side.)
template <typename V, typename SrcFn>
__device__ __forceinline__ void gather_short_rows(int nv, int64_t R, float *out, SrcFn src_of_row,
                                                  int64_t warp_global, int64_t nwarps_global, int lane) {
  V *o = reinterpret_cast<V *>(out);
  for (int64_t r0 = warp_global * 4; r0 < R; r0 += nwarps_global * 4) {
    V v[4];
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (r0 + u < R && lane < nv) v[u] = __ldg(reinterpret_cast<const V *>(src_of_row(r0 + u)) + lane);
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (r0 + u < R && lane < nv) o[(r0 + u) * nv + lane] = v[u];
  }
}
template <typename V, typename SrcFn>
__device__ __forceinline__ void gather_block_v(int nv, int dim, int64_t R, float *out, SrcFn src_of_row,
                                               int64_t warp_global, int64_t nwarps_global, int lane) {
  if (nv <= 32) {
    gather_short_rows<V>(nv, R, out, src_of_row, warp_global, nwarps_global, lane);
    return;
  }
  for (int64_t r = warp_global; r < R; r += nwarps_global)
    copy_row<V>(src_of_row(r), out + r * dim, nv, lane);
}
template <typename SrcFn>
__device__ __forceinline__ void gather_block(int dim, int64_t R, float *out, SrcFn src_of_row,
                                             int64_t warp_global, int64_t nwarps_global, int lane) {
  if ((dim & 3) == 0) gather_block_v<float4>(dim >> 2, dim, R, out, src_of_row, warp_global, nwarps_global, lane);
  else if ((dim & 1) == 0) gather_block_v<float2>(dim >> 1, dim, R, out, src_of_row, warp_global, nwarps_global, lane);
  else gather_block_v<float>(dim, dim, R, out, src_of_row, warp_global, nwarps_global, lane);
}
__device__ __forceinline__ void gather_row(const float *feat, int dim, int64_t src_row, float *out,
                                           int64_t dst_row, int lane) {
  const float *src = feat + src_row * dim;
  float *dst = out + dst_row * dim;
  if ((dim & 3) == 0) copy_row<float4>(src, dst, dim >> 2, lane);
  else if ((dim & 1) == 0) copy_row<float2>(src, dst, dim >> 1, lane);
  else copy_row<float>(src, dst, dim, lane);
}


}  // namespace
