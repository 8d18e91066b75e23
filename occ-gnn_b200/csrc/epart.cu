// epart.cu -- gs_epart_*: the edge partitioning of the Python sibling's sampler
// (Sampler.edge_partitioning, python/utils/sampler.py:64-134) as CUDA kernels:
//   every layer but the last: an edge goes to the GPU that owns its SOURCE
//     (sampler.py:74-76), owned_nds[g] = unique destinations owned by g (:80-81),
//     shuffle_nds[g][r] = unique destinations owned by r != g among g's edges (:83-87);
//   last layer: an edge whose source is resident in the cache of the destination's
//     owner stays with that owner ("natural", :93-95,105); the others ("extra") go
//     to the owner of the source (:106); part g = its natural edges followed by its
//     extra edges, both in edge order (:107-110); owned_nds as above over all
//     destinations (:112-113); shuffle_nds only if g has extra edges, over those (:114-119).
// torch.unique returns SORTED ids, torch.where keeps edge order: the lists below are
// bit-identical to the reference's.  Integer work: stable partition by class (count,
// spine, scatter) for the edges, mark + ordered compaction over the node range for
// the unique lists.  B200 (sm_100a) only; no CPU fallback.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "slicer_internal.cuh"

namespace {

constexpr int EP_TPB = 256;
constexpr int EP_NW = EP_TPB / 32;
constexpr int EP_MAXC = 2 * GS_MAX_PARTS;                        // edge classes: natural_g, extra_g
constexpr int EP_MAXL = GS_MAX_PARTS + GS_MAX_PARTS * GS_MAX_PARTS;  // node lists: owned_o, shuffle_(g -> o)
#define EP_FULL 0xFFFFFFFFu

struct EpArgs {
  const int32_t *src, *dst;   // [n] edges (device, narrowed)
  int64_t n;
  const uint8_t *owner;       // [N]
  const uint8_t *resident;    // [N][P] or nullptr
  uint32_t *marks;            // [N] bit 0: is a destination; bit 1 + g: destination of an edge part g sends away
  int32_t N, P, last;
  int32_t grid_e, grid_n;
  uint32_t *ecnt;             // [2P][grid_e] per-CTA class counts -> exclusive prefixes (+ class base)
  uint32_t *ncnt;             // [P + P*P][grid_n]
  int64_t *tot;               // [0..2P): class totals; [32..32+P+P*P): list totals; [120]: first bad edge + 1
  int32_t *osrc, *odst;       // [n] partitioned edges
  int32_t *onodes;            // unique lists, concatenated in list order
};

__device__ __forceinline__ int ep_chunk(int64_t n, int grid, int bx, int64_t &lo, int64_t &hi) {
  const int64_t per = ((n + grid - 1) / grid + EP_TPB - 1) / EP_TPB * EP_TPB;
  lo = (int64_t)bx * per < n ? (int64_t)bx * per : n;
  hi = lo + per < n ? lo + per : n;
  return (int)((hi - lo + EP_TPB - 1) / EP_TPB);
}

// class of an edge: 2 g (+ 1 if extra); -1 = out-of-range id
__device__ __forceinline__ int ep_class(const EpArgs &a, int32_t s, int32_t d, int &gs, int &gd) {
  if ((uint32_t)s >= (uint32_t)a.N || (uint32_t)d >= (uint32_t)a.N) return -1;
  gs = a.owner[s];
  gd = a.owner[d];
  if (!a.last) return 2 * gs;                                        // sampler.py:76
  const bool natural = a.resident[(size_t)s * a.P + gd] != 0;       // sampler.py:93
  return natural ? 2 * gd : 2 * gs + 1;                              // :105 / :106
}

// pass 1: class counts per CTA; destination marks
__global__ void __launch_bounds__(EP_TPB) k_ep_count(EpArgs a) {
  __shared__ uint32_t s_cnt[EP_MAXC];
  if (threadIdx.x < EP_MAXC) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  int64_t lo, hi;
  ep_chunk(a.n, a.grid_e, blockIdx.x, lo, hi);
  for (int64_t e = lo + threadIdx.x; e < hi; e += EP_TPB) {
    const int32_t s = a.src[e], d = a.dst[e];
    int gs = 0, gd = 0;
    const int c = ep_class(a, s, d, gs, gd);
    if (c < 0) {
      atomicMin((unsigned long long *)&a.tot[120], (unsigned long long)e + 1ull);
      continue;
    }
    atomicAdd(&s_cnt[c], 1u);
    uint32_t m = 1u;                                    // a destination: owned_nds of its owner (:80 / :112)
    const int g = c >> 1;                               // the part that holds the edge
    if (gd != g && (!a.last || (c & 1))) m |= 2u << g;  // :83-87 all of g's edges / :114-119 its extra edges
    if ((a.marks[d] & m) != m) atomicOr(&a.marks[d], m);
  }
  __syncthreads();
  if (threadIdx.x < 2 * a.P) a.ecnt[(size_t)threadIdx.x * a.grid_e + blockIdx.x] = s_cnt[threadIdx.x];
}

// one CTA: exclusive scan of nc counter rows over nb CTAs; row c then starts at the sum of the
// rows before it (lists / classes are stored one after the other)
__global__ void __launch_bounds__(EP_TPB) k_ep_spine(uint32_t *cnt, int nb, int nc, int64_t *tot) {
  __shared__ uint32_t s_tot[EP_MAXL], s_base[EP_MAXL];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int c = w; c < nc; c += EP_NW) {
    uint32_t *row = cnt + (size_t)c * nb;
    uint32_t carry = 0u;
    for (int b0 = 0; b0 < nb; b0 += 32) {
      const int b = b0 + lane;
      const uint32_t v = b < nb ? row[b] : 0u;
      uint32_t inc = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t x = __shfl_up_sync(EP_FULL, inc, d);
        if (lane >= d) inc += x;
      }
      if (b < nb) row[b] = carry + inc - v;
      carry += __shfl_sync(EP_FULL, inc, 31);
    }
    if (lane == 0) s_tot[c] = carry;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0u;
    for (int c = 0; c < nc; c++) {
      s_base[c] = run;
      run += s_tot[c];
      tot[c] = (int64_t)s_tot[c];
    }
  }
  __syncthreads();
  for (int c = w; c < nc; c += EP_NW)
    for (int b = lane; b < nb; b += 32) cnt[(size_t)c * nb + b] += s_base[c];
}

// pass 2: stable scatter of the edges by class (edge order kept inside a class)
__global__ void __launch_bounds__(EP_TPB) k_ep_scatter(EpArgs a) {
  __shared__ uint32_t s_run[EP_MAXC];          // next free place of every class for this CTA
  __shared__ uint32_t s_wc[EP_NW][EP_MAXC];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, NC = 2 * a.P;
  if (threadIdx.x < NC) s_run[threadIdx.x] = a.ecnt[(size_t)threadIdx.x * a.grid_e + blockIdx.x];
  int64_t lo, hi;
  const int ntiles = ep_chunk(a.n, a.grid_e, blockIdx.x, lo, hi);
  for (int t = 0; t < ntiles; t++) {
    const int64_t e = lo + (int64_t)t * EP_TPB + threadIdx.x;
    int32_t s = 0, d = 0;
    int c = -1, gs = 0, gd = 0;
    if (e < hi) {
      s = a.src[e];
      d = a.dst[e];
      c = ep_class(a, s, d, gs, gd);
    }
    const uint32_t same = __match_any_sync(EP_FULL, c);  // lanes of the same class
    const uint32_t rank = __popc(same & ((1u << lane) - 1u));
    if (lane < NC) s_wc[w][lane] = 0u;
    __syncwarp();
    if (c >= 0 && rank == 0u) s_wc[w][c] = __popc(same);
    __syncthreads();
    uint32_t before = 0u;
    if (c >= 0) {
      before = s_run[c];
      for (int k = 0; k < w; k++) before += s_wc[k][c];
    }
    __syncthreads();
    if (threadIdx.x < NC) {
      uint32_t add = 0u;
#pragma unroll
      for (int k = 0; k < EP_NW; k++) add += s_wc[k][threadIdx.x];
      s_run[threadIdx.x] += add;
    }
    if (c >= 0) {
      a.osrc[before + rank] = s;
      a.odst[before + rank] = d;
    }
    __syncthreads();
  }
}

// list of a marked node: owned_o -> o; shuffle (g -> o) -> P + g * P + o
// pass 3: list counts per CTA over the node range
__global__ void __launch_bounds__(EP_TPB) k_ep_ncount(EpArgs a) {
  __shared__ uint32_t s_cnt[EP_MAXL];
  const int NL = a.P + a.P * a.P;
  for (int i = threadIdx.x; i < NL; i += EP_TPB) s_cnt[i] = 0u;
  __syncthreads();
  int64_t lo, hi;
  ep_chunk(a.N, a.grid_n, blockIdx.x, lo, hi);
  for (int64_t v = lo + threadIdx.x; v < hi; v += EP_TPB) {
    const uint32_t m = a.marks[v];
    if (!m) continue;
    const int o = a.owner[v];
    atomicAdd(&s_cnt[o], 1u);
    for (uint32_t b = m >> 1; b; b &= b - 1u) atomicAdd(&s_cnt[a.P + (__ffs((int)b) - 1) * a.P + o], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NL; i += EP_TPB) a.ncnt[(size_t)i * a.grid_n + blockIdx.x] = s_cnt[i];
}

// pass 4: the lists, ascending ids (= torch.unique's order); the marks are cleared on the way
__global__ void __launch_bounds__(EP_TPB) k_ep_nwrite(EpArgs a) {
  __shared__ uint32_t s_run[EP_MAXL];
  __shared__ uint32_t s_wc[EP_NW][EP_MAXL];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, P = a.P, NL = P + P * P;
  for (int i = threadIdx.x; i < NL; i += EP_TPB) s_run[i] = a.ncnt[(size_t)i * a.grid_n + blockIdx.x];
  int64_t lo, hi;
  const int ntiles = ep_chunk(a.N, a.grid_n, blockIdx.x, lo, hi);
  for (int t = 0; t < ntiles; t++) {
    const int64_t v = lo + (int64_t)t * EP_TPB + threadIdx.x;
    uint32_t m = 0u;
    int o = -1;
    if (v < hi) {
      m = a.marks[v];
      if (m) {
        a.marks[v] = 0u;
        o = a.owner[v];
      }
    }
    for (int i = lane; i < NL; i += 32) s_wc[w][i] = 0u;
    __syncwarp();
    const uint32_t same = __match_any_sync(EP_FULL, o);  // marked lanes with the same owner
    uint32_t rk[1 + GS_MAX_PARTS];
#pragma unroll
    for (int b = 0; b <= GS_MAX_PARTS; b++) {
      rk[b] = 0u;
      if (b > P) continue;
      const uint32_t grp = __ballot_sync(EP_FULL, (m >> b) & 1u) & same;
      if ((m >> b) & 1u) {
        rk[b] = __popc(grp & ((1u << lane) - 1u));
        if (rk[b] == 0u) s_wc[w][b == 0 ? o : P + (b - 1) * P + o] = __popc(grp);
      }
    }
    __syncthreads();
#pragma unroll
    for (int b = 0; b <= GS_MAX_PARTS; b++) {
      if (b > P || !((m >> b) & 1u)) continue;
      const int li = b == 0 ? o : P + (b - 1) * P + o;
      uint32_t at = s_run[li] + rk[b];
      for (int k = 0; k < w; k++) at += s_wc[k][li];
      a.onodes[at] = (int32_t)v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NL; i += EP_TPB) {
      uint32_t add = 0u;
#pragma unroll
      for (int k = 0; k < EP_NW; k++) add += s_wc[k][i];
      s_run[i] += add;
    }
    __syncthreads();
  }
}

__global__ void k_ep_narrow(const int64_t *s64, const int64_t *d64, int64_t n, int32_t *s32, int32_t *d32) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = s64[i], d = d64[i];
    s32[i] = (s < 0 || s > 0x7FFFFFFF) ? -1 : (int32_t)s;
    d32[i] = (d < 0 || d > 0x7FFFFFFF) ? -1 : (int32_t)d;
  }
}

}  // namespace

struct gs_epart {
  int device = 0;
  int64_t N = 0, max_edges = 0;
  int P = 0;
  bool has_resident = false;
  cudaStream_t st = nullptr;
  std::vector<void *> dev, pinned;
  EpArgs a{};
  int64_t *s64 = nullptr, *d64 = nullptr;
  int64_t *tot_h = nullptr;
  int32_t *osrc_h = nullptr, *odst_h = nullptr, *onodes_h = nullptr;
  int64_t ebase[EP_MAXC + 1] = {0}, nbase[EP_MAXL + 1] = {0};
  bool valid = false;
};

namespace {
int ep_fail(int code, const char *fmt, long long x = 0, long long y = 0) {
  char buf[256];
  snprintf(buf, sizeof buf, fmt, x, y);
  return gs_set_error(code, buf);
}
#define EPC(call)                                                                        \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      char b_[256];                                                                      \
      snprintf(b_, sizeof b_, "%s failed: %s (epart.cu:%d)", #call, cudaGetErrorString(e_), __LINE__); \
      return gs_set_error(GS_ERR_CUDA, b_);                                              \
    }                                                                                    \
  } while (0)
template <typename T>
int ep_dev(gs_epart *e, T *&p, size_t n) {
  void *q = nullptr;
  EPC(cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
  e->dev.push_back(q);
  p = (T *)q;
  return GS_OK;
}
template <typename T>
int ep_pin(gs_epart *e, T *&p, size_t n) {
  void *q = nullptr;
  EPC(cudaHostAlloc(&q, std::max<size_t>(n, 1) * sizeof(T), cudaHostAllocDefault));
  e->pinned.push_back(q);
  p = (T *)q;
  return GS_OK;
}
}  // namespace

extern "C" int gs_epart_destroy(gs_epart *e) {
  if (!e) return GS_OK;
  cudaSetDevice(e->device);
  if (e->st) cudaStreamSynchronize(e->st);
  for (void *p : e->dev) cudaFree(p);
  for (void *p : e->pinned) cudaFreeHost(p);
  if (e->st) cudaStreamDestroy(e->st);
  delete e;
  return GS_OK;
}

extern "C" int gs_epart_create(int device, int64_t num_nodes, int32_t num_parts, const int32_t *owner_host,
                               const uint8_t *resident_host, int64_t max_edges, gs_epart **out) {
  if (!out) return ep_fail(GS_ERR_INVALID, "gs_epart_create: out is NULL");
  *out = nullptr;
  if (num_nodes < 1 || num_nodes > 0x7FFFFFFFll || max_edges < 1 || max_edges > 0x7FFFFFFFll)
    return ep_fail(GS_ERR_INVALID, "gs_epart_create: num_nodes %lld / max_edges %lld out of range", num_nodes, max_edges);
  if (num_parts < 1 || num_parts > GS_MAX_PARTS || !owner_host)
    return ep_fail(GS_ERR_INVALID, "gs_epart_create: num_parts must be in [1,%lld] and owner_host given", GS_MAX_PARTS);
  for (int64_t v = 0; v < num_nodes; v++)
    if (owner_host[v] < 0 || owner_host[v] >= num_parts)
      return ep_fail(GS_ERR_INVALID, "gs_epart_create: owner[%lld] = %lld is not a part", v, owner_host[v]);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || device < 0 || device >= ndev) {
    cudaGetLastError();
    return ep_fail(GS_ERR_CUDA, "gs_epart_create: no CUDA device %lld (a GPU is required: no CPU fallback)", device);
  }
  EPC(cudaSetDevice(device));
  gs_epart *e = new (std::nothrow) gs_epart();
  if (!e) return ep_fail(GS_ERR_INVALID, "out of host memory");
  e->device = device;
  e->N = num_nodes;
  e->P = num_parts;
  e->max_edges = max_edges;
  e->has_resident = resident_host != nullptr;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  EpArgs &a = e->a;
  a.N = (int32_t)num_nodes;
  a.P = num_parts;
  a.grid_e = (int)std::max<int64_t>(1, std::min<int64_t>(4 * sms, (max_edges + EP_TPB - 1) / EP_TPB));
  a.grid_n = (int)std::max<int64_t>(1, std::min<int64_t>(4 * sms, (num_nodes + EP_TPB - 1) / EP_TPB));
  int rc = GS_OK;
  uint8_t *own = nullptr, *res = nullptr;
  int32_t *s32 = nullptr, *d32 = nullptr;
  auto bail = [&](int r) { gs_epart_destroy(e); return r; };
  if (cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking) != cudaSuccess) return bail(ep_fail(GS_ERR_CUDA, "cudaStreamCreate failed"));
  if ((rc = ep_dev(e, own, (size_t)num_nodes))) return bail(rc);
  if (resident_host && (rc = ep_dev(e, res, (size_t)num_nodes * num_parts))) return bail(rc);
  if ((rc = ep_dev(e, a.marks, (size_t)num_nodes)) || (rc = ep_dev(e, e->s64, (size_t)max_edges)) ||
      (rc = ep_dev(e, e->d64, (size_t)max_edges)) || (rc = ep_dev(e, s32, (size_t)max_edges)) ||
      (rc = ep_dev(e, d32, (size_t)max_edges)) || (rc = ep_dev(e, a.osrc, (size_t)max_edges)) ||
      (rc = ep_dev(e, a.odst, (size_t)max_edges)) || (rc = ep_dev(e, a.onodes, 2 * (size_t)max_edges)) ||
      (rc = ep_dev(e, a.ecnt, (size_t)EP_MAXC * a.grid_e)) || (rc = ep_dev(e, a.ncnt, (size_t)EP_MAXL * a.grid_n)) ||
      (rc = ep_dev(e, a.tot, 128)))
    return bail(rc);
  if ((rc = ep_pin(e, e->tot_h, 128)) || (rc = ep_pin(e, e->osrc_h, (size_t)max_edges)) ||
      (rc = ep_pin(e, e->odst_h, (size_t)max_edges)) || (rc = ep_pin(e, e->onodes_h, 2 * (size_t)max_edges)))
    return bail(rc);
  a.src = s32;
  a.dst = d32;
  a.owner = own;
  a.resident = res;
  {
    std::vector<uint8_t> o8((size_t)num_nodes);
    for (int64_t v = 0; v < num_nodes; v++) o8[v] = (uint8_t)owner_host[v];
    if (cudaMemcpy(own, o8.data(), (size_t)num_nodes, cudaMemcpyHostToDevice) != cudaSuccess)
      return bail(ep_fail(GS_ERR_CUDA, "gs_epart_create: owner copy failed"));
  }
  if (res && cudaMemcpy(res, resident_host, (size_t)num_nodes * num_parts, cudaMemcpyHostToDevice) != cudaSuccess)
    return bail(ep_fail(GS_ERR_CUDA, "gs_epart_create: residency mask copy failed"));
  if (cudaMemset(a.marks, 0, (size_t)num_nodes * sizeof(uint32_t)) != cudaSuccess)
    return bail(ep_fail(GS_ERR_CUDA, "gs_epart_create: memset failed"));
  *out = e;
  return GS_OK;
}

extern "C" int gs_epart_run(gs_epart *e, const int64_t *src_host, const int64_t *dst_host, int64_t n_edges,
                            int32_t last_layer) {
  if (!e || n_edges < 0 || (n_edges > 0 && (!src_host || !dst_host)))
    return ep_fail(GS_ERR_INVALID, "gs_epart_run: bad argument");
  if (n_edges > e->max_edges) return ep_fail(GS_ERR_INVALID, "gs_epart_run: %lld edges exceed max_edges = %lld", n_edges, e->max_edges);
  if (last_layer && !e->has_resident)
    return ep_fail(GS_ERR_INVALID, "gs_epart_run: the last layer's rule needs the residency mask (node_gpu_mask)");
  GsRange nv("gs_epart_run");
  EPC(cudaSetDevice(e->device));
  e->valid = false;
  EpArgs a = e->a;
  a.n = n_edges;
  a.last = last_layer ? 1 : 0;
  const int NC = 2 * e->P, NL = e->P + e->P * e->P;
  cudaStream_t st = e->st;
  memset(e->tot_h, 0, 128 * sizeof(int64_t));
  e->tot_h[120] = -1;  // as unsigned: no bad edge yet
  EPC(cudaMemcpyAsync(a.tot, e->tot_h, 128 * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  if (n_edges > 0) {
    EPC(cudaMemcpyAsync(e->s64, src_host, (size_t)n_edges * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    EPC(cudaMemcpyAsync(e->d64, dst_host, (size_t)n_edges * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    k_ep_narrow<<<592, 256, 0, st>>>(e->s64, e->d64, n_edges, (int32_t *)a.src, (int32_t *)a.dst);
  }
  k_ep_count<<<a.grid_e, EP_TPB, 0, st>>>(a);
  k_ep_spine<<<1, EP_TPB, 0, st>>>(a.ecnt, a.grid_e, NC, a.tot);
  k_ep_scatter<<<a.grid_e, EP_TPB, 0, st>>>(a);
  k_ep_ncount<<<a.grid_n, EP_TPB, 0, st>>>(a);
  k_ep_spine<<<1, EP_TPB, 0, st>>>(a.ncnt, a.grid_n, NL, a.tot + 32);
  k_ep_nwrite<<<a.grid_n, EP_TPB, 0, st>>>(a);
  EPC(cudaGetLastError());
  EPC(cudaMemcpyAsync(e->tot_h, a.tot, 128 * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  EPC(cudaStreamSynchronize(st));
  if (e->tot_h[120] != -1) {
    // the marks of a failed run are stale: clear them before reporting
    cudaMemsetAsync(a.marks, 0, (size_t)e->N * sizeof(uint32_t), st);
    cudaStreamSynchronize(st);
    return ep_fail(GS_ERR_INVALID, "gs_epart_run: edge %lld has an endpoint outside [0, %lld)", e->tot_h[120] - 1, e->N);
  }
  int64_t et = 0, nt = 0;
  for (int c = 0; c < NC; c++) {
    e->ebase[c] = et;
    et += e->tot_h[c];
  }
  e->ebase[NC] = et;
  for (int l = 0; l < NL; l++) {
    e->nbase[l] = nt;
    nt += e->tot_h[32 + l];
  }
  e->nbase[NL] = nt;
  if (et != n_edges) return ep_fail(GS_ERR_STATE, "gs_epart_run: %lld of %lld edges placed", et, n_edges);
  if (et > 0) {
    EPC(cudaMemcpyAsync(e->osrc_h, a.osrc, (size_t)et * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    EPC(cudaMemcpyAsync(e->odst_h, a.odst, (size_t)et * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  }
  if (nt > 0) EPC(cudaMemcpyAsync(e->onodes_h, a.onodes, (size_t)nt * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  EPC(cudaStreamSynchronize(st));
  e->valid = true;
  return GS_OK;
}

extern "C" int gs_epart_get(gs_epart *e, int32_t part, int32_t what, int32_t remote, const int32_t **host,
                            const int32_t **device, int64_t *len) {
  if (!e || !len || part < 0 || part >= e->P) return ep_fail(GS_ERR_INVALID, "gs_epart_get: bad argument");
  if (!e->valid) return ep_fail(GS_ERR_STATE, "gs_epart_get: no completed gs_epart_run");
  const int P = e->P;
  int64_t off = 0, n = 0;
  const int32_t *h = nullptr, *d = nullptr;
  switch (what) {
    case GS_EP_SRC:
    case GS_EP_DST:
      off = e->ebase[2 * part];
      n = e->ebase[2 * part + 2] - off;
      h = (what == GS_EP_SRC ? e->osrc_h : e->odst_h) + off;
      d = (what == GS_EP_SRC ? e->a.osrc : e->a.odst) + off;
      break;
    case GS_EP_OWNED:
      off = e->nbase[part];
      n = e->nbase[part + 1] - off;
      h = e->onodes_h + off;
      d = e->a.onodes + off;
      break;
    case GS_EP_SHUFFLE:
      if (remote < 0 || remote >= P || remote == part) return ep_fail(GS_ERR_INVALID, "gs_epart_get: remote must be another part");
      off = e->nbase[P + part * P + remote];
      n = e->nbase[P + part * P + remote + 1] - off;
      h = e->onodes_h + off;
      d = e->a.onodes + off;
      break;
    case GS_EP_NUM_EXTRA:
      n = e->ebase[2 * part + 2] - e->ebase[2 * part + 1];
      break;
    default:
      return ep_fail(GS_ERR_INVALID, "gs_epart_get: unknown list %lld", what);
  }
  if (host) *host = h;
  if (device) *device = d;
  *len = n;
  return GS_OK;
}
