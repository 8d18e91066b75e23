// pslicer_api.cu -- host side of the partitioned-mode C ABI (gs_pslicer_*,
// include/gsplit_slicer.h).  Owns the rank's graph shard, its peer-visible
// arena (exported with CUDA IPC), the private workspaces, the sample slots and
// the per-slot CUDA graphs.  No slicing logic (pslicer_kernels.cu), no CPU path.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "pslicer_internal.cuh"

namespace {

int pfail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  return gs_set_error(code, buf);
}

#define CU(call)                                                                        \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess)                                                              \
      return pfail(GS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                   __FILE__, __LINE__);                                                 \
  } while (0)

enum { PSLOT_FREE = 0, PSLOT_STAGED = 1, PSLOT_BUSY = 2, PSLOT_DONE = 3 };

struct PSlot {
  PsWs ws{};
  gs_sample_header *hdr = nullptr;  // device
  int32_t *slab = nullptr;          // device
  GsInput *input = nullptr;         // pinned, mapped
  int64_t *seeds = nullptr;
  gs_sample_header *hdr_host = nullptr;
  int32_t *slab_host = nullptr;
  float *feat_out = nullptr;        // device
  cudaGraphExec_t exec = nullptr;
  cudaEvent_t done = nullptr;
  int state = PSLOT_FREE;
};

size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

// Peer-visible arena: one allocation, same layout on every rank.
size_t arena_layout(char *base, int64_t table_cap, int64_t Ccap, int64_t max_batch, int64_t stage_floats,
                    PsArena *a) {
  size_t o = 0;
  auto take = [&](size_t bytes) {
    char *p = base ? base + o : nullptr;
    o += align_up(bytes);
    return p;
  };
  a->table = (uint2 *)take((size_t)table_cap * sizeof(uint2));
  a->cand = (int32_t *)take((size_t)Ccap * sizeof(int32_t));
  a->cflag = (uint8_t *)take((size_t)Ccap);
  a->nf = (uint2 *)take((size_t)Ccap * sizeof(uint2));
  a->cnt0 = (uint32_t *)take((size_t)max_batch * sizeof(uint32_t));
  a->rowrem = nullptr;
  a->stage = (float *)take((size_t)stage_floats * sizeof(float));
  a->flags = (uint32_t *)take(GS_MAX_PARTS * sizeof(uint32_t));
  a->err = (int32_t *)take(sizeof(int32_t));
  return o;
}

}  // namespace

struct gs_pslicer {
  int device = 0;
  gs_pconfig cfg{};
  int64_t N = 0, n_local = 0, m_local = 0, table_cap = 0;
  int64_t Fcap = 0, Ccap = 0, slab_cap = 0;
  int64_t *indptr = nullptr;
  int32_t *indices = nullptr;
  float *feat = nullptr;  // [n_local][feat_dim]
  int64_t stage_floats = 0;
  int64_t feat_cap_rows = 0;
  char *arena = nullptr;
  size_t arena_bytes = 0;
  PsArena me{};
  PsArena peer[GS_MAX_PARTS]{};
  bool connected[GS_MAX_PARTS] = {false};
  void *ipc_base[GS_MAX_PARTS] = {nullptr};
  std::vector<void *> dev;  // private device allocations
  PsWs base{};
  std::vector<PSlot> slots;
  cudaStream_t stream = nullptr;
  bool own_stream = false, ready = false;
  int grid = 0, grid_rows[GS_MAX_LAYERS] = {0}, grid_cand[GS_MAX_LAYERS] = {0};
  int next_slot = 0;
  int64_t bytes = 0;
};

namespace {

template <typename T>
int palloc(gs_pslicer *s, T **p, size_t n) {
  void *q = nullptr;
  const size_t b = std::max<size_t>(n, 1) * sizeof(T);
  CU(cudaMalloc(&q, b));
  s->dev.push_back(q);
  s->bytes += (int64_t)b;
  *p = (T *)q;
  return GS_OK;
}

int num_segments(const gs_pslicer *s) { return 3 * s->cfg.num_layers + 2; }

// Segment k of a minibatch (see gsplit_slicer.h).  Every segment but the first
// starts by waiting for the peers' previous signal; every one but the last
// ends by signalling.
int enqueue_segment(gs_pslicer *s, PSlot &sl, int seg) {
  const int L = s->cfg.num_layers;
  cudaStream_t st = s->stream;
  const PsWs &ws = sl.ws;
  if (seg == 0) {
    ps_launch_begin(ws, st);
    ps_launch_signal(ws, st);
  } else {
    ps_launch_wait(ws, st);
    const int l = (seg - 1) / 3, k = (seg - 1) % 3;
    if (k == 0) {
      if (l > 0) {  // row side of the previous layer, then this layer's rows
        ps_launch_front_count(ws, l - 1, s->grid_cand[l - 1], st);
        ps_launch_front_apply(ws, l - 1, s->grid_cand[l - 1], st);
        ps_launch_relabel(ws, l - 1, s->grid_cand[l - 1], st);
      }
      if (l < L) {
        ps_launch_rows_count(ws, l, s->grid_rows[l], st);
        ps_launch_rows_scan(ws, l, s->grid_rows[l], st);
        ps_launch_signal(ws, st);
      } else {
        if (s->feat) ps_launch_gather(ws, 8 * (s->grid / std::max(1, s->cfg.grid_ctas > 0 ? s->cfg.grid_ctas : 4)), st);
        ps_launch_finish(ws, s->grid, st);
      }
    } else if (k == 1) {
      ps_launch_expand(ws, l, s->grid_cand[l], st);
      ps_launch_signal(ws, st);
    } else {
      ps_launch_cand_flags(ws, l, s->grid_cand[l], st);
      ps_launch_cand_apply(ws, l, s->grid_cand[l], st);
      ps_launch_signal(ws, st);
    }
  }
  CU(cudaGetLastError());
  return GS_OK;
}

int pick_slot(gs_pslicer *s) {
  const int ns = (int)s->slots.size();
  for (int t = 0; t < ns; t++) {
    const int c = (s->next_slot + t) % ns;
    if (s->slots[c].state == PSLOT_FREE) return c;
  }
  return -1;
}

int stage(gs_pslicer *s, const int64_t *seeds, int32_t n, uint64_t batch_index, int32_t *slot_out) {
  if (!s || !slot_out) return pfail(GS_ERR_INVALID, "NULL argument");
  if (!s->ready) return pfail(GS_ERR_STATE, "gs_pslicer_ready() has not been called");
  if (n < 0 || n > s->cfg.max_batch)
    return pfail(GS_ERR_INVALID, "batch of %d seeds exceeds max_batch=%d", n, s->cfg.max_batch);
  if (n && !seeds) return pfail(GS_ERR_INVALID, "seeds_host is NULL");
  const int k = pick_slot(s);
  if (k < 0) return pfail(GS_ERR_BUSY, "all %d sample slots are in use; release one", (int)s->slots.size());
  PSlot &sl = s->slots[k];
  memcpy(sl.seeds, seeds, (size_t)n * sizeof(int64_t));
  sl.input->n_seeds = n;
  sl.input->want_host = s->cfg.host_output;
  sl.input->batch_index = batch_index;
  sl.state = PSLOT_STAGED;
  s->next_slot = (k + 1) % (int)s->slots.size();
  *slot_out = k;
  return GS_OK;
}

}  // namespace

extern "C" {

void gs_pconfig_default(gs_pconfig *c) {
  memset(c, 0, sizeof *c);
  c->abi_version = GS_ABI_VERSION;
  c->num_layers = 3;
  for (int i = 0; i < GS_MAX_LAYERS; i++) c->fanouts[i] = 10;
  c->num_parts = 4;
  c->owner_mode = GS_OWNER_MOD;
  c->max_batch = 4096;
  c->num_slots = 2;
  c->host_output = 1;
  c->timeout_ms = 5000;
  c->use_graph = 1;
}

int gs_pslicer_destroy(gs_pslicer *s) {
  if (!s) return GS_OK;
  cudaSetDevice(s->device);
  if (s->stream) cudaStreamSynchronize(s->stream);
  for (PSlot &sl : s->slots) {
    if (sl.exec) cudaGraphExecDestroy(sl.exec);
    if (sl.done) cudaEventDestroy(sl.done);
    cudaFree(sl.hdr);
    cudaFree(sl.slab);
    cudaFree(sl.feat_out);
    if (sl.input) cudaFreeHost(sl.input);
    if (sl.seeds) cudaFreeHost(sl.seeds);
    if (sl.hdr_host) cudaFreeHost(sl.hdr_host);
    if (sl.slab_host) cudaFreeHost(sl.slab_host);
  }
  for (int q = 0; q < GS_MAX_PARTS; q++)
    if (s->ipc_base[q]) cudaIpcCloseMemHandle(s->ipc_base[q]);
  for (void *p : s->dev) cudaFree(p);
  cudaFree(s->arena);
  cudaFree(s->feat);
  cudaFree(s->indptr);
  cudaFree(s->indices);
  if (s->own_stream && s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return GS_OK;
}

int gs_pslicer_create(int device, int64_t N, int64_t n_local, int64_t m_local,
                      const int64_t *indptr_host, const int64_t *indices_host,
                      const gs_pconfig *cfg, gs_pslicer **out) {
  if (!cfg || !out || !indptr_host || (m_local && !indices_host))
    return pfail(GS_ERR_INVALID, "NULL argument");
  if (cfg->abi_version != GS_ABI_VERSION) return pfail(GS_ERR_INVALID, "gs_pconfig.abi_version mismatch");
  const int P = cfg->num_parts, L = cfg->num_layers;
  if (L < 1 || L > GS_MAX_LAYERS) return pfail(GS_ERR_INVALID, "num_layers must be in [1,%d]", GS_MAX_LAYERS);
  if (P < 1 || P > GS_MAX_PARTS) return pfail(GS_ERR_INVALID, "num_parts must be in [1,%d]", GS_MAX_PARTS);
  if (cfg->rank < 0 || cfg->rank >= P) return pfail(GS_ERR_INVALID, "rank outside [0,num_parts)");
  if (N < 1 || N >= (1ll << 31) - 1) return pfail(GS_ERR_INVALID, "num_nodes must be in [1, 2^31-2]");
  for (int l = 0; l < L; l++)
    if (cfg->fanouts[l] < -1 || cfg->fanouts[l] == 0 || cfg->fanouts[l] > 1023)
      return pfail(GS_ERR_INVALID, "fanouts[%d]=%d: use -1 or 1..1023", l, cfg->fanouts[l]);
  if (cfg->max_batch < 1 || cfg->num_slots < 1 || cfg->num_slots > 64)
    return pfail(GS_ERR_INVALID, "max_batch >= 1 and num_slots in [1,64] required");
  int64_t table_cap = 0, want_local = 0;
  if (cfg->owner_mode == GS_OWNER_MOD) {
    table_cap = (N + P - 1) / P;
    want_local = (N - cfg->rank + P - 1) / P;
  } else if (cfg->owner_mode == GS_OWNER_RANGE) {
    if (cfg->range_bounds[0] != 0 || cfg->range_bounds[P] != N)
      return pfail(GS_ERR_INVALID, "range_bounds must start at 0 and end at num_nodes");
    for (int q = 0; q < P; q++) {
      const int64_t w = (int64_t)cfg->range_bounds[q + 1] - cfg->range_bounds[q];
      if (w < 0) return pfail(GS_ERR_INVALID, "range_bounds must be non-decreasing");
      table_cap = std::max(table_cap, w);
    }
    want_local = (int64_t)cfg->range_bounds[cfg->rank + 1] - cfg->range_bounds[cfg->rank];
  } else {
    return pfail(GS_ERR_INVALID, "unknown owner_mode %d", cfg->owner_mode);
  }
  if (cfg->rng_mode != GS_RNG_PHILOX && cfg->rng_mode != GS_RNG_PHILOX_NOREPLACE)
    return pfail(GS_ERR_INVALID, "partitioned slicer: rng_mode must be GS_RNG_PHILOX or GS_RNG_PHILOX_NOREPLACE");
  if (n_local != want_local)
    return pfail(GS_ERR_INVALID, "rank %d owns %lld nodes under this owner map, shard has %lld",
                 cfg->rank, (long long)want_local, (long long)n_local);
  if (indptr_host[0] != 0 || indptr_host[n_local] != m_local)
    return pfail(GS_ERR_INVALID, "indptr_local[0] != 0 or indptr_local[n_local] != m_local");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return pfail(GS_ERR_CUDA, "no CUDA device (%s): this library has no CPU fallback",
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  CU(cudaSetDevice(device));
  gs_pslicer *s = new (std::nothrow) gs_pslicer();
  if (!s) return pfail(GS_ERR_INVALID, "out of host memory");
  s->device = device;
  s->cfg = *cfg;
  s->cfg.features_local_host = nullptr;
  if (s->cfg.timeout_ms <= 0) s->cfg.timeout_ms = 5000;
  s->N = N;
  s->n_local = n_local;
  s->m_local = m_local;
  s->table_cap = table_cap;
  // global capacities (identical on every rank): F_{l+1} <= min(N, C_l), C_l <= F_l (k_l + 1)
  int64_t F = cfg->max_batch, fmax = F, cmax = 0, slab = F + 8, Fb[GS_MAX_LAYERS], Cb[GS_MAX_LAYERS];
  for (int l = 0; l < L; l++) {
    F = std::min(F, N);
    fmax = std::max(fmax, F);
    const int64_t fan = cfg->fanouts[l];
    if (fan < 0 && cfg->num_edges_global <= 0 && cfg->cand_capacity <= 0) {
      delete s;
      return pfail(GS_ERR_INVALID, "fanout -1 needs num_edges_global or explicit capacities");
    }
    const int64_t Mg = cfg->num_edges_global > 0 ? cfg->num_edges_global : (int64_t)1 << 40;
    const int64_t E = fan >= 0 ? std::min(F * fan, Mg) : std::min(Mg, cfg->cand_capacity > 0 ? cfg->cand_capacity : Mg);
    const int64_t C = F + E;
    cmax = std::max(cmax, C);
    Fb[l] = F;
    Cb[l] = C;
    slab += std::min(N, C) + std::min(N, E) + E + 6 * F + 64;
    F = std::min(N, C);
  }
  fmax = std::max(fmax, F);
  s->Fcap = cfg->frontier_capacity > 0 ? cfg->frontier_capacity : fmax;
  s->Ccap = cfg->cand_capacity > 0 ? cfg->cand_capacity : cmax;
  s->slab_cap = cfg->slab_capacity > 0 ? cfg->slab_capacity : slab;
  if (s->Ccap >= (1ll << 31) - 2 * GS_TILE || s->Fcap >= (1ll << 31) - 2 * GS_TILE ||
      s->slab_cap >= (1ll << 31) - 16) {
    delete s;
    return pfail(GS_ERR_INVALID, "capacities exceed 2^31 positions: set frontier_capacity / "
                                 "cand_capacity / slab_capacity (required for fanout -1)");
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  s->grid = (cfg->grid_ctas > 0 ? cfg->grid_ctas : 4) * sms;
  for (int l = 0; l < L; l++) {
    const int64_t fb = std::min(Fb[l], s->Fcap), cb = std::min(Cb[l], s->Ccap);
    s->grid_rows[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (fb + GS_TILE - 1) / GS_TILE));
    s->grid_cand[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (cb + GS_TILE - 1) / GS_TILE));
  }
#define CUS(call)                                                                          \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      pfail(GS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
            __LINE__);                                                                     \
      std::string keep = gs_last_error();                                                  \
      gs_pslicer_destroy(s);                                                               \
      return gs_set_error(GS_ERR_CUDA, keep.c_str());                                      \
    }                                                                                      \
  } while (0)
#define PA(ptr, n)                                      \
  do {                                                  \
    if (palloc(s, &ptr, (size_t)(n)) != GS_OK) {        \
      std::string keep = gs_last_error();               \
      gs_pslicer_destroy(s);                            \
      return gs_set_error(GS_ERR_CUDA, keep.c_str());   \
    }                                                   \
  } while (0)
  // graph shard; validated on the device like gs_graph_create's (the kernels index the
  // shard's rows and the peers' tables with these ids)
  CUS(cudaMalloc((void **)&s->indptr, (size_t)(n_local + 1) * sizeof(int64_t)));
  CUS(cudaMemcpy(s->indptr, indptr_host, (size_t)(n_local + 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
  CUS(cudaMalloc((void **)&s->indices, (size_t)std::max<int64_t>(m_local, 1) * sizeof(int32_t)));
  {
    const int64_t chunk = 1ll << 26;
    int64_t *tmp = nullptr;
    unsigned long long *err = nullptr, bad[2] = {0, 0};
    CUS(cudaMalloc((void **)&err, 2 * sizeof(unsigned long long)));
    CUS(cudaMemset(err, 0xFF, 2 * sizeof(unsigned long long)));
    gs_launch_check_indptr(s->indptr, n_local, m_local, err, 0);
    CUS(cudaMalloc((void **)&tmp, (size_t)std::min(chunk, std::max<int64_t>(m_local, 1)) * sizeof(int64_t)));
    for (int64_t o = 0; o < m_local; o += chunk) {
      const int64_t n = std::min(chunk, m_local - o);
      CUS(cudaMemcpy(tmp, indices_host + o, (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice));
      gs_launch_check_ids(tmp, (size_t)n, o, N, err, 0);
      gs_launch_narrow_ids(tmp, s->indices + o, (size_t)n, 0);
    }
    CUS(cudaMemcpy(bad, err, sizeof bad, cudaMemcpyDeviceToHost));
    CUS(cudaDeviceSynchronize());
    cudaFree(tmp);
    cudaFree(err);
    if (bad[0] != ~0ull || bad[1] != ~0ull) {
      gs_pslicer_destroy(s);
      if (bad[0] != ~0ull)
        return pfail(GS_ERR_INVALID, "indptr_local is not a non-decreasing sequence inside [0, m_local]: first bad row %llu",
                     bad[0] - 1ull);
      return pfail(GS_ERR_INVALID, "indices_local[%llu] lies outside [0, num_nodes_global)", bad[1] - 1ull);
    }
  }
  // arena
  if (cfg->shuffle_rows_capacity < 0 || cfg->shuffle_max_dim < 0) {
    gs_pslicer_destroy(s);
    return pfail(GS_ERR_INVALID, "negative shuffle capacity");
  }
  const int64_t stage_dim = ((int64_t)cfg->shuffle_max_dim + 3) & ~3ll;  // 16-byte aligned peer blocks and rows
  s->stage_floats = (int64_t)P * cfg->shuffle_rows_capacity * stage_dim;
  PsArena dummy;
  s->arena_bytes = arena_layout(nullptr, table_cap, s->Ccap, cfg->max_batch, s->stage_floats, &dummy);
  CUS(cudaMalloc((void **)&s->arena, s->arena_bytes));
  s->bytes += (int64_t)s->arena_bytes;
  arena_layout(s->arena, table_cap, s->Ccap, cfg->max_batch, s->stage_floats, &s->me);
  CUS(cudaMemset(s->arena, 0, s->arena_bytes));
  CUS(cudaMemset(s->me.table, 0xFF, (size_t)table_cap * sizeof(uint2)));
  s->peer[cfg->rank] = s->me;
  s->connected[cfg->rank] = true;
  if (cfg->stream) {
    s->stream = (cudaStream_t)cfg->stream;
  } else {
    CUS(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    s->own_stream = true;
  }
  if (cfg->features_local_host) {
    if (cfg->feature_dim < 1) {
      gs_pslicer_destroy(s);
      return pfail(GS_ERR_INVALID, "feature_dim must be >= 1");
    }
    const size_t fb = (size_t)std::max<int64_t>(n_local, 1) * cfg->feature_dim * sizeof(float);
    CUS(cudaMalloc((void **)&s->feat, fb));
    CUS(cudaMemcpy(s->feat, cfg->features_local_host, (size_t)n_local * cfg->feature_dim * sizeof(float),
                   cudaMemcpyHostToDevice));
    s->bytes += (int64_t)fb;
    const int last = L - 1;
    s->feat_cap_rows = cfg->feature_rows_capacity > 0
                           ? cfg->feature_rows_capacity
                           : std::max<int64_t>(1, std::min(n_local, Cb[last] - Fb[last]));
  }
  // private workspaces
  PsWs &w = s->base;
  w.indptr = s->indptr;
  w.indices = s->indices;
  w.num_nodes = (int32_t)N;
  w.n_local = (int32_t)n_local;
  w.P = P;
  w.L = L;
  w.rank = cfg->rank;
  w.owner_mode = cfg->owner_mode == GS_OWNER_MOD ? PS_OWNER_MOD : PS_OWNER_RANGE;
  for (int q = 0; q <= GS_MAX_PARTS; q++) w.bounds[q] = q <= P ? cfg->range_bounds[q] : 0;
  for (int l = 0; l < GS_MAX_LAYERS; l++) w.fanouts[l] = cfg->fanouts[l];
  w.seed_lo = (uint32_t)cfg->seed;
  w.seed_hi = (uint32_t)(cfg->seed >> 32);
  w.rng_mode = cfg->rng_mode;
  w.pmask = (P & (P - 1)) == 0 ? P - 1 : -1;
  w.pshift = 0;
  while ((1 << w.pshift) < P) w.pshift++;
  w.pmagic = (uint32_t)((0x100000000ull + (uint64_t)P - 1) / (uint64_t)P);
  w.me = s->me;
  w.Fcap = (int32_t)s->Fcap;
  w.Ccap = (int32_t)s->Ccap;
  w.grid_max = s->grid;
  w.timeout_ms = s->cfg.timeout_ms;
  w.slab_cap = s->slab_cap;
  w.stage_rows = cfg->shuffle_rows_capacity;
  w.stage_dim = (int32_t)(((int64_t)cfg->shuffle_max_dim + 3) & ~3ll);
  w.feat = s->feat;
  w.feat_dim = cfg->feature_dim;
  w.feat_cap_rows = s->feat_cap_rows;
  const int ctiles = (int)((s->Ccap + GS_TILE - 1) / GS_TILE) + 1;
  PA(w.rowstart, s->Fcap);
  PA(w.rowdeg, s->Fcap);
  PA(w.cnt[0], s->Fcap);
  PA(w.cnt[1], s->Fcap);
  PA(w.off, s->Fcap + 1);
  PA(w.rowscan, s->Fcap + 1);
  PA(w.rowrem, s->Fcap);
  PA(w.tile_first, ctiles + 1);
  PA(w.ftile_sum, s->grid);
  PA(w.ftile_pre, s->grid);
  PA(w.ctr, (size_t)PS_NCTR * s->grid);
  PA(w.tickets, 8);
  PA(w.epoch, 2);
  CUS(cudaMemset(w.tickets, 0, 8 * sizeof(uint32_t)));
  CUS(cudaMemset(w.epoch, 0, 2 * sizeof(uint32_t)));
  CUS(cudaMemset(w.rowrem, 0, (size_t)s->Fcap * sizeof(uint32_t)));
  s->slots.resize(cfg->num_slots);
  for (PSlot &sl : s->slots) {
    CUS(cudaMalloc((void **)&sl.hdr, sizeof(gs_sample_header)));
    CUS(cudaMemset(sl.hdr, 0, sizeof(gs_sample_header)));
    CUS(cudaMalloc((void **)&sl.slab, (size_t)(s->slab_cap + 4) * sizeof(int32_t)));
    s->bytes += (int64_t)((s->slab_cap + 4) * sizeof(int32_t));
    CUS(cudaHostAlloc((void **)&sl.input, sizeof(GsInput), cudaHostAllocMapped));
    CUS(cudaHostAlloc((void **)&sl.seeds, (size_t)cfg->max_batch * sizeof(int64_t), cudaHostAllocMapped));
    CUS(cudaHostAlloc((void **)&sl.hdr_host, sizeof(gs_sample_header), cudaHostAllocMapped));
    memset(sl.hdr_host, 0, sizeof(gs_sample_header));
    memset(sl.input, 0, sizeof(GsInput));
    if (cfg->host_output)
      CUS(cudaHostAlloc((void **)&sl.slab_host, (size_t)(s->slab_cap + 4) * sizeof(int32_t), cudaHostAllocMapped));
    CUS(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    if (s->feat) {
      CUS(cudaMalloc((void **)&sl.feat_out, (size_t)s->feat_cap_rows * cfg->feature_dim * sizeof(float)));
      s->bytes += (int64_t)(s->feat_cap_rows * cfg->feature_dim * sizeof(float));
    }
  }
  CUS(cudaDeviceSynchronize());
#undef PA
#undef CUS
  *out = s;
  return GS_OK;
}

int gs_pslicer_ipc_handle(gs_pslicer *s, void *handle64) {
  if (!s || !handle64) return pfail(GS_ERR_INVALID, "NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  CU(cudaSetDevice(s->device));
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, s->arena));
  memcpy(handle64, &h, sizeof h);
  return GS_OK;
}

int gs_pslicer_connect_ipc(gs_pslicer *s, int32_t q, const void *handle64) {
  if (!s || !handle64) return pfail(GS_ERR_INVALID, "NULL argument");
  if (q < 0 || q >= s->cfg.num_parts || q == s->cfg.rank) return pfail(GS_ERR_INVALID, "bad peer rank %d", q);
  if (s->ready) return pfail(GS_ERR_STATE, "already frozen by gs_pslicer_ready()");
  CU(cudaSetDevice(s->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof h);
  void *base = nullptr;
  CU(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  s->ipc_base[q] = base;
  arena_layout((char *)base, s->table_cap, s->Ccap, s->cfg.max_batch, s->stage_floats, &s->peer[q]);
  s->connected[q] = true;
  return GS_OK;
}

int gs_pslicer_connect_local(gs_pslicer *s, int32_t q, gs_pslicer *peer) {
  if (!s || !peer) return pfail(GS_ERR_INVALID, "NULL argument");
  if (q < 0 || q >= s->cfg.num_parts || q == s->cfg.rank) return pfail(GS_ERR_INVALID, "bad peer rank %d", q);
  if (s->ready) return pfail(GS_ERR_STATE, "already frozen by gs_pslicer_ready()");
  if (peer->cfg.rank != q || peer->cfg.num_parts != s->cfg.num_parts || peer->Ccap != s->Ccap ||
      peer->table_cap != s->table_cap || peer->cfg.max_batch != s->cfg.max_batch ||
      peer->stage_floats != s->stage_floats)
    return pfail(GS_ERR_INVALID, "peer %d was created with a different configuration", q);
  if (peer->device != s->device) {
    CU(cudaSetDevice(s->device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
      return pfail(GS_ERR_CUDA, "no peer access from device %d to %d: %s", s->device, peer->device,
                   cudaGetErrorString(e));
    cudaGetLastError();
  }
  s->peer[q] = peer->me;
  s->connected[q] = true;
  return GS_OK;
}

int gs_pslicer_ready(gs_pslicer *s) {
  if (!s) return pfail(GS_ERR_INVALID, "NULL argument");
  if (s->ready) return GS_OK;
  for (int q = 0; q < s->cfg.num_parts; q++)
    if (!s->connected[q]) return pfail(GS_ERR_STATE, "peer %d is not connected", q);
  CU(cudaSetDevice(s->device));
  for (int q = 0; q < s->cfg.num_parts; q++) s->base.peer[q] = s->peer[q];
  for (PSlot &sl : s->slots) {
    sl.ws = s->base;
    sl.ws.hdr = sl.hdr;
    sl.ws.slab = sl.slab;
    sl.ws.feat_out = sl.feat_out;
    void *dp = nullptr;
    CU(cudaHostGetDevicePointer(&dp, sl.input, 0));
    sl.ws.input = (const GsInput *)dp;
    CU(cudaHostGetDevicePointer(&dp, sl.seeds, 0));
    sl.ws.seeds_in = (const int64_t *)dp;
    CU(cudaHostGetDevicePointer(&dp, sl.hdr_host, 0));
    sl.ws.hdr_host = (gs_sample_header *)dp;
    sl.ws.slab_host = nullptr;
    if (sl.slab_host) {
      CU(cudaHostGetDevicePointer(&dp, sl.slab_host, 0));
      sl.ws.slab_host = (int32_t *)dp;
    }
    if (s->cfg.use_graph && s->own_stream) {
      cudaGraph_t graph = nullptr;
      CU(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
      int rc = GS_OK;
      for (int k = 0; k < num_segments(s) && rc == GS_OK; k++) rc = enqueue_segment(s, sl, k);
      cudaError_t ce = cudaStreamEndCapture(s->stream, &graph);
      if (rc != GS_OK) return rc;
      if (ce != cudaSuccess) return pfail(GS_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(ce));
      CU(cudaGraphInstantiate(&sl.exec, graph, 0));
      cudaGraphDestroy(graph);
    }
  }
  s->ready = true;
  return GS_OK;
}

int gs_pslicer_num_segments(const gs_pslicer *s) { return s ? num_segments(s) : 0; }

int gs_pslicer_stage(gs_pslicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                     int32_t *slot_out) {
  return stage(s, seeds_host, n, batch_index, slot_out);
}

int gs_pslicer_enqueue(gs_pslicer *s, int32_t slot, int32_t seg_lo, int32_t seg_hi) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  PSlot &sl = s->slots[slot];
  if (sl.state != PSLOT_STAGED) return pfail(GS_ERR_STATE, "slot %d is not staged", slot);
  if (seg_lo < 0 || seg_hi > num_segments(s) || seg_lo > seg_hi) return pfail(GS_ERR_INVALID, "bad segment range");
  CU(cudaSetDevice(s->device));
  for (int k = seg_lo; k < seg_hi; k++) {
    const int rc = enqueue_segment(s, sl, k);
    if (rc != GS_OK) return rc;
  }
  return GS_OK;
}

int gs_pslicer_commit(gs_pslicer *s, int32_t slot) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  PSlot &sl = s->slots[slot];
  if (sl.state != PSLOT_STAGED) return pfail(GS_ERR_STATE, "slot %d is not staged", slot);
  CU(cudaSetDevice(s->device));
  CU(cudaEventRecord(sl.done, s->stream));
  sl.state = PSLOT_BUSY;
  return GS_OK;
}

int gs_pslicer_submit(gs_pslicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                      int32_t *slot_out) {
  const GsRange trace("gs_pslicer_submit");
  int rc = stage(s, seeds_host, n, batch_index, slot_out);
  if (rc != GS_OK) return rc;
  PSlot &sl = s->slots[*slot_out];
  CU(cudaSetDevice(s->device));
  if (sl.exec) {
    CU(cudaGraphLaunch(sl.exec, s->stream));
  } else {
    for (int k = 0; k < num_segments(s); k++)
      if ((rc = enqueue_segment(s, sl, k)) != GS_OK) return rc;
  }
  return gs_pslicer_commit(s, *slot_out);
}

int gs_pslicer_wait(gs_pslicer *s, int32_t slot, const gs_sample_header **hdr_out) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  PSlot &sl = s->slots[slot];
  if (sl.state == PSLOT_FREE || sl.state == PSLOT_STAGED)
    return pfail(GS_ERR_STATE, "slot %d holds no submitted sample", slot);
  const GsRange trace("gs_pslicer_wait");
  CU(cudaSetDevice(s->device));
  if (sl.state == PSLOT_BUSY) CU(cudaEventSynchronize(sl.done));
  sl.state = PSLOT_DONE;
  if (hdr_out) *hdr_out = sl.hdr_host;
  const int st = sl.hdr_host->status ? sl.hdr_host->status : sl.hdr_host->late_status;
  if (st == GS_ERR_OVERFLOW)
    return pfail(GS_ERR_OVERFLOW, "sample outgrew its workspace (frontier_capacity=%lld cand_capacity=%lld "
                 "slab_capacity=%lld): raise them on every rank, then gs_pslicer_reset()",
                 (long long)s->Fcap, (long long)s->Ccap, (long long)s->slab_cap);
  if (st == GS_ERR_SEED) return pfail(GS_ERR_SEED, "seed ids must be distinct and in [0, num_nodes)");
  if (st == GS_ERR_PEER)
    return pfail(GS_ERR_PEER, "a peer rank did not reach a barrier within %d ms", s->cfg.timeout_ms);
  if (st != 0) return pfail(st, "device raised status %d", st);
  return GS_OK;
}

int gs_pslicer_slab(gs_pslicer *s, int32_t slot, const int32_t **device_slab, const int32_t **host_slab) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  PSlot &sl = s->slots[slot];
  if (sl.state != PSLOT_DONE) return pfail(GS_ERR_STATE, "slot %d not waited on", slot);
  if (device_slab) *device_slab = sl.slab;
  if (host_slab) *host_slab = sl.slab_host;
  return GS_OK;
}

int gs_pslicer_feature_slab(gs_pslicer *s, int32_t slot, const float **device_rows, int32_t *dim,
                            int32_t *rows) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  if (!s->feat) return pfail(GS_ERR_STATE, "slicer was created without features_local_host");
  PSlot &sl = s->slots[slot];
  if (sl.state != PSLOT_DONE) return pfail(GS_ERR_STATE, "slot %d not waited on", slot);
  if (device_rows) *device_rows = sl.feat_out;
  if (dim) *dim = s->cfg.feature_dim;
  if (rows) *rows = sl.hdr_host->feature_rows;
  return GS_OK;
}

int gs_pslicer_release(gs_pslicer *s, int32_t slot) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  PSlot &sl = s->slots[slot];
  if (sl.state == PSLOT_BUSY) {
    CU(cudaSetDevice(s->device));
    CU(cudaEventSynchronize(sl.done));
  }
  sl.state = PSLOT_FREE;
  return GS_OK;
}

int gs_pslicer_shuffle(gs_pslicer *s, int32_t slot, int32_t layer, float *rows, int32_t dim,
                       int32_t backward, int32_t phase) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return pfail(GS_ERR_INVALID, "bad slot");
  PSlot &sl = s->slots[slot];
  if (sl.state != PSLOT_DONE) return pfail(GS_ERR_STATE, "slot %d not waited on", slot);
  const GsRange trace(backward ? "gs_pslicer_shuffle(backward)" : "gs_pslicer_shuffle(forward)");
  if (layer < 0 || layer >= s->cfg.num_layers) return pfail(GS_ERR_INVALID, "bad layer");
  if (!rows || dim < 1 || dim > s->cfg.shuffle_max_dim)
    return pfail(GS_ERR_INVALID, "rows NULL or dim %d outside [1, shuffle_max_dim=%d]", dim,
                 s->cfg.shuffle_max_dim);
  if (phase < -1 || phase > 2) return pfail(GS_ERR_INVALID, "phase must be -1, 0, 1 or 2");
  const gs_sample_header *h = sl.hdr_host;
  for (int q = 0; q < s->cfg.num_parts; q++)
    if (q != s->cfg.rank && (h->field[layer][q][GS_F_FROM_IDS].len > s->cfg.shuffle_rows_capacity ||
                             h->field[layer][q][GS_F_TO_IDS].len > s->cfg.shuffle_rows_capacity))
      return pfail(GS_ERR_OVERFLOW, "shuffle of layer %d needs more than shuffle_rows_capacity=%lld rows per peer",
                   layer, (long long)s->cfg.shuffle_rows_capacity);
  CU(cudaSetDevice(s->device));
  const PsWs &ws = sl.ws;
  cudaStream_t st = s->stream;
  if (phase == -1 || phase == 0) {
    ps_launch_shuffle_push(ws, layer, rows, dim, backward, s->grid, st);
    ps_launch_signal(ws, st);
  }
  if (phase == -1 || phase == 1) {
    ps_launch_wait(ws, st);
    for (int q = 0; q < s->cfg.num_parts; q++)
      if (q != s->cfg.rank) ps_launch_shuffle_merge(ws, layer, rows, dim, backward, q, s->grid, st);
    ps_launch_signal(ws, st);  // peers may reuse my staging arena only after I merged
  }
  if (phase == -1 || phase == 2) ps_launch_wait(ws, st);
  CU(cudaGetLastError());
  return GS_OK;
}

int gs_pslicer_sync(gs_pslicer *s) {
  if (!s) return pfail(GS_ERR_INVALID, "NULL argument");
  CU(cudaSetDevice(s->device));
  CU(cudaStreamSynchronize(s->stream));
  return GS_OK;
}

int gs_pslicer_reset(gs_pslicer *s) {
  if (!s) return pfail(GS_ERR_INVALID, "NULL argument");
  CU(cudaSetDevice(s->device));
  CU(cudaStreamSynchronize(s->stream));
  CU(cudaMemsetAsync(s->me.table, 0xFF, (size_t)s->table_cap * sizeof(uint2), s->stream));
  CU(cudaMemsetAsync(s->me.cflag, 0, (size_t)s->Ccap, s->stream));
  CU(cudaMemsetAsync(s->me.nf, 0, (size_t)s->Ccap * sizeof(uint2), s->stream));
  CU(cudaMemsetAsync(s->me.err, 0, sizeof(int32_t), s->stream));
  CU(cudaMemsetAsync(s->base.tickets, 0, 8 * sizeof(uint32_t), s->stream));
  // barrier state: after a peer timeout the ranks' epochs differ (the rank that ran alone
  // signalled, the others did not), so a reset starts every rank from epoch 0 with clean flags
  CU(cudaMemsetAsync(s->base.epoch, 0, 2 * sizeof(uint32_t), s->stream));
  CU(cudaMemsetAsync(s->me.flags, 0, GS_MAX_PARTS * sizeof(uint32_t), s->stream));
  CU(cudaStreamSynchronize(s->stream));
  for (PSlot &sl : s->slots) sl.state = PSLOT_FREE;
  return GS_OK;
}

int gs_pslicer_launches_per_sample(const gs_pslicer *s) {
  if (!s) return 0;
  const int L = s->cfg.num_layers;
  // begin + finish, 7 kernels per layer, 3L+1 barriers of two launches each
  return 2 + 7 * L + 2 * (3 * L + 1) + (s->feat ? 1 : 0);
}
int64_t gs_pslicer_device_bytes(const gs_pslicer *s) { return s ? s->bytes : 0; }

}  // extern "C"
