// slicer_api.cu -- host side of the C ABI declared in include/gsplit_slicer.h.
// Owns device memory, streams and the per-slot CUDA graphs; contains no
// slicing logic of its own (that is slicer_kernels.cu) and no CPU fallback.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "slicer_internal.cuh"

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

}  // namespace
// shared with pslicer_api.cu
int gs_set_error(int code, const char *msg) {
  g_err = msg;
  return code;
}

static bool gs_nvtx_enabled() {
  static const bool on = [] {
    const char *e = getenv("GSPLIT_NVTX");
    return e && *e && *e != '0';
  }();
  return on;
}
GsRange::GsRange(const char *name) : on(gs_nvtx_enabled()) {
  if (on) nvtxRangePushA(name);
}
GsRange::~GsRange() {
  if (on) nvtxRangePop();
}
namespace {

#define CU(call)                                                                       \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess)                                                             \
      return fail(GS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                  __FILE__, __LINE__);                                                 \
  } while (0)

// tagged dedup table: tag indices start at 1 (index 0 is the fill pattern's tag)
const uint32_t kEpochStart[2] = {1u, 1u};

enum SlotState { SLOT_FREE = 0, SLOT_BUSY = 1, SLOT_DONE = 2 };

struct Slot {
  GsWs ws{};
  std::vector<void *> dev;  // device allocations owned by the slot
  GsInput *input = nullptr; // pinned, mapped
  int64_t *seeds = nullptr; // pinned, mapped
  gs_sample_header *hdr_host = nullptr;
  int32_t *slab_host = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaEvent_t done = nullptr;
  cudaEvent_t chain_done = nullptr;  // copy-engine mirror: kernels of the sample finished
  cudaEvent_t pack_done = nullptr;   // gs_slicer_pack_parts: the part blocks are written
  int64_t copied = 0;                // ... int32 elements of the slab the DMA of this sample moved
  cudaStream_t stream = nullptr;
  int state = SLOT_FREE;
};

}  // namespace

struct gs_graph {
  int device = 0;
  int64_t num_nodes = 0, num_edges = 0;
  int64_t *indptr = nullptr;
  int32_t *indices = nullptr;
};

struct gs_features {
  int device = 0;
  int64_t num_rows = 0;
  int32_t dim = 0, id_div = 1;
  float *rows = nullptr;          // HBM: every row, or the cached rows
  int32_t *slot_of = nullptr;     // cached store: [num_rows] slot in `rows` or -1
  const float *host_dev = nullptr;// cached store: device alias of the caller's host table
  void *host_reg = nullptr;       // ... the range this library page-locked (NULL: was already pinned)
  int64_t n_cached = 0;
};

struct gs_slicer {
  gs_graph *g = nullptr;
  gs_features *feat = nullptr;
  int64_t feat_cap_rows = 0;
  gs_config cfg{};
  int grid = 0;                       // CTAs of a full-machine launch
  int grid_full = 0;                  // 8 CTAs per SM: full occupancy for the streaming gather
  int grid_rows[GS_MAX_LAYERS] = {0};  // per layer: enough CTAs for the largest possible frontier
  int grid_cand[GS_MAX_LAYERS] = {0};  // ... and candidate stream (static upper bounds)
  int64_t Fcap = 0, Ccap = 0, slab_cap = 0, fslab_cap = 0;
  uint8_t *owner = nullptr;
  uint32_t *draws = nullptr;
  uint64_t *draw_cursor = nullptr;
  std::vector<cudaStream_t> streams;
  cudaStream_t copy_stream = nullptr;  // host mirror DMAs of all slots, one after another (interleaved
                                       // DMAs lose ~12 % of the link: 50 vs 57 GB/s measured)
  int64_t mirror_est = 0;              // running estimate of slab_used (int32 elements) + margin
  std::vector<Slot> slots;
  int next_slot = 0;
  int64_t slot_bytes = 0;
};

namespace {

template <typename T>
int dev_alloc(Slot &sl, T **p, size_t n, int64_t *bytes) {
  void *q = nullptr;
  size_t b = std::max<size_t>(n, 1) * sizeof(T);
  CU(cudaMalloc(&q, b));
  sl.dev.push_back(q);
  *p = (T *)q;
  *bytes += (int64_t)b;
  return GS_OK;
}

// The static launch sequence of one minibatch.  `ev` (optional) receives one
// event after every launch, for gs_slicer_profile.
int enqueue_sample(gs_slicer *s, Slot &sl, cudaStream_t st, std::vector<cudaEvent_t> *ev,
                   std::vector<const char *> *names) {
  static const char *kNames[] = {"rows_count", "rows_scan",   "expand",      "cand_flags",
                                 "cand_apply", "front_count", "front_apply", "relabel"};
  auto mark = [&](const char *nm) -> int {
    if (ev) {
      cudaEvent_t e;
      CU(cudaEventCreate(&e));
      CU(cudaEventRecord(e, st));
      ev->push_back(e);
      names->push_back(nm);
    }
    return GS_OK;
  };
  int rc;
  if ((rc = mark("start"))) return rc;
  gs_launch_begin(sl.ws, st);
  if ((rc = mark("begin"))) return rc;
  for (int l = 0; l < s->cfg.num_layers; l++) {
    const int gr = s->grid_rows[l], gc = s->grid_cand[l];
    gs_launch_rows_count(sl.ws, l, gr, st);
    if ((rc = mark(kNames[0]))) return rc;
    gs_launch_rows_scan(sl.ws, l, gr, st);
    if ((rc = mark(kNames[1]))) return rc;
    gs_launch_expand(sl.ws, l, gc, st);
    if ((rc = mark(kNames[2]))) return rc;
    gs_launch_cand_flags(sl.ws, l, gc, st);
    if ((rc = mark(kNames[3]))) return rc;
    gs_launch_cand_apply(sl.ws, l, gc, st);
    if ((rc = mark(kNames[4]))) return rc;
    gs_launch_front_count(sl.ws, l, gr, st);
    if ((rc = mark(kNames[5]))) return rc;
    gs_launch_front_apply(sl.ws, l, gr, st);
    if ((rc = mark(kNames[6]))) return rc;
    gs_launch_relabel(sl.ws, l, gc, st);
    if ((rc = mark(kNames[7]))) return rc;
  }
  if (s->feat) {
    gs_launch_gather_sample(sl.ws, s->grid_full, st);  // streaming kernel: wants every warp slot
    if ((rc = mark("gather_features"))) return rc;
  }
  gs_launch_finish(sl.ws, s->grid, st);
  if ((rc = mark("finish"))) return rc;
  CU(cudaGetLastError());
  return GS_OK;
}

int stage_input(gs_slicer *s, Slot &sl, const int64_t *seeds, int32_t n, uint64_t batch_index) {
  if (n < 0 || n > s->cfg.max_batch)
    return fail(GS_ERR_INVALID, "batch of %d seeds exceeds max_batch=%d", n, s->cfg.max_batch);
  if (n && !seeds) return fail(GS_ERR_INVALID, "seeds_host is NULL");
  memcpy(sl.seeds, seeds, (size_t)n * sizeof(int64_t));
  sl.input->n_seeds = n;
  sl.input->want_host = s->cfg.host_output;
  sl.input->batch_index = batch_index;
  return GS_OK;
}

void compute_capacities(const gs_config &c, int64_t N, int64_t M, int64_t *Fcap, int64_t *Ccap,
                        int64_t *slab, int64_t *fslab, int64_t *Fb, int64_t *Cb) {
  int64_t F = c.max_batch, fmax = F, cmax = 0, sl = 8, fs = F + 8;
  const int64_t P = c.num_parts;
  for (int l = 0; l < c.num_layers; l++) {
    F = std::min(F, N);
    fmax = std::max(fmax, F);
    const int64_t fan = c.fanouts[l];
    const int64_t E = fan >= 0 ? std::min(F * fan, M) : M;
    const int64_t C = F + E;
    cmax = std::max(cmax, C);
    Fb[l] = F;
    Cb[l] = C;
    sl += std::min(N, E) + E + 3 * std::min(P * F, E) + P + 4 * F + 4 * GS_NUM_FIELDS * P + 8;
    fs += std::min(N, C) + 8;
    F = std::min(N, C);
  }
  *Fcap = c.frontier_capacity > 0 ? c.frontier_capacity : fmax;
  *Ccap = c.cand_capacity > 0 ? c.cand_capacity : cmax;
  *slab = c.slab_capacity > 0 ? c.slab_capacity : sl;
  *fslab = fs;
}

}  // namespace

extern "C" {

const char *gs_last_error(void) { return g_err.c_str(); }
int gs_abi_version(void) { return GS_ABI_VERSION; }

void gs_config_default(gs_config *c) {
  memset(c, 0, sizeof *c);
  c->abi_version = GS_ABI_VERSION;
  c->num_layers = 3;
  for (int i = 0; i < GS_MAX_LAYERS; i++) c->fanouts[i] = 10;
  c->num_parts = 4;
  c->rng_mode = GS_RNG_PHILOX;
  c->max_batch = 4096;
  c->num_slots = 4;
  c->num_streams = 4;
  c->host_output = 1;
  c->use_graph = 1;
  c->l2_hints = 1;
  c->host_copy_engine = 1;
}

int gs_graph_create(int device, int64_t num_nodes, int64_t num_edges,
                    const int64_t *indptr_host, const int64_t *indices_host, gs_graph **out) {
  if (!out || !indptr_host || (num_edges && !indices_host))
    return fail(GS_ERR_INVALID, "NULL argument");
  if (num_nodes < 1 || num_nodes >= (1ll << 31) - 1 || num_edges < 0)
    return fail(GS_ERR_INVALID, "num_nodes must be in [1, 2^31-2] (int32 ids on device)");
  if (indptr_host[0] != 0 || indptr_host[num_nodes] != num_edges)
    return fail(GS_ERR_INVALID, "indptr[0] != 0 or indptr[N] != num_edges");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(GS_ERR_CUDA, "no CUDA device (%s): this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  CU(cudaSetDevice(device));
  gs_graph *g = new (std::nothrow) gs_graph();
  if (!g) return fail(GS_ERR_INVALID, "out of host memory");
  g->device = device;
  g->num_nodes = num_nodes;
  g->num_edges = num_edges;
  CU(cudaMalloc((void **)&g->indptr, (size_t)(num_nodes + 1) * sizeof(int64_t)));
  CU(cudaMemcpy(g->indptr, indptr_host, (size_t)(num_nodes + 1) * sizeof(int64_t),
                cudaMemcpyHostToDevice));
  CU(cudaMalloc((void **)&g->indices, (size_t)std::max<int64_t>(num_edges, 1) * sizeof(int32_t)));
  // narrow int64 -> int32 on the device, in bounded chunks
  const int64_t chunk = 1ll << 26;
  int64_t *tmp = nullptr;
  CU(cudaMalloc((void **)&tmp, (size_t)std::min(chunk, std::max<int64_t>(num_edges, 1)) *
                                   sizeof(int64_t)));
  for (int64_t o = 0; o < num_edges; o += chunk) {
    const int64_t n = std::min(chunk, num_edges - o);
    CU(cudaMemcpy(tmp, indices_host + o, (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice));
    gs_launch_narrow_ids(tmp, g->indices + o, (size_t)n, 0);
    CU(cudaGetLastError());
  }
  CU(cudaDeviceSynchronize());
  CU(cudaFree(tmp));
  *out = g;
  return GS_OK;
}

int gs_graph_destroy(gs_graph *g) {
  if (!g) return GS_OK;
  cudaSetDevice(g->device);
  cudaFree(g->indptr);
  cudaFree(g->indices);
  delete g;
  return GS_OK;
}
int64_t gs_graph_num_nodes(const gs_graph *g) { return g ? g->num_nodes : 0; }
int64_t gs_graph_num_edges(const gs_graph *g) { return g ? g->num_edges : 0; }

}  // extern "C"

extern "C" {

int gs_slicer_destroy(gs_slicer *s) {
  if (!s) return GS_OK;
  cudaSetDevice(s->g->device);
  cudaDeviceSynchronize();
  for (Slot &sl : s->slots) {
    if (sl.exec) cudaGraphExecDestroy(sl.exec);
    if (sl.done) cudaEventDestroy(sl.done);
    if (sl.chain_done) cudaEventDestroy(sl.chain_done);
    if (sl.pack_done) cudaEventDestroy(sl.pack_done);
    for (void *p : sl.dev) cudaFree(p);
    if (sl.input) cudaFreeHost(sl.input);
    if (sl.seeds) cudaFreeHost(sl.seeds);
    if (sl.hdr_host) cudaFreeHost(sl.hdr_host);
    if (sl.slab_host) cudaFreeHost(sl.slab_host);
  }
  for (cudaStream_t st : s->streams) cudaStreamDestroy(st);
  if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
  cudaFree(s->owner);
  cudaFree(s->draws);
  cudaFree(s->draw_cursor);
  delete s;
  return GS_OK;
}

int gs_slicer_create(gs_graph *g, const gs_config *cfg, gs_slicer **out) {
  if (!g || !cfg || !out) return fail(GS_ERR_INVALID, "NULL argument");
  if (cfg->abi_version != GS_ABI_VERSION)
    return fail(GS_ERR_INVALID, "gs_config.abi_version %d != %d", cfg->abi_version, GS_ABI_VERSION);
  if (cfg->num_layers < 1 || cfg->num_layers > GS_MAX_LAYERS)
    return fail(GS_ERR_INVALID, "num_layers must be in [1,%d]", GS_MAX_LAYERS);
  if (cfg->num_parts < 1 || cfg->num_parts > GS_MAX_PARTS)
    return fail(GS_ERR_INVALID, "num_parts must be in [1,%d]", GS_MAX_PARTS);
  for (int l = 0; l < cfg->num_layers; l++)
    if (cfg->fanouts[l] < -1 || cfg->fanouts[l] == 0 || cfg->fanouts[l] > 1023)
      return fail(GS_ERR_INVALID, "fanouts[%d]=%d: use -1 (all neighbours) or 1..1023", l,
                  cfg->fanouts[l]);
  if (cfg->max_batch < 1) return fail(GS_ERR_INVALID, "max_batch must be >= 1");
  if (cfg->num_slots < 1 || cfg->num_slots > 256)
    return fail(GS_ERR_INVALID, "num_slots must be in [1,256]");
  if (cfg->rng_mode != GS_RNG_PHILOX && cfg->rng_mode != GS_RNG_DRAWS &&
      cfg->rng_mode != GS_RNG_PHILOX_NOREPLACE)
    return fail(GS_ERR_INVALID, "unknown rng_mode %d", cfg->rng_mode);
  if (cfg->rng_mode == GS_RNG_DRAWS && (!cfg->draws_host || cfg->n_draws == 0))
    return fail(GS_ERR_INVALID, "GS_RNG_DRAWS needs draws_host / n_draws");
  CU(cudaSetDevice(g->device));

  gs_slicer *s = new (std::nothrow) gs_slicer();
  if (!s) return fail(GS_ERR_INVALID, "out of host memory");
  s->g = g;
  s->cfg = *cfg;
  s->cfg.owner_host = nullptr;
  s->cfg.draws_host = nullptr;
  s->cfg.features = nullptr;
  const int64_t N = g->num_nodes, M = g->num_edges;
  int64_t Fb[GS_MAX_LAYERS] = {0}, Cb[GS_MAX_LAYERS] = {0};
  compute_capacities(*cfg, N, M, &s->Fcap, &s->Ccap, &s->slab_cap, &s->fslab_cap, Fb, Cb);
  if (s->Ccap >= (1ll << 31) - 2 * GS_TILE || s->Fcap >= (1ll << 31) - 2 * GS_TILE ||
      s->slab_cap >= (1ll << 31) - 16) {
    delete s;
    return fail(GS_ERR_INVALID, "derived capacities exceed 2^31 positions; lower max_batch or set "
                                "cand_capacity / slab_capacity");
  }
  if (cfg->features) {
    if (cfg->features->device != g->device) {
      delete s;
      return fail(GS_ERR_INVALID, "features live on device %d, graph on %d", cfg->features->device,
                  g->device);
    }
    s->feat = cfg->features;
    const int last = cfg->num_layers - 1;
    const int64_t e_last = Cb[last] - Fb[last];
    s->feat_cap_rows = cfg->feature_rows_capacity > 0 ? cfg->feature_rows_capacity
                                                      : std::max<int64_t>(1, std::min(N, e_last));
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, g->device);
  // grid_ctas = CTAs per SM of a full-machine launch.  Sizes are only known on
  // the device, so each launch gets the grid its layer's static upper bound
  // needs, capped at the full-machine grid (tile loops stride over the rest).
  s->grid = (cfg->grid_ctas > 0 ? cfg->grid_ctas : 4) * sms;
  s->grid_full = 8 * sms;
  for (int l = 0; l < cfg->num_layers; l++) {
    const int64_t fb = std::min(Fb[l], s->Fcap), cb = std::min(Cb[l], s->Ccap);
    s->grid_rows[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (fb + GS_TILE - 1) / GS_TILE));
    s->grid_cand[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (cb + GS_TILE - 1) / GS_TILE));
  }

#define CUS(call)                                                                        \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      fail(GS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
           __LINE__);                                                                    \
      std::string keep = g_err;                                                          \
      gs_slicer_destroy(s);                                                              \
      g_err = keep;                                                                      \
      return GS_ERR_CUDA;                                                                \
    }                                                                                    \
  } while (0)

  if (cfg->owner_host) {
    std::vector<uint8_t> o8((size_t)N);
    for (int64_t i = 0; i < N; i++) {
      const int32_t o = cfg->owner_host[i];
      if (o < 0 || o >= cfg->num_parts) {
        gs_slicer_destroy(s);
        return fail(GS_ERR_INVALID, "owner_host[%lld]=%d outside [0,%d)", (long long)i, o,
                    cfg->num_parts);
      }
      o8[(size_t)i] = (uint8_t)o;
    }
    CUS(cudaMalloc((void **)&s->owner, (size_t)N));
    CUS(cudaMemcpy(s->owner, o8.data(), (size_t)N, cudaMemcpyHostToDevice));
  }
  if (cfg->rng_mode == GS_RNG_DRAWS) {
    CUS(cudaMalloc((void **)&s->draws, (size_t)cfg->n_draws * sizeof(uint32_t)));
    CUS(cudaMemcpy(s->draws, cfg->draws_host, (size_t)cfg->n_draws * sizeof(uint32_t),
                   cudaMemcpyHostToDevice));
  }
  CUS(cudaMalloc((void **)&s->draw_cursor, sizeof(uint64_t)));
  CUS(cudaMemset(s->draw_cursor, 0, sizeof(uint64_t)));

  // replay mode consumes one sequential draw stream: keep every slot on one stream
  const int nstreams =
      cfg->rng_mode == GS_RNG_DRAWS ? 1 : std::max(1, std::min(cfg->num_streams, cfg->num_slots));
  s->streams.resize(nstreams);
  for (int i = 0; i < nstreams; i++) {
    s->streams[i] = nullptr;
    CUS(cudaStreamCreateWithFlags(&s->streams[i], cudaStreamNonBlocking));
  }
  if (cfg->host_output && cfg->host_copy_engine)
    CUS(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));

  const int ftiles = s->grid;  // per-CTA count arrays: one entry per CTA of the largest grid
  const int ctiles = (int)((s->Ccap + GS_TILE - 1) / GS_TILE) + 1;  // tile_first: one per candidate tile
  s->slots.resize(cfg->num_slots);
  for (int k = 0; k < cfg->num_slots; k++) {
    Slot &sl = s->slots[k];
    GsWs &w = sl.ws;
    int64_t bytes = 0;
    w.indptr = g->indptr;
    w.indices = g->indices;
    w.owner = s->owner;
    w.num_nodes = (int32_t)N;
    w.P = cfg->num_parts;
    w.pmask = (cfg->num_parts & (cfg->num_parts - 1)) == 0 ? cfg->num_parts - 1 : -1;
    w.pmagic = (uint32_t)((0x100000000ull + (uint64_t)cfg->num_parts - 1) / (uint64_t)cfg->num_parts);
    w.L = cfg->num_layers;
    for (int l = 0; l < GS_MAX_LAYERS; l++) w.fanouts[l] = cfg->fanouts[l];
    w.rng_mode = cfg->rng_mode;
    w.l2_hints = cfg->l2_hints;
    w.seed_lo = (uint32_t)cfg->seed;
    w.seed_hi = (uint32_t)(cfg->seed >> 32);
    w.draws = s->draws;
    w.n_draws = cfg->n_draws;
    w.draw_cursor = s->draw_cursor;
    w.slab_cap = s->slab_cap;
    w.fslab_cap = s->fslab_cap;
    w.Fcap = (int32_t)s->Fcap;
    w.Ccap = (int32_t)s->Ccap;
    {  // tagged table when positions leave room for >= 4 samples' worth of tags in a 32-bit entry
      int S = 1;
      while (S < 32 && (1ll << S) <= std::max<int64_t>(s->Ccap, cfg->max_batch)) S++;
      const int64_t ntags = S < 32 ? (1ll << (32 - S)) : 1;
      w.tag_shift = (GS_TABLE_TAGS && ntags >= 4 * (int64_t)(cfg->num_layers + 1)) ? S : 0;
    }
    w.ftiles_max = ftiles;
    w.ctiles_max = s->grid;
#define DA(ptr, n)                                         \
  do {                                                     \
    if (dev_alloc(sl, &ptr, (size_t)(n), &bytes) != GS_OK) { \
      std::string keep = g_err;                            \
      gs_slicer_destroy(s);                                \
      g_err = keep;                                        \
      return GS_ERR_CUDA;                                  \
    }                                                      \
  } while (0)
    DA(w.hdr, 1);
    DA(w.table, N);
    DA(w.slab, s->slab_cap + 4);
    DA(w.fslab, s->fslab_cap + 4);
    DA(w.rowstart, s->Fcap);
    DA(w.rowdeg, s->Fcap);
    DA(w.off, s->Fcap + 1);
    DA(w.samp_rank, s->Fcap);
    DA(w.rowscan, (s->Fcap + 1) * cfg->num_parts);
    DA(w.cand, s->Ccap);
    DA(w.cflag, s->Ccap);
    DA(w.kill, s->Ccap);
    DA(w.irank, s->Ccap);
    DA(w.rowa, s->Fcap);
    DA(w.rowhas, s->Fcap + 4);
    DA(w.tile_first, ctiles + 1);
    DA(w.ftile_sum, ftiles);
    DA(w.ftile_samp, ftiles);
    DA(w.ftile_pre, 2 * (size_t)ftiles);
    DA(w.ctile_cnt, (size_t)GS_NC_MAX * s->grid);
    DA(w.ftile_cnt, (size_t)GS_NF_MAX * ftiles);
    DA(w.tickets, 8);
    DA(w.epoch, 2);
    if (s->feat) {
      w.feat = s->feat->rows;
      w.feat_slot = s->feat->slot_of;
      w.feat_host = s->feat->host_dev;
      w.feat_dim = s->feat->dim;
      w.feat_div = s->feat->id_div;
      w.feat_rows = s->feat->num_rows;
      w.feat_cap_rows = s->feat_cap_rows;
      DA(w.feat_out, (size_t)s->feat_cap_rows * s->feat->dim);
    }
#undef DA
    s->slot_bytes = bytes;
    CUS(cudaMemset(w.hdr, 0, sizeof(gs_sample_header)));
    CUS(cudaMemset(w.tickets, 0, 8 * sizeof(uint32_t)));
    CUS(cudaMemset(w.table, 0xFF, (size_t)N * sizeof(uint32_t)));
    CUS(cudaMemcpy(w.epoch, kEpochStart, sizeof(kEpochStart), cudaMemcpyHostToDevice));
    CUS(cudaMemset(w.kill, 0, (size_t)s->Ccap));
    CUS(cudaHostAlloc((void **)&sl.input, sizeof(GsInput), cudaHostAllocMapped));
    CUS(cudaHostAlloc((void **)&sl.seeds, (size_t)cfg->max_batch * sizeof(int64_t),
                      cudaHostAllocMapped));
    CUS(cudaHostAlloc((void **)&sl.hdr_host, sizeof(gs_sample_header), cudaHostAllocMapped));
    memset(sl.hdr_host, 0, sizeof(gs_sample_header));
    memset(sl.input, 0, sizeof(GsInput));
    if (cfg->host_output)
      CUS(cudaHostAlloc((void **)&sl.slab_host, (size_t)(s->slab_cap + 4) * sizeof(int32_t),
                        cudaHostAllocMapped));
    void *dp = nullptr;
    CUS(cudaHostGetDevicePointer(&dp, sl.input, 0));
    w.input = (const GsInput *)dp;
    CUS(cudaHostGetDevicePointer(&dp, sl.seeds, 0));
    w.seeds_in = (const int64_t *)dp;
    CUS(cudaHostGetDevicePointer(&dp, sl.hdr_host, 0));
    w.hdr_host = (gs_sample_header *)dp;
    w.slab_host = nullptr;
    if (sl.slab_host && !cfg->host_copy_engine) {  // mirrored by k_finish's own stores
      CUS(cudaHostGetDevicePointer(&dp, sl.slab_host, 0));
      w.slab_host = (int32_t *)dp;
    }
    sl.stream = s->streams[k % nstreams];
    CUS(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    CUS(cudaEventCreateWithFlags(&sl.chain_done, cudaEventDisableTiming));
    CUS(cudaEventCreateWithFlags(&sl.pack_done, cudaEventDisableTiming));
    if (cfg->use_graph) {
      cudaGraph_t graph = nullptr;
      CUS(cudaStreamBeginCapture(sl.stream, cudaStreamCaptureModeThreadLocal));
      int rc = enqueue_sample(s, sl, sl.stream, nullptr, nullptr);
      cudaError_t ce = cudaStreamEndCapture(sl.stream, &graph);
      if (rc != GS_OK || ce != cudaSuccess) {
        if (rc == GS_OK) fail(GS_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(ce));
        std::string keep = g_err;
        gs_slicer_destroy(s);
        g_err = keep;
        return GS_ERR_CUDA;
      }
      CUS(cudaGraphInstantiate(&sl.exec, graph, 0));
      cudaGraphDestroy(graph);
    }
  }
  CUS(cudaDeviceSynchronize());
#undef CUS
  *out = s;
  return GS_OK;
}

int gs_slicer_submit(gs_slicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                     int32_t *slot_out) {
  if (!s || !slot_out) return fail(GS_ERR_INVALID, "NULL argument");
  const GsRange trace("gs_slicer_submit");
  CU(cudaSetDevice(s->g->device));
  const int ns = (int)s->slots.size();
  int k = -1;
  for (int t = 0; t < ns; t++) {
    const int c = (s->next_slot + t) % ns;
    if (s->slots[c].state == SLOT_FREE) {
      k = c;
      break;
    }
  }
  if (k < 0) return fail(GS_ERR_BUSY, "all %d sample slots are in use; release one", ns);
  Slot &sl = s->slots[k];
  int rc = stage_input(s, sl, seeds_host, n, batch_index);
  if (rc != GS_OK) return rc;
  if (sl.exec) {
    CU(cudaGraphLaunch(sl.exec, sl.stream));
  } else {
    rc = enqueue_sample(s, sl, sl.stream, nullptr, nullptr);
    if (rc != GS_OK) return rc;
  }
  sl.copied = 0;
  if (s->copy_stream && sl.slab_host) {
    // mirror with the copy engine: its size must be known now, so move the running
    // maximum of slab_used (+2 %); gs_slicer_wait tops up the rare remainder
    CU(cudaEventRecord(sl.chain_done, sl.stream));
    CU(cudaStreamWaitEvent(s->copy_stream, sl.chain_done, 0));
    sl.copied = std::min<int64_t>(s->slab_cap, s->mirror_est);
    if (sl.copied > 0)
      CU(cudaMemcpyAsync(sl.slab_host, sl.ws.slab, (size_t)sl.copied * sizeof(int32_t),
                         cudaMemcpyDeviceToHost, s->copy_stream));
    CU(cudaEventRecord(sl.done, s->copy_stream));
  } else {
    CU(cudaEventRecord(sl.done, sl.stream));
  }
  sl.state = SLOT_BUSY;
  s->next_slot = (k + 1) % ns;
  *slot_out = k;
  return GS_OK;
}

static int finish_wait(gs_slicer *s, Slot &sl, const gs_sample_header **hdr_out) {
  sl.state = SLOT_DONE;
  const int st = sl.hdr_host->status ? sl.hdr_host->status : sl.hdr_host->late_status;
  if (hdr_out) *hdr_out = sl.hdr_host;
  if (st != 0) {
    // a sample that stopped half way may have left touched table entries behind
    cudaMemsetAsync(sl.ws.table, 0xFF, (size_t)s->g->num_nodes * sizeof(uint32_t), sl.stream);
    cudaMemcpyAsync(sl.ws.epoch, kEpochStart, sizeof(kEpochStart), cudaMemcpyHostToDevice, sl.stream);
    cudaMemsetAsync(sl.ws.kill, 0, (size_t)s->Ccap, sl.stream);
    cudaStreamSynchronize(sl.stream);
    if (st == GS_ERR_OVERFLOW)
      return fail(GS_ERR_OVERFLOW,
                  "sample outgrew its workspace (frontier_capacity=%lld cand_capacity=%lld "
                  "slab_capacity=%lld feature_rows_capacity=%lld): raise them in gs_config",
                  (long long)s->Fcap, (long long)s->Ccap, (long long)s->slab_cap,
                  (long long)s->feat_cap_rows);
    if (st == GS_ERR_SEED)
      return fail(GS_ERR_SEED, "seed ids must be distinct and in [0, num_nodes)");
    return fail(st, "device raised status %d", st);
  }
  if (s->copy_stream && sl.slab_host) {
    const int64_t used = sl.hdr_host->slab_used;
    if (used > sl.copied) {  // first samples, or a sample larger than any before it
      cudaError_t e = cudaMemcpyAsync(sl.slab_host + sl.copied, sl.ws.slab + sl.copied,
                                      (size_t)(used - sl.copied) * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                      s->copy_stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(s->copy_stream);
      if (e != cudaSuccess) return fail(GS_ERR_CUDA, "host mirror failed: %s", cudaGetErrorString(e));
      sl.copied = used;
    }
    s->mirror_est = std::max(s->mirror_est, used + used / 50 + 1024);
  }
  return GS_OK;
}

int gs_slicer_wait(gs_slicer *s, int32_t slot, const gs_sample_header **hdr_out) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  Slot &sl = s->slots[slot];
  if (sl.state == SLOT_FREE) return fail(GS_ERR_STATE, "slot %d holds no sample", slot);
  const GsRange trace("gs_slicer_wait");
  CU(cudaSetDevice(s->g->device));
  if (sl.state == SLOT_BUSY) CU(cudaEventSynchronize(sl.done));
  return finish_wait(s, sl, hdr_out);
}

int gs_slicer_slab(gs_slicer *s, int32_t slot, const int32_t **device_slab,
                   const int32_t **host_slab) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  Slot &sl = s->slots[slot];
  if (sl.state != SLOT_DONE) return fail(GS_ERR_STATE, "slot %d not waited on", slot);
  if (device_slab) *device_slab = sl.ws.slab;
  if (host_slab) *host_slab = sl.slab_host;
  return GS_OK;
}

int gs_slicer_frontiers(gs_slicer *s, int32_t slot, const int32_t **device_frontiers) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  Slot &sl = s->slots[slot];
  if (sl.state != SLOT_DONE) return fail(GS_ERR_STATE, "slot %d not waited on", slot);
  if (device_frontiers) *device_frontiers = sl.ws.fslab;
  return GS_OK;
}

int gs_slicer_read_frontier(gs_slicer *s, int32_t slot, int32_t layer, int32_t *host_out) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  Slot &sl = s->slots[slot];
  if (sl.state != SLOT_DONE) return fail(GS_ERR_STATE, "slot %d not waited on", slot);
  if (layer < 0 || layer > s->cfg.num_layers) return fail(GS_ERR_INVALID, "bad layer");
  const gs_span sp = sl.hdr_host->frontier[layer];
  if (sp.len == 0) return GS_OK;
  if (!host_out) return fail(GS_ERR_INVALID, "NULL argument");
  CU(cudaSetDevice(s->g->device));
  CU(cudaMemcpy(host_out, sl.ws.fslab + sp.off, (size_t)sp.len * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return GS_OK;
}

int gs_slicer_release(gs_slicer *s, int32_t slot) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  Slot &sl = s->slots[slot];
  if (sl.state == SLOT_BUSY) {
    CU(cudaSetDevice(s->g->device));
    CU(cudaEventSynchronize(sl.done));
    finish_wait(s, sl, nullptr);
  }
  sl.state = SLOT_FREE;
  return GS_OK;
}

static int scatter_parts(gs_slicer *s, int32_t slot, const GsPartDst &dst, int64_t block_words,
                         void *after_event, void *consumer_stream) {
  Slot &sl = s->slots[slot];
  if (sl.state == SLOT_FREE) return fail(GS_ERR_STATE, "slot %d holds no sample", slot);
  const GsRange trace("gs_slicer_pack_parts");
  CU(cudaSetDevice(s->g->device));
  if (after_event) CU(cudaStreamWaitEvent(sl.stream, (cudaEvent_t)after_event, 0));
  // on the slot's own stream: after the sample's kernels, before the slot's next sample.
  // >= 4 parts (most stores remote): bulk-copy engine, one CTA per SM over all parts keeps
  // ~4.7 MB in flight (4 B200: 25.2 K minibatches/s against 24.7 K with plain stores);
  // fewer parts: plain 16-byte stores, 8 CTAs per SM (2 B200: 14.2 K against 13.2 K)
  // GSPLIT_PACK_TMA=0/1 in the environment overrides the choice (tests force both paths on 2 ranks)
  const char *ov = getenv("GSPLIT_PACK_TMA");
  const bool tma = GS_PACK_TMA && (ov ? atoi(ov) != 0 : s->cfg.num_parts >= 4);
  const int pack_ctas = tma ? std::max(2, s->grid_full / 8) : s->grid_full;
  gs_launch_pack_parts(sl.ws, dst, block_words, std::max(1, pack_ctas / std::max(1, s->cfg.num_parts)), tma,
                       sl.stream);
  CU(cudaGetLastError());
  CU(cudaEventRecord(sl.pack_done, sl.stream));
  CU(cudaStreamWaitEvent((cudaStream_t)consumer_stream, sl.pack_done, 0));  // NULL: the default stream waits
  return GS_OK;
}

int gs_slicer_pack_parts(gs_slicer *s, int32_t slot, int32_t *dst_device, int64_t block_words,
                         void *consumer_stream) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  if (!dst_device || block_words < GS_PART_HEADER_WORDS || (block_words & 3) ||
      ((uintptr_t)dst_device & 15))
    return fail(GS_ERR_INVALID, "pack_parts: dst must be 16-byte aligned, block_words a multiple of 4 and >= %d",
                (int)GS_PART_HEADER_WORDS);
  GsPartDst dst{};
  for (int g = 0; g < s->cfg.num_parts; g++) dst.blk[g] = dst_device + (int64_t)g * block_words;
  return scatter_parts(s, slot, dst, block_words, nullptr, consumer_stream);
}

int gs_slicer_scatter_parts(gs_slicer *s, int32_t slot, int32_t *const *dst_blocks, int64_t block_words,
                            void *after_event, void *consumer_stream) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  if (!dst_blocks || block_words < GS_PART_HEADER_WORDS || (block_words & 3))
    return fail(GS_ERR_INVALID, "scatter_parts: block_words must be a multiple of 4 and >= %d",
                (int)GS_PART_HEADER_WORDS);
  GsPartDst dst{};
  for (int g = 0; g < s->cfg.num_parts; g++) {
    if (!dst_blocks[g] || ((uintptr_t)dst_blocks[g] & 15))
      return fail(GS_ERR_INVALID, "scatter_parts: dst_blocks[%d] must be a 16-byte aligned device pointer", g);
    dst.blk[g] = dst_blocks[g];
  }
  return scatter_parts(s, slot, dst, block_words, after_event, consumer_stream);
}

int gs_ipc_alloc(int device, int64_t bytes, void **dev_ptr, void *handle_out) {
  if (!dev_ptr || !handle_out || bytes <= 0) return fail(GS_ERR_INVALID, "gs_ipc_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  CU(cudaSetDevice(device));
  void *p = nullptr;
  CU(cudaMalloc(&p, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(GS_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  memcpy(handle_out, &h, 64);
  *dev_ptr = p;
  return GS_OK;
}
int gs_ipc_open(int device, const void *handle, void **dev_ptr) {
  if (!dev_ptr || !handle) return fail(GS_ERR_INVALID, "gs_ipc_open: bad argument");
  CU(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  CU(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return GS_OK;
}
int gs_ipc_close(int device, void *dev_ptr) {
  if (!dev_ptr) return GS_OK;
  CU(cudaSetDevice(device));
  CU(cudaIpcCloseMemHandle(dev_ptr));
  return GS_OK;
}
int gs_ipc_free(int device, void *dev_ptr) {
  if (!dev_ptr) return GS_OK;
  CU(cudaSetDevice(device));
  CU(cudaFree(dev_ptr));
  return GS_OK;
}

int gs_slicer_profile(gs_slicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                      float *ms, const char **names, int32_t cap, int32_t *n_out) {
  if (!s || !ms || !names || !n_out) return fail(GS_ERR_INVALID, "NULL argument");
  CU(cudaSetDevice(s->g->device));
  Slot &sl = s->slots[0];
  if (sl.state != SLOT_FREE) return fail(GS_ERR_BUSY, "slot 0 must be free to profile");
  int rc = stage_input(s, sl, seeds_host, n, batch_index);
  if (rc != GS_OK) return rc;
  std::vector<cudaEvent_t> ev;
  std::vector<const char *> nm;
  rc = enqueue_sample(s, sl, sl.stream, &ev, &nm);
  if (rc == GS_OK) {
    cudaError_t e = cudaStreamSynchronize(sl.stream);
    if (e != cudaSuccess) rc = fail(GS_ERR_CUDA, "profile run failed: %s", cudaGetErrorString(e));
  }
  int cnt = 0;
  if (rc == GS_OK)
    for (size_t i = 1; i < ev.size() && cnt < cap; i++, cnt++) {
      cudaEventElapsedTime(&ms[cnt], ev[i - 1], ev[i]);
      names[cnt] = nm[i];
    }
  for (cudaEvent_t e : ev) cudaEventDestroy(e);
  *n_out = cnt;
  if (rc != GS_OK) return rc;
  sl.state = SLOT_BUSY;  // completed; let wait()/release() run the normal path
  CU(cudaEventRecord(sl.done, sl.stream));
  const gs_sample_header *h;
  rc = finish_wait(s, sl, &h);
  sl.state = SLOT_FREE;
  return rc;
}

int gs_features_create(int device, int64_t num_rows, int32_t dim, int32_t id_div,
                       const float *rows_host, gs_features **out) {
  if (!out || !rows_host) return fail(GS_ERR_INVALID, "NULL argument");
  if (num_rows < 1 || dim < 1 || id_div < 1) return fail(GS_ERR_INVALID, "num_rows, dim, id_div must be >= 1");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(GS_ERR_CUDA, "no CUDA device (%s): this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  CU(cudaSetDevice(device));
  gs_features *f = new (std::nothrow) gs_features();
  if (!f) return fail(GS_ERR_INVALID, "out of host memory");
  f->device = device;
  f->num_rows = num_rows;
  f->dim = dim;
  f->id_div = id_div;
  const size_t bytes = (size_t)num_rows * dim * sizeof(float);
  cudaError_t ce = cudaMalloc((void **)&f->rows, bytes);
  if (ce == cudaSuccess) ce = cudaMemcpy(f->rows, rows_host, bytes, cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) {
    cudaFree(f->rows);
    delete f;
    return fail(GS_ERR_CUDA, "feature upload failed: %s", cudaGetErrorString(ce));
  }
  *out = f;
  return GS_OK;
}

// Cached store: the table stays where the caller has it, in
// host memory, page-locked and mapped here; the rows named in cached_rows are
// copied to HBM.  Replaces MemoryManager.initialize + refresh_cache
// (python/utils/memory_manager.py:40-104): a miss is served by the gather
// kernel itself over PCIe, so nothing is refreshed or undone per minibatch.
int gs_features_create_cached(int device, int64_t num_rows, int32_t dim, int32_t id_div,
                              const float *rows_host, const int64_t *cached_rows, int64_t n_cached,
                              gs_features **out) {
  if (!out || !rows_host || (n_cached > 0 && !cached_rows)) return fail(GS_ERR_INVALID, "NULL argument");
  if (num_rows < 1 || dim < 1 || id_div < 1) return fail(GS_ERR_INVALID, "num_rows, dim, id_div must be >= 1");
  if (n_cached < 0 || n_cached > num_rows || num_rows >= (1ll << 31))
    return fail(GS_ERR_INVALID, "n_cached must be in [0, num_rows], num_rows < 2^31");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(GS_ERR_CUDA, "no CUDA device (%s): this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  std::vector<int32_t> slot((size_t)num_rows, -1);
  for (int64_t k = 0; k < n_cached; k++) {
    const int64_t r = cached_rows[k];
    if (r < 0 || r >= num_rows) return fail(GS_ERR_INVALID, "cached_rows[%lld]=%lld outside [0,num_rows)", (long long)k, (long long)r);
    if (slot[(size_t)r] >= 0) return fail(GS_ERR_INVALID, "cached_rows lists row %lld twice", (long long)r);
    slot[(size_t)r] = (int32_t)k;
  }
  CU(cudaSetDevice(device));
  gs_features *f = new (std::nothrow) gs_features();
  if (!f) return fail(GS_ERR_INVALID, "out of host memory");
  f->device = device;
  f->num_rows = num_rows;
  f->dim = dim;
  f->id_div = id_div;
  f->n_cached = n_cached;
  const size_t row_bytes = (size_t)dim * sizeof(float);
  auto bail = [&](cudaError_t ce, const char *what) {
    cudaGetLastError();
    gs_features_destroy(f);
    return fail(GS_ERR_CUDA, "%s failed: %s", what, cudaGetErrorString(ce));
  };
  // page-lock + map the caller's table in place (no copy); already-pinned memory is used as it is
  cudaError_t ce = cudaHostRegister((void *)rows_host, (size_t)num_rows * row_bytes,
                                    cudaHostRegisterMapped | cudaHostRegisterPortable);
  if (ce == cudaSuccess) f->host_reg = (void *)rows_host;
  else if (ce == cudaErrorHostMemoryAlreadyRegistered) cudaGetLastError();
  else return bail(ce, "cudaHostRegister of the feature table");
  void *alias = nullptr;
  if ((ce = cudaHostGetDevicePointer(&alias, (void *)rows_host, 0)) != cudaSuccess)
    return bail(ce, "cudaHostGetDevicePointer");
  f->host_dev = (const float *)alias;
  if ((ce = cudaMalloc((void **)&f->slot_of, (size_t)num_rows * sizeof(int32_t))) != cudaSuccess)
    return bail(ce, "cudaMalloc(slot map)");
  if ((ce = cudaMemcpy(f->slot_of, slot.data(), (size_t)num_rows * sizeof(int32_t), cudaMemcpyHostToDevice)) != cudaSuccess)
    return bail(ce, "slot map upload");
  if ((ce = cudaMalloc((void **)&f->rows, std::max<size_t>(1, (size_t)n_cached * row_bytes))) != cudaSuccess)
    return bail(ce, "cudaMalloc(cache rows)");
  // upload the cached rows in runs of consecutive source rows (a degree-ordered
  // cache of a degree-sorted graph is one run; an arbitrary set degrades to row copies)
  for (int64_t k = 0; k < n_cached;) {
    int64_t k1 = k + 1;
    while (k1 < n_cached && cached_rows[k1] == cached_rows[k1 - 1] + 1) k1++;
    ce = cudaMemcpyAsync(f->rows + (size_t)k * dim, rows_host + (size_t)cached_rows[k] * dim,
                         (size_t)(k1 - k) * row_bytes, cudaMemcpyHostToDevice, 0);
    if (ce != cudaSuccess) return bail(ce, "cache row upload");
    k = k1;
  }
  if ((ce = cudaStreamSynchronize(0)) != cudaSuccess) return bail(ce, "cache row upload");
  *out = f;
  return GS_OK;
}

int gs_features_info(const gs_features *f, int64_t *num_rows, int32_t *dim, int64_t *rows_in_hbm) {
  if (!f) return fail(GS_ERR_INVALID, "NULL argument");
  if (num_rows) *num_rows = f->num_rows;
  if (dim) *dim = f->dim;
  if (rows_in_hbm) *rows_in_hbm = f->slot_of ? f->n_cached : f->num_rows;
  return GS_OK;
}

int gs_features_destroy(gs_features *f) {
  if (!f) return GS_OK;
  cudaSetDevice(f->device);
  cudaFree(f->rows);
  cudaFree(f->slot_of);
  if (f->host_reg) cudaHostUnregister(f->host_reg);
  delete f;
  return GS_OK;
}

int gs_features_gather(gs_features *f, const int32_t *ids_device, int64_t n, float *out_device,
                       void *stream) {
  if (!f || (n > 0 && (!ids_device || !out_device))) return fail(GS_ERR_INVALID, "NULL argument");
  if (n <= 0) return GS_OK;
  CU(cudaSetDevice(f->device));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, f->device);
  const int grid = (int)std::min<int64_t>((n + GS_NW - 1) / GS_NW, (int64_t)sms * 8);
  gs_launch_gather_ids(f->rows, f->slot_of, f->host_dev, f->dim, f->id_div, f->num_rows, ids_device,
                       n, out_device, grid, (cudaStream_t)stream);
  CU(cudaGetLastError());
  return GS_OK;
}

int gs_slicer_feature_slab(gs_slicer *s, int32_t slot, const float **device_rows, int32_t *dim) {
  if (!s || slot < 0 || slot >= (int)s->slots.size()) return fail(GS_ERR_INVALID, "bad slot");
  if (!s->feat) return fail(GS_ERR_STATE, "slicer was created without gs_config.features");
  Slot &sl = s->slots[slot];
  if (sl.state != SLOT_DONE) return fail(GS_ERR_STATE, "slot %d not waited on", slot);
  if (device_rows) *device_rows = sl.ws.feat_out;
  if (dim) *dim = s->feat->dim;
  return GS_OK;
}

int gs_reference_shuffle(int64_t *nodes, int64_t n, int reset) {
  if (!nodes && n > 0) return fail(GS_ERR_INVALID, "NULL argument");
  if (reset) srand(1);
  for (int64_t i = 1; i < n; i++) std::swap(nodes[i], nodes[(int64_t)rand() % (i + 1)]);
  return GS_OK;
}

int gs_slicer_launches_per_sample(const gs_slicer *s) {
  return s ? 2 + 8 * s->cfg.num_layers + (s->feat ? 1 : 0) : 0;
}
int64_t gs_slicer_slot_bytes(const gs_slicer *s) { return s ? s->slot_bytes : 0; }
int64_t gs_slicer_slab_capacity(const gs_slicer *s) { return s ? s->slab_cap : 0; }

}  // extern "C"
