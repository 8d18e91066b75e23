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

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

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

// per-slot state at creation / after a reset: first tag index, first scan epoch
const GsState kStateStart = {1u, 1u, {0u, 0u}};

enum SlotState { SLOT_FREE = 0, SLOT_STAGED = 1, SLOT_BUSY = 2, SLOT_DONE = 3 };

struct Slot {
  GsWs ws{};
  std::vector<void *> dev;  // device allocations owned by the slot
  GsInput *input = nullptr; // pinned, mapped
  int64_t *seeds = nullptr; // pinned, mapped
  gs_sample_header *hdr_host = nullptr;
  int32_t *slab_host = nullptr;
  cudaEvent_t done = nullptr;
  cudaEvent_t pack_done = nullptr;   // gs_slicer_pack_parts: the part blocks are written
  int64_t copied = 0;                // ... int32 elements of the slab the DMA of this sample moved
  int group = 0;
  int state = SLOT_FREE;
  bool direct = false;               // submitted with gs_slicer_submit_direct: the lists are in the peers' blocks
};

// `members` consecutive slots share ONE launch sequence (grid.y = member): the
// kernels of small layers then fill the machine, and a minibatch costs
// (4 L + 1) / members launches.
struct Group {
  int first = 0, members = 0;        // slots [first, first + members)
  GsWs *wsv = nullptr;               // device copy of the members' GsWs
  cudaStream_t stream = nullptr;
  cudaGraphExec_t exec = nullptr;    // the full group
  cudaEvent_t chain_done = nullptr;  // kernels of the group's last launch finished
  int staged = 0;                    // members staged for the next launch
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
  int group_size = 1;
  int grid = 0;                          // CTAs per member of a full-machine launch
  int grid_gather = 0;                   // feature gather role of the last launch (streaming: wants every warp slot)
  int grid_rows[GS_MAX_LAYERS] = {0};    // per layer: enough CTAs for the largest possible frontier
  int grid_cand[GS_MAX_LAYERS] = {0};    // ... candidate stream
  int grid_relabel[GS_MAX_LAYERS] = {0}; // ... edge list (static upper bounds)
  int64_t Fcap = 0, Ccap = 0, slab_cap = 0, fslab_cap = 0;
  uint8_t *owner = nullptr;
  uint32_t *draws = nullptr;
  uint64_t *draw_cursor = nullptr;
  std::vector<cudaStream_t> streams;
  cudaStream_t copy_stream = nullptr;  // host mirror DMAs of all slots, one after another (interleaved
                                       // DMAs lose ~12 % of the link: 50 vs 57 GB/s measured)
  int64_t mirror_est = 0;              // running estimate of slab_used (int32 elements) + margin
  std::vector<Slot> slots;
  std::vector<Group> groups;
  cudaStream_t push_stream[GS_MAX_PARTS] = {nullptr};  // gs_slicer_push_parts: one copy-engine stream per part
  cudaEvent_t push_ev[GS_MAX_PARTS] = {nullptr};
  int cur_group = -1;                  // group being filled by gs_slicer_submit (-1: none)
  int next_group = 0;
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

// The static launch sequence of `members` minibatches of one group (grid.y =
// member): 4 launches per layer + 1.  `ev` (optional) receives one event after
// every launch, for gs_slicer_profile.
int enqueue_group(gs_slicer *s, Group &gr, int members, cudaStream_t st, std::vector<cudaEvent_t> *ev,
                  std::vector<const char *> *names) {
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
  const int L = s->cfg.num_layers;
  if ((rc = mark("start"))) return rc;
  const bool wide = s->cfg.num_parts > 4;
  // DIAGNOSTIC ONLY (tools/chain_cost.py): GSPLIT_DIAG_STOP=k launches the first k kernels of the
  // chain and a tail that only does the slot's housekeeping -- the samples are incomplete; the
  // difference of two such rates is the machine time kernel k costs with all slots in flight.
  static const int stop = [] {
    const char *e = getenv("GSPLIT_DIAG_STOP");
    return e && *e ? atoi(e) : 1 << 20;
  }();
  // GSPLIT_DIAG_ROLES: bit 0 drops the relabel roles, bit 1 the row-side roles (lists left unwritten)
  static const int roles_off = [] {
    const char *e = getenv("GSPLIT_DIAG_ROLES");
    return e && *e ? atoi(e) : 0;
  }();
  const bool relabel_on = !(roles_off & 1), rowside_on = !(roles_off & 2);
  int nl = 0;
  for (int l = 0; l < L; l++) {
    if (nl++ < stop) gs_launch_rows(gr.wsv, members, l, s->grid_rows[l], st);
    if ((rc = mark("rows"))) return rc;
    if (nl++ < stop)
      gs_launch_expand(gr.wsv, members, wide, s->cfg.rng_mode, s->cfg.fanouts[l] > 0, l, s->grid_cand[l],
                       l > 0 && relabel_on ? s->grid_relabel[l - 1] : 0, l > 0 && rowside_on ? s->grid_rows[l - 1] : 0, st);
    if ((rc = mark(l == 0 ? "expand" : "expand+relabel+rowside"))) return rc;
    if (nl++ < stop) gs_launch_cand(gr.wsv, members, wide, 0, l, s->grid_cand[l], s->grid_rows[l], st);
    if ((rc = mark("cand_flags+rowcount"))) return rc;
    if (nl++ < stop) gs_launch_cand(gr.wsv, members, wide, 1, l, s->grid_cand[l], 0, st);
    if ((rc = mark("cand_apply"))) return rc;
  }
  if (nl < stop)
    gs_launch_tail(gr.wsv, members, wide, relabel_on ? s->grid_relabel[L - 1] : 0, rowside_on ? s->grid_rows[L - 1] : 0,
                   s->feat ? s->grid_gather : (roles_off ? 1 : 0), st);
  else
    gs_launch_tail(gr.wsv, members, wide, 0, 0, 1, st);
  if ((rc = mark(s->feat ? "relabel+rowside+gather_features+finish" : "relabel+rowside+finish"))) return rc;
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
  sl.input->direct = 0;
  sl.direct = false;
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

// Launch the group being filled with the members staged so far.  A full group
// replays its CUDA graph; a partial one (gs_slicer_flush, or a wait on a staged
// slot) is launched eagerly with grid.y = staged, so members that are not part
// of the launch -- and whatever their slots still hold -- are not touched.
int launch_group(gs_slicer *s, Group &gr) {
  const int m = gr.staged;
  if (m <= 0) return GS_OK;
  if (gr.exec && m == gr.members) {
    CU(cudaGraphLaunch(gr.exec, gr.stream));
  } else {
    int rc = enqueue_group(s, gr, m, gr.stream, nullptr, nullptr);
    if (rc != GS_OK) return rc;
  }
  const bool mirror = s->copy_stream != nullptr;
  if (mirror) {
    CU(cudaEventRecord(gr.chain_done, gr.stream));
    CU(cudaStreamWaitEvent(s->copy_stream, gr.chain_done, 0));
  }
  CU(cudaGetLastError());
  for (int j = 0; j < m; j++) {
    Slot &sl = s->slots[gr.first + j];
    sl.copied = 0;
    if (mirror && sl.slab_host) {
      // mirror with the copy engine: its size must be known now, so move the running
      // maximum of slab_used (+2 %); gs_slicer_wait tops up the rare remainder
      sl.copied = std::min<int64_t>(s->slab_cap, s->mirror_est);
      if (sl.copied > 0)
        CU(cudaMemcpyAsync(sl.slab_host, sl.ws.slab, (size_t)sl.copied * sizeof(int32_t),
                           cudaMemcpyDeviceToHost, s->copy_stream));
      CU(cudaEventRecord(sl.done, s->copy_stream));
    } else {
      CU(cudaEventRecord(sl.done, gr.stream));
    }
    sl.state = SLOT_BUSY;
  }
  gr.staged = 0;
  s->cur_group = -1;
  return GS_OK;
}

int flush_if_staged(gs_slicer *s, Slot &sl) {
  if (sl.state != SLOT_STAGED) return GS_OK;
  return launch_group(s, s->groups[sl.group]);
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
  c->group_size = 1;
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
  // The kernels index indptr[] and the dedup table with the ids they read from
  // `indices`: a malformed CSR is rejected here, by one device pass over each
  // array BEFORE the ids are narrowed to int32 (Dataset's checksums compare sums
  // only, dataset.cpp:27,35).
  unsigned long long *err = nullptr;
  int64_t *tmp = nullptr;
  auto bail = [&](int code) {
    cudaFree(err);
    cudaFree(tmp);
    std::string keep = g_err;
    gs_graph_destroy(g);
    g_err = keep;
    return code;
  };
#define CUG(call)                                                                          \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess)                                                                 \
      return bail(fail(GS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                       __FILE__, __LINE__));                                               \
  } while (0)
  CUG(cudaMalloc((void **)&err, 2 * sizeof(unsigned long long)));
  CUG(cudaMemset(err, 0xFF, 2 * sizeof(unsigned long long)));
  CUG(cudaMalloc((void **)&g->indptr, (size_t)(num_nodes + 1) * sizeof(int64_t)));
  CUG(cudaMemcpy(g->indptr, indptr_host, (size_t)(num_nodes + 1) * sizeof(int64_t),
                 cudaMemcpyHostToDevice));
  gs_launch_check_indptr(g->indptr, num_nodes, num_edges, err, 0);
  CUG(cudaGetLastError());
  CUG(cudaMalloc((void **)&g->indices, (size_t)std::max<int64_t>(num_edges, 1) * sizeof(int32_t)));
  // narrow int64 -> int32 on the device, in bounded chunks
  const int64_t chunk = 1ll << 26;
  CUG(cudaMalloc((void **)&tmp, (size_t)std::min(chunk, std::max<int64_t>(num_edges, 1)) *
                                    sizeof(int64_t)));
  for (int64_t o = 0; o < num_edges; o += chunk) {
    const int64_t n = std::min(chunk, num_edges - o);
    CUG(cudaMemcpy(tmp, indices_host + o, (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice));
    gs_launch_check_ids(tmp, (size_t)n, o, num_nodes, err, 0);
    gs_launch_narrow_ids(tmp, g->indices + o, (size_t)n, 0);
    CUG(cudaGetLastError());
  }
  unsigned long long bad[2] = {0, 0};
  CUG(cudaMemcpy(bad, err, sizeof bad, cudaMemcpyDeviceToHost));  // synchronises the default stream
#undef CUG
  if (bad[0] != ~0ull)
    return bail(fail(GS_ERR_INVALID, "indptr is not a non-decreasing sequence inside [0, num_edges]: first bad row %llu",
                     bad[0] - 1ull));
  if (bad[1] != ~0ull)
    return bail(fail(GS_ERR_INVALID, "indices[%llu] lies outside [0, num_nodes)", bad[1] - 1ull));
  cudaFree(err);
  cudaFree(tmp);
  *out = g;
  return GS_OK;
}

// the reference's on-disk layout (dataset.cpp:18-36, written by
// convert_dgl_dataset.py:99-127) straight into HBM.  meta.txt gives the sizes and checksums;
// indptr.bin / indices.bin stream through two pinned staging buffers (pread -> H2D copy -> device
// passes), so the host never holds the graph -- 13 GB of int64 indices for papers100M -- and the
// device validates (monotone offsets, ids in range), sums (the checksums the reference compiles out
// under NDEBUG, dataset.cpp:27,35) and narrows the ids as they arrive.
int gs_graph_load(int device, const char *directory, int verify_checksums, gs_graph **out) {
  if (!out || !directory) return fail(GS_ERR_INVALID, "NULL argument");
  *out = nullptr;
  const std::string dir(directory);
  long long n = -1, m = -1, cs_off = 0, cs_edges = 0;
  bool has_off = false, has_edges = false;
  {
    FILE *f = fopen((dir + "/meta.txt").c_str(), "r");
    if (!f) return fail(GS_ERR_INVALID, "cannot open %s/meta.txt", directory);
    char line[512];
    while (fgets(line, sizeof line, f)) {
      char *eq = strchr(line, '=');
      if (!eq) continue;
      *eq = 0;
      const std::string key(line);
      char *end = nullptr;
      long long v = strtoll(eq + 1, &end, 10);
      if (end && (*end == '.' || *end == 'e' || *end == 'E')) v = (long long)strtod(eq + 1, nullptr);
      if (key == "num_nodes") n = v;
      else if (key == "num_edges") m = v;
      else if (key == "csum_offsets") { cs_off = v; has_off = true; }
      else if (key == "csum_edges") { cs_edges = v; has_edges = true; }
    }
    fclose(f);
  }
  if (n < 1 || n >= (1ll << 31) - 1 || m < 0) return fail(GS_ERR_INVALID, "meta.txt: num_nodes / num_edges missing or out of range");
  const int fd_p = open((dir + "/indptr.bin").c_str(), O_RDONLY), fd_x = open((dir + "/indices.bin").c_str(), O_RDONLY);
  auto close_fds = [&] { if (fd_p >= 0) close(fd_p); if (fd_x >= 0) close(fd_x); };
  struct stat sp, sx;
  if (fd_p < 0 || fd_x < 0 || fstat(fd_p, &sp) || fstat(fd_x, &sx) || sp.st_size != (off_t)((n + 1) * 8) ||
      sx.st_size != (off_t)(m * 8)) {
    close_fds();
    return fail(GS_ERR_INVALID, "indptr.bin / indices.bin missing or their size disagrees with meta.txt");
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1 || device < 0 || device >= ndev) {
    cudaGetLastError();
    close_fds();
    return fail(GS_ERR_CUDA, "no CUDA device %d (%s): this library has no CPU fallback", device,
                e != cudaSuccess ? cudaGetErrorString(e) : "device count");
  }
  gs_graph *g = new (std::nothrow) gs_graph();
  if (!g) { close_fds(); return fail(GS_ERR_INVALID, "out of host memory"); }
  g->device = device;
  g->num_nodes = n;
  g->num_edges = m;
  const size_t CH = 4u << 20;  // int64 elements per staging buffer (32 MiB)
  int64_t *pin[2] = {nullptr, nullptr}, *tmp[2] = {nullptr, nullptr};
  unsigned long long *err = nullptr;  // [0] bad row + 1, [1] bad edge + 1, [2] sum(indptr), [3] sum(indices)
  cudaStream_t st[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  auto cleanup = [&] {
    for (int i = 0; i < 2; i++) {
      if (pin[i]) cudaFreeHost(pin[i]);
      if (tmp[i]) cudaFree(tmp[i]);
      if (st[i]) cudaStreamDestroy(st[i]);
      if (ev[i]) cudaEventDestroy(ev[i]);
    }
    cudaFree(err);
    close_fds();
  };
  auto bail = [&](int code) {
    std::string keep = g_err;
    cleanup();
    gs_graph_destroy(g);
    g_err = keep;
    return code;
  };
#define CUL(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return bail(fail(GS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__)); \
  } while (0)
  CUL(cudaSetDevice(device));
  CUL(cudaMalloc((void **)&g->indptr, (size_t)(n + 1) * sizeof(int64_t)));
  CUL(cudaMalloc((void **)&g->indices, (size_t)std::max<long long>(m, 1) * sizeof(int32_t)));
  CUL(cudaMalloc((void **)&err, 4 * sizeof(unsigned long long)));
  {
    const unsigned long long init[4] = {~0ull, ~0ull, 0ull, 0ull};
    CUL(cudaMemcpy(err, init, sizeof init, cudaMemcpyHostToDevice));
  }
  for (int i = 0; i < 2; i++) {
    CUL(cudaHostAlloc((void **)&pin[i], CH * sizeof(int64_t), cudaHostAllocDefault));
    CUL(cudaMalloc((void **)&tmp[i], CH * sizeof(int64_t)));
    CUL(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));
    CUL(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
  }
  // one pass over a file: chunk k goes through staging buffer k & 1
  auto stream_file = [&](int fd, long long count, bool is_indices) -> int {
    int k = 0;
    for (long long o = 0; o < count; o += (long long)CH, k++) {
      const int b = k & 1;
      const size_t cnt = (size_t)std::min<long long>((long long)CH, count - o);
      CUL(cudaEventSynchronize(ev[b]));  // the copy that last read this pinned buffer is done
      size_t got = 0;
      while (got < cnt * 8) {
        const ssize_t r = pread(fd, (char *)pin[b] + got, cnt * 8 - got, (off_t)(o * 8 + (long long)got));
        if (r <= 0) return bail(fail(GS_ERR_INVALID, "short read in %s", is_indices ? "indices.bin" : "indptr.bin"));
        got += (size_t)r;
      }
      int64_t *dst = is_indices ? tmp[b] : g->indptr + o;
      CUL(cudaMemcpyAsync(dst, pin[b], cnt * 8, cudaMemcpyHostToDevice, st[b]));
      CUL(cudaEventRecord(ev[b], st[b]));
      gs_launch_sum_i64(dst, cnt, err + (is_indices ? 3 : 2), st[b]);
      if (is_indices) {
        gs_launch_check_ids(tmp[b], cnt, o, n, err, st[b]);
        gs_launch_narrow_ids(tmp[b], g->indices + o, cnt, st[b]);
        // the next use of tmp[b] is two chunks away on the same stream: ordered
      }
      CUL(cudaGetLastError());
    }
    return GS_OK;
  };
  int rc = stream_file(fd_p, n + 1, false);
  if (rc != GS_OK) return rc;
  CUL(cudaStreamSynchronize(st[0]));
  CUL(cudaStreamSynchronize(st[1]));
  gs_launch_check_indptr(g->indptr, n, m, err, st[0]);
  rc = stream_file(fd_x, m, true);
  if (rc != GS_OK) return rc;
  CUL(cudaStreamSynchronize(st[0]));
  CUL(cudaStreamSynchronize(st[1]));
  unsigned long long res[4] = {0, 0, 0, 0};
  long long first = -1, last = -1;
  CUL(cudaMemcpy(res, err, sizeof res, cudaMemcpyDeviceToHost));
  CUL(cudaMemcpy(&first, g->indptr, 8, cudaMemcpyDeviceToHost));
  CUL(cudaMemcpy(&last, g->indptr + n, 8, cudaMemcpyDeviceToHost));
#undef CUL
  if (first != 0 || last != m) return bail(fail(GS_ERR_INVALID, "indptr[0] != 0 or indptr[N] != num_edges"));
  if (res[0] != ~0ull)
    return bail(fail(GS_ERR_INVALID, "indptr is not a non-decreasing sequence inside [0, num_edges]: first bad row %llu",
                     res[0] - 1ull));
  if (res[1] != ~0ull) return bail(fail(GS_ERR_INVALID, "indices[%llu] lies outside [0, num_nodes)", res[1] - 1ull));
  if (verify_checksums) {
    if (has_off && (long long)res[2] != cs_off)
      return bail(fail(GS_ERR_INVALID, "indptr checksum mismatch: file sums to %lld, meta.txt says %lld", (long long)res[2], cs_off));
    if (has_edges && (long long)res[3] != cs_edges)
      return bail(fail(GS_ERR_INVALID, "indices checksum mismatch: file sums to %lld, meta.txt says %lld", (long long)res[3], cs_edges));
  }
  cleanup();
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

int gs_slicer_destroy(gs_slicer *s) {
  if (!s) return GS_OK;
  cudaSetDevice(s->g->device);
  cudaDeviceSynchronize();
  for (Group &gr : s->groups) {
    if (gr.exec) cudaGraphExecDestroy(gr.exec);
    if (gr.chain_done) cudaEventDestroy(gr.chain_done);
    cudaFree(gr.wsv);
  }
  for (Slot &sl : s->slots) {
    if (sl.done) cudaEventDestroy(sl.done);
    if (sl.pack_done) cudaEventDestroy(sl.pack_done);
    for (void *p : sl.dev) cudaFree(p);
    if (sl.input) cudaFreeHost(sl.input);
    if (sl.seeds) cudaFreeHost(sl.seeds);
    if (sl.hdr_host) cudaFreeHost(sl.hdr_host);
    if (sl.slab_host) cudaFreeHost(sl.slab_host);
  }
  for (cudaStream_t st : s->streams) cudaStreamDestroy(st);
  if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
  for (int q = 0; q < GS_MAX_PARTS; q++) {
    if (s->push_stream[q]) cudaStreamDestroy(s->push_stream[q]);
    if (s->push_ev[q]) cudaEventDestroy(s->push_ev[q]);
  }
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
  // replay mode consumes one sequential draw stream: one minibatch at a time, on one stream
  const int gsz = cfg->rng_mode == GS_RNG_DRAWS ? 1 : std::max(1, cfg->group_size);
  if (gsz > 16 || cfg->num_slots % gsz != 0)
    return fail(GS_ERR_INVALID, "group_size=%d must be in [1,16] and divide num_slots=%d", gsz, cfg->num_slots);
  CU(cudaSetDevice(g->device));
  gs_kernels_prepare();

  gs_slicer *s = new (std::nothrow) gs_slicer();
  if (!s) return fail(GS_ERR_INVALID, "out of host memory");
  s->g = g;
  s->cfg = *cfg;
  s->cfg.owner_host = nullptr;
  s->cfg.draws_host = nullptr;
  s->cfg.features = nullptr;
  s->group_size = gsz;
  const int64_t N = g->num_nodes, M = g->num_edges;
  int64_t Fb[GS_MAX_LAYERS] = {0}, Cb[GS_MAX_LAYERS] = {0};
  compute_capacities(*cfg, N, M, &s->Fcap, &s->Ccap, &s->slab_cap, &s->fslab_cap, Fb, Cb);
  // table entries are (tag | kind | payload): positions and row indices need a payload of
  // <= 30 bits, which leaves at least one tag
  if (s->Ccap >= (1ll << 30) - 2 * GS_TILE || s->Fcap >= (1ll << 30) - 2 * GS_TILE ||
      s->slab_cap >= (1ll << 31) - 16) {
    delete s;
    return fail(GS_ERR_INVALID, "derived capacities exceed 2^30 positions; lower max_batch or set "
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
  // grid_ctas = CTAs per SM and member of a full-machine launch.  Sizes are only
  // known on the device, so each launch gets the grid its layer's static upper
  // bound needs, capped at the full-machine grid (tile loops run over the rest).
  s->grid = (cfg->grid_ctas > 0 ? cfg->grid_ctas : 4) * sms;
  s->grid_gather = std::max(1, 8 * sms / gsz);
  for (int l = 0; l < cfg->num_layers; l++) {
    const int64_t fb = std::min(Fb[l], s->Fcap), cb = std::min(Cb[l], s->Ccap);
    const int64_t eb = std::max<int64_t>(1, cb - fb);
    s->grid_rows[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (fb + GS_TILE - 1) / GS_TILE));
    s->grid_cand[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (cb + GS_TILE - 1) / GS_TILE));
    s->grid_relabel[l] = (int)std::max<int64_t>(1, std::min<int64_t>(s->grid, (eb + GS_TILE - 1) / GS_TILE));
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

  const int ngroups = cfg->num_slots / gsz;
  const int nstreams =
      cfg->rng_mode == GS_RNG_DRAWS ? 1 : std::max(1, std::min(cfg->num_streams, ngroups));
  s->streams.resize(nstreams);
  for (int i = 0; i < nstreams; i++) {
    s->streams[i] = nullptr;
    CUS(cudaStreamCreateWithFlags(&s->streams[i], cudaStreamNonBlocking));
  }
  if (cfg->host_output) CUS(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));

  const int ftiles = (int)((s->Fcap + GS_TILE - 1) / GS_TILE) + 1;
  const int ctiles = (int)((s->Ccap + GS_TILE - 1) / GS_TILE) + 1;
  int pos_bits = 1;
  while (pos_bits < 30 && (1ll << pos_bits) <= std::max<int64_t>(std::max(s->Ccap, s->Fcap), cfg->max_batch)) pos_bits++;
  s->slots.resize(cfg->num_slots);
  s->groups.resize(ngroups);
  for (int k = 0; k < cfg->num_slots; k++) {
    Slot &sl = s->slots[k];
    GsWs &w = sl.ws;
    int64_t bytes = 0;
    sl.group = k / gsz;
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
    w.pos_bits = pos_bits;
    w.ftiles_cap = ftiles;
    w.ctiles_cap = ctiles;
#define DA(ptr, n)                                         \
  do {                                                     \
    if (dev_alloc(sl, &ptr, (size_t)(n), &bytes) != GS_OK) { \
      std::string keep = g_err;                            \
      gs_slicer_destroy(s);                                \
      g_err = keep;                                        \
      return GS_ERR_CUDA;                                  \
    }                                                      \
  } while (0)
    const int L = cfg->num_layers;
    DA(w.hdr, 1);
    DA(w.state, 1);
    DA(w.table, N);
    DA(w.slab, s->slab_cap + 4);
    DA(w.fslab, s->fslab_cap + 4);
    DA(w.rowinfo, s->Fcap);
    DA(w.off, s->Fcap + 1);
    DA(w.samp_rank, s->Fcap);
    DA(w.selfrank, 2 * (s->Fcap + 4));  // per-row arrays ping-pong by layer (rowside of layer l runs next to k_rows of l + 1)
    DA(w.ea, 2 * (s->Fcap + 4));
    DA(w.rowhas, 2 * (s->Fcap + 4));
    DA(w.rowscan, (s->Fcap + 1) * (cfg->num_parts > 4 ? 8 : 4));  // one record per row
    DA(w.cand, 2 * ((s->Ccap + 7) & ~(int64_t)3));  // ping-pong by layer
    DA(w.cflag, s->Ccap + 4);
    DA(w.ixs, s->Ccap + 64);
    DA(w.tgt, 1);
    DA(w.tile_first, ctiles + 1);
    DA(w.cnt, (size_t)L * GS_CNT_STRIDE);
    DA(w.tickets, (size_t)(L + 1) * GS_TICKETS);
    DA(w.rs_status, ftiles);
    DA(w.rs_val, 2 * (size_t)ftiles * GS_RS_NC);
    w.cgrid_cap = s->grid;
    DA(w.ccnt, (size_t)GS_CS_NC * s->grid);
    DA(w.fcnt, (size_t)GS_FS_NC * s->grid);
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
    CUS(cudaMemcpy(w.state, &kStateStart, sizeof(GsState), cudaMemcpyHostToDevice));
    CUS(cudaMemset(w.table, 0xFF, (size_t)N * sizeof(uint32_t)));
    CUS(cudaMemset(w.cnt, 0, (size_t)L * GS_CNT_STRIDE * sizeof(uint32_t)));
    CUS(cudaMemset(w.tickets, 0, (size_t)(L + 1) * GS_TICKETS * sizeof(uint32_t)));
    CUS(cudaMemset(w.rs_status, 0, (size_t)ftiles * sizeof(uint32_t)));
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
    CUS(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    CUS(cudaEventCreateWithFlags(&sl.pack_done, cudaEventDisableTiming));
  }
  for (int j = 0; j < ngroups; j++) {
    Group &gr = s->groups[j];
    gr.first = j * gsz;
    gr.members = gsz;
    gr.stream = s->streams[j % nstreams];
    CUS(cudaMalloc((void **)&gr.wsv, (size_t)gsz * sizeof(GsWs)));
    for (int m = 0; m < gsz; m++)
      CUS(cudaMemcpy(gr.wsv + m, &s->slots[gr.first + m].ws, sizeof(GsWs), cudaMemcpyHostToDevice));
    CUS(cudaEventCreateWithFlags(&gr.chain_done, cudaEventDisableTiming));
    if (cfg->use_graph) {
      cudaGraph_t graph = nullptr;
      CUS(cudaStreamBeginCapture(gr.stream, cudaStreamCaptureModeThreadLocal));
      int rc = enqueue_group(s, gr, gsz, gr.stream, nullptr, nullptr);
      cudaError_t ce = cudaStreamEndCapture(gr.stream, &graph);
      if (rc != GS_OK || ce != cudaSuccess) {
        if (rc == GS_OK) fail(GS_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(ce));
        std::string keep = g_err;
        gs_slicer_destroy(s);
        g_err = keep;
        return GS_ERR_CUDA;
      }
      CUS(cudaGraphInstantiate(&gr.exec, graph, 0));
      cudaGraphDestroy(graph);
    }
  }
  CUS(cudaDeviceSynchronize());
#undef CUS
  *out = s;
  return GS_OK;
}

struct DirectArgs {
  int32_t *const *dst_blocks;
  int64_t block_words;
  void *after_event;
};
static int submit_impl(gs_slicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                       int32_t *slot_out, const DirectArgs *da) {
  if (!s || !slot_out) return fail(GS_ERR_INVALID, "NULL argument");
  const GsRange trace("gs_slicer_submit");
  CU(cudaSetDevice(s->g->device));
  const int ng = (int)s->groups.size();
  if (s->cur_group < 0) {
    // the next group in ring order, as soon as its first slot is free again (a
    // released slot implies that the group's last launch has completed)
    int pick = -1;
    for (int t = 0; t < ng && pick < 0; t++) {
      const Group &cand = s->groups[(s->next_group + t) % ng];
      bool idle = true;
      for (int j = 0; j < cand.members; j++) {
        const int st = s->slots[cand.first + j].state;
        // a member that was launched but not waited on yet is complete as soon as slot 0 was
        // released (one launch sequence serves the whole group); it is restaged only once free
        if (st == SLOT_STAGED || (j == 0 && st != SLOT_FREE)) idle = false;
      }
      if (idle) pick = (s->next_group + t) % ng;
    }
    if (pick < 0)
      return fail(GS_ERR_BUSY, "all %d sample slots are in use; release one", (int)s->slots.size());
    s->cur_group = pick;
    s->next_group = (pick + 1) % ng;
    s->groups[pick].staged = 0;
  }
  Group &gr = s->groups[s->cur_group];
  const int k = gr.first + gr.staged;
  Slot &sl = s->slots[k];
  if (sl.state != SLOT_FREE)
    return fail(GS_ERR_BUSY, "slot %d (member %d of the group being filled) is still held; release it", k,
                gr.staged);
  int rc = stage_input(s, sl, seeds_host, n, batch_index);
  if (rc != GS_OK) return rc;
  if (da) {
    sl.input->direct = 1;
    sl.input->block_words = da->block_words;
    for (int g = 0; g < s->cfg.num_parts; g++) sl.input->part_dst[g] = da->dst_blocks[g];
    sl.direct = true;
    // the owners are done with these blocks once this event has completed
    if (da->after_event) CU(cudaStreamWaitEvent(gr.stream, (cudaEvent_t)da->after_event, 0));
  }
  sl.state = SLOT_STAGED;
  gr.staged++;
  *slot_out = k;
  if (gr.staged == gr.members) return launch_group(s, gr);
  return GS_OK;
}

int gs_slicer_submit(gs_slicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                     int32_t *slot_out) {
  return submit_impl(s, seeds_host, n, batch_index, slot_out, nullptr);
}

int gs_slicer_submit_direct(gs_slicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                            int32_t *const *dst_blocks, int64_t block_words, void *after_event,
                            int32_t *slot_out) {
  if (!s || !slot_out) return fail(GS_ERR_INVALID, "NULL argument");
  if (s->group_size != 1) return fail(GS_ERR_INVALID, "submit_direct: needs group_size = 1");
  if (s->cfg.host_output) return fail(GS_ERR_INVALID, "submit_direct: create the slicer with host_output = 0 (the lists leave the GPU)");
  if (s->feat) return fail(GS_ERR_INVALID, "submit_direct: not combined with the fused feature gather");
  if (!dst_blocks || block_words < GS_PART_HEADER_WORDS || (block_words & 3) || block_words > 0x7FFFFFF0ll)
    return fail(GS_ERR_INVALID, "submit_direct: block_words must be a multiple of 4 in [%d, 2^31)",
                (int)GS_PART_HEADER_WORDS);
  for (int g = 0; g < s->cfg.num_parts; g++)
    if (!dst_blocks[g] || ((uintptr_t)dst_blocks[g] & 15))
      return fail(GS_ERR_INVALID, "submit_direct: dst_blocks[%d] must be a 16-byte aligned device pointer", g);
  const DirectArgs da{dst_blocks, block_words, after_event};
  return submit_impl(s, seeds_host, n, batch_index, slot_out, &da);
}

int gs_slicer_flush(gs_slicer *s) {
  if (!s) return fail(GS_ERR_INVALID, "NULL argument");
  if (s->cur_group < 0) return GS_OK;
  CU(cudaSetDevice(s->g->device));
  return launch_group(s, s->groups[s->cur_group]);
}

static int finish_wait(gs_slicer *s, Slot &sl, const gs_sample_header **hdr_out) {
  sl.state = SLOT_DONE;
  {
    const cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) return fail(GS_ERR_CUDA, "sample failed on the device: %s", cudaGetErrorString(le));
  }
  const int st = sl.hdr_host->status ? sl.hdr_host->status : sl.hdr_host->late_status;
  if (hdr_out) *hdr_out = sl.hdr_host;
  if (st != 0) {
    // nothing to repair: a sample's tag, scan epochs and tickets are retired by its own
    // last launch, whatever it left in the table reads as stale from the next sample on
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
      // blocking copy on the legacy stream's behalf of THIS slot only: it must not queue
      // behind the mirrors of samples submitted later (those sit on copy_stream)
      cudaError_t e = cudaMemcpy(sl.slab_host + sl.copied, sl.ws.slab + sl.copied,
                                 (size_t)(used - sl.copied) * sizeof(int32_t), cudaMemcpyDeviceToHost);
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
  int rc = flush_if_staged(s, sl);
  if (rc != GS_OK) return rc;
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
  if (sl.state == SLOT_STAGED || sl.state == SLOT_BUSY) {
    CU(cudaSetDevice(s->g->device));
    int rc = flush_if_staged(s, sl);
    if (rc != GS_OK) return rc;
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
  int rc = flush_if_staged(s, sl);
  if (rc != GS_OK) return rc;
  cudaStream_t st = s->groups[sl.group].stream;
  if (sl.direct) {
    // gs_slicer_submit_direct: the sample's own kernels wrote the blocks; only the ordering is left
    CU(cudaEventRecord(sl.pack_done, st));
    CU(cudaStreamWaitEvent((cudaStream_t)consumer_stream, sl.pack_done, 0));
    return GS_OK;
  }
  if (after_event) CU(cudaStreamWaitEvent(st, (cudaEvent_t)after_event, 0));
  // on the group's own stream: after the sample's kernels, before the group's next launch.
  // >= 4 parts (most stores remote): bulk-copy engine, one CTA per SM over all parts keeps
  // ~4.7 MB in flight (4 B200: 25.2 K minibatches/s against 24.7 K with plain stores);
  // fewer parts: plain 16-byte stores, 8 CTAs per SM (2 B200: 14.2 K against 13.2 K)
  // GSPLIT_PACK_TMA=0/1 in the environment overrides the choice (tests force both paths on 2 ranks)
  const char *ov = getenv("GSPLIT_PACK_TMA");
  const bool tma = GS_PACK_TMA && (ov ? atoi(ov) != 0 : s->cfg.num_parts >= 4);
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->g->device);
  const int pack_ctas = tma ? std::max(2, sms) : 8 * sms;
  gs_launch_pack_parts(sl.ws, dst, block_words, std::max(1, pack_ctas / std::max(1, s->cfg.num_parts)), tma, st);
  CU(cudaGetLastError());
  CU(cudaEventRecord(sl.pack_done, st));
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

int gs_slicer_push_parts(gs_slicer *s, int32_t slot, int32_t *const *dst_blocks, void *after_event,
                         void *consumer_stream) {
  if (!s || slot < 0 || slot >= (int)s->slots.size() || !dst_blocks) return fail(GS_ERR_INVALID, "push_parts: bad argument");
  Slot &sl = s->slots[slot];
  if (!sl.direct || sl.state != SLOT_DONE)
    return fail(GS_ERR_STATE, "push_parts: slot %d must hold a sample submitted with gs_slicer_submit_direct and waited for", slot);
  const GsRange trace("gs_slicer_push_parts");
  CU(cudaSetDevice(s->g->device));
  const int P = s->cfg.num_parts;
  const int st0 = sl.hdr_host->status ? sl.hdr_host->status : sl.hdr_host->late_status;
  for (int q = 0; q < P; q++) {
    if (!dst_blocks[q]) return fail(GS_ERR_INVALID, "push_parts: dst_blocks[%d] is NULL", q);
    if (!s->push_stream[q]) {
      CU(cudaStreamCreateWithFlags(&s->push_stream[q], cudaStreamNonBlocking));
      CU(cudaEventCreateWithFlags(&s->push_ev[q], cudaEventDisableTiming));
    }
    // words of block q in use: left in the header by the sample's last launch (direct mode)
    const int64_t words = st0 == 0 ? (int64_t)sl.hdr_host->feature_part_off[q] : (int64_t)GS_PART_HEADER_WORDS;
    if (words < GS_PART_HEADER_WORDS || words > sl.input->block_words)
      return fail(GS_ERR_STATE, "push_parts: block %d reports %lld words in use", q, (long long)words);
    cudaStream_t st = s->push_stream[q];
    if (after_event) CU(cudaStreamWaitEvent(st, (cudaEvent_t)after_event, 0));
    if (dst_blocks[q] != sl.input->part_dst[q])
      CU(cudaMemcpyAsync(dst_blocks[q], sl.input->part_dst[q], (size_t)words * sizeof(int32_t), cudaMemcpyDefault, st));
    CU(cudaEventRecord(s->push_ev[q], st));
    CU(cudaStreamWaitEvent((cudaStream_t)consumer_stream, s->push_ev[q], 0));
  }
  return GS_OK;
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
  Group &gr = s->groups[0];
  for (int j = 0; j < gr.members; j++)
    if (s->slots[gr.first + j].state != SLOT_FREE)
      return fail(GS_ERR_BUSY, "the slots of group 0 must be free to profile");
  Slot &sl = s->slots[gr.first];
  int rc = stage_input(s, sl, seeds_host, n, batch_index);
  if (rc != GS_OK) return rc;
  std::vector<cudaEvent_t> ev;
  std::vector<const char *> nm;
  rc = enqueue_group(s, gr, 1, gr.stream, &ev, &nm);
  if (rc == GS_OK) {
    cudaError_t e = cudaStreamSynchronize(gr.stream);
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
  sl.state = SLOT_BUSY;  // completed; run the normal completion path
  CU(cudaEventRecord(sl.done, gr.stream));
  sl.copied = 0;
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

// Random-sector yardstick (roofline.random_access in bench.py): n_reads independent
// 4-byte reads, eight in flight per thread, from pseudo-random words of a table of
// table_bytes -- the access pattern of the neighbour picks and of the dedup-table
// look-ups.  With a table larger than L2 every read moves one 32-byte DRAM sector.
int gs_bench_sector_probe(int device, int64_t table_bytes, int64_t n_reads, double *reads_per_s) {
  if (!reads_per_s || table_bytes < 4 || n_reads < 1) return fail(GS_ERR_INVALID, "gs_bench_sector_probe: bad argument");
  CU(cudaSetDevice(device));
  uint32_t *tab = nullptr, *sink = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const uint64_t elems = (uint64_t)table_bytes / 4;
  int rc = GS_OK;
  cudaError_t e = cudaMalloc((void **)&tab, (size_t)elems * 4);
  if (e == cudaSuccess) e = cudaMalloc((void **)&sink, 4);
  if (e == cudaSuccess) e = cudaMemset(tab, 0x5A, (size_t)elems * 4);
  if (e == cudaSuccess) e = cudaEventCreate(&e0);
  if (e == cudaSuccess) e = cudaEventCreate(&e1);
  if (e == cudaSuccess) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int grid = sms * 8;
    gs_launch_sector_probe(tab, elems, (uint64_t)n_reads, 1u, sink, grid, 0);  // warm-up
    cudaEventRecord(e0, 0);
    for (uint32_t rep = 0; rep < 3; rep++) gs_launch_sector_probe(tab, elems, (uint64_t)n_reads, 2u + rep, sink, grid, 0);
    cudaEventRecord(e1, 0);
    e = cudaEventSynchronize(e1);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    if (e == cudaSuccess) *reads_per_s = 3.0 * (double)n_reads / ((double)ms * 1e-3);
  }
  if (e != cudaSuccess) rc = fail(GS_ERR_CUDA, "gs_bench_sector_probe: %s", cudaGetErrorString(e));
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(tab);
  cudaFree(sink);
  return rc;
}

int gs_slicer_launches_per_sample(const gs_slicer *s) {
  return s ? 4 * s->cfg.num_layers + 1 : 0;
}
int64_t gs_slicer_slot_bytes(const gs_slicer *s) { return s ? s->slot_bytes : 0; }
int64_t gs_slicer_slab_capacity(const gs_slicer *s) { return s ? s->slab_cap : 0; }

}  // extern "C"
