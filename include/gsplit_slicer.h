/*
 * gsplit_slicer.h -- C ABI of the B200-native minibatch slicer.
 *
 * This is the drop-in boundary for ONE path of sandeep06011991/OCC-GNN: the
 * `cslicer` module (per-minibatch neighbour sampling over CSR, frontier dedup +
 * relabel, split of every layer into per-GPU bipartites).  Each entry point
 * names the reference interface it replaces; citations are relative to
 * /root/reference/cslicer/.  Plain pointers and sizes only: no C++ or torch
 * types cross this boundary, no exceptions; every call returns a gs_status and
 * gs_last_error() describes the last failure on the calling thread.
 *
 * Execution model: a slicer owns `num_slots` sample slots (the reference's
 * bounded queue, WorkerPool.cpp:23-24), in groups of `group_size` consecutive
 * slots.  gs_slicer_submit() stages one minibatch in the next slot of the group
 * being filled; the submit that fills the group launches it -- ONE pre-captured
 * CUDA graph of 4 L + 1 kernels for all `group_size` minibatches (blockIdx.y =
 * member) on the group's stream -- and returns at once.  gs_slicer_wait() blocks
 * until that slot's sample is complete (launching a partly filled group first);
 * gs_slicer_release() hands the slot back.  group_size = 1 is the reference's
 * one-batch-at-a-time behaviour.  There is no CPU fallback: without a CUDA
 * device gs_graph_create() fails with GS_ERR_CUDA.
 */
#ifndef GSPLIT_SLICER_H
#define GSPLIT_SLICER_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GS_MAX_LAYERS 8
#define GS_MAX_PARTS 8
#define GS_ABI_VERSION 2

typedef enum {
  GS_OK = 0,
  GS_ERR_INVALID = 1,  /* bad argument / configuration */
  GS_ERR_CUDA = 2,     /* CUDA runtime failure (incl. no device) */
  GS_ERR_BUSY = 3,     /* no free sample slot: release one first */
  GS_ERR_OVERFLOW = 4, /* a sample outgrew its workspace: raise the capacities */
  GS_ERR_SEED = 5,     /* seed id out of range */
  GS_ERR_STATE = 6,    /* slot not in the state the call needs */
  GS_ERR_PEER = 7      /* partitioned mode: a peer rank failed or never reached a barrier */
} gs_status;

/* GS_RNG_PHILOX: picks WITH replacement (slicer.cpp:14-18), counter-based draws.
 * GS_RNG_DRAWS: the same rule fed from an explicit draw stream (mt19937 replay).
 * GS_RNG_PHILOX_NOREPLACE: the `fanout` picks of a row are distinct -- the rule
 * of the Python sibling's sampler (dgl sample_neighbors, replace=False,
 * python/utils/sampler.py:187); SURVEY 8(f) N1.  Parity with DGL's generator is
 * unpinned (DGL absent); the oracle restates this library's own permutation. */
typedef enum { GS_RNG_PHILOX = 0, GS_RNG_DRAWS = 1, GS_RNG_PHILOX_NOREPLACE = 2 } gs_rng_mode;

/* Per-bipartite lists.  The first seven are the reference's fields
 * (bipartite.h:9-24, exposed at pyfrontend.cpp:128-138); for from_ids/to_ids the
 * list is slot [own part], the only slot the reference ever fills
 * (slicer.cpp:41-42).  The reference's `indptr` (all ones, one per out node) and
 * `indices` (always empty) carry no information and are synthesised by the
 * host-side mirror; CSR_INDPTR / CSR_INDICES are the repaired block structure:
 * rows = out_nodes, indices = positions in in_nodes, edge multiplicity kept. */
typedef enum {
  GS_F_IN_NODES = 0,
  GS_F_OUT_NODES = 1,
  GS_F_OWNED_OUT = 2,
  GS_F_FROM_IDS = 3,
  GS_F_TO_IDS = 4,
  GS_F_SELF_IN = 5,
  GS_F_SELF_OUT = 6,
  GS_F_CSR_INDPTR = 7,
  GS_F_CSR_INDICES = 8,
  GS_NUM_FIELDS = 9
} gs_field_id;

typedef struct {
  int32_t off; /* element offset into the sample slab (int32 units) */
  int32_t len;
} gs_span;

/* One sample's table of contents.  Written by the device, mirrored to pinned
 * host memory at the end of every sample. */
typedef struct {
  int32_t status;      /* 0 or a gs_status raised on the device */
  int32_t num_layers;
  int32_t num_parts;
  int32_t n_seeds;
  uint64_t batch_index;
  uint64_t draw_cursor; /* GS_RNG_DRAWS: draws consumed so far (carried) */
  int64_t edges;        /* real (non-self) edges in this sample */
  int64_t slab_used;    /* int32 elements of the slab in use */
  int32_t frontier_len[GS_MAX_LAYERS + 1]; /* F_l: nodes entering layer l; [L] = input nodes */
  int32_t cand_len[GS_MAX_LAYERS];         /* C_l = F_l self entries + sampled entries */
  gs_span frontier[GS_MAX_LAYERS + 1];
  gs_span field[GS_MAX_LAYERS][GS_MAX_PARTS][GS_NUM_FIELDS];
  uint64_t draw_base[GS_MAX_LAYERS];
  int32_t late_status;   /* device-internal: error raised inside a tiled kernel */
  int32_t feature_rows;  /* rows written by the fused feature gather (0 without features) */
  int32_t feature_part_off[GS_MAX_PARTS + 1]; /* row offset of each part's block in the feature slab; a direct
                                               * submit (no fused gather): words in use of part q's block */
  int32_t reserved[3];   /* [0] host mirror wanted, [1] frontier slab elements in use, [2] 1: `frontier`
                          * spans index the frontier slab (gs_slicer_frontiers), 0: the sample slab */
} gs_sample_header;

typedef struct gs_graph gs_graph;
typedef struct gs_slicer gs_slicer;
typedef struct gs_features gs_features;

typedef struct {
  int32_t abi_version;             /* GS_ABI_VERSION */
  int32_t num_layers;              /* reference: 3 (slicer.cpp:75) */
  int32_t fanouts[GS_MAX_LAYERS];  /* reference: 10 (slicer.cpp:10,15); -1 = all neighbours */
  int32_t num_parts;               /* reference: 4 (bipartite.h:19) */
  int32_t rng_mode;                /* gs_rng_mode */
  uint64_t seed;                   /* philox key */
  const int32_t *owner_host;       /* [num_nodes] or NULL = id % num_parts (pyfrontend.cpp:57) */
  const uint32_t *draws_host;      /* GS_RNG_DRAWS: 32-bit draw stream (e.g. mt19937 replay) */
  uint64_t n_draws;
  int32_t max_batch;               /* minibatch_size (pyfrontend.cpp:42) */
  int32_t num_slots;               /* queue_size; samples in flight / unreleased */
  int32_t num_streams;             /* no_worker_threads; concurrent CUDA streams */
  int32_t host_output;             /* 1: mirror every sample's slab to pinned host memory */
  int64_t frontier_capacity;       /* 0 = derive from fanouts */
  int64_t cand_capacity;           /* 0 = derive */
  int64_t slab_capacity;           /* 0 = derive (int32 elements) */
  int32_t grid_ctas;               /* CTAs per SM of a full-machine launch; 0 = 4 */
  int32_t use_graph;               /* 1 (default via gs_config_default): CUDA-graph replay */
  int32_t l2_hints;                /* 1 (default): L2 eviction-priority hints, tables kept resident */
  int32_t host_copy_engine;        /* kept for ABI stability; the host_output mirror is always one DMA per sample on
                                    * a dedicated copy stream, sized from the running maximum of slab_used (the rare
                                    * remainder is topped up in gs_slicer_wait).  Measured D2H ceilings on this box:
                                    * copy engine 57.0 GB/s, stores issued by SMs or TMA 52.3 (the round-1 alternative) */
  gs_features *features;           /* optional: fuse the input-feature gather into every sample */
  int64_t feature_rows_capacity;   /* 0 = derive (rows of the last layer's in_nodes, all parts) */
  int32_t group_size;              /* minibatches per launch sequence (1..16, must divide num_slots; GS_RNG_DRAWS
                                    * forces 1): the kernels of `group_size` consecutive slots run as ONE grid each */
  int32_t reserved0;
} gs_config;


const char *gs_last_error(void);
int gs_abi_version(void);
void gs_config_default(gs_config *cfg);

/* Replaces Dataset::read_graph's in-memory product (dataset.cpp:18-36,
 * dataset.h:39-41): CSR in host memory -> device-resident graph (int64 row
 * offsets, int32 neighbour ids; num_nodes must be < 2^31). */
int gs_graph_create(int device, int64_t num_nodes, int64_t num_edges,
                    const int64_t *indptr_host, const int64_t *indices_host,
                    gs_graph **out);
/* SURVEY 8(f) row N4 -- replaces Dataset::read_graph (dataset.cpp:18-36) for the files
 * convert_dgl_dataset.py:99-127 writes: `directory`/meta.txt (num_nodes, num_edges,
 * csum_offsets, csum_edges), indptr.bin int64[N+1], indices.bin int64[M], streamed from
 * disk through two pinned staging buffers into HBM; the device validates, checksums
 * (verify_checksums != 0: the sums the reference only asserts in debug builds,
 * dataset.cpp:27,35) and narrows the ids as they arrive -- the host never holds the graph. */
int gs_graph_load(int device, const char *directory, int verify_checksums, gs_graph **out);
int gs_graph_destroy(gs_graph *g);
int64_t gs_graph_num_nodes(const gs_graph *g);
int64_t gs_graph_num_edges(const gs_graph *g);

/* Node features resident in HBM (replaces Dataset::features, dataset.cpp:38-48,
 * and the gather half of python/utils/memory_manager.py:75-104 /
 * python/data/bipartite.py:61-67).  rows_host: [num_rows x dim] fp32, row of
 * node id at index id / id_div (id_div = 1 for a full replica; = num_parts for a
 * shard that holds only the nodes with id % num_parts == part). */
int gs_features_create(int device, int64_t num_rows, int32_t dim, int32_t id_div,
                       const float *rows_host, gs_features **out);
/* The same store with the table left in HOST memory and an HBM cache in front
 * (replaces MemoryManager.initialize + refresh_cache,
 * python/utils/memory_manager.py:40-104).  rows_host is page-locked and mapped
 * in place (no copy; it must stay valid and unchanged until gs_features_destroy);
 * cached_rows: n_cached distinct row indices (id / id_div) whose rows are copied
 * to HBM.  Gathers read cached rows from HBM and every other row straight from
 * host memory over PCIe inside the same kernel -- no per-minibatch refresh. */
int gs_features_create_cached(int device, int64_t num_rows, int32_t dim, int32_t id_div,
                              const float *rows_host, const int64_t *cached_rows, int64_t n_cached,
                              gs_features **out);
int gs_features_info(const gs_features *f, int64_t *num_rows, int32_t *dim, int64_t *rows_in_hbm);
int gs_features_destroy(gs_features *f);
/* Standalone gather on device pointers: out[i, :] = rows[ids[i] / id_div, :].
 * stream: a cudaStream_t (NULL = default stream).  Asynchronous. */
int gs_features_gather(gs_features *f, const int32_t *ids_device, int64_t n,
                       float *out_device, void *stream);
/* Fused gather of a completed slot (gs_config.features set): the feature slab
 * holds, part after part, the rows of the LAST layer's in_nodes_g -- the input
 * block each GPU reads.  Offsets are in gs_sample_header.feature_part_off. */
int gs_slicer_feature_slab(gs_slicer *s, int32_t slot, const float **device_rows,
                           int32_t *dim);

/* Replaces CSlicer::CSlicer + WorkerPool::WorkerPool + Slicer::Slicer
 * (pyfrontend.cpp:41-70, WorkerPool.cpp:4-35, slicer.h:41-72): allocates the
 * per-slot dedup tables, workspaces and sample slabs and captures one CUDA graph
 * per slot. */
int gs_slicer_create(gs_graph *g, const gs_config *cfg, gs_slicer **out);
int gs_slicer_destroy(gs_slicer *s);

/* Replaces WorkerPool::run's push of one batch (WorkerPool.cpp:41-49) and the
 * Slicer::run iteration that consumes it (slicer.cpp:120-139).  seeds_host:
 * n int64 node ids, distinct.  Returns GS_ERR_BUSY when no group can take the
 * batch: every slot in use, or -- group_size > 1 -- the next member slot of the
 * group being filled is still held (groups are refilled in slot order). */
int gs_slicer_submit(gs_slicer *s, const int64_t *seeds_host, int32_t n,
                     uint64_t batch_index, int32_t *slot_out);
/* Launch the group being filled now, with the members staged so far (the stop
 * sentinel of WorkerPool.cpp:52-55: no more batches will come).  gs_slicer_wait
 * / release / pack_parts on a staged slot do this implicitly. */
int gs_slicer_flush(gs_slicer *s);

/* Replaces CSlicer::getSample / ConQueue::pop_object (pyfrontend.cpp:73-78,
 * conqueue.h:42-60) for one slot: blocks until the sample is complete, then
 * returns its header (host memory, valid until release). */
int gs_slicer_wait(gs_slicer *s, int32_t slot, const gs_sample_header **hdr_out);

/* Base pointers of a completed slot's slab: every gs_span indexes into these.
 * host_slab is NULL unless host_output was set. */
int gs_slicer_slab(gs_slicer *s, int32_t slot, const int32_t **device_slab,
                   const int32_t **host_slab);

/* Frontiers 0..L of a completed slot (gs_sample_header.frontier spans index this
 * device array).  They are by-products, not part of a reference Sample, so they are
 * not mirrored to the host; gs_slicer_read_frontier copies one out on demand. */
int gs_slicer_frontiers(gs_slicer *s, int32_t slot, const int32_t **device_frontiers);
int gs_slicer_read_frontier(gs_slicer *s, int32_t slot, int32_t layer, int32_t *host_out);

/* Replaces PySample's destructor (pybipartite.cpp:68-76). */
int gs_slicer_release(gs_slicer *s, int32_t slot);

/* Split-parallel hand-over (the consumer of Sample.layers[l][g] is GPU g,
 * python/train.py:120-160, python/data/bipartite.py:20-58): every list of part g of
 * the slot's sample is packed into one self-describing block of `block_words`
 * int32 at dst_device + g * block_words, so that the P blocks can leave in ONE
 * all-to-all over NVLink (block g -> rank g) while every rank keeps slicing whole
 * minibatches on a replicated topology.  Block layout (int32 words):
 *   [0] GS_PART_MAGIC  [1] status (0, or GS_ERR_OVERFLOW: block_words too small)
 *   [2] part  [3] num_layers  [4] num_parts  [5] n_seeds  [6,7] batch_index (lo, hi)
 *   [8] words in use  [9] edges of this part  [10..15] 0
 *   [16 + 2*(l*GS_NUM_FIELDS + f)], [.. + 1] = offset (words from the block start,
 *   16-byte aligned) and length of field f of layer l;  data from word
 *   GS_PART_HEADER_WORDS on.
 * The slot must hold a submitted sample (waiting for it is not required: the pack
 * runs on the slot's stream, after the sample's kernels).  consumer_stream (a
 * cudaStream_t; NULL = the default stream) is made to wait for the pack.  block_words % 4 == 0. */
#define GS_PART_MAGIC 0x31505347
#define GS_PART_HEADER_WORDS (16 + 2 * GS_MAX_LAYERS * GS_NUM_FIELDS)
int gs_slicer_pack_parts(gs_slicer *s, int32_t slot, int32_t *dst_device, int64_t block_words,
                         void *consumer_stream);

/* Fused form of the hand-over: block g is written STRAIGHT into dst_blocks[g] -- a
 * device pointer that may be PEER memory (a buffer of rank g mapped with
 * gs_ipc_open), so the parts cross NVLink inside the packing kernel's own 16-byte
 * stores and no send buffer or library collective touches the payload.
 * dst_blocks: host array of num_parts device pointers (16-byte aligned).
 * after_event (a recorded cudaEvent_t, may be NULL = none): the pack first waits for
 * it (the barrier after which the owners are done with these buffers). */
int gs_slicer_scatter_parts(gs_slicer *s, int32_t slot, int32_t *const *dst_blocks,
                            int64_t block_words, void *after_event, void *consumer_stream);

/* Direct hand-over: like gs_slicer_submit, but the sample's OWN kernels write every list
 * of part g straight into dst_blocks[g] (local or peer memory, e.g. a buffer of rank g
 * mapped with gs_ipc_open) in the block layout of gs_slicer_pack_parts -- each list
 * crosses NVLink once, as it is produced; nothing is written to the slot's slab and no
 * packing kernel runs.  after_event (a recorded cudaEvent_t or NULL): the sample's launch
 * sequence first waits for it (the owners are done with these blocks).  Needs group_size
 * = 1, host_output = 0, no fused feature gather.  gs_slicer_wait() then returns the
 * header only (its spans are relative to the part's block); gs_slicer_scatter_parts()
 * on such a slot moves nothing and only makes consumer_stream wait for the sample.
 * block_words too small -> the sample's status is GS_ERR_OVERFLOW (also in every block
 * header).  No counterpart in the reference (one process, host vectors). */
int gs_slicer_submit_direct(gs_slicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                            int32_t *const *dst_blocks, int64_t block_words, void *after_event,
                            int32_t *slot_out);

/* Staged hand-over: for a slot submitted with gs_slicer_submit_direct into LOCAL blocks
 * (a staging area: the sample's kernels then write each part's lists once, contiguously,
 * behind its block header) and waited for: copies block q -- only its words in use, which
 * the sample left in gs_sample_header.feature_part_off[q] -- to dst_blocks[q] (peer memory)
 * with the COPY ENGINES, one stream per part; the copies first wait for after_event (may be
 * NULL), consumer_stream waits for all of them.  No SM is involved in the exchange.  A block
 * whose destination equals its staging address is not copied. */
int gs_slicer_push_parts(gs_slicer *s, int32_t slot, int32_t *const *dst_blocks, void *after_event,
                         void *consumer_stream);

/* The hand-over's barrier without a collective (one warp on `stream`): first stores
 * signal_value into the n_signal words signal_words[i] (device pointers, local or peer
 * memory mapped with gs_ipc_open) -- ordered, system-wide, after everything the stream
 * did before -- then waits until every one of the n_wait LOCAL words wait_words[i] is
 * >= wait_value (wrap-around compare).  A wait that lasts longer than timeout_ms (0 =
 * 5000) sets *status_mapped (a pinned, mapped host word) to 1 and ends: no kernel hangs
 * on a peer that died.  n_signal, n_wait <= GS_MAX_PARTS.  No counterpart in the reference. */
int gs_flags_post(int device, uint32_t *const *signal_words, int32_t n_signal, uint32_t signal_value,
                  const uint32_t *const *wait_words, int32_t n_wait, uint32_t wait_value,
                  uint32_t timeout_ms, int32_t *status_mapped, void *stream);

/* Device buffers another process can map (cudaIpcGetMemHandle / OpenMemHandle):
 * handle = 64 bytes.  gs_ipc_open enables peer access to the owner's GPU. */
int gs_ipc_alloc(int device, int64_t bytes, void **dev_ptr, void *handle_out);
int gs_ipc_open(int device, const void *handle, void **dev_ptr);
int gs_ipc_close(int device, void *dev_ptr);
int gs_ipc_free(int device, void *dev_ptr);

/* Diagnostics: run one minibatch eagerly (no graph) on slot 0 with a CUDA event
 * between kernels; ms[i] receives the duration of launch i, names[i] a static
 * string.  Returns the number of launches through *n_out (<= cap). */
int gs_slicer_profile(gs_slicer *s, const int64_t *seeds_host, int32_t n,
                      uint64_t batch_index, float *ms, const char **names,
                      int32_t cap, int32_t *n_out);

/* Replaces the epoch shuffle of WorkerPool::run (WorkerPool.cpp:40):
 * std::random_shuffle over nodes[0..n) driven by the C library's rand(), i.e.
 * swap(nodes[i], nodes[rand() % (i+1)]) for i = 1..n-1.  reset != 0 first puts
 * rand() in the state of a process that never called srand() (the reference
 * never seeds it), so epoch e of a fresh run is reproduced by calling this e+1
 * times with reset only on the first call.  Host-only; no device work. */
int gs_reference_shuffle(int64_t *nodes, int64_t n, int reset);

/* Measurement aid ("each random 4/8-byte gather charged 32 B"): the rate
 * at which this GPU serves independent random 4-byte reads from a table of
 * table_bytes, eight in flight per thread -- one 32-byte DRAM sector each once the
 * table exceeds L2.  bench.py divides the slicer's sector-granular byte bound by it.
 * The reference has no counterpart (a CPU program). */
int gs_bench_sector_probe(int device, int64_t table_bytes, int64_t n_reads, double *reads_per_s);

/* ------------------------------------------------------------------------
 * the edge partitioning of the Python sibling's sampler
 * (Sampler.edge_partitioning, python/utils/sampler.py:64-134).  For one layer's
 * edges (src = the sampled neighbour or the self loop, dest = the node it feeds):
 *   every layer but the last  (last_layer = 0): the edge goes to the owner of src
 *     (:76); owned_nds[g] = sorted unique dest owned by g (:80-81); shuffle_nds[g][r] =
 *     sorted unique dest owned by r != g among g's edges (:83-87);
 *   the last layer (last_layer = 1): if src is resident in the feature cache of
 *     owner(dest) -- resident[src * P + owner(dest)], the reference's
 *     memory_manager.node_gpu_mask -- the edge stays with owner(dest), else it goes
 *     to owner(src) (:93-106); part g = its natural edges then its extra edges, each in
 *     edge order (:107-110); shuffle_nds over g's extra edges only (:114-119).
 * owner_host: int32[num_nodes] (workload_assignment); resident_host: uint8
 * [num_nodes * num_parts] row-major, or NULL if last_layer is never used.
 * gs_epart_run takes host int64 edge arrays (the reference's CPU tensors), runs on the
 * GPU and returns when the lists are in the handle's pinned host buffers (and still on
 * the device).  Errors: an endpoint outside [0, num_nodes) -> GS_ERR_INVALID. */
typedef struct gs_epart gs_epart;
enum { GS_EP_SRC = 0, GS_EP_DST = 1, GS_EP_OWNED = 2, GS_EP_SHUFFLE = 3, GS_EP_NUM_EXTRA = 4 };
int gs_epart_create(int device, int64_t num_nodes, int32_t num_parts, const int32_t *owner_host,
                    const uint8_t *resident_host, int64_t max_edges, gs_epart **out);
int gs_epart_destroy(gs_epart *e);
int gs_epart_run(gs_epart *e, const int64_t *src_host, const int64_t *dst_host, int64_t n_edges,
                 int32_t last_layer);
/* List `what` of part `part` (remote: the peer of GS_EP_SHUFFLE, else ignored) of the last
 * run: host / device pointers (either may be NULL) and length.  GS_EP_NUM_EXTRA: *len =
 * number of extra edges of the part (the reference's shuffle_nds is {} when it is 0). */
int gs_epart_get(gs_epart *e, int32_t part, int32_t what, int32_t remote, const int32_t **host,
                 const int32_t **device, int64_t *len);

/* Allocation-free frontier dedup.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * out_device[0..*count_host) = the distinct values of ids_device[0..n) in order of
 * first occurrence -- the rule by which the slicer appends new nodes to a frontier
 * (slicer.cpp:22-43), so for one seed whose neighbour list is `ids` it equals
 * frontier 1 of gs_slicer_submit() minus the seed.  ids in [0, num_nodes).
 * The dedup table, the per-CTA counts and the control words live in one static
 * device array of 4 MiB + 4.25 KiB that every GPU loading the library holds until
 * exit: num_nodes <= GS_NOALLOC_MAX_NODES and n <= GS_NOALLOC_MAX_IDS, larger
 * sizes return GS_ERR_INVALID.  Calls no cudaMalloc / cudaFree; the caller owns
 * ids_device and out_device (n int32 each, device memory).  Runs on the legacy
 * default stream and returns when the result is complete.  All calls on a device
 * share the array (cleared at the start of each call): calls that overlap on one
 * GPU race -- serialise them. */
#define GS_NOALLOC_MAX_NODES (1 << 20)
#define GS_NOALLOC_MAX_IDS (1 << 24)
int gs_unique_first_noalloc(int device, const int32_t *ids_device, int64_t n, int64_t num_nodes,
                            int32_t *out_device, int64_t *count_host);

/* Number of kernel launches one launch sequence performs (4 L + 1), shared by the
 * group_size minibatches of a group. */
int gs_slicer_launches_per_sample(const gs_slicer *s);
/* Device bytes held by one slot / by the slicer. */
int64_t gs_slicer_slot_bytes(const gs_slicer *s);
/* int32 elements one slot's sample slab can hold (gs_config.slab_capacity, or the
 * bound derived from max_batch and the fanouts): an upper bound for any part block. */
int64_t gs_slicer_slab_capacity(const gs_slicer *s);

/* ------------------------------------------------------------------------
 * Partitioned mode: one rank per GPU, rank g of num_parts holds
 * only the CSR rows (and a dedup table) of the nodes it owns and produces
 * B[layer][g] -- what GPU g of the reference's 4-GPU layout consumes
 * (slicer.cpp:36-43 routes an edge to the bipartite of the source's owner).
 * Ranks exchange sampled edges and new frontier nodes with their own stores /
 * reductions into peer memory over NVLink (arenas mapped with CUDA IPC, or
 * directly when several ranks share a process); no host round trip and no
 * library collective sits on the data path.  Every rank must be given the same
 * batches in the same order.  The reference has no counterpart (one process,
 * whole graph in host memory, dataset.cpp:18-36); outputs are bit-identical to
 * part g of the single-GPU slicer above and of the oracle.
 * ------------------------------------------------------------------------ */
typedef enum { GS_OWNER_MOD = 0, GS_OWNER_RANGE = 1 } gs_owner_mode;

typedef struct gs_pslicer gs_pslicer;

typedef struct {
  int32_t abi_version;
  int32_t num_layers;
  int32_t fanouts[GS_MAX_LAYERS];
  int32_t num_parts;                       /* = number of ranks */
  int32_t rank;
  int32_t owner_mode;                      /* gs_owner_mode */
  int32_t range_bounds[GS_MAX_PARTS + 1];  /* GS_OWNER_RANGE: part q owns [b[q], b[q+1]) */
  uint64_t seed;                           /* philox key (same on every rank) */
  int32_t max_batch;
  int32_t num_slots;                       /* samples that may be held unreleased */
  int32_t host_output;
  int32_t grid_ctas;                       /* 0 = 4 CTAs per SM */
  int32_t timeout_ms;                      /* barrier timeout; 0 = 5000 */
  int32_t use_graph;                       /* 1: one CUDA graph per slot (gs_pslicer_submit) */
  int64_t frontier_capacity;               /* 0 = derive; must agree on all ranks */
  int64_t cand_capacity;
  int64_t slab_capacity;
  int64_t num_edges_global;                /* upper bound of the edge count: bounds fanout -1 layers; 0 = unknown */
  void *stream;                            /* cudaStream_t to run on; NULL = own stream */
  const float *features_local_host;        /* optional [n_local x feature_dim] fp32: rows of the owned nodes, local order */
  int32_t feature_dim;
  int32_t rng_mode;                        /* GS_RNG_PHILOX (default) or GS_RNG_PHILOX_NOREPLACE */
  int64_t feature_rows_capacity;           /* 0 = derive */
  int64_t shuffle_rows_capacity;           /* rows one rank may send to one peer per shuffle; 0 = no shuffle */
  int32_t shuffle_max_dim;                 /* widest row a shuffle may carry */
  int32_t reserved1;
} gs_pconfig;

void gs_pconfig_default(gs_pconfig *cfg);

/* indptr_local_host: [n_local + 1] row offsets of the nodes this rank owns, in
 * local-index order (GS_OWNER_MOD: id / num_parts; GS_OWNER_RANGE: id - b[rank]);
 * indices_local_host: their neighbour lists, GLOBAL ids. */
int gs_pslicer_create(int device, int64_t num_nodes_global, int64_t n_local, int64_t m_local,
                      const int64_t *indptr_local_host, const int64_t *indices_local_host,
                      const gs_pconfig *cfg, gs_pslicer **out);
int gs_pslicer_destroy(gs_pslicer *s);
/* 64-byte cudaIpcMemHandle_t of this rank's arena, to be sent to every peer. */
int gs_pslicer_ipc_handle(gs_pslicer *s, void *handle64);
int gs_pslicer_connect_ipc(gs_pslicer *s, int32_t peer_rank, const void *handle64);
/* Peer living in the same process (tests / single-GPU emulation of P ranks). */
int gs_pslicer_connect_local(gs_pslicer *s, int32_t peer_rank, gs_pslicer *peer);
/* All peers connected: freeze the launch sequence (captures the CUDA graphs). */
int gs_pslicer_ready(gs_pslicer *s);

/* One minibatch = gs_pslicer_num_segments() launch segments separated by the
 * inter-rank barriers.  gs_pslicer_submit = stage + every segment + commit (one
 * process per rank).  stage / enqueue / commit expose the segments so that
 * several ranks sharing ONE stream can be interleaved segment by segment. */
int gs_pslicer_num_segments(const gs_pslicer *s);
int gs_pslicer_submit(gs_pslicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                      int32_t *slot_out);
int gs_pslicer_stage(gs_pslicer *s, const int64_t *seeds_host, int32_t n, uint64_t batch_index,
                     int32_t *slot_out);
int gs_pslicer_enqueue(gs_pslicer *s, int32_t slot, int32_t seg_lo, int32_t seg_hi);
int gs_pslicer_commit(gs_pslicer *s, int32_t slot);
int gs_pslicer_wait(gs_pslicer *s, int32_t slot, const gs_sample_header **hdr_out);
int gs_pslicer_slab(gs_pslicer *s, int32_t slot, const int32_t **device_slab,
                    const int32_t **host_slab);
int gs_pslicer_release(gs_pslicer *s, int32_t slot);
/* Fused input-feature gather (features_local_host set): rows of the LAST layer's
 * in_nodes of this rank's part -- all owned by this rank, so the gather is local
 * (python/data/bipartite.py:61-67, memory_manager.py:75-104). */
int gs_pslicer_feature_slab(gs_pslicer *s, int32_t slot, const float **device_rows, int32_t *dim,
                            int32_t *rows);
/* Split-parallel shuffle of one layer (replaces the 12 peer copies + mergeKernel per
 * layer of src/gnn/dist_sage.cu:120-160,193-199 and pull_for_remotes / push_from_remotes
 * of python/data/bipartite.py:88-99): rows_device is this rank's [len(out_nodes) x dim]
 * fp32 block of partial aggregates of the slot's sample.
 *   forward  (backward = 0): row FROM_IDS[q][k] goes to rank q, which adds it to its row
 *                            TO_IDS[me][k] (skipped when -1), peers merged in rank order;
 *   backward (backward = 1): the transpose -- row TO_IDS[q][k] of the owner (zeros when -1)
 *                            overwrites row FROM_IDS[owner][k] of the contributor.
 * The rows move by the kernel's own stores into the peers' staging arenas over NVLink.
 * Collective over the ranks (same slot order on every rank), asynchronous on the slicer's
 * stream; gs_pslicer_sync() waits for it.  phase -1 = whole shuffle; 0 / 1 / 2 = push +
 * signal / wait + merge + signal / wait, for ranks that share one stream. */
int gs_pslicer_shuffle(gs_pslicer *s, int32_t slot, int32_t layer, float *rows_device, int32_t dim,
                       int32_t backward, int32_t phase);
int gs_pslicer_sync(gs_pslicer *s);
/* After any error: COLLECTIVE -- every rank calls this with its streams idle, and no rank
 * submits again before all ranks have returned (a host-side barrier, e.g. dist.barrier()).
 * Clears the dedup table, the dense marks, the error word, the "group broken" latch and the
 * barrier state (epoch and flags restart at 0: after a peer timeout the epochs differ). */
int gs_pslicer_reset(gs_pslicer *s);
int gs_pslicer_launches_per_sample(const gs_pslicer *s);
int64_t gs_pslicer_device_bytes(const gs_pslicer *s);

#ifdef __cplusplus
}
#endif
#endif
