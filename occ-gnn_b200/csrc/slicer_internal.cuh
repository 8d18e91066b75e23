// slicer_internal.cuh -- shared declarations between the kernels and the C ABI.
// B200 (sm_100a) only.  See DESIGN.md for the pipeline and the data layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gsplit_slicer.h"

#define GS_TPB 256          // threads per CTA in every tiled kernel
#define GS_NW (GS_TPB / 32) // warps per CTA
#define GS_IPT 4            // sub-iterations (items per thread) per tile
#define GS_TILE (GS_TPB * GS_IPT)
#define GS_INF 0xFFFFFFFFu
#define GS_NC_MAX (2 * GS_MAX_PARTS + 1) // candidate-side counters
#define GS_NF_MAX (4 * GS_MAX_PARTS)     // frontier-side counters

// candidate flag byte
#define GS_CF_SELF 0x01u   // slot 0 of a row: the destination itself (slicer.cpp:7)
#define GS_CF_LOOP 0x02u   // sampled neighbour equal to the destination (treated as self, slicer.cpp:33)
#define GS_CF_FFIRST 0x04u // first occurrence in the layer's candidate stream -> next frontier
#define GS_CF_IFIRST 0x08u // first occurrence as an edge source -> in_nodes of its owner
#define GS_CF_PART_SHIFT 4

// Inputs of one minibatch, in pinned mapped host memory (read by k_begin over PCIe).
// Tagged dedup table (no per-layer / per-sample resets); A/B with -DGS_TABLE_TAGS=0
#ifndef GS_TABLE_TAGS
#define GS_TABLE_TAGS 1
#endif

// hand-over copy through the bulk-copy engine (cp.async.bulk); A/B with -DGS_PACK_TMA=0
#ifndef GS_PACK_TMA
#define GS_PACK_TMA 1
#endif

struct GsInput {
  int32_t n_seeds;
  int32_t want_host;
  uint64_t batch_index;
};

// Everything a kernel needs; passed by value, constant for the life of a slot,
// so a slot's whole minibatch is one replayable CUDA graph.
struct GsWs {
  // graph (shared by all slots)
  const int64_t *indptr;  // [N+1]
  const int32_t *indices; // [M]
  const uint8_t *owner;   // [N] or nullptr -> id % P
  int32_t num_nodes;
  int32_t P, L;
  int32_t pmask;     // P - 1 when P is a power of two, else -1
  uint32_t pmagic;   // ceil(2^32 / P): id / P by multiply-high (+ one correction step)
  int32_t fanouts[GS_MAX_LAYERS];
  // rng
  int32_t rng_mode;
  int32_t l2_hints;      // 1: keep the dedup table in L2 (evict_last), stream the rest (evict_first)
  uint32_t seed_lo, seed_hi;
  const uint32_t *draws;
  uint64_t n_draws;
  uint64_t *draw_cursor; // shared by all slots (GS_RNG_DRAWS runs on one stream)
  // per-slot state
  gs_sample_header *hdr; // device
  uint32_t *table;       // [N] minimum edge-source position of the node in the current layer (GS_INF: none)
  uint32_t *epoch;       // [2] tagged table: [0] next free tag index, [1] first tag index of the current sample
  int32_t tag_shift;     // S > 0: table value = (tag << S) | position, newer layers use SMALLER tags, so a
                         // stale entry loses every red.min and reads as "none": no per-layer / per-sample
                         // resets, one dense fill every 2^(32-S) / (L+1) samples.  0: untagged (reset each time)
  int32_t *slab;         // sample slab: every bipartite list (mirrored to the host in e2e mode)
  int64_t slab_cap;
  int32_t *fslab;        // frontier slab: frontiers 0..L (device only; read back on demand)
  int64_t fslab_cap;
  int64_t *rowstart;     // [Fcap]   indptr[front[i]]
  int32_t *rowdeg;       // [Fcap]
  uint32_t *off;         // [Fcap+1] first candidate position of row i
  uint32_t *samp_rank;   // [Fcap]   # earlier rows that draw (replay mode)
  uint32_t *rowscan;     // [P][Fcap+1] edges of part g emitted before row i
  int32_t *cand;         // [Ccap]   candidate stream (slot 0 of a row holds the row index)
  uint8_t *cflag;        // [Ccap]
  uint8_t *kill;         // [Ccap] 1: the edge candidate here is NOT a frontier first (its node's self entry came earlier); zero between layers
  uint32_t *irank;       // [Ccap] rank in in_nodes_g of the first edge occurrence at this position
  uint8_t *rowhas;       // [Fcap] bit q: row emitted >= 1 edge into part q (k_front_count -> k_front_apply)
  uint32_t *rowa;        // [Fcap] table value (min edge-source position) of every row's node
  uint32_t *tile_first;  // [ctiles_max+1] row containing candidate t*GS_TILE
  unsigned long long *ftile_sum; // [ftiles_max] candidates per frontier tile
  uint32_t *ftile_samp;  // [ftiles_max] drawing rows per frontier tile
  uint32_t *ftile_pre;   // [2][ftiles_max] exclusive prefixes of the two above
  uint32_t *ctile_cnt;   // [GS_NC_MAX][ctiles_max]
  uint32_t *ftile_cnt;   // [GS_NF_MAX][ftiles_max]
  uint32_t *tickets;     // [8] last-block tickets, one per spine kernel
  int32_t Fcap, Ccap, ftiles_max, ctiles_max;
  // fused input-feature gather (optional)
  const float *feat;     // [rows][feat_dim] fp32 (cached store: the cached rows)
  const int32_t *feat_slot; // cached store: [feat_rows] slot of every row or -1; else NULL
  const float *feat_host;   // cached store: device alias of the host-resident table
  float *feat_out;       // per-slot feature slab [feat_cap_rows][feat_dim]
  int32_t feat_dim, feat_div;
  int64_t feat_rows, feat_cap_rows;
  // host mirrors (pinned, mapped)
  const int64_t *seeds_in;
  const GsInput *input;
  gs_sample_header *hdr_host;
  int32_t *slab_host;
};

// error string of the calling thread (slicer_api.cu); returns `code`
int gs_set_error(int code, const char *msg);

// Tracing (the reference marks its phases with NVTX ranges,
// src/samplers/neigh_sample.h:98, python/train.py:68).  GSPLIT_NVTX=1 in the
// environment turns on one range per submit / wait / shuffle of the C ABI; off
// (the default) it costs one predictable branch.  Header-only NVTX v3.
struct GsRange {
  bool on;
  explicit GsRange(const char *name);
  ~GsRange();
};

// launchers (slicer_kernels.cu); every one enqueues exactly one kernel on `st`
void gs_launch_begin(const GsWs &ws, cudaStream_t st);
void gs_launch_rows_count(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_rows_scan(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_expand(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_cand_flags(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_cand_apply(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_front_count(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_front_apply(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_relabel(const GsWs &ws, int layer, int grid, cudaStream_t st);
void gs_launch_finish(const GsWs &ws, int grid, cudaStream_t st);
void gs_launch_gather_sample(const GsWs &ws, int grid, cudaStream_t st);
void gs_launch_gather_ids(const float *feat, const int32_t *slot_of, const float *host_rows, int dim,
                          int div, int64_t rows, const int32_t *ids, int64_t n, float *out, int grid,
                          cudaStream_t st);
struct GsPartDst { int32_t *blk[GS_MAX_PARTS]; };  // block g of a hand-over (local or peer memory)
void gs_launch_pack_parts(const GsWs &ws, const GsPartDst &dst, int64_t block_words, int grid_x, bool tma,
                          cudaStream_t st);
void gs_launch_fill_u32(uint32_t *p, size_t n, uint32_t v, cudaStream_t st);
void gs_launch_narrow_ids(const int64_t *src, int32_t *dst, size_t n, cudaStream_t st);
