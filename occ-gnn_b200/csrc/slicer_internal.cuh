// slicer_internal.cuh -- shared declarations between the kernels and the C ABI.
// B200 (sm_100a) only.  See DESIGN.md for the pipeline and the data layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gsplit_slicer.h"

#define GS_TPB 256          // threads per CTA in every tiled kernel
#define GS_NW (GS_TPB / 32) // warps per CTA
#define GS_IPT 4            // consecutive items per thread and tile
#define GS_TILE (GS_TPB * GS_IPT)
#define GS_INF 0xFFFFFFFFu

// candidate flag byte: kind in bits 0-1, part in bits 4-6 (edge: owner of the
// source; self entry: owner of the row's node)
#define GS_CK_EDGE 0u  // cand = source id; its table entry holds the minimum position
#define GS_CK_SELF 1u  // slot 0 of a row: the destination itself (slicer.cpp:7); cand = row index
#define GS_CK_DEAD 2u  // sampled neighbour equal to the destination (slicer.cpp:33), or past the end
#define GS_CK_FHIT 3u  // edge whose source is a node of the CURRENT frontier; cand = its row index
#define GS_CK_MASK 3u
#define GS_CF_PART_SHIFT 4

// per-layer totals, exact once k_expand has run (index into GsWs::cnt[layer])
#define GS_CNT_STRIDE 64
#define GS_CNT_NEW 0                      // nodes new to the frontier
#define GS_CNT_IN(P, q) (1 + (q))         // |in_nodes_q|
#define GS_CNT_E(P, q) (1 + (P) + (q))    // edges of part q
#define GS_CNT_OUT(P, q) (1 + 2 * (P) + (q))
#define GS_CNT_OWN(P, q) (1 + 3 * (P) + (q))
#define GS_CNT_TO(P, q) (1 + 4 * (P) + (q))
#define GS_CNT_SELF(P, q) (1 + 5 * (P) + (q))  // rows owned by q (k_rows)
#define GS_TICKETS 4                      // per layer: [0] k_rows tiles, [1] k_expand done, [2] k_cand_flags done, [3] row counts done

// decoupled look-back descriptors
#define GS_RS_NC 16   // k_rows scan: candidates, drawing rows, rows per owner (uint64 each; 2 + P used)
#define GS_CS_NC 24   // k_cand scan: 1 + 2 * 8 = 17 counters (uint32), padded
#define GS_FS_NC 24   // rowside scan: 3 * 8 counters (uint32)
#define GS_SKIPPED 1000  // header status of a group member that holds no minibatch

// hand-over copy through the bulk-copy engine (cp.async.bulk); A/B with -DGS_PACK_TMA=0
#ifndef GS_PACK_TMA
#define GS_PACK_TMA 1
#endif

// Inputs of one minibatch, in pinned mapped host memory (read by k_rows<FIRST> over PCIe).
struct GsInput {
  int32_t n_seeds;   // < 0: this member of the group holds no minibatch
  int32_t want_host;
  uint64_t batch_index;
  // direct hand-over (gs_slicer_submit_direct): the lists of part q are written straight into
  // block q -- local or PEER memory -- behind a block header, nothing into the slot's slab
  int32_t direct, pad0;
  int64_t block_words;
  int32_t *part_dst[GS_MAX_PARTS];
};

// Where the lists of every part of the slot's current sample go (device memory; written by the
// sample's first launch from GsInput, read by every later one).
struct GsTargets {
  int32_t *base[GS_MAX_PARTS];    // the slot's slab for every part, or the P hand-over blocks
  int64_t cap;                    // words available behind a base
  int64_t cursor[GS_MAX_PARTS];   // direct: words of block q in use (header included)
  int32_t direct, pad0;
};

// Per-slot state carried from sample to sample; prepared for sample k + 1 by the
// last CTA of sample k's k_tail, so no kernel of a sample ever has to reset anything.
struct GsState {
  uint32_t tag_idx;     // tag index of the current sample; stored tag = ntags - 1 - tag_idx (index 0 = fill pattern)
  uint32_t scan_epoch;  // scan descriptors of layer l: epoch scan_epoch + 3 l (k_rows), + 1 (k_cand), + 2 (rowside)
  uint32_t pad[2];
};

// Everything a kernel needs about one slot.  The kernels receive an ARRAY of
// these (one per minibatch of a group, blockIdx.y selects) and stage their own
// in shared memory.
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
  int32_t l2_hints;      // 1: keep the dedup table in L2 (evict_last), stream the picks (evict_first)
  uint32_t seed_lo, seed_hi;
  const uint32_t *draws;
  uint64_t n_draws;
  uint64_t *draw_cursor; // shared by all slots (GS_RNG_DRAWS runs on one stream, groups of one)
  // per-slot state
  gs_sample_header *hdr; // device
  GsState *state;
  uint32_t *table;       // [N] (tag | kind | payload): kind 0 = node of the current frontier, payload = its row;
                         //     kind 1 = minimum edge-source position of a node new to this layer
  int32_t pos_bits;      // S: payload bits; the kind bit is 1 << S, the tag sits above it
  GsTargets *tgt;        // list bases of the current sample (slab, or hand-over blocks)
  int32_t *ixs;          // [Ccap] csr_indices before relabel: rank, or ~position of the first occurrence
  int32_t *slab;         // sample slab: every bipartite list (mirrored to the host in e2e mode)
  int64_t slab_cap;
  int32_t *fslab;        // frontier slab: frontiers 0..L (device only; read back on demand)
  int64_t fslab_cap;
  int4 *rowinfo;         // [Fcap] {indptr lo, indptr hi, degree, node id} of every frontier row
  uint32_t *off;         // [Fcap+1] first candidate position of row i
  uint32_t *samp_rank;   // [Fcap]   # earlier rows that draw (replay mode)
  uint32_t *selfrank;    // [2][Fcap] # earlier rows with the same owner = rank in self_ids_* of that part
  uint32_t *ea;          // [2][Fcap] minimum edge-source position of the row's node in this layer (GS_INF: none)
  uint32_t *rowhas;      // [2][Fcap] (ping-pong by layer) bit q: the row emitted >= 1 edge into part q
  uint32_t *rowscan;     // [Fcap+1][4 or 8] edges of part q emitted before row i (k_cand -> rowside)
  int32_t *cand;         // [2][Ccap] (ping-pong by layer) candidate stream; after k_cand: in_nodes rank at first-occurrence positions
  uint8_t *cflag;        // [Ccap]
  uint32_t *tile_first;  // [ctiles_cap] row containing candidate t*GS_TILE
  uint32_t *cnt;         // [L][GS_CNT_STRIDE] per-layer totals
  uint32_t *tickets;     // [L][GS_TICKETS] + [GS_TICKETS] (tail)
  uint32_t *rs_status;   // [ftiles_cap]
  unsigned long long *rs_val;  // [2][ftiles_cap][GS_RS_NC]  aggregates, inclusive prefixes
  uint32_t *ccnt;        // [GS_CS_NC][cgrid_cap] k_cand_flags: class counts per CTA -> exclusive prefixes
  int32_t cgrid_cap;
  uint32_t *fcnt;        // [GS_FS_NC][cgrid_cap] row side: class counts per CTA run of rows -> exclusive prefixes
  int32_t Fcap, Ccap, ftiles_cap, ctiles_cap;
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

// launchers (slicer_kernels.cu); every one enqueues exactly one kernel on `st`.
// wsv: device array of `members` GsWs (one group); grid.y = members.
struct GsGrids {
  int rows, cand, relabel;  // CTAs of the three roles for one layer (static upper bounds)
};
void gs_kernels_prepare();  // once per device, before the first launch / graph capture
// wide: more than 4 parts (64-bit packed per-part counters)
void gs_launch_rows(const GsWs *wsv, int members, int layer, int grid_rows, cudaStream_t st);
// by_rows: sampled layer (fan > 0) -> one thread per (row, four picks); else candidate-parallel (fan = -1).
// grid_relabel / grid_rowside: extra CTAs that finish the previous layer's lists.
void gs_launch_expand(const GsWs *wsv, int members, bool wide, int rng_mode, bool by_rows, int layer, int grid,
                      int grid_relabel, int grid_rowside, cudaStream_t st);
// pass 0: flags (+ grid_rowcount CTAs that count the row-side classes), pass 1: apply
void gs_launch_cand(const GsWs *wsv, int members, bool wide, int pass, int layer, int grid, int grid_rowcount,
                    cudaStream_t st);
void gs_launch_tail(const GsWs *wsv, int members, bool wide, int grid_relabel, int grid_rowside, int grid_gather,
                    cudaStream_t st);
void gs_launch_gather_ids(const float *feat, const int32_t *slot_of, const float *host_rows, int dim,
                          int div, int64_t rows, const int32_t *ids, int64_t n, float *out, int grid,
                          cudaStream_t st);
struct GsPartDst { int32_t *blk[GS_MAX_PARTS]; };  // block g of a hand-over (local or peer memory)
void gs_launch_pack_parts(const GsWs &ws, const GsPartDst &dst, int64_t block_words, int grid_x, bool tma,
                          cudaStream_t st);
void gs_launch_fill_u32(uint32_t *p, size_t n, uint32_t v, cudaStream_t st);
void gs_launch_narrow_ids(const int64_t *src, int32_t *dst, size_t n, cudaStream_t st);
// CSR validation on the device (gs_graph_create): err[0] = first offending index + 1 (0: none), by kind
void gs_launch_check_indptr(const int64_t *indptr, int64_t n, int64_t m, unsigned long long *err, cudaStream_t st);
void gs_launch_check_ids(const int64_t *ids, size_t n, int64_t base, int64_t num_nodes, unsigned long long *err,
                         cudaStream_t st);
// wrapping sum of an int64 array into *out (atomicAdd; zero it first): the dataset checksums
void gs_launch_sum_i64(const int64_t *src, size_t n, unsigned long long *out, cudaStream_t st);
// random-sector yardstick (bench): n reads of one 4-byte word each from a table of `elems` words
void gs_launch_sector_probe(const uint32_t *table, uint64_t elems, uint64_t n, uint32_t seed, uint32_t *sink,
                            int grid, cudaStream_t st);
