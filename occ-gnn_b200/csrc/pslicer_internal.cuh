// pslicer_internal.cuh -- partitioned (one rank per GPU) slicer: shared
// declarations between pslicer_kernels.cu and pslicer_api.cu.  B200 (sm_100a).
//
// every rank g of P stores only the CSR rows of the nodes it owns
// and a dedup table over those nodes.  The frontier of a layer is replicated
// (ids + per-row candidate counts), so all ranks agree on the global candidate
// positions p of the sequential scan.  A rank expands the
// rows it owns and hands every sampled edge (p, src) to owner(src) with plain
// stores / red.min over NVLink into that peer's arena; after a device-side
// barrier each rank holds exactly the candidates whose source it owns, at their
// global positions in a dense stream, and builds its own bipartite B[layer][g]
// with the same position-ordered scans as the single-GPU path.  New frontier
// nodes are published to every peer at their position; a compaction scan over
// the dense marks rebuilds the replicated frontier.
#pragma once
#include "slicer_internal.cuh"

#define PS_OWNER_MOD 0    // owner = id % P, local index = id / P   (pyfrontend.cpp:57)
#define PS_OWNER_RANGE 1  // owner = contiguous id range, local index = id - bounds[owner]

// dense candidate flag byte (0 = empty position on this rank)
#define PS_CF_SELF 0x01u   // slot 0 of a row (cand = row index)
#define PS_CF_OWN 0x02u    // ... of a row this rank owns
#define PS_CF_FFIRST 0x04u
#define PS_CF_IFIRST 0x08u
#define PS_CF_EDGE 0x10u   // edge whose source this rank owns (cand = source id)

#define PS_NCTR (5 + 2 * GS_MAX_PARTS)  // per-kernel counter arrays: 5 + per-peer from / to lists

// What a peer may touch of a rank's memory: one cudaMalloc arena, exported
// with cudaIpcGetMemHandle; identical layout on every rank.
struct PsArena {
  uint2 *table;      // [table_cap] {x: min edge-source position | in_nodes rank, y: self position}
  int32_t *cand;     // [Ccap] dense candidate stream
  uint8_t *cflag;    // [Ccap]
  uint2 *nf;         // [Ccap] new frontier node at position p: {id, candidates it will own next layer}; y == 0: none
  uint32_t *cnt0;    // [max_batch] candidate counts of the seeds
  uint8_t *rowrem;   // unused by peers (kept here so the arena is one allocation)
  float *stage;      // [P][shuffle_rows][shuffle_dim] rows received from every peer in a shuffle
  uint32_t *flags;   // [GS_MAX_PARTS] barrier: flags[q] = last epoch rank q signalled
  int32_t *err;      // [1] status raised by any peer
};

struct PsWs {
  // local graph shard: rows of the nodes this rank owns, global neighbour ids
  const int64_t *indptr;  // [n_local+1]
  const int32_t *indices; // [m_local]
  int32_t num_nodes;      // global N
  int32_t n_local;
  int32_t P, L, rank;
  int32_t owner_mode;
  int32_t pmask, pshift;  // PS_OWNER_MOD, P a power of two: id & pmask, id >> pshift (else pmask = -1)
  uint32_t pmagic;        // ceil(2^32 / P): id / P by multiply-high + one correction step
  int32_t bounds[GS_MAX_PARTS + 1];  // PS_OWNER_RANGE
  int32_t fanouts[GS_MAX_LAYERS];
  uint32_t seed_lo, seed_hi;
  int32_t rng_mode;       // GS_RNG_PHILOX or GS_RNG_PHILOX_NOREPLACE
  PsArena me;
  PsArena peer[GS_MAX_PARTS];  // peer[rank] == me
  // private workspaces (one set: minibatches run one after another on one stream)
  int64_t *rowstart;   // [Fcap] owned rows only
  int32_t *rowdeg;     // [Fcap]
  uint32_t *cnt[2];    // [Fcap] candidate count of every frontier row (replicated), ping-pong by layer
  uint32_t *off;       // [Fcap+1]
  uint32_t *rowscan;   // [Fcap+1] edges of MY part emitted before row i
  uint32_t *rowrem;    // [Fcap] owned row: bit q = it sampled a source owned by rank q != this rank
  uint32_t *tile_first;// [ctiles+1]
  unsigned long long *ftile_sum; // [grid]
  uint32_t *ftile_pre; // [grid]
  uint32_t *ctr;       // [PS_NCTR][grid] per-CTA counters of the current spine kernel
  uint32_t *tickets;   // [8]
  uint32_t *epoch;     // [2] barriers executed so far; [1] != 0: a barrier timed out (group broken until reset)
  int32_t Fcap, Ccap, grid_max, timeout_ms;
  int64_t stage_rows;  // shuffle capacity: rows per peer
  int32_t stage_dim;
  const float *feat;   // [n_local][feat_dim] rows of the owned nodes (optional)
  int32_t feat_dim;
  int64_t feat_cap_rows;
  // per-slot
  float *feat_out;     // [feat_cap_rows][feat_dim]
  gs_sample_header *hdr;
  int32_t *slab;
  int64_t slab_cap;
  const int64_t *seeds_in;
  const GsInput *input;
  gs_sample_header *hdr_host;
  int32_t *slab_host;
};

void ps_launch_begin(const PsWs &ws, cudaStream_t st);
void ps_launch_signal(const PsWs &ws, cudaStream_t st);
void ps_launch_wait(const PsWs &ws, cudaStream_t st);
void ps_launch_rows_count(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_rows_scan(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_expand(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_cand_flags(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_cand_apply(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_front_count(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_front_apply(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_relabel(const PsWs &ws, int layer, int grid, cudaStream_t st);
void ps_launch_gather(const PsWs &ws, int grid, cudaStream_t st);
void ps_launch_shuffle_push(const PsWs &ws, int layer, float *rows, int dim, int backward, int grid, cudaStream_t st);
void ps_launch_shuffle_merge(const PsWs &ws, int layer, float *rows, int dim, int backward, int peer, int grid, cudaStream_t st);
void ps_launch_finish(const PsWs &ws, int grid, cudaStream_t st);
