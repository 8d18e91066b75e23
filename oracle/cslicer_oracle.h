/*
 * oracle/cslicer_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's `cslicer` hot path (sequential neighbour
 * sampling over CSR, frontier dedup, owner split into per-GPU bipartites and
 * relabel).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (occ-gnn_b200/) never
 * does.
 *
 * Parity status: PINNED.  oracle/ref_shim.cpp drives the unmodified reference
 * sources (compiled from /root/reference/cslicer into oracle/_ref/) and
 * tests/test_oracle_vs_reference.py + tests/golden/ hold the bit-for-bit
 * comparison at the one configuration the reference supports (3 layers,
 * fanout 10 with replacement, 4 parts, owner = id % 4).
 *
 * Reference lines restated (relative to /root/reference/cslicer):
 *   slicer.cpp:6-22    neighbour_sample      -> orc_neighbour_sample
 *   slicer.cpp:25-64   slice_layer           -> orc_slice_layer
 *   slicer.cpp:69-81   get_sample            -> orc_slice
 *   bipartite.h:33-66  add_self_edge/add_from_node/add_to_node/add_edge
 *   bipartite.cpp:3-17 reorder               -> orc_reorder
 *   util/duplicate.cpp:14-39 order_and_remove_duplicates/replace/clear
 *   WorkerPool.cpp:37-51 epoch shuffle + batch cut -> orc_epoch_batches
 *   pybipartite.cpp:10-43 deep copy per sample -> orc_pool_run (baseline only)
 */
#ifndef CSLICER_ORACLE_H
#define CSLICER_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_LAYERS 8
#define ORC_MAX_PARTS 16

/* ORC_RNG_PHILOX_NOREPLACE: distinct picks per row (the product's
 * GS_RNG_PHILOX_NOREPLACE; semantics of python/utils/sampler.py:187,
 * dgl sample_neighbors replace=False).  DGL's own generator is not available
 * here: this mode restates the product's permutation -- parity unpinned. */
enum { ORC_RNG_MT19937 = 0, ORC_RNG_PHILOX = 1, ORC_RNG_DRAWS = 2, ORC_RNG_PHILOX_NOREPLACE = 3 };

enum {
  ORC_F_IN_NODES = 0,
  ORC_F_OUT_NODES = 1,
  ORC_F_OWNED_OUT = 2,
  ORC_F_FROM_IDS = 3,      /* slot [own part] -- the only slot the reference fills */
  ORC_F_TO_IDS = 4,        /* slot [own part] */
  ORC_F_SELF_IN = 5,
  ORC_F_SELF_OUT = 6,
  ORC_F_INDPTR_COMPAT = 7, /* reference's broken indptr: ones, len == #out pushes */
  ORC_F_CSR_INDPTR = 8,    /* restatement-defined clean CSR */
  ORC_F_CSR_INDICES = 9,
  ORC_NUM_FIELDS = 10
};

typedef struct {
  int32_t num_layers;
  int32_t fanouts[ORC_MAX_LAYERS]; /* -1 = all neighbours */
  int32_t num_parts;
  int32_t rng_mode;
  uint64_t seed;          /* philox key */
  const int32_t *owner;   /* NULL -> id % num_parts (pyfrontend.cpp:57) */
  const uint32_t *draws;  /* ORC_RNG_DRAWS: explicit 32-bit draw stream */
  uint64_t n_draws;
} orc_config;

typedef struct orc_slicer orc_slicer;

orc_slicer *orc_create(int64_t num_nodes, const int64_t *indptr,
                       const int64_t *indices, const orc_config *cfg);
void orc_destroy(orc_slicer *s);

/* Slice one minibatch.  batch_index only feeds the philox counter. */
int orc_slice(orc_slicer *s, const int64_t *seeds, int64_t n_seeds,
              uint64_t batch_index);

/* Field access on the last sliced sample (pointers valid until next orc_slice). */
int64_t orc_field_len(const orc_slicer *s, int layer, int part, int field);
/* pick j (0-based) of a row of degree deg under ORC_RNG_PHILOX_NOREPLACE */
uint32_t orc_noreplace_pick(uint64_t seed, uint64_t batch_index, int layer, int64_t node,
                            uint32_t j, uint32_t deg);
const int64_t *orc_field_ptr(const orc_slicer *s, int layer, int part, int field);
int64_t orc_frontier_len(const orc_slicer *s, int layer); /* layer in [0, L] */
const int64_t *orc_frontier_ptr(const orc_slicer *s, int layer);
int64_t orc_edges_last(const orc_slicer *s);  /* real (non-self) edges of last sample */
uint64_t orc_draws_used(const orc_slicer *s); /* rng draws consumed so far */

/* WorkerPool.cpp:37-51: libstdc++ random_shuffle with unseeded rand().
 * Writes the permuted node ids of `epochs` epochs into out[epochs*num_nodes]. */
void orc_epoch_permutations(int64_t num_nodes, int epochs, int64_t *out);

/* Multithreaded baseline: `workers` slicers, one per thread, each popping
 * batches from a shared cursor and deep-copying every produced sample
 * (WorkerPool.cpp:29-33, slicer.cpp:120-139, pybipartite.cpp:10-66).
 * batches: concatenated seed ids; batch_off[nb+1] offsets.  Returns seconds;
 * *edges_out = total real edges produced. */
double orc_pool_run(int64_t num_nodes, const int64_t *indptr,
                    const int64_t *indices, const orc_config *cfg, int workers,
                    const int64_t *batches, const int64_t *batch_off, int64_t nb,
                    int64_t *edges_out);

/* philox4x32-10, exposed for known-answer tests. */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
uint32_t orc_philox_pick(uint64_t seed, uint64_t batch_index, int layer,
                         int64_t node, int slot);
/* mt19937 (default seed 5489) raw stream, for replay tests. */
void orc_mt19937_stream(uint32_t seed, uint64_t n, uint32_t *out);

#ifdef __cplusplus
}
#endif
#endif
