/*
 * oracle/cslicer_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * See cslicer_oracle.h for scope, parity status and the reference lines restated.
 * Plain C11, sequential per slicer, parameterised in layers / parts / fanouts /
 * owner map / rng, and asserted equal to the compiled reference at L=3, k=10, P=4.
 */
#define _GNU_SOURCE
#include "cslicer_oracle.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ---------------------------------------------------------------- vectors */
typedef struct {
  int64_t *d;
  int64_t n, cap;
} vec;

static inline void vec_push(vec *v, int64_t x) {
  if (v->n == v->cap) {
    v->cap = v->cap ? v->cap * 2 : 16;
    v->d = (int64_t *)realloc(v->d, (size_t)v->cap * sizeof(int64_t));
  }
  v->d[v->n++] = x;
}
static inline void vec_clear(vec *v) { v->n = 0; }
static inline void vec_free(vec *v) {
  free(v->d);
  v->d = NULL;
  v->n = v->cap = 0;
}
static inline int vec_back_is(const vec *v, int64_t x) {
  return v->n != 0 && v->d[v->n - 1] == x;
}

/* ---------------------------------------------------------------- mt19937 */
/* std::mt19937 as libstdc++ implements it (slicer.h:33, default seed 5489). */
typedef struct {
  uint32_t mt[624];
  int idx;
} mt19937;

static void mt_seed(mt19937 *m, uint32_t seed) {
  m->mt[0] = seed;
  for (int i = 1; i < 624; i++)
    m->mt[i] = 1812433253u * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t)i;
  m->idx = 624;
}
static uint32_t mt_next(mt19937 *m) {
  if (m->idx >= 624) {
    for (int i = 0; i < 624; i++) {
      uint32_t y = (m->mt[i] & 0x80000000u) | (m->mt[(i + 1) % 624] & 0x7fffffffu);
      m->mt[i] = m->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    m->idx = 0;
  }
  uint32_t y = m->mt[m->idx++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}
void orc_mt19937_stream(uint32_t seed, uint64_t n, uint32_t *out) {
  mt19937 m;
  mt_seed(&m, seed);
  for (uint64_t i = 0; i < n; i++) out[i] = mt_next(&m);
}

/* ----------------------------------------------------------------- philox */
/* Philox4x32-10 (Salmon et al., SC'11), the counter-based generator the
 * product's native mode uses; restated independently here. */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* draw for (seed, batch, layer, destination node id, pick slot) */
uint32_t orc_philox_pick(uint64_t seed, uint64_t batch_index, int layer,
                         int64_t node, int slot) {
  uint32_t ctr[4], key[2], out[4];
  ctr[0] = (uint32_t)(uint64_t)node;
  ctr[1] = (uint32_t)((uint64_t)node >> 32);
  ctr[2] = (uint32_t)batch_index;
  ctr[3] = ((uint32_t)layer << 24) | (uint32_t)(slot >> 2);
  key[0] = (uint32_t)seed;
  key[1] = (uint32_t)(seed >> 32);
  orc_philox4x32_10(ctr, key, out);
  return out[slot & 3];
}

/* Without-replacement picks: pick j of a row = perm(j), perm a keyed bijection
 * of [0, deg); semantics follow python/utils/sampler.py:187 (distinct
 * neighbours per destination).  Keys: the Philox block of (seed, batch, layer,
 * node) whose block id is 0xFFFFFF (pick blocks use ids slot/4, far below).
 *   deg <= 16: the first j+1 steps of a Fisher-Yates shuffle of 0..15 (step i
 *              swaps entry i with entry i + floor(draw_i * (deg-i) / 2^32));
 *   deg  > 16: an unbalanced Feistel network over ceil(log2 deg) bits (12 rounds
 *              when that is 5 bits, else 8), re-applied until the value < deg. */
static uint32_t nr_mix(uint32_t right, uint32_t key) {
  uint32_t f = right * 0x9E3779B1u + key;
  f ^= f >> 16;
  f *= 0x85EBCA6Bu;
  f ^= f >> 13;
  f *= 0xC2B2AE35u;
  f ^= f >> 16;
  return f;
}

uint32_t orc_noreplace_pick(uint64_t seed, uint64_t batch_index, int layer, int64_t node,
                            uint32_t j, uint32_t deg) {
  if (deg <= 1u) return 0u;
  uint32_t ctr[4], key[2], rk[4];
  ctr[0] = (uint32_t)(uint64_t)node;
  ctr[1] = (uint32_t)((uint64_t)node >> 32);
  ctr[2] = (uint32_t)batch_index;
  ctr[3] = ((uint32_t)layer << 24) | 0x00FFFFFFu;
  key[0] = (uint32_t)seed;
  key[1] = (uint32_t)(seed >> 32);
  orc_philox4x32_10(ctr, key, rk);
  if (deg <= 16u) {
    uint8_t a[16];
    for (int i = 0; i < 16; i++) a[i] = (uint8_t)i;
    for (uint32_t i = 0; i <= j; i++) {
      uint32_t draw = nr_mix(i, rk[i % 4]);
      uint32_t t = i + (uint32_t)(((uint64_t)draw * (deg - i)) >> 32);
      uint8_t tmp = a[i]; a[i] = a[t]; a[t] = tmp;
    }
    return a[j];
  }
  int bits = 0;
  while (((uint64_t)1 << bits) < deg) bits++; /* smallest bits with 2^bits >= deg */
  const int rounds = bits <= 5 ? 12 : 8;
  uint32_t x = j;
  for (;;) {
    int wl = bits / 2, wr = bits - wl;
    for (int r = 0; r < rounds; r++) {
      uint32_t left = x >> wr;
      uint32_t right = x & (uint32_t)(((uint64_t)1 << wr) - 1);
      uint32_t k = rk[r % 4] + (uint32_t)(r / 4) * 0x632BE5ABu;
      uint32_t f = nr_mix(right, k) & (uint32_t)(((uint64_t)1 << wl) - 1);
      x = (right << wl) | (left ^ f);
      int t = wl; wl = wr; wr = t;
    }
    if (x < deg) return x;
  }
}

/* ------------------------------------------------------ duplicate remover */
/* util/duplicate.cpp:8-45: dense int mask over all nodes + touched list. */
typedef struct {
  int32_t *mask;
  vec used;
} dupr;

static void dupr_init(dupr *d, int64_t n) {
  d->mask = (int32_t *)calloc((size_t)n, sizeof(int32_t));
  memset(&d->used, 0, sizeof(vec));
}
static void dupr_free(dupr *d) {
  free(d->mask);
  vec_free(&d->used);
}
/* duplicate.cpp:14-26: stable first-occurrence unique, mask[v] = 1 + position */
static void dupr_order_and_remove(dupr *d, vec *nodes) {
  int64_t i = 0;
  for (int64_t k = 0; k < nodes->n; k++) {
    int64_t nd = nodes->d[k];
    if (d->mask[nd] == 0) {
      nodes->d[i] = nd;
      i++;
      d->mask[nd] = (int32_t)i;
      vec_push(&d->used, nd);
    }
  }
  nodes->n = i;
}
/* duplicate.cpp:35-39: v -> mask[v]-1, i.e. -1 when absent */
static void dupr_replace(const dupr *d, vec *v) {
  for (int64_t i = 0; i < v->n; i++) v->d[i] = (int64_t)d->mask[v->d[i]] - 1;
}
/* duplicate.cpp:28-33 */
static void dupr_clear(dupr *d) {
  for (int64_t i = 0; i < d->used.n; i++) d->mask[d->used.d[i]] = 0;
  vec_clear(&d->used);
}

/* -------------------------------------------------------------- bipartite */
typedef struct {
  vec f[ORC_NUM_FIELDS];
  vec edge_dst; /* per pushed edge: destination id (restatement-only, for clean CSR) */
  vec edge_src; /* per pushed edge: source id before dedup */
} bipart;

static void bp_clear(bipart *b) {
  for (int i = 0; i < ORC_NUM_FIELDS; i++) vec_clear(&b->f[i]);
  vec_clear(&b->edge_dst);
  vec_clear(&b->edge_src);
}
static void bp_free(bipart *b) {
  for (int i = 0; i < ORC_NUM_FIELDS; i++) vec_free(&b->f[i]);
  vec_free(&b->edge_dst);
  vec_free(&b->edge_src);
}
/* bipartite.h:33-39 */
static inline void bp_add_self_edge(bipart *b, int64_t nd1) {
  if (vec_back_is(&b->f[ORC_F_SELF_IN], nd1)) return;
  vec_push(&b->f[ORC_F_SELF_IN], nd1);
  vec_push(&b->f[ORC_F_SELF_OUT], nd1);
}
/* bipartite.h:41-46 (slot is always the bipartite's own index, slicer.cpp:42) */
static inline void bp_add_from_node(bipart *b, int64_t nd1) {
  if (vec_back_is(&b->f[ORC_F_FROM_IDS], nd1)) return;
  vec_push(&b->f[ORC_F_FROM_IDS], nd1);
}
/* bipartite.h:48-53 */
static inline void bp_add_to_node(bipart *b, int64_t nd1) {
  if (vec_back_is(&b->f[ORC_F_TO_IDS], nd1)) return;
  vec_push(&b->f[ORC_F_TO_IDS], nd1);
}
/* bipartite.h:55-66; the indptr[l] += 1 there lands one past the end and is
 * overwritten by the next push_back(1), so indptr stays all ones. */
static inline void bp_add_edge(bipart *b, int64_t nd1, int64_t nd2, int islocal) {
  if (islocal && !vec_back_is(&b->f[ORC_F_OWNED_OUT], nd1))
    vec_push(&b->f[ORC_F_OWNED_OUT], nd1);
  if (!vec_back_is(&b->f[ORC_F_OUT_NODES], nd1)) {
    vec_push(&b->f[ORC_F_OUT_NODES], nd1);
    vec_push(&b->f[ORC_F_INDPTR_COMPAT], 1);
  }
  vec_push(&b->f[ORC_F_IN_NODES], nd2);
  vec_push(&b->edge_dst, nd1);
  vec_push(&b->edge_src, nd2);
}

/* bipartite.cpp:3-17, plus the restatement-defined clean CSR (A7): rows are
 * out_nodes, indices[e] = position of the edge's source in in_nodes, edges in
 * emission order, multiplicity kept. */
static void bp_reorder(bipart *b, dupr *dr) {
  dupr_order_and_remove(dr, &b->f[ORC_F_IN_NODES]);
  dupr_replace(dr, &b->f[ORC_F_SELF_IN]);
  vec *ci = &b->f[ORC_F_CSR_INDICES];
  for (int64_t e = 0; e < b->edge_src.n; e++)
    vec_push(ci, (int64_t)dr->mask[b->edge_src.d[e]] - 1);
  dupr_clear(dr);

  dupr_order_and_remove(dr, &b->f[ORC_F_OUT_NODES]);
  dupr_replace(dr, &b->f[ORC_F_OWNED_OUT]);
  dupr_replace(dr, &b->f[ORC_F_SELF_OUT]);
  dupr_replace(dr, &b->f[ORC_F_FROM_IDS]);
  dupr_replace(dr, &b->f[ORC_F_TO_IDS]);
  /* clean indptr: count edges per out row (rows are contiguous runs in the
   * frontier order, but count generally so duplicate seeds stay well defined) */
  vec *cp = &b->f[ORC_F_CSR_INDPTR];
  int64_t rows = b->f[ORC_F_OUT_NODES].n;
  for (int64_t r = 0; r <= rows; r++) vec_push(cp, 0);
  for (int64_t e = 0; e < b->edge_dst.n; e++)
    cp->d[(int64_t)dr->mask[b->edge_dst.d[e]]] += 1; /* mask = 1 + row */
  for (int64_t r = 0; r < rows; r++) cp->d[r + 1] += cp->d[r];
  dupr_clear(dr);
}

/* ----------------------------------------------------------------- slicer */
struct orc_slicer {
  int64_t num_nodes;
  const int64_t *indptr, *indices;
  orc_config cfg;
  dupr dr, out_dr;
  mt19937 mt;
  uint64_t draws_used;
  vec frontier[ORC_MAX_LAYERS + 1];
  vec nbrs;
  bipart bp[ORC_MAX_LAYERS][ORC_MAX_PARTS];
  int64_t edges_last;
  uint64_t batch_index;
};

orc_slicer *orc_create(int64_t num_nodes, const int64_t *indptr,
                       const int64_t *indices, const orc_config *cfg) {
  if (cfg->num_layers < 1 || cfg->num_layers > ORC_MAX_LAYERS) return NULL;
  if (cfg->num_parts < 1 || cfg->num_parts > ORC_MAX_PARTS) return NULL;
  orc_slicer *s = (orc_slicer *)calloc(1, sizeof(orc_slicer));
  s->num_nodes = num_nodes;
  s->indptr = indptr;
  s->indices = indices;
  s->cfg = *cfg;
  dupr_init(&s->dr, num_nodes);
  dupr_init(&s->out_dr, num_nodes);
  mt_seed(&s->mt, 5489u); /* std::mt19937 default ctor, slicer.h:33 */
  return s;
}

void orc_destroy(orc_slicer *s) {
  if (!s) return;
  dupr_free(&s->dr);
  dupr_free(&s->out_dr);
  for (int l = 0; l <= ORC_MAX_LAYERS; l++) vec_free(&s->frontier[l]);
  vec_free(&s->nbrs);
  for (int l = 0; l < ORC_MAX_LAYERS; l++)
    for (int p = 0; p < ORC_MAX_PARTS; p++) bp_free(&s->bp[l][p]);
  free(s);
}

static inline int orc_owner(const orc_slicer *s, int64_t nd) {
  return s->cfg.owner ? s->cfg.owner[nd] : (int)(nd % s->cfg.num_parts);
}

static inline uint32_t orc_draw(orc_slicer *s, int layer, int64_t nd1, int slot) {
  uint32_t r;
  switch (s->cfg.rng_mode) {
    case ORC_RNG_MT19937:
      r = mt_next(&s->mt);
      break;
    case ORC_RNG_DRAWS:
      r = s->draws_used < s->cfg.n_draws ? s->cfg.draws[s->draws_used] : 0u;
      break;
    default:
      r = orc_philox_pick(s->cfg.seed, s->batch_index, layer, nd1, slot);
      break;
  }
  s->draws_used++;
  return r;
}

/* slicer.cpp:6-22: self entry first; all neighbours in CSR order when
 * deg < fanout (or fanout == -1); else `fanout` picks WITH replacement,
 * one 32-bit draw per pick, draw % deg. */
static void orc_neighbour_sample(orc_slicer *s, int layer, int64_t nd1, vec *out) {
  vec_push(out, nd1);
  int64_t off = s->indptr[nd1];
  int64_t deg = s->indptr[nd1 + 1] - off;
  int fan = s->cfg.fanouts[layer];
  if (fan < 0 || deg < fan) {
    for (int64_t i = 0; i < deg; i++) vec_push(out, s->indices[off + i]);
  } else {
    for (int i = 0; i < fan; i++) {
      int64_t r;
      if (s->cfg.rng_mode == ORC_RNG_PHILOX_NOREPLACE) {
        r = (int64_t)orc_noreplace_pick(s->cfg.seed, s->batch_index, layer, nd1, (uint32_t)i, (uint32_t)deg);
        s->draws_used++;
      } else {
        r = (int64_t)(orc_draw(s, layer, nd1, i) % (uint64_t)deg);
      }
      vec_push(out, s->indices[off + r]);
    }
  }
}

/* slicer.cpp:25-64 */
static void orc_slice_layer(orc_slicer *s, int layer, const vec *in, vec *out) {
  const int P = s->cfg.num_parts;
  bipart *B = s->bp[layer];
  for (int64_t a = 0; a < in->n; a++) {
    int64_t nd1 = in->d[a];
    vec_clear(&s->nbrs);
    orc_neighbour_sample(s, layer, nd1, &s->nbrs);
    int to = orc_owner(s, nd1);
    for (int64_t b = 0; b < s->nbrs.n; b++) {
      int64_t nd2 = s->nbrs.d[b];
      if (nd1 == nd2) {
        bp_add_self_edge(&B[to], nd1);
      } else {
        int from = orc_owner(s, nd2);
        s->edges_last++;
        if (to == from) {
          bp_add_edge(&B[to], nd1, nd2, 1);
        } else {
          bp_add_edge(&B[from], nd1, nd2, 0);
          bp_add_to_node(&B[to], nd1);
          bp_add_from_node(&B[from], nd1);
        }
      }
      if (s->out_dr.mask[nd2] == 0) {
        vec_push(out, nd2);
        s->out_dr.mask[nd2] = 1;
        vec_push(&s->out_dr.used, nd2);
      }
    }
  }
  for (int g = 0; g < P; g++) bp_reorder(&B[g], &s->dr);
  dupr_clear(&s->out_dr);
}

/* slicer.cpp:69-81 (+ clear(), slicer.cpp:114-118) */
int orc_slice(orc_slicer *s, const int64_t *seeds, int64_t n_seeds,
              uint64_t batch_index) {
  const int L = s->cfg.num_layers, P = s->cfg.num_parts;
  for (int l = 0; l < L; l++)
    for (int g = 0; g < P; g++) bp_clear(&s->bp[l][g]);
  for (int l = 0; l <= L; l++) vec_clear(&s->frontier[l]);
  s->edges_last = 0;
  s->batch_index = batch_index;
  for (int64_t i = 0; i < n_seeds; i++) {
    if (seeds[i] < 0 || seeds[i] >= s->num_nodes) return -1;
    vec_push(&s->frontier[0], seeds[i]);
  }
  for (int l = 0; l < L; l++)
    orc_slice_layer(s, l, &s->frontier[l], &s->frontier[l + 1]);
  return 0;
}

int64_t orc_field_len(const orc_slicer *s, int layer, int part, int field) {
  return s->bp[layer][part].f[field].n;
}
const int64_t *orc_field_ptr(const orc_slicer *s, int layer, int part, int field) {
  return s->bp[layer][part].f[field].d;
}
int64_t orc_frontier_len(const orc_slicer *s, int layer) { return s->frontier[layer].n; }
const int64_t *orc_frontier_ptr(const orc_slicer *s, int layer) {
  return s->frontier[layer].d;
}
int64_t orc_edges_last(const orc_slicer *s) { return s->edges_last; }
uint64_t orc_draws_used(const orc_slicer *s) { return s->draws_used; }

/* WorkerPool.cpp:40: std::random_shuffle(first,last) in libstdc++ is
 *   for (i = first+1; i != last; ++i) iter_swap(i, first + rand() % ((i-first)+1));
 * with the process-global, never-seeded rand(). */
void orc_epoch_permutations(int64_t num_nodes, int epochs, int64_t *out) {
  int64_t *cur = (int64_t *)malloc((size_t)num_nodes * sizeof(int64_t));
  for (int64_t i = 0; i < num_nodes; i++) cur[i] = i;
  srand(1); /* == the state of a process that never called srand */
  for (int e = 0; e < epochs; e++) {
    for (int64_t i = 1; i < num_nodes; i++) {
      int64_t j = (int64_t)(rand() % (i + 1));
      int64_t t = cur[i];
      cur[i] = cur[j];
      cur[j] = t;
    }
    memcpy(out + (size_t)e * num_nodes, cur, (size_t)num_nodes * sizeof(int64_t));
  }
  free(cur);
}

/* ------------------------------------------------- multithreaded baseline */
/* Same threading model as WorkerPool.cpp:29-33 + slicer.cpp:120-139: one
 * slicer (own masks, own rng) per thread, batches handed out in order, every
 * finished sample deep-copied into a fresh owning object
 * (pybipartite.cpp:10-66) and then dropped by the consumer. */
typedef struct {
  orc_slicer *s;
  const int64_t *batches, *batch_off;
  int64_t nb;
  int64_t *cursor;
  pthread_mutex_t *mtx;
  int64_t edges;
  int64_t sink;
} pool_arg;

static void *pool_worker(void *p) {
  pool_arg *a = (pool_arg *)p;
  const int L = a->s->cfg.num_layers, P = a->s->cfg.num_parts;
  for (;;) {
    pthread_mutex_lock(a->mtx);
    int64_t b = (*a->cursor)++;
    pthread_mutex_unlock(a->mtx);
    if (b >= a->nb) break;
    orc_slice(a->s, a->batches + a->batch_off[b], a->batch_off[b + 1] - a->batch_off[b],
              (uint64_t)b);
    a->edges += a->s->edges_last;
    /* deep copy of every field, as `new PySample(&sample)` does */
    for (int l = 0; l < L; l++)
      for (int g = 0; g < P; g++)
        for (int f = 0; f < ORC_NUM_FIELDS; f++) {
          const vec *v = &a->s->bp[l][g].f[f];
          if (v->n == 0) continue;
          int64_t *c = (int64_t *)malloc((size_t)v->n * sizeof(int64_t));
          memcpy(c, v->d, (size_t)v->n * sizeof(int64_t));
          a->sink += c[v->n - 1];
          free(c);
        }
  }
  return NULL;
}

double orc_pool_run(int64_t num_nodes, const int64_t *indptr,
                    const int64_t *indices, const orc_config *cfg, int workers,
                    const int64_t *batches, const int64_t *batch_off, int64_t nb,
                    int64_t *edges_out) {
  if (workers < 1) workers = 1;
  pool_arg *args = (pool_arg *)calloc((size_t)workers, sizeof(pool_arg));
  pthread_t *th = (pthread_t *)calloc((size_t)workers, sizeof(pthread_t));
  pthread_mutex_t mtx = PTHREAD_MUTEX_INITIALIZER;
  int64_t cursor = 0;
  for (int w = 0; w < workers; w++) {
    args[w].s = orc_create(num_nodes, indptr, indices, cfg);
    args[w].batches = batches;
    args[w].batch_off = batch_off;
    args[w].nb = nb;
    args[w].cursor = &cursor;
    args[w].mtx = &mtx;
  }
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int w = 0; w < workers; w++) pthread_create(&th[w], NULL, pool_worker, &args[w]);
  int64_t edges = 0;
  for (int w = 0; w < workers; w++) {
    pthread_join(th[w], NULL);
    edges += args[w].edges;
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  for (int w = 0; w < workers; w++) orc_destroy(args[w].s);
  free(args);
  free(th);
  if (edges_out) *edges_out = edges;
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
