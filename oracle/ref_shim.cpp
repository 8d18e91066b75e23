// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Drives the UNMODIFIED reference slicer.  This file contains no reference
// code: it #includes the reference headers where they lie
// (/root/reference/cslicer, via -I) and is linked against the reference's own
// dataset.cpp / bipartite.cpp / util/duplicate.cpp / slicer.cpp /
// pybipartite.cpp by oracle/Makefile into oracle/_ref/libcslicer_ref.so.
//
// Why a shim and not the reference's pybind module: CSlicer hard-codes the
// dataset root "/data/sandeep/" (pyfrontend.cpp:24) and chooses batches itself;
// parity tests need caller-chosen batches and a directory under the repo.  The
// shim builds the same objects CSlicer/WorkerPool build (pyfrontend.cpp:51-68,
// WorkerPool.cpp:23-33) and feeds Slicer::run through its own two ConQueues.
#include <fcntl.h>
#include <unistd.h>

#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

#include "slicer.h"  // reference header (pulls dataset.h, pybipartite.h, conqueue.h)

namespace {

// The reference prints unconditionally from its hot loop (slicer.cpp:123,132,
// conqueue.h:44, pybipartite.cpp:46,75); silence fd 1 for the duration.
struct StdoutSilencer {
  int saved;
  StdoutSilencer() {
    std::cout.flush();
    fflush(stdout);
    saved = dup(1);
    int devnull = open("/dev/null", O_WRONLY);
    dup2(devnull, 1);
    close(devnull);
  }
  ~StdoutSilencer() {
    std::cout.flush();
    fflush(stdout);
    dup2(saved, 1);
    close(saved);
  }
};

struct Maps {
  std::vector<int> workload;
  std::vector<int> *storage[4];
  explicit Maps(long n) {
    // pyfrontend.cpp:53-65: owner = id % 4
    for (int k = 0; k < 4; k++) storage[k] = new std::vector<int>();
    for (long j = 0; j < n; j++) {
      workload.push_back(j % 4);
      for (int k = 0; k < 4; k++) storage[k]->push_back(k == j % 4 ? 1 : 0);
    }
  }
  ~Maps() {
    for (int k = 0; k < 4; k++) delete storage[k];
  }
};

void put(std::vector<int64_t> &buf, const std::vector<long> &v) {
  buf.push_back((int64_t)v.size());
  for (long x : v) buf.push_back((int64_t)x);
}

}  // namespace

extern "C" {

// Slice nb caller-chosen batches with ONE reference Slicer (one mt19937 stream,
// state carried across batches as in the reference).  Output: malloc'd int64
// stream; per sample, per layer (3), per part (4):
//   in_nodes, indptr, out_nodes, owned_out_nodes, indices, from_ids[0..3],
//   to_ids[0..3], self_ids_in, self_ids_out  (each as len, data...), gpu_id.
// Returns the number of int64 written, or -1.
int64_t ref_slice_batches(const char *dataset_dir, const int64_t *batches,
                          const int64_t *batch_off, int64_t nb, int64_t **out) {
  StdoutSilencer quiet;
  Dataset *ds = new Dataset(std::string(dataset_dir));
  Maps maps(ds->num_nodes);
  auto *work = new ConQueue<std::vector<long> *>(10);
  auto *done = new ConQueue<PySample *>(10);
  Slicer *sl = new Slicer(ds, &maps.workload, maps.storage, 0, done, work);
  std::thread th(&Slicer::run, sl);
  std::vector<int64_t> buf;
  for (int64_t b = 0; b < nb; b++) {
    auto *batch = new std::vector<long>(batches + batch_off[b], batches + batch_off[b + 1]);
    if (batch->empty()) {  // an empty vector is the stop sentinel (slicer.cpp:124)
      delete batch;
      continue;
    }
    work->push_object(batch);
    PySample *s = done->pop_object();
    for (auto *layer : s->layers)
      for (PyBipartite *bp : *layer) {
        put(buf, bp->in_nodes);
        put(buf, bp->indptr);
        put(buf, bp->out_nodes);
        put(buf, bp->owned_out_nodes);
        put(buf, bp->indices);
        for (int i = 0; i < 4; i++) put(buf, bp->from_ids[i]);
        for (int i = 0; i < 4; i++) put(buf, bp->to_ids[i]);
        put(buf, bp->self_ids_in);
        put(buf, bp->self_ids_out);
        buf.push_back(bp->gpu_id);
      }
    delete s;
  }
  work->push_object(new std::vector<long>());
  th.join();
  *out = (int64_t *)malloc(sizeof(int64_t) * (buf.size() + 1));
  memcpy(*out, buf.data(), sizeof(int64_t) * buf.size());
  return (int64_t)buf.size();
}

void ref_free(int64_t *p) { free(p); }

// WorkerPool.cpp:40's call, verbatim in effect: libstdc++ std::random_shuffle
// over 0..n-1 with the process-global rand() in its never-seeded state
// (srand(1) restores that state), repeated `epochs` times on the same array.
void ref_epoch_permutations(int64_t n, int epochs, int64_t *out) {
  std::vector<long> nodes(n);
  for (int64_t i = 0; i < n; i++) nodes[i] = i;
  srand(1);
  for (int e = 0; e < epochs; e++) {
    std::random_shuffle(&nodes[0], &nodes[0] + n);
    for (int64_t i = 0; i < n; i++) out[(int64_t)e * n + i] = nodes[i];
  }
}

// Timed baseline: `workers` reference Slicers on `workers` std::threads
// (WorkerPool.cpp:29-33) fed nb caller-chosen batches through the reference's
// ConQueues; the consumer pops and drops each PySample as test_py.py does.
// Returns seconds from first push to last pop; dataset load is outside.
double ref_pool_time(const char *dataset_dir, int workers, const int64_t *batches,
                     const int64_t *batch_off, int64_t nb) {
  StdoutSilencer quiet;
  Dataset *ds = new Dataset(std::string(dataset_dir));
  Maps maps(ds->num_nodes);
  auto *work = new ConQueue<std::vector<long> *>(10);
  auto *done = new ConQueue<PySample *>(10);
  std::vector<Slicer *> sl;
  std::vector<std::thread> th;
  for (int w = 0; w < workers; w++)
    sl.push_back(new Slicer(ds, &maps.workload, maps.storage, 0, done, work));
  auto t0 = std::chrono::steady_clock::now();
  for (int w = 0; w < workers; w++) th.emplace_back(&Slicer::run, sl[w]);
  std::thread producer([&] {
    for (int64_t b = 0; b < nb; b++)
      work->push_object(
          new std::vector<long>(batches + batch_off[b], batches + batch_off[b + 1]));
    for (int w = 0; w < workers; w++) work->push_object(new std::vector<long>());
  });
  for (int64_t b = 0; b < nb; b++) delete done->pop_object();
  auto t1 = std::chrono::steady_clock::now();
  producer.join();
  for (auto &t : th) t.join();
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
