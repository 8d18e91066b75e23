// noalloc.cu -- gs_unique_first_noalloc(): the slicer's frontier dedup (first
// occurrence = smallest position, list order = ascending position; the rule
// k_expand / k_cand implement for slicer.cpp:22-43) as a stand-alone entry point
// that allocates NOTHING: its dedup table, its per-CTA counts and its control
// words live in one file-scope device array.  No counterpart in the reference;
// added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// Every call on a device shares the array: calls that overlap on that GPU race.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "slicer_internal.cuh"

namespace {

constexpr int NA_TPB = 256;
constexpr int NA_MAX_GRID = 1024;
constexpr int NA_MAX_DEVICES = 64;
// scratch layout (uint32 words): [table: MAX_NODES][per-CTA counts: MAX_GRID][control: 64]
constexpr size_t NA_TABLE = 0;
constexpr size_t NA_BCNT = (size_t)GS_NOALLOC_MAX_NODES;
constexpr size_t NA_CTL = NA_BCNT + NA_MAX_GRID;
constexpr size_t NA_WORDS = NA_CTL + 64;
enum { CTL_TICKET = 0, CTL_TOTAL = 1, CTL_BAD = 2 };  // CTL_BAD: first out-of-range position + 1

// 4 MiB + 4.25 KiB, held by every GPU that loads the library until the process exits.
// In HOST code this name is a stand-in, not a device address: na_scratch() asks the
// runtime for the address once per device.
static __device__ __align__(256) uint32_t g_na_scratch[NA_WORDS];

// table entry: 0 = empty, else 0xFFFFFFFF - (smallest position seen) -- atomicMax keeps the first
__global__ void __launch_bounds__(NA_TPB) k_na_claim(const int32_t *__restrict__ ids, uint32_t n,
                                                     uint32_t num_nodes, uint32_t *scr) {
  for (uint32_t i = blockIdx.x * NA_TPB + threadIdx.x; i < n; i += gridDim.x * NA_TPB) {
    uint32_t id = (uint32_t)ids[i];
    if (id >= num_nodes)
      atomicMin(&scr[NA_CTL + CTL_BAD], i + 1);
    else
      atomicMax(&scr[NA_TABLE + id], GS_INF - i);
  }
}

__device__ __forceinline__ bool na_first(const int32_t *ids, uint32_t i, uint32_t n, uint32_t num_nodes,
                                         const uint32_t *scr, int32_t &id) {
  if (i >= n) return false;
  id = ids[i];
  return (uint32_t)id < num_nodes && scr[NA_TABLE + (uint32_t)id] == GS_INF - i;
}

// CTA b owns positions [b * chunk, (b + 1) * chunk); the last CTA to finish turns the
// per-CTA counts into exclusive prefixes and publishes the total.
__global__ void __launch_bounds__(NA_TPB) k_na_count(const int32_t *__restrict__ ids, uint32_t n,
                                                     uint32_t num_nodes, uint32_t chunk, uint32_t *scr) {
  __shared__ uint32_t s_cnt;
  __shared__ bool s_last;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  uint32_t lo = blockIdx.x * chunk, hi = min(n, lo + chunk), mine = 0;
  int32_t id;
  for (uint32_t i = lo + threadIdx.x; i < hi; i += NA_TPB) mine += na_first(ids, i, n, num_nodes, scr, id);
  mine = __reduce_add_sync(0xFFFFFFFFu, mine);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_cnt, mine);
  __syncthreads();
  if (threadIdx.x == 0) {
    scr[NA_BCNT + blockIdx.x] = s_cnt;
    __threadfence();
    s_last = atomicAdd(&scr[NA_CTL + CTL_TICKET], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    volatile uint32_t *bc = scr + NA_BCNT;
    uint32_t run = 0;
    for (uint32_t b = 0; b < gridDim.x; ++b) {
      uint32_t c = bc[b];
      bc[b] = run;
      run += c;
    }
    scr[NA_CTL + CTL_TOTAL] = run;
    scr[NA_CTL + CTL_TICKET] = 0;
  }
}

__global__ void __launch_bounds__(NA_TPB) k_na_write(const int32_t *__restrict__ ids, uint32_t n,
                                                     uint32_t num_nodes, uint32_t chunk, const uint32_t *scr,
                                                     int32_t *__restrict__ out) {
  __shared__ uint32_t s_w[NA_TPB / 32];
  uint32_t lo = blockIdx.x * chunk, hi = min(n, lo + chunk);
  uint32_t base = scr[NA_BCNT + blockIdx.x];
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (uint32_t t = lo; t < hi; t += NA_TPB) {
    uint32_t i = t + threadIdx.x;
    int32_t id = 0;
    bool f = i < hi && na_first(ids, i, n, num_nodes, scr, id);
    uint32_t m = __ballot_sync(0xFFFFFFFFu, f);
    if (lane == 0) s_w[w] = __popc(m);
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < NA_TPB / 32; ++k) {
      uint32_t c = s_w[k];
      before += k < (int)w ? c : 0;
      total += c;
    }
    if (f) out[base + before + __popc(m & ((1u << lane) - 1))] = id;
    base += total;
    __syncthreads();
  }
}

std::mutex g_na_mu;
uint32_t *g_na_ptr[NA_MAX_DEVICES];  // device address of g_na_scratch, per device

int na_fail(int code, const char *what, cudaError_t e) {
  char buf[256];
  snprintf(buf, sizeof buf, "gs_unique_first_noalloc: %s failed: %s", what, cudaGetErrorString(e));
  return gs_set_error(code, buf);
}

// device `device` is current
int na_scratch(int device, uint32_t **out) {
  std::lock_guard<std::mutex> lk(g_na_mu);
  if (!g_na_ptr[device]) {
    uint32_t *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_na_scratch);
    if (e != cudaSuccess) return na_fail(GS_ERR_CUDA, "cudaGetSymbolAddress(scratch)", e);
    g_na_ptr[device] = p;
  }
  *out = g_na_ptr[device];
  return GS_OK;
}

}  // namespace

extern "C" int gs_unique_first_noalloc(int device, const int32_t *ids_device, int64_t n, int64_t num_nodes,
                                       int32_t *out_device, int64_t *count_host) {
  char buf[256];
  if (!count_host || n < 0 || num_nodes <= 0 || (n > 0 && (!ids_device || !out_device)))
    return gs_set_error(GS_ERR_INVALID, "gs_unique_first_noalloc: null pointer or negative size");
  if (device < 0 || device >= NA_MAX_DEVICES)
    return gs_set_error(GS_ERR_INVALID, "gs_unique_first_noalloc: device ordinal out of range");
  if (num_nodes > GS_NOALLOC_MAX_NODES || n > GS_NOALLOC_MAX_IDS) {
    snprintf(buf, sizeof buf,
             "gs_unique_first_noalloc: num_nodes %lld / n %lld beyond the static scratch "
             "(GS_NOALLOC_MAX_NODES %d, GS_NOALLOC_MAX_IDS %d)",
             (long long)num_nodes, (long long)n, GS_NOALLOC_MAX_NODES, GS_NOALLOC_MAX_IDS);
    return gs_set_error(GS_ERR_INVALID, buf);
  }
  *count_host = 0;
  if (n == 0) return GS_OK;
  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e == cudaSuccess) e = cudaSetDevice(device);
  if (e != cudaSuccess) return na_fail(GS_ERR_CUDA, "cudaSetDevice (a GPU is required: no CPU fallback)", e);
  uint32_t *scr = nullptr;
  int rc = na_scratch(device, &scr);
  uint32_t ctl[3] = {0, 0, 0};
  if (rc == GS_OK) {
    const uint32_t un = (uint32_t)n, nn = (uint32_t)num_nodes;
    int grid = (int)((n + NA_TPB - 1) / NA_TPB);
    if (grid > NA_MAX_GRID) grid = NA_MAX_GRID;
    uint32_t chunk = (uint32_t)(((n + grid - 1) / grid + NA_TPB - 1) / NA_TPB * NA_TPB);
    cudaStream_t st = cudaStreamLegacy;
    // the array keeps the last call's contents: clear the table range in use and the control words
    e = cudaMemsetAsync(scr + NA_TABLE, 0, (size_t)nn * sizeof(uint32_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(scr + NA_CTL, 0, 2 * sizeof(uint32_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(scr + NA_CTL + CTL_BAD, 0xFF, sizeof(uint32_t), st);
    if (e == cudaSuccess) {
      k_na_claim<<<grid, NA_TPB, 0, st>>>(ids_device, un, nn, scr);
      k_na_count<<<grid, NA_TPB, 0, st>>>(ids_device, un, nn, chunk, scr);
      k_na_write<<<grid, NA_TPB, 0, st>>>(ids_device, un, nn, chunk, scr, out_device);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(ctl, scr + NA_CTL, sizeof ctl, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) rc = na_fail(GS_ERR_CUDA, "dedup launch sequence", e);
  }
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  if (rc != GS_OK) return rc;
  if (ctl[CTL_BAD] != GS_INF) {
    snprintf(buf, sizeof buf, "gs_unique_first_noalloc: ids[%u] lies outside [0, num_nodes)", ctl[CTL_BAD] - 1);
    return gs_set_error(GS_ERR_INVALID, buf);
  }
  *count_host = (int64_t)ctl[CTL_TOTAL];
  return GS_OK;
}
