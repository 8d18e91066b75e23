// Probe: how fast can a kernel mirror a device buffer into pinned, mapped host
// memory?  (a) 16-byte SM stores (what k_finish does), (b) TMA bulk stores
// shared -> global (cp.async.bulk), (c) the copy engine.  Diagnostic only.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void k_st16(const int4 *src, int4 *dst, int64_t n4) {
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += nth) dst[i] = src[i];
}

template <int CHUNK>  // bytes per bulk store, two buffers per CTA
__global__ void __launch_bounds__(256) k_tma(const char *src, char *dst, int64_t bytes) {
  extern __shared__ __align__(128) char smem[];
  const int64_t nchunks = bytes / CHUNK;
  int buf = 0;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x, buf ^= 1) {
    char *s = smem + buf * CHUNK;
    // the bulk store that last read this buffer (two iterations ago) must have finished reading it
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    __syncthreads();
    const int4 *g = (const int4 *)(src + c * CHUNK);
    int4 *s4 = (int4 *)s;
    for (int i = threadIdx.x; i < CHUNK / 16; i += blockDim.x) s4[i] = g[i];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t sa = (uint32_t)__cvta_generic_to_shared(s);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * CHUNK), "r"(sa), "r"(CHUNK) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main() {
  const int64_t bytes = 40ll << 20;
  char *d, *h, *hd;
  CK(cudaMalloc(&d, bytes));
  CK(cudaMemset(d, 0x5a, bytes));
  CK(cudaHostAlloc(&h, bytes, cudaHostAllocMapped));
  CK(cudaHostGetDevicePointer((void **)&hd, h, 0));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  auto time = [&](const char *name, auto fn) {
    memset(h, 0, bytes);
    for (int i = 0; i < 2; i++) fn();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 10; i++) fn();
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    bool ok = true;
    for (int64_t i = 0; i < bytes; i += 4093) ok &= h[i] == 0x5a;
    ok &= h[bytes - 1] == 0x5a;
    printf("%-34s %6.1f GB/s %s\n", name, bytes * 10 / (ms * 1e-3) / 1e9, ok ? "ok" : "MISMATCH");
  };
  time("copy engine", [&] { CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, 0)); });
  for (int g : {148, 296, 592, 1184}) {
    char nm[64]; snprintf(nm, 64, "st.v4 grid %d x 256", g);
    time(nm, [&] { k_st16<<<g, 256>>>((const int4 *)d, (int4 *)hd, bytes / 16); });
  }
  for (int g : {32, 74, 148, 296}) {
    char nm[64];
    snprintf(nm, 64, "TMA bulk 4 KB grid %d", g);
    time(nm, [&] { k_tma<4096><<<g, 256, 2 * 4096>>>(d, hd, bytes); });
    snprintf(nm, 64, "TMA bulk 16 KB grid %d", g);
    time(nm, [&] { k_tma<16384><<<g, 256, 2 * 16384>>>(d, hd, bytes); });
  }
  CK(cudaFuncSetAttribute(k_tma<65536>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 65536));
  for (int g : {16, 74, 148})  {
    char nm[64]; snprintf(nm, 64, "TMA bulk 64 KB grid %d", g);
    time(nm, [&] { k_tma<65536><<<g, 256, 2 * 65536>>>(d, hd, bytes); });
  }
  return 0;
}
