// flags.cu -- gs_flags_post: the hand-over's barrier without a collective.  A rank tells its
// peers "my blocks of round r are complete" by storing r + 1 into a flag word in each peer's
// memory (CUDA IPC over NVLink) and learns that a round has fully arrived by polling its OWN
// flag words; buffer releases travel the other way by the same means.  One warp per call;
// no rank waits for a rank it does not need (a library all-reduce per round cost the
// hand-over 15 % at two GPUs: lock-step plus a collective kernel per round).
// No counterpart in the reference (one process).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "slicer_internal.cuh"

namespace {

struct FlagArgs {
  uint32_t *sig[GS_MAX_PARTS];        // words to store `sig_value` into (local or peer memory)
  const uint32_t *wait[GS_MAX_PARTS]; // local words to watch until each is >= wait_value
  int32_t n_sig, n_wait;
  uint32_t sig_value, wait_value;
  unsigned long long timeout_ns;
  int32_t *status;                    // pinned mapped host word: set to 1 on a timeout
};

__global__ void __launch_bounds__(32) k_flags(FlagArgs a) {
  const int t = threadIdx.x;
  if (t < a.n_sig) {
    // everything this stream did before -- the kernels that wrote the blocks -- is ordered before
    // the flag for an observer anywhere in the system
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.sig[t]), "r"(a.sig_value) : "memory");
  }
  if (t < a.n_wait) {
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      uint32_t v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.wait[t]) : "memory");
      if ((int32_t)(v - a.wait_value) >= 0) break;
      __nanosleep(500);
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (now - t0 > a.timeout_ns) {  // a peer never got there: report, do not hang the GPU
        *a.status = 1;
        __threadfence_system();
        break;
      }
    }
  }
}

}  // namespace

extern "C" int gs_flags_post(int device, uint32_t *const *signal_words, int32_t n_signal, uint32_t signal_value,
                             const uint32_t *const *wait_words, int32_t n_wait, uint32_t wait_value,
                             uint32_t timeout_ms, int32_t *status_mapped, void *stream) {
  if (n_signal < 0 || n_signal > GS_MAX_PARTS || n_wait < 0 || n_wait > GS_MAX_PARTS ||
      (n_signal && !signal_words) || (n_wait && (!wait_words || !status_mapped)))
    return gs_set_error(GS_ERR_INVALID, "gs_flags_post: at most GS_MAX_PARTS words to signal / to wait for, and a status word");
  FlagArgs a{};
  for (int i = 0; i < n_signal; i++) a.sig[i] = signal_words[i];
  for (int i = 0; i < n_wait; i++) a.wait[i] = wait_words[i];
  a.n_sig = n_signal;
  a.n_wait = n_wait;
  a.sig_value = signal_value;
  a.wait_value = wait_value;
  a.timeout_ns = (unsigned long long)(timeout_ms ? timeout_ms : 5000u) * 1000000ull;
  a.status = status_mapped;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) {
    k_flags<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    char buf[160];
    snprintf(buf, sizeof buf, "gs_flags_post: %s", cudaGetErrorString(e));
    return gs_set_error(GS_ERR_CUDA, buf);
  }
  return GS_OK;
}
