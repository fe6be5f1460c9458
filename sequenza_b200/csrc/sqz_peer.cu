// sqz_peer.cu -- result buffers shared between the GPUs of one box (one process per GPU).
//
// The reference assembles result$LPP from its mclapply workers over pipes
// (R/bayes.R:114-116).  Here every rank's grid kernels store their results straight into
// every rank's whole-grid buffer over NVLink (sqz_baf_grid_dev_units with n_out = world),
// so the last kernel of the fit is also its all-gather; what is left of the collective is
// ordering: a rank raises its flag in every peer once its stores are out (sqz_peer_signal)
// and waits until every peer's flag is up in its own buffer (sqz_peer_wait).
//
// A peer buffer is  [ 64 flags of 8 bytes | reserved to kPeerHeader ][ data ... ], allocated
// with cudaMalloc and exported with CUDA IPC; peers open the handle and get an address that
// is valid in their process (NVLink peer mapping).
#include "sqz_common.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

constexpr size_t kPeerHeader = 1024;
constexpr int kPeerMaxRanks = 16;

struct PeerPtrs {
  unsigned long long* flags[kPeerMaxRanks];
};

// one thread per peer: all earlier writes of this GPU (the grid kernels before this one in
// the stream) become visible system-wide, then the flag goes up in that peer's header
__global__ void peer_signal_kernel(PeerPtrs p, int n, int my_rank, unsigned long long epoch) {
  const int k = threadIdx.x;
  if (k >= n) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p.flags[k] + my_rank), "l"(epoch)
               : "memory");
}

// one thread per rank: spin on this GPU's own header until that rank's flag shows `epoch`
// (flags only grow).  A peer that never arrives is a broken job: after ten seconds the
// kernel traps instead of returning data that is not there.
__global__ void peer_wait_kernel(const unsigned long long* flags, int n,
                                 unsigned long long epoch) {
  const int r = threadIdx.x;
  if (r < n) {
    unsigned long long t0, t1, v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + r) : "memory");
      if (v >= epoch) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 10000000000ull) __trap();
      __nanosleep(200);
    }
  }
  __syncthreads();
  __threadfence_system();
}

}  // namespace sqz

using namespace sqz;

extern "C" {

size_t sqz_peer_data_offset(void) { return kPeerHeader; }

int sqz_peer_alloc(size_t data_bytes, void** d_base, unsigned char* handle64,
                   size_t* total_bytes) {
  if (!d_base || !handle64) return fail(SQZ_ERR_ARG, "null pointer argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  void* p = nullptr;
  const size_t total = kPeerHeader + ((data_bytes + 255) & ~(size_t)255);
  SQZ_CUDA(cudaMalloc(&p, total));
  cudaError_t e = cudaMemset(p, 0, total);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    cudaGetLastError();
    return fail(SQZ_ERR_CUDA, std::string("peer buffer: ") + cudaGetErrorString(e));
  }
  memcpy(handle64, &h, 64);
  *d_base = p;
  if (total_bytes) *total_bytes = total;
  return SQZ_OK;
}

int sqz_peer_open(const unsigned char* handle64, void** d_base) {
  if (!d_base || !handle64) return fail(SQZ_ERR_ARG, "null pointer argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(SQZ_ERR_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
  }
  *d_base = p;
  return SQZ_OK;
}

int sqz_peer_close(void* d_base) {
  if (d_base) SQZ_CUDA(cudaIpcCloseMemHandle(d_base));
  return SQZ_OK;
}

int sqz_peer_free(void* d_base) {
  if (d_base) SQZ_CUDA(cudaFree(d_base));
  return SQZ_OK;
}

int sqz_peer_signal(void* const* d_bases, int n, int my_rank, uint64_t epoch, void* stream) {
  if (!d_bases || n < 1 || n > kPeerMaxRanks || my_rank < 0 || my_rank >= n)
    return fail(SQZ_ERR_ARG, "sqz_peer_signal: between 1 and 16 ranks");
  PeerPtrs p;
  for (int k = 0; k < kPeerMaxRanks; ++k)
    p.flags[k] = k < n ? (unsigned long long*)d_bases[k] : nullptr;
  peer_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(p, n, my_rank,
                                                         (unsigned long long)epoch);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

int sqz_peer_wait(void* d_base, int n, uint64_t epoch, void* stream) {
  if (!d_base || n < 1 || n > kPeerMaxRanks)
    return fail(SQZ_ERR_ARG, "sqz_peer_wait: between 1 and 16 ranks");
  peer_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((const unsigned long long*)d_base, n,
                                                       (unsigned long long)epoch);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

}  // extern "C"
