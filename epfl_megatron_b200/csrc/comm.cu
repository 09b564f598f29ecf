// Peer-mapped ("symmetric") memory for the GPUs of one NVSwitch domain plus the small stream-
// ordered signalling kernels of the fused tensor-parallel GEMMs (gemm_comm_tcgen05.cu).
//
// One process per GPU: every rank cudaMalloc's a region, exports it with cudaIpcGetMemHandle,
// the handles travel through torch.distributed (parallel/symm.py), and each rank maps its
// peers' regions with cudaIpcOpenMemHandle. After that kernels read and write peer memory with
// ordinary ld/st on the mapped addresses (NVLink P2P). The reference has no counterpart: all its
// tensor-parallel traffic goes through NCCL calls.
#include "common.cuh"
#include <string.h>

namespace b200 {

__global__ void signal_kernel(unsigned* a0, unsigned* a1, unsigned* a2, unsigned* a3, unsigned* a4,
                              unsigned* a5, unsigned* a6, unsigned* a7, int n, unsigned value) {
  unsigned* a[8] = {a0, a1, a2, a3, a4, a5, a6, a7};
  if (threadIdx.x < n && a[threadIdx.x] != nullptr) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a[threadIdx.x]), "r"(value) : "memory");
  }
}

// Spin until flags[i] (i < n, i != skip) have reached `value` (wrap-safe); bounded.
__global__ void wait_kernel(const unsigned* flags, int n, int skip, unsigned value) {
  const int i = threadIdx.x;
  if (i >= n || i == skip) return;
  const long long t0 = clock64();
  for (;;) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + i) : "memory");
    if (static_cast<int>(v - value) >= 0) break;
    __nanosleep(128);
    if (clock64() - t0 > 10000000000LL) asm volatile("trap;");
  }
}

// dst0 (and dst1 when non-null) <- src, 16 bytes per thread step.
__global__ void copy2_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst0,
                             uint4* __restrict__ dst1, long long n16) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
    const uint4 v = src[i];
    dst0[i] = v;
    if (dst1) dst1[i] = v;
  }
}

constexpr int kStageBlocks = 148 * 4;

// Stand-alone staging for kernels that pull rows themselves (parallel/symm.py gather_begin, e.g. the
// embedding backward; the all-gather GEMM stages inside its own kernel): (1) wait until every peer has
// pulled the previous content of this staging shard, (2) copy the local rows into the staging shard
// (peers pull from there) and, when asked, into a second destination, (3) the last block to finish
// tells the peers "epoch e ready".
__global__ void stage_shard_kernel(const uint4* __restrict__ src, uint4* __restrict__ stage,
                                   uint4* __restrict__ own_rows, long long n16,
                                   const unsigned* consumed, int tp, int rank, unsigned wait_value,
                                   int do_wait, unsigned* done_counter, unsigned done_target,
                                   unsigned* s0, unsigned* s1, unsigned* s2, unsigned* s3, unsigned* s4,
                                   unsigned* s5, unsigned* s6, int n_signal, unsigned epoch) {
  if (do_wait) {
    if (threadIdx.x < tp && threadIdx.x != rank) {
      const long long t0 = clock64();
      for (;;) {
        unsigned v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(consumed + threadIdx.x) : "memory");
        if (static_cast<int>(v - wait_value) >= 0) break;
        __nanosleep(128);
        if (clock64() - t0 > 10000000000LL) asm volatile("trap;");
      }
    }
    __syncthreads();
  }
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
    const uint4 v = src[i];
    stage[i] = v;
    own_rows[i] = v;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned old = atomicAdd(done_counter, 1u);
    if (old + 1u == done_target) {
      unsigned* sig[7] = {s0, s1, s2, s3, s4, s5, s6};
      __threadfence_system();
      for (int i = 0; i < n_signal; ++i)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sig[i]), "r"(epoch) : "memory");
    }
  }
}

}  // namespace b200

using namespace b200;

// See stage_shard_kernel. consumed: local flags [tp]; wait_value: epoch whose pull must be over
// (ignored when do_wait == 0); done_counter: a local u32 used only by this entry point, call_index
// = 1, 2, ... per workspace; signal_addrs: the peers' shard_ready[<me>] words.
B200_API int b200_stage_shard(const void* src, void* stage, void* own_rows, long long bytes,
                              const unsigned* consumed, int tp, int rank, unsigned wait_value,
                              int do_wait, unsigned* done_counter, unsigned call_index,
                              unsigned* const* signal_addrs, int n_signal, unsigned epoch,
                              cudaStream_t stream) {
  if (bytes % 16 || n_signal < 0 || n_signal > 7 || tp > 32) return B200_ERR_BAD_ARG;
  unsigned* s[7] = {};
  for (int i = 0; i < n_signal; ++i) s[i] = signal_addrs[i];
  stage_shard_kernel<<<kStageBlocks, 256, 0, stream>>>(
      (const uint4*)src, (uint4*)stage, (uint4*)own_rows, bytes / 16, consumed, tp, rank, wait_value,
      do_wait, done_counter, call_index * (unsigned)kStageBlocks, s[0], s[1], s[2], s[3], s[4], s[5], s[6],
      n_signal, epoch);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_symm_alloc(long long bytes, void** out) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemset(p, 0, (size_t)bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return (int)e;
  *out = p;
  return 0;
}

B200_API int b200_symm_free(void* p) { return (int)cudaFree(p); }

// handle: 64 bytes
B200_API int b200_symm_export(void* p, void* handle_out) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) return (int)e;
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

B200_API int b200_symm_import(const void* handle, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return (int)e;
  *out = p;
  return 0;
}

B200_API int b200_symm_unmap(void* p) { return (int)cudaIpcCloseMemHandle(p); }

// Store `value` (release, system scope) to up to 8 addresses, ordered after the stream's
// previous work: "my shard is ready" / "I have consumed yours".
B200_API int b200_symm_signal(unsigned* const* addrs, int n, unsigned value, cudaStream_t stream) {
  if (n < 1 || n > 8) return B200_ERR_BAD_ARG;
  unsigned* a[8] = {};
  for (int i = 0; i < n; ++i) a[i] = addrs[i];
  signal_kernel<<<1, 32, 0, stream>>>(a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], n, value);
  B200_CHECK_LAUNCH();
  return 0;
}

// Block the stream until local flags[0..n) except `skip` have reached `value`.
B200_API int b200_symm_wait(const unsigned* flags, int n, int skip, unsigned value, cudaStream_t stream) {
  if (n < 1 || n > 32) return B200_ERR_BAD_ARG;
  wait_kernel<<<1, 32, 0, stream>>>(flags, n, skip, value);
  B200_CHECK_LAUNCH();
  return 0;
}

// src -> dst0 and (optionally) dst1; bytes must be a multiple of 16.
B200_API int b200_copy2(const void* src, void* dst0, void* dst1, long long bytes, cudaStream_t stream) {
  if (bytes % 16) return B200_ERR_UNSUPPORTED;
  const long long n16 = bytes / 16;
  long long blocks = (n16 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  copy2_kernel<<<(unsigned)blocks, 256, 0, stream>>>((const uint4*)src, (uint4*)dst0, (uint4*)dst1, n16);
  B200_CHECK_LAUNCH();
  return 0;
}
