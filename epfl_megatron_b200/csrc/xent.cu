// Vocab-parallel cross-entropy on bf16 logits with fp32 math (K14 building block).
//
// Reference behaviour: megatron/core/tensor_parallel/cross_entropy.py:16-126 runs on an
// fp32 copy of the [s, b, V/tp] logits and issues three TP all-reduces (max, target
// logit, sum-exp). Here each rank makes ONE pass over its bf16 shard producing the
// triple (local max, local sum-exp relative to that max, target logit if owned); the
// triples are exchanged once across the TP group and folded by `xent_combine`; the
// backward rewrites the logits buffer in place with (softmax - onehot) * dloss.
#include "common.cuh"

namespace b200 {

constexpr int kXentThreads = 256;

// stats[row] = {max, sumexp, target_logit_or_0}
__global__ void __launch_bounds__(kXentThreads)
xent_stats_kernel(const __nv_bfloat16* __restrict__ logits, const long long* __restrict__ labels,
                  float* __restrict__ stats, int T, int V, long long vocab_start) {
  __shared__ float red[33];
  const int tid = threadIdx.x;
  const int nvec = V >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const bf16x8* lr = reinterpret_cast<const bf16x8*>(logits + (size_t)row * V);
    float m = -INFINITY, s = 0.f;
    for (int v = tid; v < nvec; v += kXentThreads) {
      float f[8];
      unpack8(lr[v], f);
      float vm = f[0];
#pragma unroll
      for (int k = 1; k < 8; ++k) vm = fmaxf(vm, f[k]);
      if (vm > m) {
        s *= __expf(m - vm);
        m = vm;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) s += __expf(f[k] - m);
    }
    const float gm = block_max<kXentThreads>(m, red);
    s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    s = block_sum<kXentThreads>(s, red);
    if (tid == 0) {
      long long lab = labels[row] - vocab_start;
      float tgt = 0.f;
      if (lab >= 0 && lab < V) tgt = __bfloat162float(logits[(size_t)row * V + lab]);
      stats[3 * (size_t)row] = gm;
      stats[3 * (size_t)row + 1] = s;
      stats[3 * (size_t)row + 2] = tgt;
    }
  }
}

// Tensor-parallel exchange inside the kernels (K14): the statistics pass stores its triple straight
// into slot [rank][row] of EVERY rank's peer-mapped table (12 bytes per token per peer over NVLink)
// and each CTA then bumps every rank's arrival counter (release, system scope); the combine kernel
// waits for the counter and folds the table. No collective call: the reference's three all-reduces
// (cross_entropy.py:21-23, :52-54, :60-62) become peer stores issued by the producer itself.
struct XentPeers {
  float* stats[8];      // rank r's table [tp][T][3] of this call's set
  unsigned* cnt[8];     // rank r's arrival counter of this set
};

__global__ void __launch_bounds__(kXentThreads)
xent_stats_push_kernel(const __nv_bfloat16* __restrict__ logits, const long long* __restrict__ labels,
                       XentPeers peers, int tp, int rank, int T, int V, long long vocab_start) {
  __shared__ float red[33];
  const int tid = threadIdx.x;
  const int nvec = V >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const bf16x8* lr = reinterpret_cast<const bf16x8*>(logits + (size_t)row * V);
    float m = -INFINITY, s = 0.f;
    for (int v = tid; v < nvec; v += kXentThreads) {
      float f[8];
      unpack8(lr[v], f);
      float vm = f[0];
#pragma unroll
      for (int k = 1; k < 8; ++k) vm = fmaxf(vm, f[k]);
      if (vm > m) {
        s *= __expf(m - vm);
        m = vm;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) s += __expf(f[k] - m);
    }
    const float gm = block_max<kXentThreads>(m, red);
    s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    s = block_sum<kXentThreads>(s, red);
    if (tid < tp) {
      long long lab = labels[row] - vocab_start;
      float tgt = 0.f;
      if (lab >= 0 && lab < V) tgt = __bfloat162float(logits[(size_t)row * V + lab]);
      float* dst = peers.stats[tid] + ((size_t)rank * T + row) * 3;   // thread q serves rank q
      dst[0] = gm;
      dst[1] = s;
      dst[2] = tgt;
    }
  }
  __syncthreads();
  if (tid < tp) {
    __threadfence_system();
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(peers.cnt[tid]), "r"(1u) : "memory");
  }
}

// Waits until `target` CTAs (of all ranks, all calls on this set so far) have delivered, then folds.
__global__ void xent_combine_wait_kernel(const float* all_stats, const unsigned* cnt, unsigned target,
                                         float* __restrict__ loss, float* __restrict__ lse, int T, int tp) {
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
      if (static_cast<int>(v - target) >= 0) break;
      __nanosleep(64);
      if (clock64() - t0 > 120000000000LL) {
        printf("b200 cross entropy exchange: wait timed out (saw %u, wanted %u)\n", v, target);
        asm volatile("trap;");
      }
    }
  }
  __syncthreads();
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const volatile float* st = all_stats;
  float m = -INFINITY;
  for (int r = 0; r < tp; ++r) m = fmaxf(m, st[((size_t)r * T + t) * 3]);
  float s = 0.f, tgt = 0.f;
  for (int r = 0; r < tp; ++r) {
    const volatile float* p = st + ((size_t)r * T + t) * 3;
    s += p[1] * __expf(p[0] - m);
    tgt += p[2];
  }
  float l = m + __logf(s);
  lse[t] = l;
  loss[t] = l - tgt;
}

// all_stats: [tp, T, 3] gathered triples -> loss[T], lse[T]
__global__ void xent_combine_kernel(const float* __restrict__ all_stats, float* __restrict__ loss,
                                    float* __restrict__ lse, int T, int tp) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  float m = -INFINITY;
  for (int r = 0; r < tp; ++r) m = fmaxf(m, all_stats[((size_t)r * T + t) * 3]);
  float s = 0.f, tgt = 0.f;
  for (int r = 0; r < tp; ++r) {
    const float* p = all_stats + ((size_t)r * T + t) * 3;
    s += p[1] * __expf(p[0] - m);
    tgt += p[2];
  }
  float l = m + __logf(s);
  lse[t] = l;
  loss[t] = l - tgt;
}

// logits <- (exp(logit - lse) - [j == label]) * dloss[row]   (in place)
__global__ void __launch_bounds__(kXentThreads)
xent_bwd_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ labels,
                const float* __restrict__ lse, const float* __restrict__ dloss, int T, int V,
                long long vocab_start) {
  const int tid = threadIdx.x;
  const int nvec = V >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    bf16x8* lr = reinterpret_cast<bf16x8*>(logits + (size_t)row * V);
    const float l = lse[row], g = dloss[row];
    const long long lab = labels[row] - vocab_start;
    for (int v = tid; v < nvec; v += kXentThreads) {
      float f[8];
      unpack8(lr[v], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float p = __expf(f[k] - l);
        if ((long long)(v * 8 + k) == lab) p -= 1.f;
        f[k] = p * g;
      }
      lr[v] = pack8(f);
    }
  }
}

}  // namespace b200

using namespace b200;

B200_API int b200_xent_stats(const void* logits, const long long* labels, float* stats, int T,
                             int V, long long vocab_start, cudaStream_t stream) {
  if (V % 8) return B200_ERR_UNSUPPORTED;
  if (T == 0) return 0;
  int grid = T < kNumSMs * 8 ? T : kNumSMs * 8;
  xent_stats_kernel<<<grid, kXentThreads, 0, stream>>>((const __nv_bfloat16*)logits, labels, stats,
                                                       T, V, vocab_start);
  B200_CHECK_LAUNCH();
  return 0;
}

// CTAs b200_xent_stats_push launches for T rows (every rank's counter receives tp times this per call).
B200_API int b200_xent_push_grid(int T) { return T < kNumSMs * 8 ? T : kNumSMs * 8; }

// peer_stats[r] / peer_cnt[r]: rank r's table [tp][T][3] and arrival counter of the set this call uses.
B200_API int b200_xent_stats_push(const void* logits, const long long* labels, float* const* peer_stats,
                                  unsigned* const* peer_cnt, int tp, int rank, int T, int V,
                                  long long vocab_start, cudaStream_t stream) {
  if (V % 8) return B200_ERR_UNSUPPORTED;
  if (tp < 2 || tp > 8 || rank < 0 || rank >= tp) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  XentPeers p = {};
  for (int r = 0; r < tp; ++r) {
    p.stats[r] = peer_stats[r];
    p.cnt[r] = peer_cnt[r];
  }
  xent_stats_push_kernel<<<b200_xent_push_grid(T), kXentThreads, 0, stream>>>(
      (const __nv_bfloat16*)logits, labels, p, tp, rank, T, V, vocab_start);
  B200_CHECK_LAUNCH();
  return 0;
}

// target: running total of arrivals expected on `cnt` (wrap-safe compare), see b200_xent_push_grid.
B200_API int b200_xent_combine_wait(const float* all_stats, const unsigned* cnt, unsigned target,
                                    float* loss, float* lse, int T, int tp, cudaStream_t stream) {
  if (T == 0) return 0;
  xent_combine_wait_kernel<<<(T + 255) / 256, 256, 0, stream>>>(all_stats, cnt, target, loss, lse, T, tp);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_xent_combine(const float* all_stats, float* loss, float* lse, int T, int tp,
                               cudaStream_t stream) {
  if (T == 0) return 0;
  xent_combine_kernel<<<(T + 255) / 256, 256, 0, stream>>>(all_stats, loss, lse, T, tp);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_xent_bwd(void* logits, const long long* labels, const float* lse,
                           const float* dloss, int T, int V, long long vocab_start,
                           cudaStream_t stream) {
  if (V % 8) return B200_ERR_UNSUPPORTED;
  if (T == 0) return 0;
  int grid = T < kNumSMs * 8 ? T : kNumSMs * 8;
  xent_bwd_kernel<<<grid, kXentThreads, 0, stream>>>((__nv_bfloat16*)logits, labels, lse, dloss, T,
                                                     V, vocab_start);
  B200_CHECK_LAUNCH();
  return 0;
}
