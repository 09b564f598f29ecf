// Data-parallel gradient reduce-scatter and parameter all-gather over peer-mapped memory, fused
// with the scale and with the partitioned Adam update (the distributed optimizer's two
// collectives, megatron/optimizer/distrib_optimizer.py:553-567 and :592-608, which the reference
// issues as one blocking NCCL reduce_scatter_tensor + one blocking _all_gather_base per step).
//
//   dp_reduce_kernel   the owner of a slice of the fp32 gradient buffer waits for its peers'
//                      "bucket ready" flags, PULLS the slice out of every peer's buffer over NVLink
//                      (16-byte loads, all ranks in flight at once), sums in rank order, scales by
//                      1/dp and overwrites its own copy. Launched per bucket on a side stream while
//                      the backward pass is still producing the next buckets.
//   adam_push_kernel   Adam/AdamW on the owned shard (fp32 master weights and moments) that writes
//                      the new bf16 parameters straight into the parameter buffer of EVERY rank
//                      (16-byte peer stores): the all-gather is the kernel's own epilogue.
//
// Flags are monotonic step numbers written with st.release.sys by the producer and polled with
// ld.acquire.sys; waits are bounded (about a minute): a timed-out wait records {site, value, target,
// index} in g_dp_debug, prints one line and traps, so a lost peer ends the job with a launch error
// instead of a hung device or a step computed from half-reduced gradients.
#include "common.cuh"

namespace b200 {

constexpr int kDpMax = 8;

__device__ unsigned g_dp_debug[8];

struct DpSrc {
  const float4* g[kDpMax];   // rank r's copy of the slice (g[rank] is the local one)
};
struct DpDst {
  uint4* p[kDpMax];          // rank r's bf16 parameter slice
};

__device__ __forceinline__ void dp_spin(const unsigned* p, unsigned target, unsigned site, unsigned index) {
  const long long t0 = clock64();
  for (;;) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    if (static_cast<int>(v - target) >= 0) return;
    __nanosleep(256);
    if (clock64() - t0 > 120000000000LL) {
      if (atomicCAS(&g_dp_debug[0], 0u, site) == 0u) {
        g_dp_debug[1] = v;
        g_dp_debug[2] = target;
        g_dp_debug[3] = index;
        g_dp_debug[4] = blockIdx.x;
        __threadfence_system();
        printf("b200 fused DP path: wait timed out (site %u, cta %u, saw %u, wanted %u, index %u)\n", site,
               blockIdx.x, v, target, index);
      }
      asm volatile("trap;");
    }
  }
}

__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  // .cv: never served from a stale line, the peer rewrites this buffer every step
  asm volatile("ld.global.cv.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// out[i] = scale * sum_r src.g[r][i]   (out == src.g[rank], in place)
template <int kDp>
__global__ void __launch_bounds__(256)
dp_reduce_kernel(DpSrc src, int rank, float4* out, long long n_vec, float scale,
                 const unsigned* __restrict__ ready, unsigned epoch) {
  if (threadIdx.x < kDp && threadIdx.x != rank) dp_spin(ready + threadIdx.x, epoch, 1u, threadIdx.x);
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kU = kDp <= 2 ? 8 : (kDp <= 4 ? 4 : 2);   // 16-byte loads in flight per thread: kU * kDp
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  for (; i + (kU - 1) * stride < n_vec; i += kU * stride) {
    float4 v[kU][kDp];
#pragma unroll
    for (int u = 0; u < kU; ++u)
#pragma unroll
      for (int r = 0; r < kDp; ++r) v[u][r] = ld_peer_f4(src.g[r] + i + u * stride);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      float4 a = v[u][0];
#pragma unroll
      for (int r = 1; r < kDp; ++r) {
        a.x += v[u][r].x; a.y += v[u][r].y; a.z += v[u][r].z; a.w += v[u][r].w;
      }
      out[i + u * stride] = make_float4(a.x * scale, a.y * scale, a.z * scale, a.w * scale);
    }
  }
  for (; i < n_vec; i += stride) {
    float4 a = ld_peer_f4(src.g[0] + i);
#pragma unroll
    for (int r = 1; r < kDp; ++r) {
      const float4 b = ld_peer_f4(src.g[r] + i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    out[i] = make_float4(a.x * scale, a.y * scale, a.z * scale, a.w * scale);
  }
}

// Adam on 8 elements per thread step; the bf16 result goes to n_dst parameter buffers.
__global__ void __launch_bounds__(256)
adam_push_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, DpDst dst, int n_dst, long long n_oct, float lr, float beta1,
                 float beta2, float eps, float wd, float inv_bc1, float inv_bc2, int adamw,
                 const float* __restrict__ coef, const float* __restrict__ skip) {
  if (skip && skip[0] != 0.f) return;
  const float gs = coef ? coef[0] : 1.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_oct;
       i += (long long)gridDim.x * blockDim.x) {
    float pa[8], ga[8], ma[8], va[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float4 pv = reinterpret_cast<const float4*>(p)[2 * i + h];
      const float4 gv = reinterpret_cast<const float4*>(g)[2 * i + h];
      const float4 mv = reinterpret_cast<const float4*>(m)[2 * i + h];
      const float4 vv = reinterpret_cast<const float4*>(v)[2 * i + h];
      pa[4 * h] = pv.x; pa[4 * h + 1] = pv.y; pa[4 * h + 2] = pv.z; pa[4 * h + 3] = pv.w;
      ga[4 * h] = gv.x; ga[4 * h + 1] = gv.y; ga[4 * h + 2] = gv.z; ga[4 * h + 3] = gv.w;
      ma[4 * h] = mv.x; ma[4 * h + 1] = mv.y; ma[4 * h + 2] = mv.z; ma[4 * h + 3] = mv.w;
      va[4 * h] = vv.x; va[4 * h + 1] = vv.y; va[4 * h + 2] = vv.z; va[4 * h + 3] = vv.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float gr = ga[k] * gs;
      if (!adamw) gr += wd * pa[k];
      ma[k] = beta1 * ma[k] + (1.f - beta1) * gr;
      va[k] = beta2 * va[k] + (1.f - beta2) * gr * gr;
      float upd = (ma[k] * inv_bc1) / (sqrtf(va[k] * inv_bc2) + eps);
      if (adamw) upd += wd * pa[k];
      pa[k] -= lr * upd;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      reinterpret_cast<float4*>(p)[2 * i + h] = make_float4(pa[4 * h], pa[4 * h + 1], pa[4 * h + 2], pa[4 * h + 3]);
      reinterpret_cast<float4*>(m)[2 * i + h] = make_float4(ma[4 * h], ma[4 * h + 1], ma[4 * h + 2], ma[4 * h + 3]);
      reinterpret_cast<float4*>(v)[2 * i + h] = make_float4(va[4 * h], va[4 * h + 1], va[4 * h + 2], va[4 * h + 3]);
    }
    uint4 o;
    o.x = pack_bf16x2(pa[0], pa[1]);
    o.y = pack_bf16x2(pa[2], pa[3]);
    o.z = pack_bf16x2(pa[4], pa[5]);
    o.w = pack_bf16x2(pa[6], pa[7]);
#pragma unroll
    for (int r = 0; r < kDpMax; ++r)
      if (r < n_dst) dst.p[r][i] = o;
  }
}

// Soft-bounded version of comm.cu's wait: flags[i] >= value for i < n, i != skip.
__global__ void dp_wait_kernel(const unsigned* flags, int n, int skip, unsigned value, unsigned site) {
  const int i = threadIdx.x;
  if (i < n && i != skip) dp_spin(flags + i, value, site, i);
}

static inline int dp_flat_grid(long long n_items) {
  long long g = (n_items + 255) / 256;
  long long cap = (long long)kNumSMs * 8;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

}  // namespace b200

using namespace b200;

// In-place reduce of one slice: grads[r] = rank r's address of the slice (grads[rank] local),
// n floats (multiple of 4), ready = local flag words [dp] that peers set to `epoch` once their
// copy of the slice is final. blocks: grid size (small under the backward pass, large after it).
B200_API int b200_dp_reduce(const void* const* grads, int dp, int rank, long long n, float scale,
                            const unsigned* ready, unsigned epoch, int blocks, cudaStream_t stream) {
  if (dp < 2 || dp > kDpMax || rank < 0 || rank >= dp || n % 4 || blocks < 1) return B200_ERR_BAD_ARG;
  if (n == 0) return 0;
  DpSrc s = {};
  for (int r = 0; r < dp; ++r) s.g[r] = (const float4*)grads[r];
  float4* out = (float4*)grads[rank];
  const long long nv = n >> 2;
  long long need = (nv + 255) / 256;
  if (need < blocks) blocks = (int)need;
#define B200_DP_CASE(D)                                                                         \
  case D:                                                                                        \
    dp_reduce_kernel<D><<<blocks, 256, 0, stream>>>(s, rank, out, nv, scale, ready, epoch);       \
    break;
  switch (dp) {
    B200_DP_CASE(2) B200_DP_CASE(3) B200_DP_CASE(4) B200_DP_CASE(5) B200_DP_CASE(6) B200_DP_CASE(7)
    B200_DP_CASE(8)
  }
#undef B200_DP_CASE
  B200_CHECK_LAUNCH();
  return 0;
}

// Adam on [n] (multiple of 8) with the bf16 copy written to n_dst buffers (own + peers).
B200_API int b200_adam_step_push(float* p, const float* g, float* m, float* v, void* const* p16_dst,
                                 int n_dst, long long n, float lr, float beta1, float beta2, float eps,
                                 float wd, int step, int adamw, int bias_correction, const float* coef,
                                 const float* skip, cudaStream_t stream) {
  if (n % 8 || n_dst < 1 || n_dst > kDpMax) return B200_ERR_UNSUPPORTED;
  if (n == 0) return 0;
  float inv_bc1 = 1.f, inv_bc2 = 1.f;
  if (bias_correction) {
    inv_bc1 = 1.f / (1.f - powf(beta1, (float)step));
    inv_bc2 = 1.f / (1.f - powf(beta2, (float)step));
  }
  DpDst d = {};
  for (int r = 0; r < n_dst; ++r) d.p[r] = (uint4*)p16_dst[r];
  adam_push_kernel<<<dp_flat_grid(n >> 3), 256, 0, stream>>>(p, g, m, v, d, n_dst, n >> 3, lr, beta1, beta2,
                                                          eps, wd, inv_bc1, inv_bc2, adamw, coef, skip);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_dp_wait(const unsigned* flags, int n, int skip, unsigned value, unsigned site,
                          cudaStream_t stream) {
  if (n < 1 || n > 32) return B200_ERR_BAD_ARG;
  dp_wait_kernel<<<1, 32, 0, stream>>>(flags, n, skip, value, site);
  B200_CHECK_LAUNCH();
  return 0;
}

// Stream-ordered copy through the copy engines (src may be a peer-mapped address): moves a peer's
// gradient slice into local staging without occupying SMs while the backward pass computes.
B200_API int b200_copy_async(void* dst, const void* src, long long bytes, cudaStream_t stream) {
  return (int)cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, stream);
}

// Copies and clears the first timed-out wait (all zero: none). Synchronises the device.
B200_API int b200_dp_debug_read(unsigned* out8) {
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return (int)e;
  e = cudaMemcpyFromSymbol(out8, g_dp_debug, sizeof(unsigned) * 8);
  if (e != cudaSuccess) return (int)e;
  unsigned zero[8] = {};
  return (int)cudaMemcpyToSymbol(g_dp_debug, zero, sizeof(zero));
}
