// RMSNorm / LayerNorm forward + backward for sm_100a (K1, K2 of SURVEY.md section 8).
//
// Behaviour parity (not code) with the reference's eager RMSNorm
// (megatron/model/fused_layer_norm.py:125-139: fp32 x*rsqrt(mean(x^2)+eps), cast to
// the input dtype, THEN multiply by the weight) and with its mixed-precision
// LayerNorm extension (megatron/fused_kernels/layer_norm_cuda_kernel.cu: Welford
// mean/var per row, two-stage gamma/beta gradient reduction).
//
// Design: one CTA walks rows grid-stride, the whole row lives in registers as 16-byte
// bf16 vectors (<= 4 vectors per thread at 256 threads -> H <= 8192), statistics are
// fp32. An optional residual-add prologue (h = x + residual, h written back) removes a
// separate elementwise pass. The weight/bias gradients are accumulated per CTA in
// registers across all of that CTA's rows and written once to a [grid, H] fp32
// scratch; a second tiny kernel folds the scratch into the (optionally accumulated)
// fp32 gradient.
#include <cstdlib>
#include "common.cuh"

namespace b200 {

constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 4;  // 256 thr * 4 vec * 8 elem = 8192 columns

template <bool kHasResidual>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                   const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y,
                   __nv_bfloat16* __restrict__ h_out, float* __restrict__ rstd_out, int M, int H,
                   float eps) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  const int tid = threadIdx.x;
  bf16x8 wv[kNormMaxVec];
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    int v = tid + i * kNormThreads;
    if (v < nvec) wv[i] = reinterpret_cast<const bf16x8*>(w)[v];
  }
  for (int row = blockIdx.x; row < M; row += gridDim.x) {
    const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * H);
    float f[kNormMaxVec][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        bf16x8 p = xr[v];
        unpack8(p, f[i]);
        if (kHasResidual) {
          bf16x8 q = reinterpret_cast<const bf16x8*>(res + (size_t)row * H)[v];
          float g[8];
          unpack8(q, g);
#pragma unroll
          for (int k = 0; k < 8; ++k) f[i][k] = bf16_round(f[i][k] + g[k]);
          reinterpret_cast<bf16x8*>(h_out + (size_t)row * H)[v] = pack8(f[i]);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) ss += f[i][k] * f[i][k];
      }
    }
    ss = block_sum<kNormThreads>(ss, red);
    const float rstd = rsqrtf(ss / (float)H + eps);
    if (tid == 0 && rstd_out) rstd_out[row] = rstd;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        float wf[8], o[8];
        unpack8(wv[i], wf);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = bf16_round(f[i][k] * rstd) * wf[k];
        reinterpret_cast<bf16x8*>(y + (size_t)row * H)[v] = pack8(o);
      }
    }
  }
}

// dx = rstd * (dy*w - xhat * mean(dy*w*xhat)); optionally dx += dres (gradient arriving
// through the residual branch). Per-CTA dw partials go to `partial[blockIdx.x]`.
template <bool kHasDres>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                   const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                   float* __restrict__ partial, int M, int H) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  const int tid = threadIdx.x;
  float wf[kNormMaxVec][8], dwacc[kNormMaxVec][8];
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    int v = tid + i * kNormThreads;
#pragma unroll
    for (int k = 0; k < 8; ++k) dwacc[i][k] = 0.f;
    if (v < nvec) unpack8(reinterpret_cast<const bf16x8*>(w)[v], wf[i]);
  }
  for (int row = blockIdx.x; row < M; row += gridDim.x) {
    const float rstd = rstd_in[row];
    float xh[kNormMaxVec][8], g[kNormMaxVec][8];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        unpack8(reinterpret_cast<const bf16x8*>(x + (size_t)row * H)[v], xh[i]);
        unpack8(reinterpret_cast<const bf16x8*>(dy + (size_t)row * H)[v], g[i]);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          xh[i][k] *= rstd;
          // forward multiplied the bf16-rounded xhat by w, so dw uses the rounded value
          dwacc[i][k] += g[i][k] * bf16_round(xh[i][k]);
          g[i][k] *= wf[i][k];
          dot += g[i][k] * xh[i][k];
        }
      }
    }
    dot = block_sum<kNormThreads>(dot, red) / (float)H;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rstd * (g[i][k] - xh[i][k] * dot);
        if (kHasDres) {
          float r[8];
          unpack8(reinterpret_cast<const bf16x8*>(dres + (size_t)row * H)[v], r);
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] += r[k];
        }
        reinterpret_cast<bf16x8*>(dx + (size_t)row * H)[v] = pack8(o);
      }
    }
  }
  float* prow = partial + (size_t)blockIdx.x * H;
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    int v = tid + i * kNormThreads;
    if (v < nvec) {
      reinterpret_cast<float4*>(prow)[2 * v] =
          make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      reinterpret_cast<float4*>(prow)[2 * v + 1] =
          make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
    }
  }
}

// out[c] (+)= sum_b partial[b][c]
__global__ void colsum_kernel(const float* __restrict__ partial, float* __restrict__ out, int nblk,
                              int H, int accumulate) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= H) return;
  float s = 0.f;
  for (int b = 0; b < nblk; ++b) s += partial[(size_t)b * H + c];
  out[c] = accumulate ? out[c] + s : s;
}

// ------------------------------- LayerNorm ----------------------------------
__global__ void __launch_bounds__(kNormThreads)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                     const __nv_bfloat16* __restrict__ b, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int H,
                     float eps) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  const int tid = threadIdx.x;
  for (int row = blockIdx.x; row < M; row += gridDim.x) {
    float f[kNormMaxVec][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        unpack8(reinterpret_cast<const bf16x8*>(x + (size_t)row * H)[v], f[i]);
#pragma unroll
        for (int k = 0; k < 8; ++k) s += f[i][k];
      }
    }
    const float mean = block_sum<kNormThreads>(s, red) / (float)H;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          f[i][k] -= mean;
          ss += f[i][k] * f[i][k];
        }
      }
    }
    const float rstd = rsqrtf(block_sum<kNormThreads>(ss, red) / (float)H + eps);
    if (tid == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        float wf[8], bf[8], o[8];
        unpack8(reinterpret_cast<const bf16x8*>(w)[v], wf);
        if (b) unpack8(reinterpret_cast<const bf16x8*>(b)[v], bf);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = f[i][k] * rstd * wf[k] + (b ? bf[k] : 0.f);
        reinterpret_cast<bf16x8*>(y + (size_t)row * H)[v] = pack8(o);
      }
    }
  }
}

// partial layout: [grid][2][H]  (dgamma then dbeta)
__global__ void __launch_bounds__(kNormThreads)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ w, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dx,
                     float* __restrict__ partial, int M, int H) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  const int tid = threadIdx.x;
  float dg[kNormMaxVec][8], db[kNormMaxVec][8];
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k) dg[i][k] = db[i][k] = 0.f;
  for (int row = blockIdx.x; row < M; row += gridDim.x) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[kNormMaxVec][8], g[kNormMaxVec][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        float wf[8];
        unpack8(reinterpret_cast<const bf16x8*>(w)[v], wf);
        unpack8(reinterpret_cast<const bf16x8*>(x + (size_t)row * H)[v], xh[i]);
        unpack8(reinterpret_cast<const bf16x8*>(dy + (size_t)row * H)[v], g[i]);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          xh[i][k] = (xh[i][k] - mean) * rstd;
          dg[i][k] += g[i][k] * xh[i][k];
          db[i][k] += g[i][k];
          g[i][k] *= wf[k];
          s1 += g[i][k];
          s2 += g[i][k] * xh[i][k];
        }
      }
    }
    s1 = block_sum<kNormThreads>(s1, red) / (float)H;
    s2 = block_sum<kNormThreads>(s2, red) / (float)H;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      int v = tid + i * kNormThreads;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rstd * (g[i][k] - s1 - xh[i][k] * s2);
        reinterpret_cast<bf16x8*>(dx + (size_t)row * H)[v] = pack8(o);
      }
    }
  }
  float* pg = partial + (size_t)blockIdx.x * 2 * H;
  float* pb = pg + H;
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    int v = tid + i * kNormThreads;
    if (v < nvec) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        pg[v * 8 + k] = dg[i][k];
        pb[v * 8 + k] = db[i][k];
      }
    }
  }
}

__global__ void colsum2_kernel(const float* __restrict__ partial, float* __restrict__ dgamma,
                               float* __restrict__ dbeta, int nblk, int H, int accumulate) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= H) return;
  float s = 0.f, t = 0.f;
  for (int b = 0; b < nblk; ++b) {
    s += partial[(size_t)b * 2 * H + c];
    t += partial[(size_t)b * 2 * H + H + c];
  }
  dgamma[c] = accumulate ? dgamma[c] + s : s;
  if (dbeta) dbeta[c] = accumulate ? dbeta[c] + t : t;
}

bool rmsnorm_warp_ok(int H);
void rmsnorm_fwd_warp(const void* x, const void* residual, const void* w, void* y, void* h_out,
                      float* rstd, int M, int H, float eps, cudaStream_t stream);
void rmsnorm_bwd_warp(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                      void* dx, float* dw, float* scratch, int max_strips, int M, int H, int accumulate,
                      cudaStream_t stream);

bool rmsnorm_tma_ok(int M, int H);
int rmsnorm_fwd_tma(const void* x, const void* residual, const void* w, void* y, void* h_out,
                    float* rstd, int M, int H, float eps, cudaStream_t stream);
int rmsnorm_bwd_tma(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                    void* dx, float* dw, float* scratch, int M, int H, int accumulate,
                    cudaStream_t stream);

int layernorm_fwd_tma(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int M,
                      int H, float eps, cudaStream_t stream);
int layernorm_bwd_tma(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                      void* dx, float* dgamma, float* dbeta, float* scratch, int M, int H, int accumulate,
                      cudaStream_t stream);

// B200_NORM_IMPL = tma (default: bulk-copy ring, csrc/norm_tma.cu) | warp | block: A/B switch for
// tools/kernel_bench.py, read once.
static int norm_impl() {
  static int impl = -1;
  if (impl < 0) {
    const char* e = getenv("B200_NORM_IMPL");
    impl = (e && e[0] == 'w') ? 1 : (e && e[0] == 'b') ? 2 : 0;
  }
  return impl;
}

static inline bool norm_shape_ok(int M, int H) {
  return M > 0 && H > 0 && (H % 8) == 0 && H <= kNormThreads * kNormMaxVec * 8;
}

}  // namespace b200

using namespace b200;

// Number of CTAs the backward kernels use == rows of the [grid, H] scratch they need.
B200_API int b200_norm_bwd_grid(int M) { return M < 2 * kNumSMs ? M : 2 * kNumSMs; }

B200_API int b200_rmsnorm_fwd(const void* x, const void* residual, const void* w, void* y,
                              void* h_out, float* rstd, int M, int H, float eps,
                              cudaStream_t stream) {
  if (!norm_shape_ok(M, H)) return B200_ERR_UNSUPPORTED;
  if (norm_impl() == 0 && rmsnorm_tma_ok(M, H))
    return rmsnorm_fwd_tma(x, residual, w, y, h_out, rstd, M, H, eps, stream);
  if (norm_impl() != 2 && rmsnorm_warp_ok(H)) {   // one warp per row (csrc/norm_warp.cu)
    rmsnorm_fwd_warp(x, residual, w, y, h_out, rstd, M, H, eps, stream);
    B200_CHECK_LAUNCH();
    return 0;
  }
  int grid = M < 8 * kNumSMs ? M : 8 * kNumSMs;
  if (residual)
    rmsnorm_fwd_kernel<true><<<grid, kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (const __nv_bfloat16*)w,
        (__nv_bfloat16*)y, (__nv_bfloat16*)h_out, rstd, M, H, eps);
  else
    rmsnorm_fwd_kernel<false><<<grid, kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)x, nullptr, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, nullptr,
        rstd, M, H, eps);
  B200_CHECK_LAUNCH();
  return 0;
}

// scratch: fp32 [b200_norm_bwd_grid(M), H]; dw: fp32 [H] (+= when accumulate != 0).
B200_API int b200_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                              const void* dres, void* dx, float* dw, float* scratch, int M, int H,
                              int accumulate, cudaStream_t stream) {
  if (!norm_shape_ok(M, H)) return B200_ERR_UNSUPPORTED;
  int grid = b200_norm_bwd_grid(M);
  if (norm_impl() == 0 && rmsnorm_tma_ok(M, H))
    return rmsnorm_bwd_tma(dy, x, w, rstd, dres, dx, dw, scratch, M, H, accumulate, stream);
  if (norm_impl() != 2 && rmsnorm_warp_ok(H)) {
    rmsnorm_bwd_warp(dy, x, w, rstd, dres, dx, dw, scratch, grid, M, H, accumulate, stream);
    B200_CHECK_LAUNCH();
    return 0;
  }
  if (dres)
    rmsnorm_bwd_kernel<true><<<grid, kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, rstd,
        (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, M, H);
  else
    rmsnorm_bwd_kernel<false><<<grid, kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, rstd, nullptr,
        (__nv_bfloat16*)dx, scratch, M, H);
  B200_CHECK_LAUNCH();
  colsum_kernel<<<(H + 255) / 256, 256, 0, stream>>>(scratch, dw, grid, H, accumulate);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean,
                                float* rstd, int M, int H, float eps, cudaStream_t stream) {
  if (!norm_shape_ok(M, H)) return B200_ERR_UNSUPPORTED;
  if (norm_impl() == 0 && rmsnorm_tma_ok(M, H)) return layernorm_fwd_tma(x, w, b, y, mean, rstd, M, H, eps, stream);
  int grid = M < 8 * kNumSMs ? M : 8 * kNumSMs;
  layernorm_fwd_kernel<<<grid, kNormThreads, 0, stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (const __nv_bfloat16*)b,
      (__nv_bfloat16*)y, mean, rstd, M, H, eps);
  B200_CHECK_LAUNCH();
  return 0;
}

// scratch: fp32 [b200_norm_bwd_grid(M), 2, H]
B200_API int b200_layernorm_bwd(const void* dy, const void* x, const void* w, const float* mean,
                                const float* rstd, void* dx, float* dgamma, float* dbeta,
                                float* scratch, int M, int H, int accumulate,
                                cudaStream_t stream) {
  if (!norm_shape_ok(M, H)) return B200_ERR_UNSUPPORTED;
  if (norm_impl() == 0 && rmsnorm_tma_ok(M, H))
    return layernorm_bwd_tma(dy, x, w, mean, rstd, dx, dgamma, dbeta, scratch, M, H, accumulate, stream);
  int grid = b200_norm_bwd_grid(M);
  layernorm_bwd_kernel<<<grid, kNormThreads, 0, stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, mean, rstd,
      (__nv_bfloat16*)dx, scratch, M, H);
  B200_CHECK_LAUNCH();
  colsum2_kernel<<<(H + 255) / 256, 256, 0, stream>>>(scratch, dgamma, dbeta, grid, H, accumulate);
  B200_CHECK_LAUNCH();
  return 0;
}

// =============================================================================
// RMSNorm backward for callers that cannot allocate device memory
// =============================================================================
// Who this is for: code that has to run the RMSNorm backward from a place where asking
// the CUDA allocator for memory is not an option -- a launch sequence that is being
// recorded or replayed and therefore must not contain cudaMalloc/cudaFree, a host program
// that was handed a fixed memory budget and has already spent it, or a plug-in that only
// receives the caller's input/output pointers and has no allocator of its own. The regular
// entry point (b200_rmsnorm_bwd) needs a [grid, H] fp32 scratch buffer from its caller for
// the per-CTA partial weight gradients; this variant brings that buffer with it as a
// fixed-size array that lives in the library image.
//
// Cost and limits: the array below occupies kNoAllocMaxRows * kNoAllocMaxH * 4 bytes
// (148 * 8192 * 4 = 4.6 MiB) of device memory on EVERY GPU that loads this library, from
// load until process exit, whether or not the entry point is ever called. H above
// kNoAllocMaxH is refused with B200_ERR_TOO_LARGE (-3). At most kNoAllocMaxRows CTAs are
// used (the regular entry point uses up to twice as many), so very tall inputs run
// somewhat slower than through b200_rmsnorm_bwd.
//
// Sharing: there is ONE array per GPU and every call uses it. The kernel overwrites every
// partial row it later reads, so nothing has to be cleared between calls and stale
// contents of an earlier call are harmless. Two calls that execute at the same time on
// one GPU (different streams, or different host threads) would write the same rows: they
// race. Keep calls to this entry point ordered on one stream per GPU.
//
// Set-up: the array's name is only a handle in host code; its device address is looked up
// with cudaGetSymbolAddress the first time the entry point runs on a given GPU (with that
// GPU current) and cached per device. Make that first call outside of any stream capture.
constexpr int kNoAllocMaxRows = b200::kNumSMs;  // CTAs == partial rows
constexpr int kNoAllocMaxH = 8192;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_rmsnorm_bwd_noalloc_scratch[kNoAllocMaxRows * kNoAllocMaxH];

#include <mutex>

namespace {
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kNoAllocMaxDevices] = {};

// Device address of the scratch array on the CURRENT device (resolved once per device).
int noalloc_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return B200_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_rmsnorm_bwd_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return B200_ERR_UNSUPPORTED;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

// Largest hidden size b200_rmsnorm_bwd_noalloc accepts.
B200_API int b200_rmsnorm_bwd_noalloc_max_h() { return kNoAllocMaxH; }

// Same result as b200_rmsnorm_bwd, without a scratch argument and without any allocation.
B200_API int b200_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w,
                                      const float* rstd, const void* dres, void* dx, float* dw,
                                      int M, int H, int accumulate, cudaStream_t stream) {
  if (H > kNoAllocMaxH) return B200_ERR_TOO_LARGE;
  if (!b200::norm_shape_ok(M, H)) return B200_ERR_UNSUPPORTED;
  float* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc != 0) return rc;
  const int grid = M < kNoAllocMaxRows ? M : kNoAllocMaxRows;
  if (dres)
    b200::rmsnorm_bwd_kernel<true><<<grid, b200::kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, rstd,
        (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, M, H);
  else
    b200::rmsnorm_bwd_kernel<false><<<grid, b200::kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, rstd, nullptr,
        (__nv_bfloat16*)dx, scratch, M, H);
  B200_CHECK_LAUNCH();
  b200::colsum_kernel<<<(H + 255) / 256, 256, 0, stream>>>(scratch, dw, grid, H, accumulate);
  B200_CHECK_LAUNCH();
  return 0;
}
