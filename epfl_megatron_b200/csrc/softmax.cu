// Scaled softmax family for the non-flash attention path (K6 of SURVEY.md section 8):
//   mode 0: softmax(scale * x)                              x [rows, sk]
//   mode 1: causal  -- row i of each [sq, sk] matrix only sees keys 0..i (+ sk - sq),
//                      everything above the diagonal is written as exact zero
//   mode 2: explicit uint8 mask [mb, 1, sq, sk] (mb = b or 1); masked scores become
//                      -10000 after scaling and a fully masked row is written as zeros
// plus the shared backward dx = scale * y * (dy - sum(dy * y)).
//
// Semantics follow the reference's three extensions
// (megatron/fused_kernels/scaled_upper_triang_masked_softmax.h,
//  scaled_masked_softmax.h, scaled_softmax_cuda.cu) and their test gate
// (fused_kernels/tests/test_fused_kernels.py:295-364) but lift the sk <= 2048/4096
// limits: one 128-thread CTA owns a row and streams it with 16-byte vectors, keeping up
// to 4096 keys in registers and re-reading (L2-resident) beyond that.
#include "common.cuh"

namespace b200 {

constexpr int kSmThreads = 128;
constexpr int kSmMaxVec = 4;  // register-cached vectors per thread: 128*4*8 = 4096 keys

template <int kMode>
__global__ void __launch_bounds__(kSmThreads)
softmax_fwd_kernel(const __nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ mask,
                   __nv_bfloat16* __restrict__ y, long long rows, int sq, int sk, int np,
                   int mask_batch, float scale) {
  __shared__ float red[33];
  const int tid = threadIdx.x;
  const int nvec = sk >> 3;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const __nv_bfloat16* xr = x + row * sk;
    __nv_bfloat16* yr = y + row * sk;
    const int qi = (int)(row % sq);
    const int limit = kMode == 1 ? qi + 1 + (sk - sq) : sk;  // number of visible keys
    const uint8_t* mr = nullptr;
    if (kMode == 2) {
      long long bidx = row / ((long long)sq * np);
      mr = mask + ((mask_batch == 1 ? 0 : bidx) * sq + qi) * (long long)sk;
    }
    float f[kSmMaxVec][8];
    float mx = -INFINITY;
    for (int it = 0, v = tid; v < nvec; ++it, v += kSmThreads) {
      float t[8];
      unpack8(reinterpret_cast<const bf16x8*>(xr)[v], t);
      uint2 mb = make_uint2(0, 0);
      if (kMode == 2) mb = reinterpret_cast<const uint2*>(mr)[v];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float s = t[k] * scale;
        if (kMode == 1 && v * 8 + k >= limit) s = -INFINITY;
        if (kMode == 2) {
          uint32_t w = k < 4 ? mb.x : mb.y;
          if ((w >> ((k & 3) * 8)) & 0xFFu) s = -10000.f;
        }
        t[k] = s;
        mx = fmaxf(mx, s);
      }
      if (it < kSmMaxVec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) f[it][k] = t[k];
      }
    }
    mx = block_max<kSmThreads>(mx, red);
    const bool dead = (kMode == 2) && (mx <= -10000.f);
    float sum = 0.f;
    for (int it = 0, v = tid; v < nvec; ++it, v += kSmThreads) {
      float t[8];
      if (it < kSmMaxVec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = f[it][k];
      } else {
        unpack8(reinterpret_cast<const bf16x8*>(xr)[v], t);
        uint2 mb = make_uint2(0, 0);
        if (kMode == 2) mb = reinterpret_cast<const uint2*>(mr)[v];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float s = t[k] * scale;
          if (kMode == 1 && v * 8 + k >= limit) s = -INFINITY;
          if (kMode == 2) {
            uint32_t w = k < 4 ? mb.x : mb.y;
            if ((w >> ((k & 3) * 8)) & 0xFFu) s = -10000.f;
          }
          t[k] = s;
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        t[k] = __expf(t[k] - mx);
        sum += t[k];
      }
      if (it < kSmMaxVec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) f[it][k] = t[k];
      } else {
        reinterpret_cast<bf16x8*>(yr)[v] = pack8(t);  // unnormalised, fixed below
      }
    }
    sum = block_sum<kSmThreads>(sum, red);
    const float inv = dead ? 0.f : 1.f / sum;
    for (int it = 0, v = tid; v < nvec; ++it, v += kSmThreads) {
      float t[8];
      if (it < kSmMaxVec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = f[it][k] * inv;
      } else {
        unpack8(reinterpret_cast<const bf16x8*>(yr)[v], t);
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] *= inv;
      }
      reinterpret_cast<bf16x8*>(yr)[v] = pack8(t);
    }
  }
}

// dx = scale * y * (dy - sum_k dy_k * y_k). Zeros in y (masked / above the diagonal)
// give zero gradients, so one kernel serves all three forward modes.
__global__ void __launch_bounds__(kSmThreads)
softmax_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                   __nv_bfloat16* __restrict__ dx, long long rows, int sk, float scale) {
  __shared__ float red[33];
  const int tid = threadIdx.x;
  const int nvec = sk >> 3;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const bf16x8* gr = reinterpret_cast<const bf16x8*>(dy + row * sk);
    const bf16x8* yr = reinterpret_cast<const bf16x8*>(y + row * sk);
    float dot = 0.f;
    for (int v = tid; v < nvec; v += kSmThreads) {
      float a[8], b[8];
      unpack8(gr[v], a);
      unpack8(yr[v], b);
#pragma unroll
      for (int k = 0; k < 8; ++k) dot += a[k] * b[k];
    }
    dot = block_sum<kSmThreads>(dot, red);
    for (int v = tid; v < nvec; v += kSmThreads) {
      float a[8], b[8], o[8];
      unpack8(gr[v], a);
      unpack8(yr[v], b);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = scale * b[k] * (a[k] - dot);
      reinterpret_cast<bf16x8*>(dx + row * sk)[v] = pack8(o);
    }
  }
}

}  // namespace b200

using namespace b200;

// x, y: [b, np, sq, sk] bf16 (y may alias x). mode: 0 plain, 1 causal, 2 masked.
B200_API int b200_softmax_fwd(const void* x, const uint8_t* mask, void* y, int b, int np, int sq,
                              int sk, int mask_batch, float scale, int mode, cudaStream_t stream) {
  if (sk % 8 || mode < 0 || mode > 2) return B200_ERR_UNSUPPORTED;
  if (mode == 2 && (!mask || (mask_batch != 1 && mask_batch != b))) return B200_ERR_BAD_ARG;
  long long rows = (long long)b * np * sq;
  if (rows == 0) return 0;
  int grid = (int)(rows < (long long)kNumSMs * 16 ? rows : (long long)kNumSMs * 16);
  const __nv_bfloat16* xi = (const __nv_bfloat16*)x;
  __nv_bfloat16* yo = (__nv_bfloat16*)y;
  if (mode == 0)
    softmax_fwd_kernel<0><<<grid, kSmThreads, 0, stream>>>(xi, mask, yo, rows, sq, sk, np,
                                                           mask_batch, scale);
  else if (mode == 1)
    softmax_fwd_kernel<1><<<grid, kSmThreads, 0, stream>>>(xi, mask, yo, rows, sq, sk, np,
                                                           mask_batch, scale);
  else
    softmax_fwd_kernel<2><<<grid, kSmThreads, 0, stream>>>(xi, mask, yo, rows, sq, sk, np,
                                                           mask_batch, scale);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int sk,
                              float scale, cudaStream_t stream) {
  if (sk % 8) return B200_ERR_UNSUPPORTED;
  if (rows == 0) return 0;
  int grid = (int)(rows < (long long)kNumSMs * 16 ? rows : (long long)kNumSMs * 16);
  softmax_bwd_kernel<<<grid, kSmThreads, 0, stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (__nv_bfloat16*)dx, rows, sk, scale);
  B200_CHECK_LAUNCH();
  return 0;
}
