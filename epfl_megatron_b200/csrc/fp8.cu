// FP8 casts for the 8-bit GEMM path (csrc/gemm_tcgen05.cu, b200_gemm_fp8).
//
// The reference reaches FP8 only through Transformer Engine's layers
// (megatron/model/transformer.py:931-950, arguments.py --fp8_e4m3 / --fp8_hybrid / --fp8_margin /
// --fp8_amax_history_len / --fp8_amax_compute_algo); its cast_transpose + amax kernels are what this
// file restates for the delayed-scaling recipe: ONE pass over a bf16 matrix produces
//   y  [rows, cols] = fp8(x * scale)            (the K-major operand of the forward / dgrad GEMM),
//   yT [cols, rows] = the same values transposed (the K-major operand of the other GEMM), optional,
//   amax = max(amax, max |x|)                   (feeds the scale of a LATER step),
// with `scale` read from device memory (no host round trip) and saturation to the finite range.
#include <cuda_fp8.h>

#include "common.cuh"

namespace b200 {

constexpr int kQT = 64;     // tile edge

template <bool kE5M2>
__device__ __forceinline__ uint32_t cvt4(float a, float b, float c, float d) {
  constexpr __nv_fp8_interpretation_t kind = kE5M2 ? __NV_E5M2 : __NV_E4M3;
  const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, kind);
  const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(c, d), __NV_SATFINITE, kind);
  return lo | (hi << 16);
}

template <bool kE5M2, bool kTranspose>
__global__ void __launch_bounds__(256)
fp8_quantize_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ y, uint8_t* __restrict__ yT,
                    int rows, int cols, long long ldx, const float* __restrict__ scale_ptr,
                    float* __restrict__ amax_out) {
  __shared__ uint8_t tile[kQT][kQT + 16];
  __shared__ float red[33];
  const int tid = threadIdx.x;
  const int r0 = blockIdx.y * kQT, c0 = blockIdx.x * kQT;
  const float scale = scale_ptr ? *scale_ptr : 1.f;
  float amax = 0.f;
  // 64 x 64 bf16 = 512 vectors of 8: two per thread
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int v = tid + i * 256;
    const int r = v >> 3, cv = (v & 7) * 8;
    const int gr = r0 + r, gc = c0 + cv;
    uint2 out = make_uint2(0u, 0u);
    if (gr < rows && gc < cols) {
      float f[8];
      unpack8(*reinterpret_cast<const bf16x8*>(x + (size_t)gr * ldx + gc), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(f[k]));
      out.x = cvt4<kE5M2>(f[0] * scale, f[1] * scale, f[2] * scale, f[3] * scale);
      out.y = cvt4<kE5M2>(f[4] * scale, f[5] * scale, f[6] * scale, f[7] * scale);
      *reinterpret_cast<uint2*>(y + (size_t)gr * cols + gc) = out;
    }
    if (kTranspose) *reinterpret_cast<uint2*>(&tile[r][cv]) = out;
  }
  if (amax_out) {
    const float m = block_max<256>(amax, red);
    // non-negative floats order like their bit patterns
    if (tid == 0 && m > 0.f) atomicMax(reinterpret_cast<int*>(amax_out), __float_as_int(m));
  } else if (kTranspose) {
    __syncthreads();
  }
  if (kTranspose) {
    // thread -> 16 consecutive rows of one column: a 16-byte piece of row (c0 + c) of yT
    const int c = tid >> 2, rr = (tid & 3) * 16;
    if (c0 + c < cols && r0 + rr < rows) {
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        w[q] = (uint32_t)tile[rr + 4 * q][c] | ((uint32_t)tile[rr + 4 * q + 1][c] << 8) |
               ((uint32_t)tile[rr + 4 * q + 2][c] << 16) | ((uint32_t)tile[rr + 4 * q + 3][c] << 24);
      *reinterpret_cast<uint4*>(yT + (size_t)(c0 + c) * rows + r0 + rr) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

}  // namespace b200

using namespace b200;

// x: bf16 [rows, cols] with row stride ldx elements; y: [rows, cols] bytes; yT: [cols, rows] bytes or
// NULL. rows % 16 == 0 and cols % 16 == 0 (the GEMM's 16-byte row strides). scale: device pointer (NULL
// = 1); amax: device pointer to a float that receives max(*amax, max |x|) (NULL = not tracked);
// e5m2: 0 = E4M3 (activations, weights), 1 = E5M2 (gradients under --fp8_hybrid).
B200_API int b200_fp8_quantize(const void* x, void* y, void* yT, int rows, int cols, long long ldx,
                               const float* scale, float* amax, int e5m2, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return B200_ERR_BAD_ARG;
  if ((rows % 16) || (cols % 16) || (ldx % 8) || (reinterpret_cast<uintptr_t>(x) & 15)) return B200_ERR_UNSUPPORTED;
  dim3 grid((cols + kQT - 1) / kQT, (rows + kQT - 1) / kQT);
  auto xb = (const __nv_bfloat16*)x;
  auto yb = (uint8_t*)y, tb = (uint8_t*)yT;
  if (e5m2) {
    if (yT) fp8_quantize_kernel<true, true><<<grid, 256, 0, stream>>>(xb, yb, tb, rows, cols, ldx, scale, amax);
    else fp8_quantize_kernel<true, false><<<grid, 256, 0, stream>>>(xb, yb, tb, rows, cols, ldx, scale, amax);
  } else {
    if (yT) fp8_quantize_kernel<false, true><<<grid, 256, 0, stream>>>(xb, yb, tb, rows, cols, ldx, scale, amax);
    else fp8_quantize_kernel<false, false><<<grid, 256, 0, stream>>>(xb, yb, tb, rows, cols, ldx, scale, amax);
  }
  B200_CHECK_LAUNCH();
  return 0;
}
