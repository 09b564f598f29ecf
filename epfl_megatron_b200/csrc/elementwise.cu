// Memory-bound elementwise kernels (K3, K4, K5, K18 of SURVEY.md section 8):
//   * RoPE applied in place on the fused QKV projection output (GQA-layout aware)
//   * GLU family (LiGLU / GEGLU / ReGLU / SwiGLU) forward + backward
//   * (bias +) GeLU, tanh or erf flavour, forward + backward
//   * (bias +) dropout + residual add with a counter-based Philox mask that the
//     backward pass regenerates instead of storing
//   * dtype casts and scaled accumulation used by the gradient buffers
//
// Behavioural references (semantics only): megatron/model/positional_embeddings.py:24-51
// (interleaved-pair rotation in fp32), megatron/model/glu_activations.py:11-54
// (x1 * act(x2), the SECOND half is the gate), megatron/model/fused_bias_gelu.py:14-43,
// megatron/model/transformer.py:532-578 (bias-dropout-add).
#include "common.cuh"

namespace b200 {

// ------------------------------- RoPE ---------------------------------------
// x: [T, G, NSLOT, HN] bf16 where T = S*B tokens in [s, b] order. The first NROT slots
// of every group are rotated (q heads + the k head for the fused QKV layout), the
// rest (v) are left untouched. table: float2 (cos, sin) [P, HN/2].
__global__ void __launch_bounds__(256)
rope_kernel(__nv_bfloat16* __restrict__ x, const float2* __restrict__ table,
            const long long* __restrict__ pos_ids, int S, int B, int G, int nslot, int nrot,
            int HN, int pos_offset, int max_pos, float sign, long long total_vec) {
  // One thread owns one 16-byte column (4 rotation pairs) of one token and walks a share of the rotated
  // heads of that token (gridDim.y shares): position lookup and the (cos, sin) values are loaded once
  // and reused for every head of the share; four heads are in flight per step. Splitting the heads
  // over gridDim.y quadruples the resident warps: with one share the kernel kept ~28 KB in flight
  // per SM and ran at 36 % of DRAM peak (profiles/ncu_all_kernels_r2_table.md).
  const int vec_per_head = HN >> 3;
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total_vec) return;
  const int v = (int)(i % vec_per_head);
  const long long t = i / vec_per_head;
  const int b = (int)(t % B);
  const int s = (int)(t / B);
  long long pos = pos_ids ? pos_ids[(long long)b * S + s] : (long long)(s + pos_offset);
  if (pos >= max_pos) pos = max_pos - 1;
  const float2* cs = table + pos * (HN >> 1) + v * 4;
  float cx[4], sn[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 c = cs[k];
    cx[k] = c.x;
    sn[k] = c.y * sign;
  }
  bf16x8* base = reinterpret_cast<bf16x8*>(x + t * (long long)G * nslot * HN) + v;
  const int heads_all = G * nrot;
  const int per_share = (heads_all + (int)gridDim.y - 1) / (int)gridDim.y;
  const int h_begin = (int)blockIdx.y * per_share;
  const int heads = min(heads_all, h_begin + per_share);
  auto head_ptr = [&](int h) {
    const int g = h / nrot, slot = h - g * nrot;
    return base + (long long)(g * nslot + slot) * vec_per_head;
  };
  auto rotate = [&](bf16x8 in) {
    float f[8], o[8];
    unpack8(in, f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[2 * k] = f[2 * k] * cx[k] - f[2 * k + 1] * sn[k];
      o[2 * k + 1] = f[2 * k] * sn[k] + f[2 * k + 1] * cx[k];
    }
    return pack8(o);
  };
  int h = h_begin;
  for (; h + 4 <= heads; h += 4) {
    bf16x8* p[4];
    bf16x8 in[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      p[u] = head_ptr(h + u);
      in[u] = *p[u];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) *p[u] = rotate(in[u]);
  }
  for (; h < heads; ++h) {
    bf16x8* p = head_ptr(h);
    *p = rotate(*p);
  }
}

// ------------------------------- activations --------------------------------
enum GluKind { kLiGLU = 0, kGEGLU = 1, kReGLU = 2, kSwiGLU = 3 };

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.f + erff(x * 0.70710678118654752f));
}
__device__ __forceinline__ float gelu_erf_grad(float x) {
  float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752f));
  float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}
__device__ __forceinline__ float gelu_tanh(float x) {
  return x * 0.5f * (1.f + tanhf(0.79788456f * x * (1.f + 0.044715f * x * x)));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  float t = tanhf(0.79788456f * x * (1.f + 0.044715f * x * x));
  return 0.5f * x * ((1.f - t * t) * (0.79788456f + 0.1070322243f * x * x)) + 0.5f * (1.f + t);
}

template <int kKind>
__device__ __forceinline__ float glu_act(float g) {
  if (kKind == kLiGLU) return g;
  if (kKind == kGEGLU) return gelu_erf(g);
  if (kKind == kReGLU) return fmaxf(g, 0.f);
  return __fdividef(g, 1.f + __expf(-g));
}
template <int kKind>
__device__ __forceinline__ float glu_act_grad(float g) {
  if (kKind == kLiGLU) return 1.f;
  if (kKind == kGEGLU) return gelu_erf_grad(g);
  if (kKind == kReGLU) return g > 0.f ? 1.f : 0.f;
  float sg = __fdividef(1.f, 1.f + __expf(-g));
  return sg * (1.f + g * (1.f - sg));
}

// in: [M, 2F] (x1 | x2), out: [M, F] = x1 * act(x2).
// Grid (ceil(F/8 / 128), ceil(M / kGluRows)): a thread owns one 16-byte column of kGluRows rows,
// issues all its loads first, and never divides.
constexpr int kGluRows = 4;
template <int kKind>
__global__ void __launch_bounds__(128)
glu_fwd_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int F,
               long long M) {
  const int vpr = F >> 3;
  const int v = blockIdx.x * 128 + threadIdx.x;
  if (v >= vpr) return;
  const long long row0 = (long long)blockIdx.y * kGluRows;
  bf16x8 A[kGluRows], Gt[kGluRows];
#pragma unroll
  for (int u = 0; u < kGluRows; ++u) {
    if (row0 + u < M) {
      const bf16x8* r = reinterpret_cast<const bf16x8*>(in + (row0 + u) * 2 * F);
      A[u] = r[v];
      Gt[u] = r[v + vpr];
    }
  }
#pragma unroll
  for (int u = 0; u < kGluRows; ++u) {
    if (row0 + u < M) {
      float a[8], g[8], o[8];
      unpack8(A[u], a);
      unpack8(Gt[u], g);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = a[k] * glu_act<kKind>(g[k]);
      reinterpret_cast<bf16x8*>(out + (row0 + u) * F)[v] = pack8(o);
    }
  }
}
template <int kKind>
__global__ void __launch_bounds__(128)
glu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ in,
               __nv_bfloat16* __restrict__ din, int F, long long M) {
  const int vpr = F >> 3;
  const int v = blockIdx.x * 128 + threadIdx.x;
  if (v >= vpr) return;
  const long long row0 = (long long)blockIdx.y * kGluRows;
  bf16x8 A[kGluRows], Gt[kGluRows], D[kGluRows];
#pragma unroll
  for (int u = 0; u < kGluRows; ++u) {
    if (row0 + u < M) {
      const bf16x8* r = reinterpret_cast<const bf16x8*>(in + (row0 + u) * 2 * F);
      A[u] = r[v];
      Gt[u] = r[v + vpr];
      D[u] = reinterpret_cast<const bf16x8*>(dy + (row0 + u) * F)[v];
    }
  }
#pragma unroll
  for (int u = 0; u < kGluRows; ++u) {
    if (row0 + u < M) {
      float a[8], g[8], d[8], da[8], dg[8];
      unpack8(A[u], a);
      unpack8(Gt[u], g);
      unpack8(D[u], d);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        da[k] = d[k] * glu_act<kKind>(g[k]);
        dg[k] = d[k] * a[k] * glu_act_grad<kKind>(g[k]);
      }
      bf16x8* w = reinterpret_cast<bf16x8*>(din + (row0 + u) * 2 * F);
      w[v] = pack8(da);
      w[v + vpr] = pack8(dg);
    }
  }
}

// y = gelu(x + bias); bias may be null. H = row length (for the bias broadcast).
template <bool kTanh>
__global__ void __launch_bounds__(256)
bias_gelu_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ bias,
                     __nv_bfloat16* __restrict__ y, int H, long long total_vec) {
  const int vpr = H >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8], o[8];
    unpack8(reinterpret_cast<const bf16x8*>(x)[i], a);
    if (bias) unpack8(reinterpret_cast<const bf16x8*>(bias)[i % vpr], b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = bias ? a[k] + b[k] : a[k];
      o[k] = kTanh ? gelu_tanh(t) : gelu_erf(t);
    }
    reinterpret_cast<bf16x8*>(y)[i] = pack8(o);
  }
}
template <bool kTanh>
__global__ void __launch_bounds__(256)
bias_gelu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ bias, __nv_bfloat16* __restrict__ dx, int H,
                     long long total_vec) {
  const int vpr = H >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8], d[8], o[8];
    unpack8(reinterpret_cast<const bf16x8*>(x)[i], a);
    unpack8(reinterpret_cast<const bf16x8*>(dy)[i], d);
    if (bias) unpack8(reinterpret_cast<const bf16x8*>(bias)[i % vpr], b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = bias ? a[k] + b[k] : a[k];
      o[k] = d[k] * (kTanh ? gelu_tanh_grad(t) : gelu_erf_grad(t));
    }
    reinterpret_cast<bf16x8*>(dx)[i] = pack8(o);
  }
}

// ------------------------------- Philox dropout-add --------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  constexpr uint32_t kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(kM0, ctr.x), lo0 = kM0 * ctr.x;
    uint32_t hi1 = __umulhi(kM1, ctr.z), lo1 = kM1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += kW0;
    key.y += kW1;
  }
  return ctr;
}

// out = residual + dropout(x + bias, p) * 1/(1-p). One Philox call yields 128 random
// bits = 8 x 16-bit lanes, i.e. one keep decision per element of a 16-byte vector.
// kBackward: out = dropout_mask(dy) * 1/(1-p) (x holds dy, residual/bias ignored).
template <bool kBackward>
__global__ void __launch_bounds__(256)
dropout_add_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ bias,
                   const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ out,
                   int H, long long total_vec, float p, unsigned long long seed,
                   unsigned long long offset) {
  const int vpr = H >> 3;
  const uint32_t thresh = (uint32_t)(p * 65536.f);
  const float scale = 1.f / (1.f - p);
  const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    float a[8], o[8];
    unpack8(reinterpret_cast<const bf16x8*>(x)[i], a);
    if (!kBackward && bias) {
      float b[8];
      unpack8(reinterpret_cast<const bf16x8*>(bias)[i % vpr], b);
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] += b[k];
    }
    uint4 rnd = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), (uint32_t)offset,
                                         (uint32_t)(offset >> 32)),
                              key);
    uint32_t w[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      uint32_t bits = (w[k >> 1] >> ((k & 1) * 16)) & 0xFFFFu;
      o[k] = bits >= thresh ? a[k] * scale : 0.f;
    }
    if (!kBackward && residual) {
      float r[8];
      unpack8(reinterpret_cast<const bf16x8*>(residual)[i], r);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] += r[k];
    }
    reinterpret_cast<bf16x8*>(out)[i] = pack8(o);
  }
}

// ------------------------------- casts / accumulate --------------------------
// dst_fp32[i] (+)= scale * src_bf16[i]
__global__ void __launch_bounds__(256)
accum_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst,
                         long long n_vec, float scale, int accumulate) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(reinterpret_cast<const bf16x8*>(src)[i], f);
    float4* d = reinterpret_cast<float4*>(dst) + 2 * i;
    float4 d0 = accumulate ? d[0] : make_float4(0, 0, 0, 0);
    float4 d1 = accumulate ? d[1] : make_float4(0, 0, 0, 0);
    d0.x += scale * f[0]; d0.y += scale * f[1]; d0.z += scale * f[2]; d0.w += scale * f[3];
    d1.x += scale * f[4]; d1.y += scale * f[5]; d1.z += scale * f[6]; d1.w += scale * f[7];
    d[0] = d0;
    d[1] = d1;
  }
}
__global__ void __launch_bounds__(256)
cast_f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                        long long n_vec) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec;
       i += (long long)gridDim.x * blockDim.x) {
    const float4* s = reinterpret_cast<const float4*>(src) + 2 * i;
    float4 a = s[0], b = s[1];
    float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    reinterpret_cast<bf16x8*>(dst)[i] = pack8(f);
  }
}

static inline int ew_grid(long long n_vec) {
  long long g = (n_vec + 255) / 256;
  long long cap = (long long)kNumSMs * 16;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

}  // namespace b200

using namespace b200;

B200_API int b200_rope(void* x, const float* cos_sin, const long long* pos_ids, int S, int B, int G,
                       int nslot, int nrot, int HN, int pos_offset, int max_pos, int inverse,
                       cudaStream_t stream) {
  if (HN % 8 || nrot > nslot || nrot <= 0) return B200_ERR_UNSUPPORTED;
  long long total = (long long)S * B * (HN >> 3);      // one thread per (token, 16-byte column)
  if (total == 0 || G == 0) return 0;
  if ((total + 127) / 128 > 0x7fffffffLL) return B200_ERR_TOO_LARGE;
  const int shares = G * nrot >= 16 ? 4 : 1;
  rope_kernel<<<dim3((unsigned)((total + 127) / 128), shares), 128, 0, stream>>>((__nv_bfloat16*)x, (const float2*)cos_sin,
                                                  pos_ids, S, B, G, nslot, nrot, HN, pos_offset,
                                                  max_pos, inverse ? -1.f : 1.f, total);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_glu_fwd(const void* in, void* out, long long M, int F, int kind,
                          cudaStream_t stream) {
  if (F % 8 || kind < 0 || kind > 3) return B200_ERR_UNSUPPORTED;
  long long total = M * (F >> 3);
  if (total == 0) return 0;
  const __nv_bfloat16* i = (const __nv_bfloat16*)in;
  __nv_bfloat16* o = (__nv_bfloat16*)out;
  // gridDim.y is limited to 65535: very tall inputs go in slabs
  const long long slab = 65535LL * kGluRows;
  for (long long m0 = 0; m0 < M; m0 += slab) {
    const long long m = M - m0 < slab ? M - m0 : slab;
    dim3 g((unsigned)(((F >> 3) + 127) / 128), (unsigned)((m + kGluRows - 1) / kGluRows));
    const __nv_bfloat16* ii = i + m0 * 2 * F;
    __nv_bfloat16* oo = o + m0 * F;
    switch (kind) {
      case kLiGLU: glu_fwd_kernel<kLiGLU><<<g, 128, 0, stream>>>(ii, oo, F, m); break;
      case kGEGLU: glu_fwd_kernel<kGEGLU><<<g, 128, 0, stream>>>(ii, oo, F, m); break;
      case kReGLU: glu_fwd_kernel<kReGLU><<<g, 128, 0, stream>>>(ii, oo, F, m); break;
      default: glu_fwd_kernel<kSwiGLU><<<g, 128, 0, stream>>>(ii, oo, F, m); break;
    }
  }
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_glu_bwd(const void* dy, const void* in, void* din, long long M, int F, int kind,
                          cudaStream_t stream) {
  if (F % 8 || kind < 0 || kind > 3) return B200_ERR_UNSUPPORTED;
  long long total = M * (F >> 3);
  if (total == 0) return 0;
  const __nv_bfloat16* d = (const __nv_bfloat16*)dy;
  const __nv_bfloat16* i = (const __nv_bfloat16*)in;
  __nv_bfloat16* o = (__nv_bfloat16*)din;
  const long long slab = 65535LL * kGluRows;
  for (long long m0 = 0; m0 < M; m0 += slab) {
    const long long m = M - m0 < slab ? M - m0 : slab;
    dim3 g((unsigned)(((F >> 3) + 127) / 128), (unsigned)((m + kGluRows - 1) / kGluRows));
    const __nv_bfloat16* dd = d + m0 * F;
    const __nv_bfloat16* ii = i + m0 * 2 * F;
    __nv_bfloat16* oo = o + m0 * 2 * F;
    switch (kind) {
      case kLiGLU: glu_bwd_kernel<kLiGLU><<<g, 128, 0, stream>>>(dd, ii, oo, F, m); break;
      case kGEGLU: glu_bwd_kernel<kGEGLU><<<g, 128, 0, stream>>>(dd, ii, oo, F, m); break;
      case kReGLU: glu_bwd_kernel<kReGLU><<<g, 128, 0, stream>>>(dd, ii, oo, F, m); break;
      default: glu_bwd_kernel<kSwiGLU><<<g, 128, 0, stream>>>(dd, ii, oo, F, m); break;
    }
  }
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_bias_gelu_fwd(const void* x, const void* bias, void* y, long long M, int H,
                                int tanh_approx, cudaStream_t stream) {
  if (H % 8) return B200_ERR_UNSUPPORTED;
  long long total = M * (H >> 3);
  if (total == 0) return 0;
  if (tanh_approx)
    bias_gelu_fwd_kernel<true><<<ew_grid(total), 256, 0, stream>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, H, total);
  else
    bias_gelu_fwd_kernel<false><<<ew_grid(total), 256, 0, stream>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, H, total);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_bias_gelu_bwd(const void* dy, const void* x, const void* bias, void* dx,
                                long long M, int H, int tanh_approx, cudaStream_t stream) {
  if (H % 8) return B200_ERR_UNSUPPORTED;
  long long total = M * (H >> 3);
  if (total == 0) return 0;
  if (tanh_approx)
    bias_gelu_bwd_kernel<true><<<ew_grid(total), 256, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias,
        (__nv_bfloat16*)dx, H, total);
  else
    bias_gelu_bwd_kernel<false><<<ew_grid(total), 256, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias,
        (__nv_bfloat16*)dx, H, total);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_dropout_add_fwd(const void* x, const void* bias, const void* residual, void* out,
                                  long long M, int H, float p, unsigned long long seed,
                                  unsigned long long offset, cudaStream_t stream) {
  if (H % 8 || p < 0.f || p >= 1.f) return B200_ERR_UNSUPPORTED;
  long long total = M * (H >> 3);
  if (total == 0) return 0;
  dropout_add_kernel<false><<<ew_grid(total), 256, 0, stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)residual,
      (__nv_bfloat16*)out, H, total, p, seed, offset);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_dropout_bwd(const void* dy, void* dx, long long M, int H, float p,
                              unsigned long long seed, unsigned long long offset,
                              cudaStream_t stream) {
  if (H % 8 || p < 0.f || p >= 1.f) return B200_ERR_UNSUPPORTED;
  long long total = M * (H >> 3);
  if (total == 0) return 0;
  dropout_add_kernel<true><<<ew_grid(total), 256, 0, stream>>>(
      (const __nv_bfloat16*)dy, nullptr, nullptr, (__nv_bfloat16*)dx, H, total, p, seed, offset);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_accum_bf16_to_f32(const void* src, float* dst, long long n, float scale,
                                    int accumulate, cudaStream_t stream) {
  if (n % 8) return B200_ERR_UNSUPPORTED;
  if (n == 0) return 0;
  accum_bf16_to_f32_kernel<<<ew_grid(n >> 3), 256, 0, stream>>>((const __nv_bfloat16*)src, dst,
                                                                n >> 3, scale, accumulate);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_cast_f32_to_bf16(const float* src, void* dst, long long n, cudaStream_t stream) {
  if (n % 8) return B200_ERR_UNSUPPORTED;
  if (n == 0) return 0;
  cast_f32_to_bf16_kernel<<<ew_grid(n >> 3), 256, 0, stream>>>(src, (__nv_bfloat16*)dst, n >> 3);
  B200_CHECK_LAUNCH();
  return 0;
}
