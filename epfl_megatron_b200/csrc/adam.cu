// Fused optimizer kernels on flat buffers (K15 phase 2, K17 of SURVEY.md section 8).
//
// The framework keeps every model chunk's parameters in ONE contiguous bf16 buffer, the
// gradients in ONE contiguous fp32 buffer and the optimizer state (fp32 master copy,
// exp_avg, exp_avg_sq) in flat buffers of the same indexing, so Adam is a single
// streaming kernel per (lr, weight-decay) range instead of the chunked multi-tensor
// launches of apex FusedAdam / amp_C that the reference relies on
// (megatron/optimizer/__init__.py:3,75-80, megatron/optimizer/clip_grads.py:75-105).
// Gradient unscale, global-norm clip and the "skip step on inf/nan" decision are read
// from DEVICE scalars, so the step needs no host synchronisation (the reference syncs
// on .item(), clip_grads.py:96).
#include "common.cuh"

namespace b200 {

// p, m, v: fp32 state; g: fp32 grads; p16: optional bf16 model copy written back.
// coef[0] multiplies the gradient (1/loss_scale * clip coefficient); skip[0] != 0 means
// "found inf": the kernel leaves everything untouched.
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
            float* __restrict__ v, __nv_bfloat16* __restrict__ p16, long long n_vec, float lr,
            float beta1, float beta2, float eps, float wd, float inv_bc1, float inv_bc2,
            int adamw, const float* __restrict__ coef, const float* __restrict__ skip) {
  if (skip && skip[0] != 0.f) return;
  const float gs = coef ? coef[0] : 1.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec;
       i += (long long)gridDim.x * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float pa[4] = {pv.x, pv.y, pv.z, pv.w}, ga[4] = {gv.x, gv.y, gv.z, gv.w};
    float ma[4] = {mv.x, mv.y, mv.z, mv.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gr = ga[k] * gs;
      if (!adamw) gr += wd * pa[k];
      ma[k] = beta1 * ma[k] + (1.f - beta1) * gr;
      va[k] = beta2 * va[k] + (1.f - beta2) * gr * gr;
      float upd = (ma[k] * inv_bc1) / (sqrtf(va[k] * inv_bc2) + eps);
      if (adamw) upd += wd * pa[k];
      pa[k] -= lr * upd;
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(va[0], va[1], va[2], va[3]);
    if (p16) {
      uint2 o;
      o.x = pack_bf16x2(pa[0], pa[1]);
      o.y = pack_bf16x2(pa[2], pa[3]);
      reinterpret_cast<uint2*>(p16)[i] = o;
    }
  }
}

// SGD with momentum on the same flat layout (reference: apex FusedSGD option,
// megatron/optimizer/__init__.py:81-86).
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ mom,
           __nv_bfloat16* __restrict__ p16, long long n, float lr, float momentum, float wd,
           int first_step, const float* __restrict__ coef, const float* __restrict__ skip) {
  if (skip && skip[0] != 0.f) return;
  const float gs = coef ? coef[0] : 1.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float gr = g[i] * gs + wd * p[i];
    float b = first_step ? gr : momentum * mom[i] + gr;
    mom[i] = b;
    float np_ = p[i] - lr * (momentum != 0.f ? b : gr);
    p[i] = np_;
    if (p16) p16[i] = __float2bfloat16_rn(np_);
  }
}

constexpr int kNormBlocks = 1024;

// partial[b] = sum over this block's grid-stride slice of x^2 ; also flags non-finite.
__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const float* __restrict__ x, long long n, float* __restrict__ partial) {
  __shared__ float red[33];
  float s = 0.f;
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 t = reinterpret_cast<const float4*>(x)[i];
    s += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
  }
  for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    s += x[i] * x[i];
  s = block_sum<256>(s, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}
// out[0] (+)= sum(partial[0..nb)) ; out[1] = 1 if the sum is not finite
__global__ void __launch_bounds__(256)
sumsq_final_kernel(const float* __restrict__ partial, int nb, float* __restrict__ out,
                   int accumulate) {
  __shared__ float red[33];
  float s = 0.f;
  for (int i = threadIdx.x; i < nb; i += 256) s += partial[i];
  s = block_sum<256>(s, red);
  if (threadIdx.x == 0) {
    float t = accumulate ? out[0] + s : s;
    out[0] = t;
    out[1] = isfinite(t) ? (accumulate ? out[1] : 0.f) : 1.f;
  }
}

__global__ void __launch_bounds__(256)
scale_kernel(float* __restrict__ x, long long n, const float* __restrict__ coef, float host_coef) {
  const float c = (coef ? coef[0] : 1.f) * host_coef;
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 t = reinterpret_cast<float4*>(x)[i];
    t.x *= c; t.y *= c; t.z *= c; t.w *= c;
    reinterpret_cast<float4*>(x)[i] = t;
  }
  for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    x[i] *= c;
}

static inline int flat_grid(long long n_items) {
  long long g = (n_items + 255) / 256;
  long long cap = (long long)kNumSMs * 8;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

}  // namespace b200

using namespace b200;

B200_API int b200_adam_step(float* p, const float* g, float* m, float* v, void* p16, long long n,
                            float lr, float beta1, float beta2, float eps, float wd, int step,
                            int adamw, int bias_correction, const float* coef, const float* skip,
                            cudaStream_t stream) {
  if (n % 4) return B200_ERR_UNSUPPORTED;
  if (n == 0) return 0;
  float inv_bc1 = 1.f, inv_bc2 = 1.f;
  if (bias_correction) {
    inv_bc1 = 1.f / (1.f - powf(beta1, (float)step));
    inv_bc2 = 1.f / (1.f - powf(beta2, (float)step));
  }
  adam_kernel<<<flat_grid(n >> 2), 256, 0, stream>>>(p, g, m, v, (__nv_bfloat16*)p16, n >> 2, lr,
                                                     beta1, beta2, eps, wd, inv_bc1, inv_bc2,
                                                     adamw, coef, skip);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_sgd_step(float* p, const float* g, float* mom, void* p16, long long n, float lr,
                           float momentum, float wd, int first_step, const float* coef,
                           const float* skip, cudaStream_t stream) {
  if (n == 0) return 0;
  sgd_kernel<<<flat_grid(n), 256, 0, stream>>>(p, g, mom, (__nv_bfloat16*)p16, n, lr, momentum, wd,
                                               first_step, coef, skip);
  B200_CHECK_LAUNCH();
  return 0;
}

// Number of fp32 scratch slots b200_sumsq needs.
B200_API int b200_sumsq_scratch_elems() { return kNormBlocks; }

// out[0] (+)= sum(x^2), out[1] = non-finite flag. scratch: fp32 [b200_sumsq_scratch_elems()].
B200_API int b200_sumsq(const float* x, long long n, float* out, float* scratch, int accumulate,
                        cudaStream_t stream) {
  long long items = (n + 3) / 4;
  int nb = flat_grid(items);
  if (nb > kNormBlocks) nb = kNormBlocks;
  sumsq_partial_kernel<<<nb, 256, 0, stream>>>(x, n, scratch);
  B200_CHECK_LAUNCH();
  sumsq_final_kernel<<<1, 256, 0, stream>>>(scratch, nb, out, accumulate);
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_scale(float* x, long long n, const float* coef, float host_coef,
                        cudaStream_t stream) {
  if (n == 0) return 0;
  scale_kernel<<<flat_grid((n + 3) / 4), 256, 0, stream>>>(x, n, coef, host_coef);
  B200_CHECK_LAUNCH();
  return 0;
}
