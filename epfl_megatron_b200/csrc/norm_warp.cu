// Warp-per-row RMSNorm forward / backward for rows that fit one warp's registers
// (H % 256 == 0, H <= 5120). One warp owns a row: 16-byte vectors strided by lane, so every
// access is a full 512-byte line burst, the reduction is five shuffles, and no block barrier
// sits between a row's load and its store: with 8-12 independent warps per SM the loads of many
// rows are in flight at once, which is what an HBM-bound kernel needs. The weight-gradient is a
// separate column reduction (each thread owns 8 columns of a strip of rows), so the row kernels
// carry no per-column accumulators. Same math and rounding order as csrc/norm.cu (the reference's
// eager RMSNorm, megatron/model/fused_layer_norm.py:125-139: normalise in fp32, round to bf16,
// then scale by the weight).
#include "common.cuh"

namespace b200 {

template <int VPL, bool kHasResidual>
__global__ void __launch_bounds__(256)
rmsnorm_fwd_warp_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                        const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y,
                        __nv_bfloat16* __restrict__ h_out, float* __restrict__ rstd_out, int M, int H,
                        float eps) {
  const int lane = threadIdx.x & 31;
  const int nvec = H >> 3;
  const int wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  const bf16x8* wv = reinterpret_cast<const bf16x8*>(w);
  for (int row = wid; row < M; row += nwarps) {
    const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * H);
    bf16x8 xv[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < nvec) xv[i] = xr[v];
    }
    if (kHasResidual) {
      const bf16x8* rr = reinterpret_cast<const bf16x8*>(res + (size_t)row * H);
      bf16x8 rv[VPL];
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        const int v = lane + 32 * i;
        if (v < nvec) rv[i] = rr[v];
      }
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        const int v = lane + 32 * i;
        if (v < nvec) {
          float a[8], b[8];
          unpack8(xv[i], a);
          unpack8(rv[i], b);
#pragma unroll
          for (int k = 0; k < 8; ++k) a[k] += b[k];
          xv[i] = pack8(a);     // h = bf16(x + residual) is what the norm sees
          reinterpret_cast<bf16x8*>(h_out + (size_t)row * H)[v] = xv[i];
        }
      }
    }
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      if (lane + 32 * i < nvec) {
        float a[8];
        unpack8(xv[i], a);
#pragma unroll
        for (int k = 0; k < 8; ++k) ss += a[k] * a[k];
      }
    }
    ss = warp_sum(ss);
    const float rstd = rsqrtf(ss / (float)H + eps);
    if (lane == 0 && rstd_out) rstd_out[row] = rstd;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < nvec) {
        float a[8], wf[8], o[8];
        unpack8(xv[i], a);
        unpack8(wv[v], wf);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = bf16_round(a[k] * rstd) * wf[k];
        reinterpret_cast<bf16x8*>(y + (size_t)row * H)[v] = pack8(o);
      }
    }
  }
}

// dx = rstd * (dy*w - xhat * mean(dy*w*xhat)) (+ dres)
template <int VPL, bool kHasDres>
__global__ void __launch_bounds__(128)
rmsnorm_bwd_dx_warp_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                           const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                           const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx, int M,
                           int H) {
  const int lane = threadIdx.x & 31;
  const int nvec = H >> 3;
  const int wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  const bf16x8* wv = reinterpret_cast<const bf16x8*>(w);
  for (int row = wid; row < M; row += nwarps) {
    const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * H);
    const bf16x8* gr = reinterpret_cast<const bf16x8*>(dy + (size_t)row * H);
    bf16x8 xv[VPL], gv[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < nvec) {
        xv[i] = xr[v];
        gv[i] = gr[v];
      }
    }
    const float rstd = rstd_in[row];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < nvec) {
        float a[8], g[8], wf[8];
        unpack8(xv[i], a);
        unpack8(gv[i], g);
        unpack8(wv[v], wf);
#pragma unroll
        for (int k = 0; k < 8; ++k) dot += g[k] * wf[k] * (a[k] * rstd);
      }
    }
    dot = warp_sum(dot) / (float)H;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < nvec) {
        float a[8], g[8], wf[8], o[8];
        unpack8(xv[i], a);
        unpack8(gv[i], g);
        unpack8(wv[v], wf);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rstd * (g[k] * wf[k] - a[k] * rstd * dot);
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
}

// partial[strip][c] = sum over the strip's rows of dy[r][c] * bf16(x[r][c] * rstd[r])
__global__ void __launch_bounds__(256)
rmsnorm_dw_strip_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                        const float* __restrict__ rstd_in, float* __restrict__ partial, int M, int H,
                        int rows_per_strip) {
  const int nvec = H >> 3;
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nvec) return;
  const int r0 = blockIdx.y * rows_per_strip;
  const int r1 = min(M, r0 + rows_per_strip);
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  int r = r0;
  for (; r + 4 <= r1; r += 4) {
    bf16x8 gq[4], xq[4];
    float rs[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      gq[u] = reinterpret_cast<const bf16x8*>(dy + (size_t)(r + u) * H)[v];
      xq[u] = reinterpret_cast<const bf16x8*>(x + (size_t)(r + u) * H)[v];
      rs[u] = rstd_in[r + u];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float g[8], a[8];
      unpack8(gq[u], g);
      unpack8(xq[u], a);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += g[k] * bf16_round(a[k] * rs[u]);
    }
  }
  for (; r < r1; ++r) {
    float g[8], a[8];
    unpack8(reinterpret_cast<const bf16x8*>(dy + (size_t)r * H)[v], g);
    unpack8(reinterpret_cast<const bf16x8*>(x + (size_t)r * H)[v], a);
    const float rs = rstd_in[r];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += g[k] * bf16_round(a[k] * rs);
  }
  float* p = partial + (size_t)blockIdx.y * H + (size_t)v * 8;
  reinterpret_cast<float4*>(p)[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

__global__ void colsum_kernel(const float* __restrict__ partial, float* __restrict__ out, int nblk,
                              int H, int accumulate);

bool rmsnorm_warp_ok(int H) { return (H % 256) == 0 && H <= 5120; }

template <int VPL>
static void launch_fwd(const void* x, const void* residual, const void* w, void* y, void* h_out,
                       float* rstd, int M, int H, float eps, cudaStream_t stream) {
  const int blocks = min((M + 7) / 8, kNumSMs * 2);
  if (residual)
    rmsnorm_fwd_warp_kernel<VPL, true><<<blocks, 256, 0, stream>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (const __nv_bfloat16*)w,
        (__nv_bfloat16*)y, (__nv_bfloat16*)h_out, rstd, M, H, eps);
  else
    rmsnorm_fwd_warp_kernel<VPL, false><<<blocks, 256, 0, stream>>>(
        (const __nv_bfloat16*)x, nullptr, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, nullptr, rstd, M,
        H, eps);
}

void rmsnorm_fwd_warp(const void* x, const void* residual, const void* w, void* y, void* h_out,
                      float* rstd, int M, int H, float eps, cudaStream_t stream) {
  const int vpl = (H / 8 + 31) / 32;
  if (vpl <= 4) launch_fwd<4>(x, residual, w, y, h_out, rstd, M, H, eps, stream);
  else if (vpl <= 8) launch_fwd<8>(x, residual, w, y, h_out, rstd, M, H, eps, stream);
  else if (vpl <= 16) launch_fwd<16>(x, residual, w, y, h_out, rstd, M, H, eps, stream);
  else launch_fwd<20>(x, residual, w, y, h_out, rstd, M, H, eps, stream);
}

template <int VPL>
static void launch_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                       void* dx, int M, int H, cudaStream_t stream) {
  const int blocks = min((M + 3) / 4, kNumSMs * 4);
  if (dres)
    rmsnorm_bwd_dx_warp_kernel<VPL, true><<<blocks, 128, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, rstd,
        (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, M, H);
  else
    rmsnorm_bwd_dx_warp_kernel<VPL, false><<<blocks, 128, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, rstd, nullptr,
        (__nv_bfloat16*)dx, M, H);
}

// scratch: fp32 [max_strips, H]; dw (+)= column sums
void rmsnorm_bwd_warp(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                      void* dx, float* dw, float* scratch, int max_strips, int M, int H, int accumulate,
                      cudaStream_t stream) {
  const int vpl = (H / 8 + 31) / 32;
  if (vpl <= 4) launch_bwd<4>(dy, x, w, rstd, dres, dx, M, H, stream);
  else if (vpl <= 8) launch_bwd<8>(dy, x, w, rstd, dres, dx, M, H, stream);
  else if (vpl <= 16) launch_bwd<16>(dy, x, w, rstd, dres, dx, M, H, stream);
  else launch_bwd<20>(dy, x, w, rstd, dres, dx, M, H, stream);
  const int rps = (M + max_strips - 1) / max_strips;
  const int strips = (M + rps - 1) / rps;
  dim3 grid((H / 8 + 255) / 256, strips);
  rmsnorm_dw_strip_kernel<<<grid, 256, 0, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, rstd,
                                                    scratch, M, H, rps);
  colsum_kernel<<<(H + 255) / 256, 256, 0, stream>>>(scratch, dw, strips, H, accumulate);
}

}  // namespace b200
