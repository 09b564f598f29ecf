// RMSNorm forward / backward as persistent kernels fed by bulk-copy (TMA) rings.
//
// Why: a norm moves 2-4 bytes per FLOP, so the only thing that matters is how many bytes are in
// flight per SM for the whole duration of the kernel. The register-resident variants
// (csrc/norm.cu, csrc/norm_warp.cu) issue a row's loads, wait, compute, store, and only then
// issue the next row's loads; with 1-2 rows per warp the memory system idles between phases and
// they reach 20-40 % of HBM bandwidth on [4096, 4096]. Here one elected thread per CTA streams rows
// into a shared-memory ring with cp.async.bulk (completion on an mbarrier, no registers, no
// address math in the consumers) up to ~190 KB ahead of the consumer warps, so the loads of the
// next rows are always outstanding while the current rows are reduced, normalised and stored.
//
//   grid   = min(148, stages of work)   one CTA per SM, rows dealt round-robin in pairs
//   block  = 16 consumer warps + 1 producer warp
//   stage  = kNtRows rows of every input array (x [, residual] / dy, x [, dres])
//   The consumer warps form independent groups (4 groups of 4 warps for H <= 4096, 2 groups of 8
//   up to H = 8192); ring stage i belongs to group i % groups. The warps of a group share the rows
//   of a stage (thread t owns 16-byte vectors t, t + group size, ...): per-thread state is
//   O(H / group size), and the weight-gradient partials of ALL rows a group sees stay in registers
//   (the separate column-sum pass that re-read x and dy is gone). One named barrier per stage
//   combines the per-warp sums of a group; the groups run out of phase, so one group's reduction
//   latency is covered by the others' loads and stores.
//
// Same math and rounding order as the reference's eager RMSNorm
// (megatron/model/fused_layer_norm.py:125-139: fp32 x * rsqrt(mean(x^2) + eps), round to the input
// dtype, then multiply by the weight).
#include "common.cuh"

namespace b200 {

constexpr int kNtConsumerWarps = 16;
constexpr int kNtConsumers = kNtConsumerWarps * 32;
constexpr int kNtThreads = kNtConsumers + 32;
constexpr int kNtRows = 2;        // rows per ring stage
constexpr int kNtVpt = 4;         // 16-byte vectors per thread and row: H <= group threads * 4 * 8
constexpr int kNtMaxStages = 12;
constexpr int kNtHeader = 2048;   // barriers + reduction scratch
constexpr int kNtSmemBudget = 200 * 1024;

__device__ __forceinline__ void nt_bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes),
      "r"(smem_u32(bar))
      : "memory");
}
// barrier `id` over `n` threads (ids 1.. are the consumer groups, kNtAllBar all consumers)
__device__ __forceinline__ void nt_bar(int id, int n) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
constexpr int kNtAllBar = 8;

struct NtSmem {
  uint64_t full[kNtMaxStages];
  uint64_t empty[kNtMaxStages];
  float red[4][2][kNtRows][8];    // [group][parity][row][warp of the group]
  float red2[4][2][kNtRows][8];   // second per-row sum of the LayerNorm kernels
};
static_assert(sizeof(NtSmem) <= kNtHeader, "header too small");

template <int kGW>
__device__ __forceinline__ void nt_init(NtSmem* sm, int S) {
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&sm->full[s], 1);
      mbar_init(&sm->empty[s], kGW);
    }
    mbar_fence_init();
  }
  __syncthreads();
}

// Producer loop shared by both kernels: stage `it` of this CTA holds rows [t * kNtRows, ...) of
// each of the `kArrays` inputs, one after the other.
template <int kArrays>
__device__ __forceinline__ void nt_produce(NtSmem* sm, uint8_t* stages, const __nv_bfloat16* a0,
                                           const __nv_bfloat16* a1, const __nv_bfloat16* a2, int M, int H,
                                           int T, int S, int stage_bytes) {
  const uint32_t row_bytes = (uint32_t)H * 2u;
  int it = 0;
  for (int t = blockIdx.x; t < T; t += gridDim.x, ++it) {
    const int slot = it % S;
    const uint32_t ph = (uint32_t)(it / S) & 1u;
    mbar_wait(&sm->empty[slot], ph ^ 1u);
    const int r0 = t * kNtRows;
    const int nr = min(kNtRows, M - r0);
    const uint32_t bytes = (uint32_t)nr * row_bytes;
    mbar_expect_tx(&sm->full[slot], bytes * kArrays);
    uint8_t* dst = stages + (size_t)slot * stage_bytes;
    nt_bulk_load(dst, a0 + (size_t)r0 * H, bytes, &sm->full[slot]);
    if (kArrays > 1) nt_bulk_load(dst + kNtRows * row_bytes, a1 + (size_t)r0 * H, bytes, &sm->full[slot]);
    if (kArrays > 2) nt_bulk_load(dst + 2 * kNtRows * row_bytes, a2 + (size_t)r0 * H, bytes, &sm->full[slot]);
  }
}

// Fold the consumer groups' per-column partial sums in group order (fixed order: deterministic)
// through `fold` (idle ring memory, >= H floats) and write the CTA's row of the [grid, .., H] scratch.
// Called by all consumer threads after their last stage.
template <int kGW>
__device__ __forceinline__ void nt_fold_groups(float (&acc)[kNtVpt][8], float* fold, float* dst, int grp,
                                               int gtid, int nvec) {
  constexpr int kGroups = kNtConsumerWarps / kGW;
  constexpr int kGT = kGW * 32;
  nt_bar(kNtAllBar, kNtConsumers);
#pragma unroll 1
  for (int gsel = 0; gsel < kGroups; ++gsel) {
    if (grp == gsel) {
#pragma unroll
      for (int i = 0; i < kNtVpt; ++i) {
        const int v = gtid + i * kGT;
        if (v < nvec) {
          float4* f = reinterpret_cast<float4*>(fold + (size_t)v * 8);
          float4 lo = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
          float4 hi = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
          if (gsel > 0) {
            const float4 a = f[0], b = f[1];
            lo.x += a.x; lo.y += a.y; lo.z += a.z; lo.w += a.w;
            hi.x += b.x; hi.y += b.y; hi.z += b.z; hi.w += b.w;
          }
          if (gsel == kGroups - 1) {
            float4* p = reinterpret_cast<float4*>(dst + (size_t)v * 8);
            p[0] = lo;
            p[1] = hi;
          } else {
            f[0] = lo;
            f[1] = hi;
          }
        }
      }
    }
    if (gsel + 1 < kGroups) nt_bar(kNtAllBar, kNtConsumers);
  }
}

// kGW = warps per consumer group (4: H <= 4096, 8: H <= 8192)
template <int kGW, bool kHasResidual>
__global__ void __launch_bounds__(kNtThreads, 1)
rmsnorm_fwd_tma_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                       const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y,
                       __nv_bfloat16* __restrict__ h_out, float* __restrict__ rstd_out, int M, int H,
                       float eps, int S, int stage_bytes) {
  constexpr int kGroups = kNtConsumerWarps / kGW;
  constexpr int kGT = kGW * 32;
  extern __shared__ __align__(1024) uint8_t nt_smem[];
  NtSmem* sm = reinterpret_cast<NtSmem*>(nt_smem);
  uint8_t* stages = nt_smem + kNtHeader;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = H >> 3;
  const int row_bytes = H * 2;
  const int T = (M + kNtRows - 1) / kNtRows;
  nt_init<kGW>(sm, S);
  if (warp == kNtConsumerWarps) {
    if (lane == 0)
      nt_produce<kHasResidual ? 2 : 1>(sm, stages, x, res, nullptr, M, H, T, S, stage_bytes);
    return;
  }
  const int grp = warp / kGW, gwarp = warp % kGW;
  const int gtid = threadIdx.x - grp * kGT;
  const bf16x8* wvec = reinterpret_cast<const bf16x8*>(w);   // 2 * H bytes, stays in L1
  int it = grp, par = 0;
  for (int t = blockIdx.x + grp * gridDim.x; t < T; t += kGroups * gridDim.x, it += kGroups, par ^= 1) {
    const int slot = it % S;
    const uint32_t ph = (uint32_t)(it / S) & 1u;
    const int r0 = t * kNtRows;
    const int nr = min(kNtRows, M - r0);
    const uint8_t* st = stages + (size_t)slot * stage_bytes;
    mbar_wait(&sm->full[slot], ph);
    bf16x8 hv[kNtRows][kNtVpt];
    float ss[kNtRows][2];
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      ss[r][0] = ss[r][1] = 0.f;
      if (r < nr) {
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(st + r * row_bytes);
        const bf16x8* rr = reinterpret_cast<const bf16x8*>(st + (kNtRows + r) * row_bytes);
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            hv[r][i] = xr[v];
            float a[8];
            unpack8(hv[r][i], a);
            if (kHasResidual) {
              float b[8];
              unpack8(rr[v], b);
#pragma unroll
              for (int k = 0; k < 8; ++k) a[k] += b[k];
              hv[r][i] = pack8(a);                       // h = bf16(x + residual) is what the norm sees
              unpack8(hv[r][i], a);
              reinterpret_cast<bf16x8*>(h_out + (size_t)(r0 + r) * H)[v] = hv[r][i];
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) ss[r][k & 1] += a[k] * a[k];
          }
        }
      }
    }
    // the stage's bytes are in registers: hand the slot back before the reduction
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm->empty[slot]);
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      const float sr = warp_sum(ss[r][0] + ss[r][1]);
      if (lane == 0) sm->red[grp][par][r][gwarp] = sr;
    }
    nt_bar(1 + grp, kGT);
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      if (r < nr) {
        float tot = 0.f;
#pragma unroll
        for (int q = 0; q < kGW; ++q) tot += sm->red[grp][par][r][q];
        const float rstd = rsqrtf(tot / (float)H + eps);
        if (gtid == 0 && rstd_out) rstd_out[r0 + r] = rstd;
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            // bf16(x * rstd) * w with the product formed by the packed bf16 multiply: the fp32 product of
            // two bf16 values is exact, so one rounding either way -- same bits, a third of the instructions
            float a[8];
            unpack8(hv[r][i], a);
            const uint4 wq = __ldg(reinterpret_cast<const uint4*>(wvec + v));
            const __nv_bfloat162* w2 = reinterpret_cast<const __nv_bfloat162*>(&wq);
            bf16x8 o;
#pragma unroll
            for (int k = 0; k < 4; ++k)
              o.v[k] = __hmul2(__floats2bfloat162_rn(a[2 * k] * rstd, a[2 * k + 1] * rstd), w2[k]);
            reinterpret_cast<bf16x8*>(y + (size_t)(r0 + r) * H)[v] = o;
          }
        }
      }
    }
  }
}

// dx = rstd * (dy*w - xhat * mean(dy*w*xhat)) (+ dres);  partial[cta][c] = sum over the CTA's rows
// of dy[r][c] * bf16(x[r][c] * rstd[r])  (the forward multiplied the ROUNDED xhat by w).
template <int kGW, bool kHasDres>
__global__ void __launch_bounds__(kNtThreads, 1)
rmsnorm_bwd_tma_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                       const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                       const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                       float* __restrict__ partial, int M, int H, int S, int stage_bytes) {
  constexpr int kGroups = kNtConsumerWarps / kGW;
  constexpr int kGT = kGW * 32;
  extern __shared__ __align__(1024) uint8_t nt_smem[];
  NtSmem* sm = reinterpret_cast<NtSmem*>(nt_smem);
  uint8_t* stages = nt_smem + kNtHeader;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = H >> 3;
  const int row_bytes = H * 2;
  const int T = (M + kNtRows - 1) / kNtRows;
  nt_init<kGW>(sm, S);
  if (warp == kNtConsumerWarps) {
    if (lane == 0)
      nt_produce<kHasDres ? 3 : 2>(sm, stages, dy, x, dres, M, H, T, S, stage_bytes);
    return;
  }
  const int grp = warp / kGW, gwarp = warp % kGW;
  const int gtid = threadIdx.x - grp * kGT;
  bf16x8 wv[kNtVpt];
  float dwacc[kNtVpt][8];
#pragma unroll
  for (int i = 0; i < kNtVpt; ++i) {
    const int v = gtid + i * kGT;
    if (v < nvec) wv[i] = reinterpret_cast<const bf16x8*>(w)[v];
#pragma unroll
    for (int k = 0; k < 8; ++k) dwacc[i][k] = 0.f;
  }
  int it = grp, par = 0;
  for (int t = blockIdx.x + grp * gridDim.x; t < T; t += kGroups * gridDim.x, it += kGroups, par ^= 1) {
    const int slot = it % S;
    const uint32_t ph = (uint32_t)(it / S) & 1u;
    const int r0 = t * kNtRows;
    const int nr = min(kNtRows, M - r0);
    const uint8_t* st = stages + (size_t)slot * stage_bytes;
    float rstd[kNtRows];
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) rstd[r] = r < nr ? rstd_in[r0 + r] : 0.f;
    mbar_wait(&sm->full[slot], ph);
    float dot[kNtRows][2];
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      dot[r][0] = dot[r][1] = 0.f;
      if (r < nr) {
        const bf16x8* gr = reinterpret_cast<const bf16x8*>(st + r * row_bytes);
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(st + (kNtRows + r) * row_bytes);
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            float g[8], a[8], wf[8];
            unpack8(gr[v], g);
            unpack8(xr[v], a);
            unpack8(wv[i], wf);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float xh = a[k] * rstd[r];
              dwacc[i][k] += g[k] * bf16_round(xh);
              dot[r][k & 1] += g[k] * wf[k] * xh;
            }
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      const float dr = warp_sum(dot[r][0] + dot[r][1]);
      if (lane == 0) sm->red[grp][par][r][gwarp] = dr;
    }
    nt_bar(1 + grp, kGT);
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      if (r < nr) {
        float tot = 0.f;
#pragma unroll
        for (int q = 0; q < kGW; ++q) tot += sm->red[grp][par][r][q];
        const float mean_dot = tot / (float)H;
        const bf16x8* gr = reinterpret_cast<const bf16x8*>(st + r * row_bytes);
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(st + (kNtRows + r) * row_bytes);
        const bf16x8* dr = reinterpret_cast<const bf16x8*>(st + (2 * kNtRows + r) * row_bytes);
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            float g[8], a[8], wf[8], o[8];
            unpack8(gr[v], g);
            unpack8(xr[v], a);
            unpack8(wv[i], wf);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = rstd[r] * (g[k] * wf[k] - a[k] * rstd[r] * mean_dot);
            if (kHasDres) {
              float d[8];
              unpack8(dr[v], d);
#pragma unroll
              for (int k = 0; k < 8; ++k) o[k] += d[k];
            }
            reinterpret_cast<bf16x8*>(dx + (size_t)(r0 + r) * H)[v] = pack8(o);
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm->empty[slot]);
  }
  nt_fold_groups<kGW>(dwacc, reinterpret_cast<float*>(stages), partial + (size_t)blockIdx.x * H, grp, gtid, nvec);
}

// ------------------------------- LayerNorm on the same ring -------------------------------
// Behaviour of the reference's mixed-precision LayerNorm extension
// (megatron/fused_kernels/layer_norm_cuda_kernel.cu: fp32 mean / variance per row, output in the
// parameter dtype, mean and 1/std saved): y = (x - mean) * rstd * w + b, two-pass variance.
template <int kGW>
__global__ void __launch_bounds__(kNtThreads, 1)
layernorm_fwd_tma_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                         const __nv_bfloat16* __restrict__ b, __nv_bfloat16* __restrict__ y,
                         float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int H,
                         float eps, int S, int stage_bytes) {
  constexpr int kGroups = kNtConsumerWarps / kGW;
  constexpr int kGT = kGW * 32;
  extern __shared__ __align__(1024) uint8_t nt_smem[];
  NtSmem* sm = reinterpret_cast<NtSmem*>(nt_smem);
  uint8_t* stages = nt_smem + kNtHeader;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = H >> 3;
  const int row_bytes = H * 2;
  const int T = (M + kNtRows - 1) / kNtRows;
  nt_init<kGW>(sm, S);
  if (warp == kNtConsumerWarps) {
    if (lane == 0) nt_produce<1>(sm, stages, x, nullptr, nullptr, M, H, T, S, stage_bytes);
    return;
  }
  const int grp = warp / kGW, gwarp = warp % kGW;
  const int gtid = threadIdx.x - grp * kGT;
  const uint4* wvec = reinterpret_cast<const uint4*>(w);
  const uint4* bvec = reinterpret_cast<const uint4*>(b);
  int it = grp, par = 0;
  for (int t = blockIdx.x + grp * gridDim.x; t < T; t += kGroups * gridDim.x, it += kGroups, par ^= 1) {
    const int slot = it % S;
    const uint32_t ph = (uint32_t)(it / S) & 1u;
    const int r0 = t * kNtRows;
    const int nr = min(kNtRows, M - r0);
    const uint8_t* st = stages + (size_t)slot * stage_bytes;
    mbar_wait(&sm->full[slot], ph);
    bf16x8 hv[kNtRows][kNtVpt];
    float s1[kNtRows];
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      s1[r] = 0.f;
      if (r < nr) {
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(st + r * row_bytes);
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            hv[r][i] = xr[v];
            float a[8];
            unpack8(hv[r][i], a);
#pragma unroll
            for (int k = 0; k < 8; ++k) s1[r] += a[k];
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm->empty[slot]);     // the rows are in registers
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      const float sr = warp_sum(s1[r]);
      if (lane == 0) sm->red[grp][par][r][gwarp] = sr;
    }
    nt_bar(1 + grp, kGT);
    float mean[kNtRows], s2[kNtRows];
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      float tot = 0.f;
#pragma unroll
      for (int q = 0; q < kGW; ++q) tot += sm->red[grp][par][r][q];
      mean[r] = tot / (float)H;
      s2[r] = 0.f;
      if (r < nr) {
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          if (gtid + i * kGT < nvec) {
            float a[8];
            unpack8(hv[r][i], a);
#pragma unroll
            for (int k = 0; k < 8; ++k) s2[r] += (a[k] - mean[r]) * (a[k] - mean[r]);
          }
        }
      }
      const float sr = warp_sum(s2[r]);
      if (lane == 0) sm->red2[grp][par][r][gwarp] = sr;
    }
    nt_bar(1 + grp, kGT);
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      if (r < nr) {
        float tot = 0.f;
#pragma unroll
        for (int q = 0; q < kGW; ++q) tot += sm->red2[grp][par][r][q];
        const float rstd = rsqrtf(tot / (float)H + eps);
        if (gtid == 0) {
          mean_out[r0 + r] = mean[r];
          rstd_out[r0 + r] = rstd;
        }
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            float a[8], wf[8], bf[8], o[8];
            unpack8(hv[r][i], a);
            unpack8(__ldg(wvec + v), wf);
            if (b) unpack8(__ldg(bvec + v), bf);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = (a[k] - mean[r]) * rstd * wf[k] + (b ? bf[k] : 0.f);
            reinterpret_cast<bf16x8*>(y + (size_t)(r0 + r) * H)[v] = pack8(o);
          }
        }
      }
    }
  }
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)), g = dy * w, xhat = (x - mean) * rstd;
// partial[cta][0][c] = sum dy * xhat (dgamma), partial[cta][1][c] = sum dy (dbeta).
template <int kGW>
__global__ void __launch_bounds__(kNtThreads, 1)
layernorm_bwd_tma_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                         const __nv_bfloat16* __restrict__ w, const float* __restrict__ mean_in,
                         const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dx,
                         float* __restrict__ partial, int M, int H, int S, int stage_bytes) {
  constexpr int kGroups = kNtConsumerWarps / kGW;
  constexpr int kGT = kGW * 32;
  extern __shared__ __align__(1024) uint8_t nt_smem[];
  NtSmem* sm = reinterpret_cast<NtSmem*>(nt_smem);
  uint8_t* stages = nt_smem + kNtHeader;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = H >> 3;
  const int row_bytes = H * 2;
  const int T = (M + kNtRows - 1) / kNtRows;
  nt_init<kGW>(sm, S);
  if (warp == kNtConsumerWarps) {
    if (lane == 0) nt_produce<2>(sm, stages, dy, x, nullptr, M, H, T, S, stage_bytes);
    return;
  }
  const int grp = warp / kGW, gwarp = warp % kGW;
  const int gtid = threadIdx.x - grp * kGT;
  const uint4* wvec = reinterpret_cast<const uint4*>(w);
  float dg[kNtVpt][8], db[kNtVpt][8];
#pragma unroll
  for (int i = 0; i < kNtVpt; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k) dg[i][k] = db[i][k] = 0.f;
  int it = grp, par = 0;
  for (int t = blockIdx.x + grp * gridDim.x; t < T; t += kGroups * gridDim.x, it += kGroups, par ^= 1) {
    const int slot = it % S;
    const uint32_t ph = (uint32_t)(it / S) & 1u;
    const int r0 = t * kNtRows;
    const int nr = min(kNtRows, M - r0);
    const uint8_t* st = stages + (size_t)slot * stage_bytes;
    float mean[kNtRows], rstd[kNtRows];
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      mean[r] = r < nr ? mean_in[r0 + r] : 0.f;
      rstd[r] = r < nr ? rstd_in[r0 + r] : 0.f;
    }
    mbar_wait(&sm->full[slot], ph);
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      float s1 = 0.f, s2 = 0.f;
      if (r < nr) {
        const bf16x8* gr = reinterpret_cast<const bf16x8*>(st + r * row_bytes);
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(st + (kNtRows + r) * row_bytes);
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            float g[8], a[8], wf[8];
            unpack8(gr[v], g);
            unpack8(xr[v], a);
            unpack8(__ldg(wvec + v), wf);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float xh = (a[k] - mean[r]) * rstd[r];
              dg[i][k] += g[k] * xh;
              db[i][k] += g[k];
              const float gw = g[k] * wf[k];
              s1 += gw;
              s2 += gw * xh;
            }
          }
        }
      }
      s1 = warp_sum(s1);
      s2 = warp_sum(s2);
      if (lane == 0) {
        sm->red[grp][par][r][gwarp] = s1;
        sm->red2[grp][par][r][gwarp] = s2;
      }
    }
    nt_bar(1 + grp, kGT);
#pragma unroll
    for (int r = 0; r < kNtRows; ++r) {
      if (r < nr) {
        float t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int q = 0; q < kGW; ++q) {
          t1 += sm->red[grp][par][r][q];
          t2 += sm->red2[grp][par][r][q];
        }
        t1 /= (float)H;
        t2 /= (float)H;
        const bf16x8* gr = reinterpret_cast<const bf16x8*>(st + r * row_bytes);
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(st + (kNtRows + r) * row_bytes);
#pragma unroll
        for (int i = 0; i < kNtVpt; ++i) {
          const int v = gtid + i * kGT;
          if (v < nvec) {
            float g[8], a[8], wf[8], o[8];
            unpack8(gr[v], g);
            unpack8(xr[v], a);
            unpack8(__ldg(wvec + v), wf);
#pragma unroll
            for (int k = 0; k < 8; ++k)
              o[k] = rstd[r] * (g[k] * wf[k] - t1 - (a[k] - mean[r]) * rstd[r] * t2);
            reinterpret_cast<bf16x8*>(dx + (size_t)(r0 + r) * H)[v] = pack8(o);
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm->empty[slot]);
  }
  float* fold = reinterpret_cast<float*>(stages);
  float* prow = partial + (size_t)blockIdx.x * 2 * H;
  nt_fold_groups<kGW>(dg, fold, prow, grp, gtid, nvec);
  nt_fold_groups<kGW>(db, fold + H, prow + H, grp, gtid, nvec);
}

// dgamma / dbeta (+)= column sums of partial [nblk][2][H]; same tree as colsum_tree_kernel.
__global__ void __launch_bounds__(256)
colsum2_tree_kernel(const float* __restrict__ partial, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    int nblk, int H, int accumulate) {
  __shared__ float red[2][8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float s = 0.f, t = 0.f;
  if (c < H)
    for (int b = ty; b < nblk; b += 8) {
      s += partial[(size_t)b * 2 * H + c];
      t += partial[(size_t)b * 2 * H + H + c];
    }
  red[0][ty][tx] = s;
  red[1][ty][tx] = t;
  __syncthreads();
  if (ty == 0 && c < H) {
    float u = 0.f, v = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      u += red[0][q][tx];
      v += red[1][q][tx];
    }
    dgamma[c] = accumulate ? dgamma[c] + u : u;
    if (dbeta) dbeta[c] = accumulate ? dbeta[c] + v : v;
  }
}

// out[c] (+)= sum_b partial[b][c]; 32 columns x 8 row groups per block, rows summed in a fixed
// order (deterministic), [nblk, H] fp32.
__global__ void __launch_bounds__(256)
colsum_tree_kernel(const float* __restrict__ partial, float* __restrict__ out, int nblk, int H,
                   int accumulate) {
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float s = 0.f;
  if (c < H)
    for (int b = ty; b < nblk; b += 8) s += partial[(size_t)b * H + c];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < H) {
    float t = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += red[q][tx];
    out[c] = accumulate ? out[c] + t : t;
  }
}

bool rmsnorm_tma_ok(int M, int H) {
  return M > 0 && H >= 256 && (H % 8) == 0 && H <= 8 * 32 * kNtVpt * 8;
}

static int nt_stages(int arrays, int H, int* stage_bytes) {
  *stage_bytes = arrays * kNtRows * H * 2;
  int s = (kNtSmemBudget - kNtHeader) / *stage_bytes;
  return s < 2 ? 2 : (s > kNtMaxStages ? kNtMaxStages : s);
}

int rmsnorm_tma_grid(int M) {
  const int T = (M + kNtRows - 1) / kNtRows;
  return T < kNumSMs ? T : kNumSMs;
}

template <typename K, typename... Args>
static int nt_launch(K kernel, int grid, int smem, cudaStream_t stream, Args... args) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  kernel<<<grid, kNtThreads, smem, stream>>>(args...);
  return (int)cudaGetLastError();
}

int rmsnorm_fwd_tma(const void* x, const void* residual, const void* w, void* y, void* h_out,
                    float* rstd, int M, int H, float eps, cudaStream_t stream) {
  int stage_bytes;
  const int S = nt_stages(residual ? 2 : 1, H, &stage_bytes);
  const int smem = kNtHeader + S * stage_bytes;
  const int grid = rmsnorm_tma_grid(M);
  const bool narrow = H <= 4 * 32 * kNtVpt * 8;
  auto xb = (const __nv_bfloat16*)x, rb = (const __nv_bfloat16*)residual, wb = (const __nv_bfloat16*)w;
  auto yb = (__nv_bfloat16*)y, hb = (__nv_bfloat16*)h_out;
  if (residual)
    return narrow ? nt_launch(rmsnorm_fwd_tma_kernel<4, true>, grid, smem, stream, xb, rb, wb, yb, hb, rstd, M,
                              H, eps, S, stage_bytes)
                  : nt_launch(rmsnorm_fwd_tma_kernel<8, true>, grid, smem, stream, xb, rb, wb, yb, hb, rstd, M,
                              H, eps, S, stage_bytes);
  return narrow ? nt_launch(rmsnorm_fwd_tma_kernel<4, false>, grid, smem, stream, xb, rb, wb, yb, hb, rstd, M, H,
                            eps, S, stage_bytes)
                : nt_launch(rmsnorm_fwd_tma_kernel<8, false>, grid, smem, stream, xb, rb, wb, yb, hb, rstd, M, H,
                            eps, S, stage_bytes);
}

// scratch: fp32 [>= rmsnorm_tma_grid(M), H]
int rmsnorm_bwd_tma(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                    void* dx, float* dw, float* scratch, int M, int H, int accumulate,
                    cudaStream_t stream) {
  int stage_bytes;
  const int S = nt_stages(dres ? 3 : 2, H, &stage_bytes);
  const int smem = kNtHeader + S * stage_bytes;
  const int grid = rmsnorm_tma_grid(M);
  const bool narrow = H <= 4 * 32 * kNtVpt * 8;
  auto gb = (const __nv_bfloat16*)dy, xb = (const __nv_bfloat16*)x, wb = (const __nv_bfloat16*)w;
  auto db = (const __nv_bfloat16*)dres;
  auto ob = (__nv_bfloat16*)dx;
  int rc;
  if (dres)
    rc = narrow ? nt_launch(rmsnorm_bwd_tma_kernel<4, true>, grid, smem, stream, gb, xb, wb, rstd, db, ob,
                            scratch, M, H, S, stage_bytes)
                : nt_launch(rmsnorm_bwd_tma_kernel<8, true>, grid, smem, stream, gb, xb, wb, rstd, db, ob,
                            scratch, M, H, S, stage_bytes);
  else
    rc = narrow ? nt_launch(rmsnorm_bwd_tma_kernel<4, false>, grid, smem, stream, gb, xb, wb, rstd, db, ob,
                            scratch, M, H, S, stage_bytes)
                : nt_launch(rmsnorm_bwd_tma_kernel<8, false>, grid, smem, stream, gb, xb, wb, rstd, db, ob,
                            scratch, M, H, S, stage_bytes);
  if (rc) return rc;
  colsum_tree_kernel<<<(H + 31) / 32, 256, 0, stream>>>(scratch, dw, grid, H, accumulate);
  return (int)cudaGetLastError();
}

int layernorm_fwd_tma(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int M,
                      int H, float eps, cudaStream_t stream) {
  int stage_bytes;
  const int S = nt_stages(1, H, &stage_bytes);
  const int smem = kNtHeader + S * stage_bytes;
  const int grid = rmsnorm_tma_grid(M);
  auto xb = (const __nv_bfloat16*)x, wb = (const __nv_bfloat16*)w, bb = (const __nv_bfloat16*)b;
  auto yb = (__nv_bfloat16*)y;
  return H <= 4 * 32 * kNtVpt * 8
             ? nt_launch(layernorm_fwd_tma_kernel<4>, grid, smem, stream, xb, wb, bb, yb, mean, rstd, M, H, eps, S,
                         stage_bytes)
             : nt_launch(layernorm_fwd_tma_kernel<8>, grid, smem, stream, xb, wb, bb, yb, mean, rstd, M, H, eps, S,
                         stage_bytes);
}

// scratch: fp32 [>= rmsnorm_tma_grid(M), 2, H]
int layernorm_bwd_tma(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                      void* dx, float* dgamma, float* dbeta, float* scratch, int M, int H, int accumulate,
                      cudaStream_t stream) {
  int stage_bytes;
  const int S = nt_stages(2, H, &stage_bytes);
  const int smem = kNtHeader + S * stage_bytes;
  const int grid = rmsnorm_tma_grid(M);
  // two accumulators per column: always the 8-warp groups (half the columns per thread)
  int rc = nt_launch(layernorm_bwd_tma_kernel<8>, grid, smem, stream, (const __nv_bfloat16*)dy,
                     (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, mean, rstd, (__nv_bfloat16*)dx, scratch, M,
                     H, S, stage_bytes);
  if (rc) return rc;
  colsum2_tree_kernel<<<(H + 31) / 32, 256, 0, stream>>>(scratch, dgamma, dbeta, grid, H, accumulate);
  return (int)cudaGetLastError();
}

}  // namespace b200
