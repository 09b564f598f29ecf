// Vocab-parallel embedding over peer memory (K13 of SURVEY.md section 8).
//
// Reference: megatron/core/tensor_parallel/layers.py:187-210 looks up the rows a rank owns, zeroes
// the others and ALL-REDUCES [b, s, h] over the tensor-parallel group (then, under sequence
// parallelism, throws (tp-1)/tp of it away, language_model.py:255-256); the backward is an
// all-gather of the output gradient followed by aten::embedding_dense_backward and a bf16 -> fp32
// add into main_grad.
//
// Here nothing is reduced. The embedding tables of the group are mapped into each other's address
// space (CUDA IPC), and
//   forward:  every output row is COPIED from the table shard of the rank that owns the token
//             (16-byte loads over NVLink when that is another GPU), already in [s, b, h] order and,
//             under sequence parallelism, only for this rank's part of the sequence;
//   backward: every rank walks all tokens, and for the ones it owns adds the output-gradient row
//             -- read from its own tensor or from the staging buffer of the rank that holds that
//             part of the sequence -- into the fp32 main_grad row with vector atomics.
// The order of the atomic adds is not fixed, so repeated tokens make the embedding gradient
// reproducible only up to fp32 rounding (as with the reference's index_add).
#include "common.cuh"
#include <string.h>

namespace b200 {

struct EmbedPeers {
  const uint4* p[8];
};

__device__ __forceinline__ uint4 ld_cv_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.cv.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// out[(si - s_off) * B + bi][:] = table_{owner}[id - owner * Vp][:], id = ids[bi * S + si]
__global__ void __launch_bounds__(128)
embed_fwd_kernel(const long long* __restrict__ ids, EmbedPeers w, uint4* __restrict__ out, int S, int B,
                 int s_off, int s_out, int nvec, int Vp, int tp) {
  const int lane = threadIdx.x & 31;
  const int wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  const int rows = s_out * B;
  for (int r = wid; r < rows; r += nwarps) {
    const int si = r / B + s_off, bi = r % B;
    long long id = ids[(size_t)bi * S + si];
    int owner = (int)(id / Vp);
    if (owner >= tp) owner = tp - 1;                 // ids are < tp * Vp by construction
    const uint4* src = w.p[owner] + (size_t)(id - (long long)owner * Vp) * nvec;
    uint4* dst = out + (size_t)r * nvec;
    for (int v0 = lane; v0 < nvec; v0 += 32 * 8) {
      uint4 q[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (v0 + 32 * u < nvec) q[u] = ld_cv_u4(src + v0 + 32 * u);
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (v0 + 32 * u < nvec) dst[v0 + 32 * u] = q[u];
    }
  }
}

// main_grad[id - lo][:] += dout_row(si, bi) for every token with lo <= id < lo + Vp.
// src.p[c] holds rows [c * s_chunk, (c + 1) * s_chunk) x B of the output gradient ([s, b, h] order).
__global__ void __launch_bounds__(128)
embed_bwd_kernel(const long long* __restrict__ ids, EmbedPeers src, float* __restrict__ main_grad, int S,
                 int B, int s_chunk, int nvec, long long lo, int Vp) {
  const int lane = threadIdx.x & 31;
  const int wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  const int rows = S * B;
  for (int r = wid; r < rows; r += nwarps) {
    const int si = r / B, bi = r % B;
    const long long id = ids[(size_t)bi * S + si] - lo;
    if (id < 0 || id >= Vp) continue;
    const int c = si / s_chunk;
    const uint4* g = src.p[c] + ((size_t)(si - c * s_chunk) * B + bi) * nvec;
    float* dst = main_grad + (size_t)id * nvec * 8;
    for (int v = lane; v < nvec; v += 32) {
      float f[8];
      unpack8(ld_cv_u4(g + v), f);
      atomicAdd(reinterpret_cast<float4*>(dst + (size_t)v * 8), make_float4(f[0], f[1], f[2], f[3]));
      atomicAdd(reinterpret_cast<float4*>(dst + (size_t)v * 8 + 4), make_float4(f[4], f[5], f[6], f[7]));
    }
  }
}

}  // namespace b200

using namespace b200;

// CUDA IPC handle of the ALLOCATION that contains `ptr` (any device pointer: a torch tensor inside a
// caching-allocator segment, or an own cudaMalloc) + the offset of `ptr` in it. The peer opens the
// handle (b200_symm_import) and adds the offset.
B200_API int b200_ipc_export_range(const void* ptr, void* handle_out, long long* offset_out) {
  typedef CUresult (*range_fn)(CUdeviceptr*, size_t*, CUdeviceptr);
  static range_fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) != cudaSuccess || !p)
      return B200_ERR_UNSUPPORTED;
    fn = reinterpret_cast<range_fn>(p);
  }
  CUdeviceptr base = 0;
  size_t size = 0;
  if (fn(&base, &size, (CUdeviceptr)ptr) != CUDA_SUCCESS) return B200_ERR_BAD_ARG;
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, (void*)base);
  if (e != cudaSuccess) return (int)e;
  memcpy(handle_out, &h, sizeof(h));
  *offset_out = (long long)((CUdeviceptr)ptr - base);
  return 0;
}

// ids [B, S] int64; tables[r] = rank r's [Vp, H] bf16 shard (peer-mapped); out [s_out * B, H] bf16 =
// rows of sequence positions [s_off, s_off + s_out) in [s, b] order.
B200_API int b200_embed_fwd(const long long* ids, const void* const* tables, void* out, int S, int B,
                            int s_off, int s_out, int H, int Vp, int tp, cudaStream_t stream) {
  if (H % 8 || tp < 1 || tp > 8 || s_off < 0 || s_off + s_out > S) return B200_ERR_BAD_ARG;
  if (s_out * B == 0) return 0;
  EmbedPeers w = {};
  for (int r = 0; r < tp; ++r) w.p[r] = (const uint4*)tables[r];
  const int rows = s_out * B;
  const int grid = rows < kNumSMs * 32 ? (rows + 3) / 4 : kNumSMs * 8;
  embed_fwd_kernel<<<grid, 128, 0, stream>>>(ids, w, (uint4*)out, S, B, s_off, s_out, H / 8, Vp, tp);
  B200_CHECK_LAUNCH();
  return 0;
}

// main_grad [Vp, H] fp32 += rows of the output gradient of the tokens in [lo, lo + Vp).
// srcs[c] (c < n_chunks): rows [c * S / n_chunks, ...) x B of the gradient, bf16, [s, b, h] order.
B200_API int b200_embed_bwd(const long long* ids, const void* const* srcs, int n_chunks, float* main_grad,
                            int S, int B, int H, long long lo, int Vp, cudaStream_t stream) {
  if (H % 8 || n_chunks < 1 || n_chunks > 8 || S % n_chunks) return B200_ERR_BAD_ARG;
  if (S * B == 0) return 0;
  EmbedPeers s = {};
  for (int c = 0; c < n_chunks; ++c) s.p[c] = (const uint4*)srcs[c];
  const int rows = S * B;
  const int grid = rows < kNumSMs * 32 ? (rows + 3) / 4 : kNumSMs * 8;
  embed_bwd_kernel<<<grid, 128, 0, stream>>>(ids, s, main_grad, S, B, S / n_chunks, H / 8, lo, Vp);
  B200_CHECK_LAUNCH();
  return 0;
}
