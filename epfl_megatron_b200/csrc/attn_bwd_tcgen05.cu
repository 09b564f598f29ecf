// Causal / full GQA-native attention backward on tcgen05 + TMEM (K12 of SURVEY.md
// section 8). Replaces the FlashAttention-2 backward the reference reaches through
// flash_attn_func at megatron/model/transformer.py:514-522 (and the K/V head expansion in
// front of it, :448-455, whose backward is a reduction over the expanded heads).
//
// Two passes over the same 128 x 128 score tiles, each a CTA with one STATIONARY operand
// pair kept in shared memory and one STREAMED pair in a 2-stage TMA ring:
//
//   pass KV (template kIsKV = true): stationary K_j, V_j; streams Q_i, dO_i of every query
//     head of the KV group and every query block that sees block j.
//       S^T  = K_j Q_i^T          tcgen05.mma  -> TMEM region A
//       dP^T = V_j dO_i^T         tcgen05.mma  -> TMEM region B
//       P^T  = exp2(S^T * scale*log2e - LSE)          (8 warps, one kv row per thread)
//       dS^T = P^T o (dP^T - D)
//       dV_j += P^T dO_i          A operand read from TMEM (bf16 P^T written over region A)
//       dK_j += dS^T Q_i          A operand read from TMEM (bf16 dS^T written over region B)
//     The GQA reduction of dK / dV over the query heads of a group happens in the TMEM
//     accumulators, so K/V are never expanded and no atomics are needed.
//   pass Q (kIsKV = false): stationary Q_i, dO_i; streams K_j, V_j.
//       S = Q_i K_j^T (double-buffered in TMEM), dP = dO_i V_j^T, dS = P o (dP - D),
//       dQ_i += dS K_j            A operand read from TMEM
//     Recomputing S/dP costs two extra MMAs per tile but removes the fp32 dQ atomics and the
//     conversion pass of the single-pass formulation.
//
// All operands come straight out of the fused projection output [s, b, g, q_per_g+2, hn]
// (rank-5 / rank-4 TMA maps, any strides) and dQ/dK/dV are written straight into the
// gradient of that tensor. A small pre-pass computes D = rowsum(dO o O) and LSE*log2e
// into caller-provided scratch padded to a multiple of 128 rows.
#include "common.cuh"

namespace b200 {

constexpr int kBwdT = 128;        // tile edge (query rows, kv rows, head dim)
constexpr int kBwdThreads = 384;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4-11 math
constexpr int kTileBytes = kBwdT * kBwdT * 2;  // 32 KB: two [128 x 128 B] swizzled boxes
constexpr int kBoxBytes = kBwdT * 128;         // 16 KB

// Shared memory (exactly 1024-aligned, no slack: the dynamic segment starts at the base of the
// CTA's window because the kernel has no static shared memory; checked at run time):
//   Xa, Xb            stationary tiles                      2 x 32 KB
//   Ya ring (3)       streamed tile used FIRST and LAST in an iteration (K_j / Q_i): needs the
//                     extra stage so its reload is issued a full iteration before it is needed
//   Yb ring (2)       streamed tile used only early in the iteration (V_j / dO_i)
//   vec (2)           LSE*log2e and D of the streamed query block (KV pass)
struct BwdSmem {
  static constexpr int kXa = 0;
  static constexpr int kXb = kXa + kTileBytes;
  static constexpr int kYa = kXb + kTileBytes;          // 3 stages
  static constexpr int kYb = kYa + 3 * kTileBytes;      // 2 stages
  static constexpr int kVec = kYb + 2 * kTileBytes;     // 2 x (lse2[128], D[128]) fp32
  static constexpr int kBar = kVec + 2 * 1024;
  static constexpr int kTotal = kBar + 192;
};
static_assert(BwdSmem::kTotal <= 232448, "shared memory budget");

__device__ __forceinline__ void tma_load_5d_b(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                              int c0, int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0),
      "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes),
      "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]),
      "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]),
      "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait_b() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// 2^x: one MUFU.EX2 (masked-out entries may overflow to +inf here; they are replaced by 0 below)
__device__ __forceinline__ float bwd_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// D[tmem] (+)= A[tmem] * B[smem]: the A operand (bf16 pairs, one row per TMEM lane) is read
// from tensor memory.
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ---------------------------------------------------------------------------------
// pre-pass: D[b, head, s] = sum_d O * dO, LSE2 = LSE * log2(e); rows s >= sq get
// D = 0, LSE2 = +inf (so P = 0 there). One warp per (b, head, s) row.
// ---------------------------------------------------------------------------------
__global__ void attn_bwd_prep_kernel(const __nv_bfloat16* __restrict__ o,
                                     const __nv_bfloat16* __restrict__ dout,
                                     const float* __restrict__ lse, float* __restrict__ lse2,
                                     float* __restrict__ dvec, int sq, int s_pad, int B, int np) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int total = B * np * s_pad;
  if (row >= total) return;
  const int s = row % s_pad;
  const int bh = row / s_pad;
  const int head = bh % np, b = bh / np;
  if (s >= sq) {
    if (lane == 0) {
      lse2[row] = INFINITY;
      dvec[row] = 0.f;
    }
    return;
  }
  const size_t off = ((size_t)s * B + b) * ((size_t)np * kBwdT) + (size_t)head * kBwdT + lane * 4;
  const uint2 ov = *reinterpret_cast<const uint2*>(o + off);
  const uint2 dv = *reinterpret_cast<const uint2*>(dout + off);
  const float2 o0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ov.x));
  const float2 o1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ov.y));
  const float2 d0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&dv.x));
  const float2 d1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&dv.y));
  float acc = o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y;
  acc = warp_sum(acc);
  if (lane == 0) {
    dvec[row] = acc;
    lse2[row] = lse[((size_t)b * np + head) * sq + s] * 1.4426950408889634f;
  }
}

// fp32 partial dK / dV sums [2][sk][B][G][128] -> bf16 at their strided destinations.
__global__ void attn_bwd_kv_convert_kernel(const float* __restrict__ kv_acc,
                                           __nv_bfloat16* __restrict__ dk,
                                           __nv_bfloat16* __restrict__ dv, long long o_ss,
                                           long long o_sb, long long o_sg, int sk, int B, int G) {
  const long long rows = (long long)sk * B * G;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one per 8 elements
  if (idx >= 2 * rows * (kBwdT / 8)) return;
  const int c8 = idx % (kBwdT / 8);
  long long row = idx / (kBwdT / 8);
  const int which = row >= rows;
  row -= which * rows;
  const int g = row % G;
  const int b = (row / G) % B;
  const long long s = row / ((long long)G * B);
  const float4* src = reinterpret_cast<const float4*>(kv_acc + (which * rows + row) * kBwdT + c8 * 8);
  const float4 x = src[0], y = src[1];
  uint4 w;
  w.x = pack_bf16x2(x.x, x.y);
  w.y = pack_bf16x2(x.z, x.w);
  w.z = pack_bf16x2(y.x, y.y);
  w.w = pack_bf16x2(y.z, y.w);
  __nv_bfloat16* dst = (which ? dv : dk) + s * o_ss + b * o_sb + g * o_sg + c8 * 8;
  *reinterpret_cast<uint4*>(dst) = w;
}

// ---------------------------------------------------------------------------------
// main kernel, both passes
// ---------------------------------------------------------------------------------
// maps: tmQ rank 5 {hn, s, j, g, b}; tmK / tmV rank 4 {hn, s, g, b}; tmdO rank 4 {hn, s, head, b}
template <bool kIsKV>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                const float* __restrict__ lse2, const float* __restrict__ dvec,
                __nv_bfloat16* __restrict__ g0, __nv_bfloat16* __restrict__ g1, long long o_ss,
                long long o_sb, long long o_sg, long long o_sj, int sq, int sk, int s_pad, int B,
                int G, int qpg, float scale, int causal, int nsplit, float* __restrict__ kv_acc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u)) asm volatile("trap;");   // swizzled tiles need it
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BwdSmem::kBar);
  uint64_t* x_full = bars + 0;
  uint64_t* ya_full = bars + 1;    // [3]
  uint64_t* ya_empty = bars + 4;   // [3]
  uint64_t* yb_full = bars + 7;    // [2]
  uint64_t* yb_empty = bars + 9;   // [2]
  uint64_t* vec_empty = bars + 11; // [2] math warps are done with the stage's LSE / D vectors
  uint64_t* s_full = bars + 13;    // [2]
  uint64_t* dp_full = bars + 15;
  uint64_t* p_ready = bars + 16;
  uint64_t* ds_ready = bars + 17;
  uint64_t* acc_done = bars + 18;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int np = G * qpg;
  const int nq = (sq + kBwdT - 1) / kBwdT, nkv = (sk + kBwdT - 1) / kBwdT;
  const int shift = sk - sq;  // query i sees keys <= i + shift when causal
  const int b = blockIdx.z;
  // KV pass with nsplit > 1: CTA `split` of a (kv block, group) owns query heads
  // [split * qpg/nsplit, (split+1) * qpg/nsplit) and adds its partial dK/dV into kv_acc.
  const int jj0 = kIsKV ? (blockIdx.y % nsplit) * (qpg / nsplit) : 0;
  const float scale_log2 = scale * 1.4426950408889634f;

  // iteration space of this CTA
  int g, head = 0, xblk, y_first, y_count, n_iter;
  if (kIsKV) {
    xblk = blockIdx.x;  // kv block; low blocks see the most queries, so they go first
    g = blockIdx.y / nsplit;
    // first query block with any visible pair: q + shift >= xblk*128
    y_first = causal ? max(0, (xblk * kBwdT - shift) / kBwdT) : 0;
    y_count = max(0, nq - y_first);
    n_iter = y_count * (qpg / nsplit);  // this CTA's share of the group's query heads
  } else {
    xblk = nq - 1 - blockIdx.x;  // query block; late blocks see the most keys
    head = blockIdx.y;
    g = head / qpg;
    y_first = 0;
    y_count = causal ? min(nkv, (xblk * kBwdT + kBwdT - 1 + shift) / kBwdT + 1) : nkv;
    n_iter = y_count;
  }

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmdO);
    mbar_init(x_full, 1);
    for (int i = 0; i < 3; ++i) {
      mbar_init(&ya_full[i], 1);
      mbar_init(&ya_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&yb_full[i], 1);
      mbar_init(&yb_empty[i], 1);
      mbar_init(&vec_empty[i], 256);
      mbar_init(&s_full[i], 1);
    }
    mbar_init(dp_full, 1);
    mbar_init(p_ready, 256);
    mbar_init(ds_ready, 256);
    mbar_init(acc_done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // TMEM columns. KV pass: A = S^T/P^T, B = dP^T/dS^T, ACC1 = dV, ACC2 = dK.
  // Q pass: S double-buffered at 0 / 128, B = dP/dS at 256, ACC = dQ at 384.
  const uint32_t tm_A = tmem_base;
  const uint32_t tm_B = tmem_base + (kIsKV ? 128 : 256);
  const uint32_t tm_acc1 = tmem_base + 256;  // KV pass only
  const uint32_t tm_acc2 = tmem_base + 384;

  uint8_t* sXa = smem + BwdSmem::kXa;
  uint8_t* sXb = smem + BwdSmem::kXb;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0 && n_iter > 0) {
      mbar_expect_tx(x_full, 2 * kTileBytes);
      if (kIsKV) {
        tma_load_4d(sXa, &tmK, x_full, 0, xblk * kBwdT, g, b);
        tma_load_4d(sXa + kBoxBytes, &tmK, x_full, 64, xblk * kBwdT, g, b);
        tma_load_4d(sXb, &tmV, x_full, 0, xblk * kBwdT, g, b);
        tma_load_4d(sXb + kBoxBytes, &tmV, x_full, 64, xblk * kBwdT, g, b);
      } else {
        const int j = head - g * qpg;
        tma_load_5d_b(sXa, &tmQ, x_full, 0, xblk * kBwdT, j, g, b);
        tma_load_5d_b(sXa + kBoxBytes, &tmQ, x_full, 64, xblk * kBwdT, j, g, b);
        tma_load_4d(sXb, &tmdO, x_full, 0, xblk * kBwdT, head, b);
        tma_load_4d(sXb + kBoxBytes, &tmdO, x_full, 64, xblk * kBwdT, head, b);
      }
      for (int it = 0; it < n_iter; ++it) {
        const int sa = it % 3, sb = it & 1;
        const uint32_t pha = ((it / 3) & 1) ^ 1, phb = ((it >> 1) & 1) ^ 1;
        uint8_t* sYa = smem + BwdSmem::kYa + sa * kTileBytes;
        uint8_t* sYb = smem + BwdSmem::kYb + sb * kTileBytes;
        if (kIsKV) {
          const int jl = it / y_count, yb = y_first + it - jl * y_count;
          const int jj = jj0 + jl;
          const int hd = g * qpg + jj;
          float* vec = reinterpret_cast<float*>(smem + BwdSmem::kVec + sb * 1024);
          const size_t voff = ((size_t)b * np + hd) * s_pad + (size_t)yb * kBwdT;
          mbar_wait(&ya_empty[sa], pha);
          mbar_expect_tx(&ya_full[sa], kTileBytes);
          tma_load_5d_b(sYa, &tmQ, &ya_full[sa], 0, yb * kBwdT, jj, g, b);
          tma_load_5d_b(sYa + kBoxBytes, &tmQ, &ya_full[sa], 64, yb * kBwdT, jj, g, b);
          mbar_wait(&yb_empty[sb], phb);
          mbar_wait(&vec_empty[sb], phb);
          mbar_expect_tx(&yb_full[sb], kTileBytes + 1024);
          tma_load_4d(sYb, &tmdO, &yb_full[sb], 0, yb * kBwdT, hd, b);
          tma_load_4d(sYb + kBoxBytes, &tmdO, &yb_full[sb], 64, yb * kBwdT, hd, b);
          bulk_load_1d(vec, lse2 + voff, 512, &yb_full[sb]);
          bulk_load_1d(vec + 128, dvec + voff, 512, &yb_full[sb]);
        } else {
          const int yb = y_first + it;
          mbar_wait(&ya_empty[sa], pha);
          mbar_expect_tx(&ya_full[sa], kTileBytes);
          tma_load_4d(sYa, &tmK, &ya_full[sa], 0, yb * kBwdT, g, b);
          tma_load_4d(sYa + kBoxBytes, &tmK, &ya_full[sa], 64, yb * kBwdT, g, b);
          mbar_wait(&yb_empty[sb], phb);
          mbar_expect_tx(&yb_full[sb], kTileBytes);
          tma_load_4d(sYb, &tmV, &yb_full[sb], 0, yb * kBwdT, g, b);
          tma_load_4d(sYb + kBoxBytes, &tmV, &yb_full[sb], 64, yb * kBwdT, g, b);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (lane == 0 && n_iter > 0) {
      constexpr uint32_t idesc_ss = make_idesc_bf16(kBwdT, kBwdT, 0, 0);
      constexpr uint32_t idesc_ts = make_idesc_bf16(kBwdT, kBwdT, 0, 1);
      const uint32_t aXa = smem_u32(sXa), aXb = smem_u32(sXb);
      const uint32_t aYa = smem_u32(smem + BwdSmem::kYa), aYb = smem_u32(smem + BwdSmem::kYb);
      // D = X * Y^T, both K-major with the head dim split into two 64-wide boxes
      auto mma_xyT = [&](uint32_t d, uint32_t x, uint32_t y) {
#pragma unroll
        for (int k = 0; k < kBwdT / 16; ++k) {
          const uint32_t off = (k >> 2) * kBoxBytes + (k & 3) * 32;
          umma_bf16_ss(d, make_sw128_desc(x + off, 16, 1024), make_sw128_desc(y + off, 16, 1024),
                       idesc_ss, k != 0);
        }
      };
      // D += A(tmem, bf16 pairs) * Y with Y [128 rows (K) x 128 (N)] as stored: MN-major B,
      // 16 rows per MMA = 2 KB, the two 64-wide N chunks one box apart.
      auto mma_tmem_y = [&](uint32_t d, uint32_t a, uint32_t y, bool first) {
#pragma unroll
        for (int k = 0; k < kBwdT / 16; ++k) {
          const uint32_t acol = k < 4 ? k * 8 : 64 + (k - 4) * 8;
          umma_bf16_ts(d, a + acol, make_sw128_desc(y + k * 2048, kBoxBytes, 1024), idesc_ts,
                       (first && k == 0) ? 0u : 1u);
        }
      };
      auto ya_addr = [&](int it) { return aYa + (it % 3) * kTileBytes; };
      auto yb_addr = [&](int it) { return aYb + (it & 1) * kTileBytes; };
      auto wait_ya = [&](int it) { mbar_wait(&ya_full[it % 3], (it / 3) & 1); };
      auto wait_yb = [&](int it) { mbar_wait(&yb_full[it & 1], (it >> 1) & 1); };
      mbar_wait(x_full, 0);
      wait_ya(0);
      tc_fence_after();
      mma_xyT(tm_A, aXa, ya_addr(0));
      umma_commit(&s_full[0]);
      wait_yb(0);
      tc_fence_after();
      mma_xyT(tm_B, aXb, yb_addr(0));
      umma_commit(dp_full);
      if (kIsKV) {
        for (int it = 0; it < n_iter; ++it) {
          mbar_wait(p_ready, it & 1);
          tc_fence_after();
          mma_tmem_y(tm_acc1, tm_A, yb_addr(it), it == 0);   // dV += P^T dO
          umma_commit(&yb_empty[it & 1]);                     // dO_i: last use
          if (it + 1 < n_iter) {
            wait_ya(it + 1);
            tc_fence_after();
            mma_xyT(tm_A, aXa, ya_addr(it + 1));
            umma_commit(&s_full[0]);
          }
          mbar_wait(ds_ready, it & 1);
          tc_fence_after();
          mma_tmem_y(tm_acc2, tm_B, ya_addr(it), it == 0);   // dK += dS^T Q
          umma_commit(&ya_empty[it % 3]);                     // Q_i: last use
          if (it + 1 < n_iter) {
            wait_yb(it + 1);
            tc_fence_after();
            mma_xyT(tm_B, aXb, yb_addr(it + 1));
            umma_commit(dp_full);
          }
        }
      } else {
        umma_commit(&yb_empty[0]);                            // V_0 is only read by dP(0)
        if (n_iter > 1) {
          wait_ya(1);
          tc_fence_after();
          mma_xyT(tm_A + 128, aXa, ya_addr(1));
          umma_commit(&s_full[1]);
        }
        for (int it = 0; it < n_iter; ++it) {
          mbar_wait(ds_ready, it & 1);
          tc_fence_after();
          mma_tmem_y(tm_acc2, tm_B, ya_addr(it), it == 0);   // dQ += dS K
          umma_commit(&ya_empty[it % 3]);                     // K_j: last use
          if (it + 1 < n_iter) {
            wait_yb(it + 1);
            tc_fence_after();
            mma_xyT(tm_B, aXb, yb_addr(it + 1));
            umma_commit(dp_full);
            umma_commit(&yb_empty[(it + 1) & 1]);             // V_{j+1}: only read by dP
          }
          if (it + 2 < n_iter) {
            wait_ya(it + 2);
            tc_fence_after();
            mma_xyT(tm_A + (it & 1) * 128, aXa, ya_addr(it + 2));
            umma_commit(&s_full[it & 1]);
          }
        }
      }
      umma_commit(acc_done);
    }
  } else if (warp >= 4) {
    // ------------------------------ math warps --------------------------------
    const int quarter = warp & 3;         // TMEM lane quarter this warp may touch
    const int half = (warp - 4) >> 2;     // which 64 of the 128 columns
    const int r = quarter * 32 + lane;    // row inside the tile == TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const int xrow = xblk * kBwdT + r;    // kv index (KV pass) or query index (Q pass)
    float my_lse2 = 0.f, my_d = 0.f;
    if (!kIsKV) {
      const size_t voff = ((size_t)b * np + head) * s_pad + xrow;
      my_lse2 = lse2[voff];
      my_d = dvec[voff];
    }
    for (int it = 0; it < n_iter; ++it) {
      const int st = it & 1;
      int yb;
      if (kIsKV) {
        const int jj = it / y_count;
        yb = y_first + it - jj * y_count;
        mbar_wait(&yb_full[st], (it >> 1) & 1);  // LSE / D vectors travel with the dO stage
      } else {
        yb = y_first + it;
      }
      const float* vec = reinterpret_cast<const float*>(smem + BwdSmem::kVec + st * 1024);
      const int ycol0 = yb * kBwdT + half * 64;
      // does this tile need the causal test at all?
      bool need_mask;
      if (kIsKV) need_mask = causal && (xblk * kBwdT + kBwdT - 1 > yb * kBwdT + shift);
      else need_mask = (causal && (yb * kBwdT + kBwdT - 1 > xblk * kBwdT + shift)) ||
                       (yb * kBwdT + kBwdT > sk);
      const uint32_t tS = kIsKV ? tm_A : tm_A + st * 128;
      mbar_wait(&s_full[kIsKV ? 0 : st], kIsKV ? (it & 1) : ((it >> 1) & 1));
      tc_fence_after();
      float P[64];
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t sr[32];
        tmem_ld_32x32(tS + lane_off + half * 64 + ch * 32, sr);
        tmem_ld_wait();
#pragma unroll
        for (int t = 0; t < 32; ++t) {
          const int c = half * 64 + ch * 32 + t;
          const float l2 = kIsKV ? vec[c] : my_lse2;
          float p = bwd_ex2(fmaf(__uint_as_float(sr[t]), scale_log2, -l2));
          if (need_mask) {
            const int ycol = ycol0 + ch * 32 + t;
            bool vis;
            if (kIsKV) vis = xrow <= ycol + shift;                 // kv <= q + shift
            else vis = (!causal || ycol <= xrow + shift) && ycol < sk;
            p = vis ? p : 0.f;
          }
          P[ch * 32 + t] = p;
        }
        if (kIsKV) {
          uint32_t pk[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) pk[t] = pack_bf16x2(P[ch * 32 + 2 * t], P[ch * 32 + 2 * t + 1]);
          tmem_st_32x16(tm_A + lane_off + half * 64 + ch * 16, pk);
        }
      }
      if (kIsKV) {
        tmem_st_wait_b();
        tc_fence_before();
        mbar_arrive(p_ready);
      }
      mbar_wait(dp_full, it & 1);
      tc_fence_after();
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t dr[32];
        tmem_ld_32x32(tm_B + lane_off + half * 64 + ch * 32, dr);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) {
          const int c = half * 64 + ch * 32 + 2 * t;
          const float d0 = kIsKV ? vec[128 + c] : my_d;
          const float d1 = kIsKV ? vec[128 + c + 1] : my_d;
          const float a = P[ch * 32 + 2 * t] * (__uint_as_float(dr[2 * t]) - d0);
          const float bb = P[ch * 32 + 2 * t + 1] * (__uint_as_float(dr[2 * t + 1]) - d1);
          pk[t] = pack_bf16x2(a, bb);
        }
        tmem_st_32x16(tm_B + lane_off + half * 64 + ch * 16, pk);
      }
      tmem_st_wait_b();
      tc_fence_before();
      mbar_arrive(ds_ready);
      if (kIsKV) mbar_arrive(&vec_empty[st]);      // this stage's LSE / D vectors may be overwritten
    }
    // ------------------------------ epilogue ----------------------------------
    if (n_iter > 0) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
    }
    const bool row_ok = xrow < (kIsKV ? sk : sq);
    const int jq = head - g * qpg;
    // KV pass: g0 = dK (scaled), g1 = dV; Q pass: g0 = dQ (scaled)
    const size_t rowoff = (size_t)xrow * o_ss + (size_t)b * o_sb + (size_t)g * o_sg +
                          (kIsKV ? 0 : (size_t)jq * o_sj) + half * 64;
    const bool split = kIsKV && nsplit > 1;
    // fp32 partial sums: kv_acc[which][sk][B][G][128]
    const size_t accoff = (((size_t)xrow * B + b) * G + g) * kBwdT + half * 64;
    const size_t acc_plane = (size_t)sk * B * G * kBwdT;
#pragma unroll 1
    for (int which = 0; which < (kIsKV ? 2 : 1); ++which) {
      const uint32_t acc = which == 0 ? tm_acc2 : tm_acc1;
      __nv_bfloat16* dst = (which == 0 ? g0 : g1) + rowoff;
      const float mul = which == 0 ? scale : 1.f;
#pragma unroll 1
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t a[32];
        if (n_iter > 0) {
          tmem_ld_32x32(acc + lane_off + half * 64 + ch * 32, a);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int t = 0; t < 32; ++t) a[t] = 0u;
        }
        if (!row_ok) continue;
        if (split) {
          if (n_iter > 0) {
            float* ap = kv_acc + which * acc_plane + accoff + ch * 32;
#pragma unroll
            for (int t = 0; t < 32; t += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ap + t),
                           "f"(__uint_as_float(a[t]) * mul), "f"(__uint_as_float(a[t + 1]) * mul),
                           "f"(__uint_as_float(a[t + 2]) * mul), "f"(__uint_as_float(a[t + 3]) * mul)
                           : "memory");
          }
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(a[8 * q + 0]) * mul, __uint_as_float(a[8 * q + 1]) * mul);
            w.y = pack_bf16x2(__uint_as_float(a[8 * q + 2]) * mul, __uint_as_float(a[8 * q + 3]) * mul);
            w.z = pack_bf16x2(__uint_as_float(a[8 * q + 4]) * mul, __uint_as_float(a[8 * q + 5]) * mul);
            w.w = pack_bf16x2(__uint_as_float(a[8 * q + 6]) * mul, __uint_as_float(a[8 * q + 7]) * mul);
            reinterpret_cast<uint4*>(dst + ch * 32)[q] = w;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}


// ---------------------------------------------------------------------------------
// Second generation of the two passes: the streamed operand comes in 64-row blocks, so the
// score tile of one step is 128 x 64 and S / dP each fit TWICE in tensor memory next to the
// two 128-column accumulators (2 x 64 + 2 x 64 + 128 + 128 = 512 columns). The MMA thread
// issues S(i+1) and dP(i+1) into the other pair of buffers BEFORE it waits for the math warps
// of step i, so the exponentials of step i run under the tensor-core work of step i+1 and the
// P -> dV -> next S chain of the first generation (one buffer, everything serial) is gone:
//
//   MMA thread:   S(i+1) dP(i+1) | wait P(i)  | dV += P(i)^T dO | wait dS(i) | dK += dS(i)^T Q
//   math warps:   P(i) = exp2(S(i)..) ........ | dS(i) = P o (dP(i) - D) ...  | P(i+1) ...
//
// Streamed blocks travel through one 4-stage ring (Ya block, Yb block, LSE / D of the block).
// Same inputs, outputs, scratch and grid as attn_bwd_kernel.
// ---------------------------------------------------------------------------------
constexpr int kB2Y = 64;                       // rows of a streamed block
constexpr int kB2YBox = kB2Y * 128;            // 8 KB: [64 rows x 128 B]
constexpr int kB2YTile = 2 * kB2YBox;          // 16 KB: 64 rows x 128 head dims
constexpr int kB2Stages = 4;
struct Bwd2Smem {
  static constexpr int kXa = 0;
  static constexpr int kXb = kXa + kTileBytes;
  static constexpr int kRing = kXb + kTileBytes;
  static constexpr int kStage = 2 * kB2YTile + 1024;   // Ya, Yb, (lse2[64], D[64]) fp32 + pad
  static constexpr int kBar = kRing + kB2Stages * kStage;
  static constexpr int kTotal = kBar + 256;
};
static_assert(Bwd2Smem::kTotal <= 232448, "shared memory budget");
static_assert(Bwd2Smem::kStage % 1024 == 0, "swizzled tiles need 1024-byte alignment");

template <bool kIsKV>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                 const float* __restrict__ lse2, const float* __restrict__ dvec,
                 __nv_bfloat16* __restrict__ g0, __nv_bfloat16* __restrict__ g1, long long o_ss,
                 long long o_sb, long long o_sg, long long o_sj, int sq, int sk, int s_pad, int B,
                 int G, int qpg, float scale, int causal, int nsplit, float* __restrict__ kv_acc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u)) asm volatile("trap;");
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Bwd2Smem::kBar);
  uint64_t* x_full = bars + 0;
  uint64_t* y_full = bars + 1;     // [4]
  uint64_t* y_empty = bars + 5;    // [4]
  uint64_t* s_full = bars + 9;     // [2]
  uint64_t* dp_full = bars + 11;   // [2]
  uint64_t* p_ready = bars + 13;   // [2]
  uint64_t* ds_ready = bars + 15;  // [2]
  uint64_t* acc_done = bars + 17;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int np = G * qpg;
  const int nq = (sq + kBwdT - 1) / kBwdT;
  const int nq64 = (sq + kB2Y - 1) / kB2Y, nkv64 = (sk + kB2Y - 1) / kB2Y;
  const int shift = sk - sq;
  const int b = blockIdx.z;
  const int jj0 = kIsKV ? (blockIdx.y % nsplit) * (qpg / nsplit) : 0;
  const float scale_log2 = scale * 1.4426950408889634f;

  int g, head = 0, xblk, y_first, y_count, n_iter;
  if (kIsKV) {
    xblk = blockIdx.x;
    g = blockIdx.y / nsplit;
    y_first = causal ? max(0, (xblk * kBwdT - shift) / kB2Y) : 0;
    y_count = max(0, nq64 - y_first);
    n_iter = y_count * (qpg / nsplit);
  } else {
    xblk = nq - 1 - blockIdx.x;
    head = blockIdx.y;
    g = head / qpg;
    y_first = 0;
    y_count = causal ? min(nkv64, (xblk * kBwdT + kBwdT - 1 + shift) / kB2Y + 1) : nkv64;
    n_iter = y_count;
  }

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmdO);
    mbar_init(x_full, 1);
    for (int i = 0; i < kB2Stages; ++i) {
      mbar_init(&y_full[i], 1);
      mbar_init(&y_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&dp_full[i], 1);
      mbar_init(&p_ready[i], 256);
      mbar_init(&ds_ready[i], 256);
    }
    mbar_init(acc_done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // TMEM columns: S buffers at 0 / 64, dP buffers at 128 / 192, accumulators at 256 (dV) and 384.
  const uint32_t tm_acc1 = tmem_base + 256;
  const uint32_t tm_acc2 = tmem_base + 384;

  uint8_t* sXa = smem + Bwd2Smem::kXa;
  uint8_t* sXb = smem + Bwd2Smem::kXb;
  uint8_t* ring = smem + Bwd2Smem::kRing;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0 && n_iter > 0) {
      mbar_expect_tx(x_full, 2 * kTileBytes);
      // a 128-row stationary tile = per 64-column box two 64-row TMA boxes, one below the other
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int off = c * kBoxBytes + h * kB2YBox, row = xblk * kBwdT + h * kB2Y;
          if (kIsKV) {
            tma_load_4d(sXa + off, &tmK, x_full, c * 64, row, g, b);
            tma_load_4d(sXb + off, &tmV, x_full, c * 64, row, g, b);
          } else {
            tma_load_5d_b(sXa + off, &tmQ, x_full, c * 64, row, head - g * qpg, g, b);
            tma_load_4d(sXb + off, &tmdO, x_full, c * 64, row, head, b);
          }
        }
      for (int it = 0; it < n_iter; ++it) {
        const int st = it % kB2Stages;
        const uint32_t ph = ((it / kB2Stages) & 1) ^ 1;
        uint8_t* sYa = ring + st * Bwd2Smem::kStage;
        uint8_t* sYb = sYa + kB2YTile;
        mbar_wait(&y_empty[st], ph);
        if (kIsKV) {
          const int jl = it / y_count, yb = y_first + it - jl * y_count;
          const int jj = jj0 + jl;
          const int hd = g * qpg + jj;
          float* vec = reinterpret_cast<float*>(sYb + kB2YTile);
          const size_t voff = ((size_t)b * np + hd) * s_pad + (size_t)yb * kB2Y;
          mbar_expect_tx(&y_full[st], 2 * kB2YTile + 512);
          tma_load_5d_b(sYa, &tmQ, &y_full[st], 0, yb * kB2Y, jj, g, b);
          tma_load_5d_b(sYa + kB2YBox, &tmQ, &y_full[st], 64, yb * kB2Y, jj, g, b);
          tma_load_4d(sYb, &tmdO, &y_full[st], 0, yb * kB2Y, hd, b);
          tma_load_4d(sYb + kB2YBox, &tmdO, &y_full[st], 64, yb * kB2Y, hd, b);
          bulk_load_1d(vec, lse2 + voff, 256, &y_full[st]);
          bulk_load_1d(vec + kB2Y, dvec + voff, 256, &y_full[st]);
        } else {
          const int yb = y_first + it;
          mbar_expect_tx(&y_full[st], 2 * kB2YTile);
          tma_load_4d(sYa, &tmK, &y_full[st], 0, yb * kB2Y, g, b);
          tma_load_4d(sYa + kB2YBox, &tmK, &y_full[st], 64, yb * kB2Y, g, b);
          tma_load_4d(sYb, &tmV, &y_full[st], 0, yb * kB2Y, g, b);
          tma_load_4d(sYb + kB2YBox, &tmV, &y_full[st], 64, yb * kB2Y, g, b);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    // The whole warp walks the loop in uniform control flow (descriptors are warp-uniform values
    // computed once per operand and advanced by constants); only the tcgen05 instructions are
    // predicated on the elected lane. With 128 x 64 x 16 MMAs (32 tensor-pipe cycles each) the
    // issue path must stay at a handful of instructions per MMA.
    if (n_iter > 0) {
      const bool leader = elect_one_sync();
      constexpr uint32_t idesc_ss = make_idesc_bf16(kBwdT, kB2Y, 0, 0);    // [128 x 64] += X Y^T
      constexpr uint32_t idesc_ts = make_idesc_bf16(kBwdT, kBwdT, 0, 1);   // [128 x 128] += A(tmem) Y
      const uint64_t dXa = make_sw128_desc(smem_u32(sXa), 16, 1024);
      const uint64_t dXb = make_sw128_desc(smem_u32(sXb), 16, 1024);
      const uint32_t aRing = smem_u32(ring);
      // D[128 x 64] = X[128 x 128] * Y[64 x 128]^T, both K-major, head dim in two 64-wide boxes
      auto mma_xyT = [&](uint32_t d, uint64_t dx, uint64_t dy) {
#pragma unroll
        for (int k = 0; k < kBwdT / 16; ++k) {
          const uint64_t ox = static_cast<uint64_t>(((k >> 2) * kBoxBytes + (k & 3) * 32) >> 4);
          const uint64_t oy = static_cast<uint64_t>(((k >> 2) * kB2YBox + (k & 3) * 32) >> 4);
          if (leader) umma_bf16_ss(d, dx + ox, dy + oy, idesc_ss, k != 0);
        }
      };
      // D[128 x 128] += A(tmem: 64 bf16 per lane, 16 packed columns per 32-column half) * Y[64 x 128]
      auto mma_tmem_y = [&](uint32_t d, uint32_t a, uint64_t dy_mn, bool first) {
#pragma unroll
        for (int k = 0; k < kB2Y / 16; ++k) {
          const uint32_t acol = (k >> 1) * 32 + (k & 1) * 8;
          if (leader)
            umma_bf16_ts(d, a + acol, dy_mn + static_cast<uint64_t>((k * 2048) >> 4), idesc_ts,
                         (first && k == 0) ? 0u : 1u);
        }
      };
      auto ya_addr = [&](int it) { return aRing + (it % kB2Stages) * Bwd2Smem::kStage; };
      auto issue_scores = [&](int it) {      // S(it), dP(it) into buffer pair it & 1
        mbar_wait(&y_full[it % kB2Stages], (it / kB2Stages) & 1);
        tc_fence_after();
        const int bf = it & 1;
        const uint32_t ya = ya_addr(it);
        mma_xyT(tmem_base + bf * kB2Y, dXa, make_sw128_desc(ya, 16, 1024));
        if (leader) umma_commit(&s_full[bf]);
        mma_xyT(tmem_base + 128 + bf * kB2Y, dXb, make_sw128_desc(ya + kB2YTile, 16, 1024));
        if (leader) umma_commit(&dp_full[bf]);
      };
      mbar_wait(x_full, 0);
      issue_scores(0);
      for (int it = 0; it < n_iter; ++it) {
        const int bf = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        if (it + 1 < n_iter) issue_scores(it + 1);
        const uint32_t ya = ya_addr(it);
        if (kIsKV) {
          mbar_wait(&p_ready[bf], ph);
          tc_fence_after();
          mma_tmem_y(tm_acc1, tmem_base + bf * kB2Y, make_sw128_desc(ya + kB2YTile, kB2YBox, 1024),
                     it == 0);                                                       // dV += P^T dO
        }
        mbar_wait(&ds_ready[bf], ph);
        tc_fence_after();
        mma_tmem_y(tm_acc2, tmem_base + 128 + bf * kB2Y, make_sw128_desc(ya, kB2YBox, 1024),
                   it == 0);                                     // dK += dS^T Q / dQ += dS K
        if (leader) umma_commit(&y_empty[it % kB2Stages]);
      }
      if (leader) umma_commit(acc_done);
    }
  } else if (warp >= 4) {
    // ------------------------------ math warps --------------------------------
    const int quarter = warp & 3;         // TMEM lane quarter this warp may touch
    const int half = (warp - 4) >> 2;     // which 32 of the block's 64 columns (64 of 128 in the epilogue)
    const int r = quarter * 32 + lane;    // row inside the tile == TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const int xrow = xblk * kBwdT + r;    // kv index (KV pass) or query index (Q pass)
    float my_lse2 = 0.f, my_d = 0.f;
    if (!kIsKV) {
      const size_t voff = ((size_t)b * np + head) * s_pad + xrow;
      my_lse2 = lse2[voff];
      my_d = dvec[voff];
    }
    for (int it = 0; it < n_iter; ++it) {
      const int bf = it & 1;
      const uint32_t ph = (it >> 1) & 1;
      int yb;
      if (kIsKV) {
        const int jj = it / y_count;
        yb = y_first + it - jj * y_count;
        mbar_wait(&y_full[it % kB2Stages], (it / kB2Stages) & 1);   // LSE / D of the block
      } else {
        yb = y_first + it;
      }
      const float* vec = reinterpret_cast<const float*>(ring + (it % kB2Stages) * Bwd2Smem::kStage +
                                                        2 * kB2YTile) + half * 32;
      const int ycol0 = yb * kB2Y + half * 32;
      bool need_mask;
      if (kIsKV) need_mask = causal && (xblk * kBwdT + kBwdT - 1 > yb * kB2Y + shift);
      else need_mask = (causal && (yb * kB2Y + kB2Y - 1 > xblk * kBwdT + shift)) ||
                       (yb * kB2Y + kB2Y > sk);
      const uint32_t tS = tmem_base + bf * kB2Y + lane_off + half * 32;
      const uint32_t tD = tmem_base + 128 + bf * kB2Y + lane_off + half * 32;
      mbar_wait(&s_full[bf], ph);
      tc_fence_after();
      float P[32];
      {
        uint32_t sr[32];
        tmem_ld_32x32(tS, sr);
        tmem_ld_wait();
#pragma unroll
        for (int t4 = 0; t4 < 8; ++t4) {
          float4 l4 = make_float4(my_lse2, my_lse2, my_lse2, my_lse2);
          if (kIsKV) l4 = reinterpret_cast<const float4*>(vec)[t4];
          const float l[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int t = t4 * 4 + u;
            float p = bwd_ex2(fmaf(__uint_as_float(sr[t]), scale_log2, -l[u]));
            if (need_mask) {
              const int ycol = ycol0 + t;
              bool vis;
              if (kIsKV) vis = xrow <= ycol + shift;
              else vis = (!causal || ycol <= xrow + shift) && ycol < sk;
              p = vis ? p : 0.f;
            }
            P[t] = p;
          }
        }
        if (kIsKV) {
          uint32_t pk[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) pk[t] = pack_bf16x2(P[2 * t], P[2 * t + 1]);
          tmem_st_32x16(tS, pk);
          tmem_st_wait_b();
          tc_fence_before();
          mbar_arrive(&p_ready[bf]);
        }
      }
      mbar_wait(&dp_full[bf], ph);
      tc_fence_after();
      {
        uint32_t dr[32];
        tmem_ld_32x32(tD, dr);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int t4 = 0; t4 < 8; ++t4) {
          float4 d4 = make_float4(my_d, my_d, my_d, my_d);
          if (kIsKV) d4 = reinterpret_cast<const float4*>(vec + kB2Y)[t4];   // D follows the 64 LSE values
          const float a0 = P[4 * t4 + 0] * (__uint_as_float(dr[4 * t4 + 0]) - d4.x);
          const float a1 = P[4 * t4 + 1] * (__uint_as_float(dr[4 * t4 + 1]) - d4.y);
          const float a2 = P[4 * t4 + 2] * (__uint_as_float(dr[4 * t4 + 2]) - d4.z);
          const float a3 = P[4 * t4 + 3] * (__uint_as_float(dr[4 * t4 + 3]) - d4.w);
          pk[2 * t4] = pack_bf16x2(a0, a1);
          pk[2 * t4 + 1] = pack_bf16x2(a2, a3);
        }
        tmem_st_32x16(tD, pk);
        tmem_st_wait_b();
        tc_fence_before();
        mbar_arrive(&ds_ready[bf]);
      }
    }
    // ------------------------------ epilogue (as in the first generation) -----
    if (n_iter > 0) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
    }
    const bool row_ok = xrow < (kIsKV ? sk : sq);
    const int jq = head - g * qpg;
    const size_t rowoff = (size_t)xrow * o_ss + (size_t)b * o_sb + (size_t)g * o_sg +
                          (kIsKV ? 0 : (size_t)jq * o_sj) + half * 64;
    const bool split = kIsKV && nsplit > 1;
    const size_t accoff = (((size_t)xrow * B + b) * G + g) * kBwdT + half * 64;
    const size_t acc_plane = (size_t)sk * B * G * kBwdT;
#pragma unroll 1
    for (int which = 0; which < (kIsKV ? 2 : 1); ++which) {
      const uint32_t acc = which == 0 ? tm_acc2 : tm_acc1;
      __nv_bfloat16* dst = (which == 0 ? g0 : g1) + rowoff;
      const float mul = which == 0 ? scale : 1.f;
#pragma unroll 1
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t a[32];
        if (n_iter > 0) {
          tmem_ld_32x32(acc + lane_off + half * 64 + ch * 32, a);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int t = 0; t < 32; ++t) a[t] = 0u;
        }
        if (!row_ok) continue;
        if (split) {
          if (n_iter > 0) {
            float* ap = kv_acc + which * acc_plane + accoff + ch * 32;
#pragma unroll
            for (int t = 0; t < 32; t += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ap + t),
                           "f"(__uint_as_float(a[t]) * mul), "f"(__uint_as_float(a[t + 1]) * mul),
                           "f"(__uint_as_float(a[t + 2]) * mul), "f"(__uint_as_float(a[t + 3]) * mul)
                           : "memory");
          }
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(a[8 * q + 0]) * mul, __uint_as_float(a[8 * q + 1]) * mul);
            w.y = pack_bf16x2(__uint_as_float(a[8 * q + 2]) * mul, __uint_as_float(a[8 * q + 3]) * mul);
            w.z = pack_bf16x2(__uint_as_float(a[8 * q + 4]) * mul, __uint_as_float(a[8 * q + 5]) * mul);
            w.w = pack_bf16x2(__uint_as_float(a[8 * q + 6]) * mul, __uint_as_float(a[8 * q + 7]) * mul);
            reinterpret_cast<uint4*>(dst + ch * 32)[q] = w;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace b200

// ------------------------------- host side -----------------------------------
namespace b200 {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_fn();

static int make_tmap_bwd(CUtensorMap* map, const void* base, int rank, const long long* dims,
                         const long long* strides, int box_rows = kBwdT) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_UNSUPPORTED;
  cuuint64_t d[5];
  cuuint64_t s[4];
  cuuint32_t box[5], es[5];
  for (int i = 0; i < rank; ++i) {
    d[i] = (cuuint64_t)dims[i];
    box[i] = 1;
    es[i] = 1;
    if (i > 0) {
      if (strides[i - 1] % 8) return B200_ERR_UNSUPPORTED;
      s[i - 1] = (cuuint64_t)strides[i - 1] * 2;
    }
  }
  box[0] = 64;
  box[1] = box_rows;
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                  d, s, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : B200_ERR_BAD_ARG;
}
}  // namespace b200

using namespace b200;

// How many CTAs share the query heads of one (kv block, kv group) in the KV pass: enough to
// fill the machine about twice when there are few KV heads (MQA / GQA under tensor parallelism).
static int kv_pass_split(int sk, int B, int G, int qpg) {
  const int nkv = (sk + kBwdT - 1) / kBwdT;
  const long long ctas = (long long)nkv * G * B;
  int want = (int)((2 * kNumSMs + ctas - 1) / ctas);
  if (want <= 1) return 1;
  int best = 1;
  for (int d = 1; d <= qpg; ++d)
    if (qpg % d == 0 && d <= want) best = d;
  if (best < want)  // next divisor above `want`, if any
    for (int d = best + 1; d <= qpg; ++d)
      if (qpg % d == 0) { best = d; break; }
  return best;
}

// fp32 elements of scratch b200_attn_bwd needs: LSE*log2e and D, each [B, np, ceil128(sq)],
// plus the partial dK / dV sums [2, sk, B, G, 128] when the KV pass is split over query heads.
B200_API long long b200_attn_bwd_scratch_elems(int sq, int sk, int B, int G, int qpg) {
  const long long s_pad = (sq + kBwdT - 1) / kBwdT * kBwdT;
  long long n = 2 * s_pad * B * G * qpg;
  if (kv_pass_split(sk, B, G, qpg) > 1) n += 2LL * sk * B * G * kBwdT;
  return n;
}

// q/k/v as in b200_attn_fwd. out, dout: [sq, B, G*qpg*128] bf16 contiguous; lse: [B, np, sq].
// dq: element (s,b,g,j,d) at dq + s*dq_ss + b*dq_sb + g*dq_sg + j*dq_sj + d;
// dk/dv: element (s,b,g,d) at dk + s*dk_ss + b*dk_sb + g*dk_sg + d.
B200_API int b200_attn_bwd(const void* q, const void* k, const void* v, const void* out,
                           const void* dout, const float* lse, void* dq, void* dk, void* dv,
                           float* scratch, int sq, int sk, int B, int G, int qpg, int hn,
                           long long q_ss, long long q_sb, long long q_sg, long long q_sj,
                           long long k_ss, long long k_sb, long long k_sg, long long dq_ss,
                           long long dq_sb, long long dq_sg, long long dq_sj, long long dk_ss,
                           long long dk_sb, long long dk_sg, float softmax_scale, int causal,
                           cudaStream_t stream) {
  if (hn != kBwdT) return B200_ERR_UNSUPPORTED;
  if (sq <= 0 || sk <= 0 || B <= 0 || G <= 0 || qpg <= 0 || !scratch) return B200_ERR_BAD_ARG;
  if (causal && sk < sq) return B200_ERR_UNSUPPORTED;
  if ((dq_ss | dq_sb | dq_sg | dq_sj | dk_ss | dk_sb | dk_sg) % 8) return B200_ERR_UNSUPPORTED;
  const int np = G * qpg;
  const int s_pad = (sq + kBwdT - 1) / kBwdT * kBwdT;
  float* lse2 = scratch;
  float* dvec = scratch + (size_t)s_pad * B * np;
  float* kv_acc = scratch + 2 * (size_t)s_pad * B * np;
  const int nsplit = kv_pass_split(sk, B, G, qpg);
  CUtensorMap tq, tk, tv, tdo;
  long long qd[5] = {hn, sq, qpg, G, B}, qs[4] = {q_ss, q_sj, q_sg, q_sb};
  long long kd[4] = {hn, sk, G, B}, ks[3] = {k_ss, k_sg, k_sb};
  long long od[4] = {hn, sq, np, B}, os[3] = {(long long)B * np * hn, hn, (long long)np * hn};
  int rc = make_tmap_bwd(&tq, q, 5, qd, qs);
  if (rc) return rc;
  rc = make_tmap_bwd(&tk, k, 4, kd, ks);
  if (rc) return rc;
  rc = make_tmap_bwd(&tv, v, 4, kd, ks);
  if (rc) return rc;
  rc = make_tmap_bwd(&tdo, dout, 4, od, os);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, BwdSmem::kTotal);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(attn_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             BwdSmem::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  {
    const int rows = B * np * s_pad;
    attn_bwd_prep_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(
        (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, lse, lse2, dvec, sq, s_pad, B, np);
    B200_CHECK_LAUNCH();
  }
  const int nq = (sq + kBwdT - 1) / kBwdT, nkv = (sk + kBwdT - 1) / kBwdT;
  if (nsplit > 1) {
    cudaError_t e = cudaMemsetAsync(kv_acc, 0, 2 * (size_t)sk * B * G * kBwdT * sizeof(float), stream);
    if (e != cudaSuccess) return (int)e;
  }
  attn_bwd_kernel<true><<<dim3(nkv, G * nsplit, B), kBwdThreads, BwdSmem::kTotal, stream>>>(
      tq, tk, tv, tdo, lse2, dvec, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, dk_ss, dk_sb, dk_sg, 0,
      sq, sk, s_pad, B, G, qpg, softmax_scale, causal, nsplit, kv_acc);
  B200_CHECK_LAUNCH();
  if (nsplit > 1) {
    const long long n8 = 2LL * sk * B * G * (kBwdT / 8);
    attn_bwd_kv_convert_kernel<<<(unsigned)((n8 + 255) / 256), 256, 0, stream>>>(
        kv_acc, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, dk_ss, dk_sb, dk_sg, sk, B, G);
    B200_CHECK_LAUNCH();
  }
  attn_bwd_kernel<false><<<dim3(nq, np, B), kBwdThreads, BwdSmem::kTotal, stream>>>(
      tq, tk, tv, tdo, lse2, dvec, (__nv_bfloat16*)dq, nullptr, dq_ss, dq_sb, dq_sg, dq_sj, sq, sk,
      s_pad, B, G, qpg, softmax_scale, causal, 1, nullptr);
  B200_CHECK_LAUNCH();
  return 0;
}

// Second-generation kernels (attn_bwd2_kernel): same contract, scratch and launch geometry.
B200_API int b200_attn_bwd_v2(const void* q, const void* k, const void* v, const void* out,
                              const void* dout, const float* lse, void* dq, void* dk, void* dv,
                              float* scratch, int sq, int sk, int B, int G, int qpg, int hn,
                              long long q_ss, long long q_sb, long long q_sg, long long q_sj,
                              long long k_ss, long long k_sb, long long k_sg, long long dq_ss,
                              long long dq_sb, long long dq_sg, long long dq_sj, long long dk_ss,
                              long long dk_sb, long long dk_sg, float softmax_scale, int causal,
                              cudaStream_t stream) {
  if (hn != kBwdT) return B200_ERR_UNSUPPORTED;
  if (sq <= 0 || sk <= 0 || B <= 0 || G <= 0 || qpg <= 0 || !scratch) return B200_ERR_BAD_ARG;
  if (causal && sk < sq) return B200_ERR_UNSUPPORTED;
  if ((dq_ss | dq_sb | dq_sg | dq_sj | dk_ss | dk_sb | dk_sg) % 8) return B200_ERR_UNSUPPORTED;
  const int np = G * qpg;
  const int s_pad = (sq + kBwdT - 1) / kBwdT * kBwdT;
  float* lse2 = scratch;
  float* dvec = scratch + (size_t)s_pad * B * np;
  float* kv_acc = scratch + 2 * (size_t)s_pad * B * np;
  const int nsplit = kv_pass_split(sk, B, G, qpg);
  CUtensorMap tq, tk, tv, tdo;
  long long qd[5] = {hn, sq, qpg, G, B}, qs[4] = {q_ss, q_sj, q_sg, q_sb};
  long long kd[4] = {hn, sk, G, B}, ks[3] = {k_ss, k_sg, k_sb};
  long long od[4] = {hn, sq, np, B}, os[3] = {(long long)B * np * hn, hn, (long long)np * hn};
  int rc = make_tmap_bwd(&tq, q, 5, qd, qs, kB2Y);
  if (rc) return rc;
  rc = make_tmap_bwd(&tk, k, 4, kd, ks, kB2Y);
  if (rc) return rc;
  rc = make_tmap_bwd(&tv, v, 4, kd, ks, kB2Y);
  if (rc) return rc;
  rc = make_tmap_bwd(&tdo, dout, 4, od, os, kB2Y);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd2_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, Bwd2Smem::kTotal);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(attn_bwd2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Bwd2Smem::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  {
    const int rows = B * np * s_pad;
    attn_bwd_prep_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(
        (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, lse, lse2, dvec, sq, s_pad, B, np);
    B200_CHECK_LAUNCH();
  }
  const int nq = (sq + kBwdT - 1) / kBwdT, nkv = (sk + kBwdT - 1) / kBwdT;
  if (nsplit > 1) {
    cudaError_t e = cudaMemsetAsync(kv_acc, 0, 2 * (size_t)sk * B * G * kBwdT * sizeof(float), stream);
    if (e != cudaSuccess) return (int)e;
  }
  attn_bwd2_kernel<true><<<dim3(nkv, G * nsplit, B), kBwdThreads, Bwd2Smem::kTotal, stream>>>(
      tq, tk, tv, tdo, lse2, dvec, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, dk_ss, dk_sb, dk_sg, 0,
      sq, sk, s_pad, B, G, qpg, softmax_scale, causal, nsplit, kv_acc);
  B200_CHECK_LAUNCH();
  if (nsplit > 1) {
    const long long n8 = 2LL * sk * B * G * (kBwdT / 8);
    attn_bwd_kv_convert_kernel<<<(unsigned)((n8 + 255) / 256), 256, 0, stream>>>(
        kv_acc, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, dk_ss, dk_sb, dk_sg, sk, B, G);
    B200_CHECK_LAUNCH();
  }
  attn_bwd2_kernel<false><<<dim3(nq, np, B), kBwdThreads, Bwd2Smem::kTotal, stream>>>(
      tq, tk, tv, tdo, lse2, dvec, (__nv_bfloat16*)dq, nullptr, dq_ss, dq_sb, dq_sg, dq_sj, sq, sk,
      s_pad, B, G, qpg, softmax_scale, causal, 1, nullptr);
  B200_CHECK_LAUNCH();
  return 0;
}
