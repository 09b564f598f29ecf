// Attention forward, second generation (K11 of SURVEY.md section 8): one CTA owns TWO 128-row
// query tiles of one head and walks the keys in blocks of 128, ping-ponging the two tiles through
// the tensor core so that the softmax of one tile runs under the MMAs of the other:
//
//   MMA thread:  S0(0) S1(0) | PV0(0) S0(1) | PV1(0) S1(1) | PV0(1) S0(2) | ...
//   softmax WG:        P0(0)      P1(0)          P0(1)          P1(1)
//
//   * S_t = Q_t K_i^T  (tcgen05.mma, M = N = K = 128, both operands from 128B-swizzled smem)
//     into its own 128 TMEM columns; the 4 softmax warps own one query row per thread, read the
//     row with tcgen05.ld, and write bf16 P back over the first 64 columns of the same region;
//   * O_t += P_t V_i reads P straight from TMEM (A operand in tensor memory), V as stored
//     (MN-major B), accumulating in 128 more TMEM columns per tile: 2 x (128 + 128) = all 512;
//   * running max with lazy rescale (O is only touched when the max grew by more than 2^8);
//   * K/V stream through a 2-stage TMA ring shared by both tiles (each K/V byte is used twice).
// Same inputs / outputs as attn_fwd_tcgen05.cu (fused-QKV strides via rank-5/4 TMA maps, O as
// [s, b, np*hn], LSE [b, np, s]); replaces flash_attn_func at megatron/model/transformer.py:514-522.
#include "common.cuh"

namespace b200 {

constexpr int kF2T = 128;                       // tile edge: query rows per tile, kv rows per block, head dim
constexpr int kF2Threads = 256;                 // warp0 TMA, warp1 MMA, warp2 TMEM, warps 4-7 softmax
constexpr int kF2Tile = kF2T * kF2T * 2;        // 32 KB = two [128 x 128 B] swizzled boxes
constexpr int kF2Box = kF2T * 128;

struct F2Smem {
  static constexpr int kQ = 0;                  // Q0, Q1
  static constexpr int kK = kQ + 2 * kF2Tile;   // K stage 0, 1
  static constexpr int kV = kK + 2 * kF2Tile;   // V stage 0, 1
  static constexpr int kBar = kV + 2 * kF2Tile;
  static constexpr int kTotal = kBar + 256 + 1024;
};

__device__ __forceinline__ void f2_tma_5d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                          int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
      "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void f2_tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]),
      "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]),
      "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]),
      "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]),
      "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void f2_tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// 2^x on the SFU: one MUFU.EX2, no range fix-ups (arguments are <= 8 here, -inf gives 0)
__device__ __forceinline__ float f2_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void f2_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// q map: rank 5 {hn, s, j, g, b};  k/v maps: rank 4 {hn, s, g, b}; boxes {64, 128, 1, ...}
__global__ void __launch_bounds__(kF2Threads, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, __nv_bfloat16* __restrict__ out,
                 float* __restrict__ lse, int sq, int sk, int B, int G, int qpg, float scale_log2,
                 int causal) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + F2Smem::kBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* k_empty = bars + 3;   // [2]
  uint64_t* v_full = bars + 5;    // [2]
  uint64_t* v_empty = bars + 7;   // [2]
  uint64_t* s_full = bars + 9;    // [2] per query tile
  uint64_t* p_ready = bars + 11;  // [2]
  uint64_t* o_done = bars + 13;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pairs_m = (sq + 2 * kF2T - 1) / (2 * kF2T);
  const int mp = pairs_m - 1 - blockIdx.x;       // heaviest tile pairs first
  const int head = blockIdx.y, b = blockIdx.z;
  const int g = head / qpg, j = head - g * qpg;
  const int m0 = mp * 2 * kF2T;
  const int shift = sk - sq;
  const int nkv = (sk + kF2T - 1) / kF2T;
  // kv blocks each query tile needs (tile 1 never needs fewer than tile 0)
  int nb[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int last_q = m0 + t * kF2T + kF2T - 1 + shift;
    nb[t] = causal ? max(1, min(nkv, last_q / kF2T + 1)) : nkv;
  }
  if (nb[1] < nb[0]) nb[1] = nb[0];
  const int n_max = nb[1];

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_ready[i], 128);
      mbar_init(&o_done[i], 1);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // S_t at t * 128, O_t at 256 + t * 128

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, 2 * kF2Tile);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        uint8_t* q = smem + F2Smem::kQ + t * kF2Tile;
        f2_tma_5d(q, &tmQ, q_full, 0, m0 + t * kF2T, j, g, b);
        f2_tma_5d(q + kF2Box, &tmQ, q_full, 64, m0 + t * kF2T, j, g, b);
      }
      for (int i = 0; i < n_max; ++i) {
        const int st = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        uint8_t* k = smem + F2Smem::kK + st * kF2Tile;
        uint8_t* v = smem + F2Smem::kV + st * kF2Tile;
        mbar_wait(&k_empty[st], ph ^ 1);
        mbar_expect_tx(&k_full[st], kF2Tile);
        tma_load_4d(k, &tmK, &k_full[st], 0, i * kF2T, g, b);
        tma_load_4d(k + kF2Box, &tmK, &k_full[st], 64, i * kF2T, g, b);
        mbar_wait(&v_empty[st], ph ^ 1);
        mbar_expect_tx(&v_full[st], kF2Tile);
        tma_load_4d(v, &tmV, &v_full[st], 0, i * kF2T, g, b);
        tma_load_4d(v + kF2Box, &tmV, &v_full[st], 64, i * kF2T, g, b);
      }
    }
  } else if (warp == 1) {
    // MMA issuer: the warp runs the loop in uniform control flow, descriptors live in uniform
    // registers and advance by constants, only the tcgen05 instructions are predicated on the
    // elected lane (a 128 x 128 x 16 MMA occupies the tensor pipe for 64 cycles: the issue path
    // has to be shorter than that).
    {
      const bool leader = elect_one_sync();
      constexpr uint32_t idesc_s = make_idesc_bf16(kF2T, kF2T, 0, 0);
      constexpr uint32_t idesc_o = make_idesc_bf16(kF2T, kF2T, 0, 1);
      const uint64_t dQ = make_sw128_desc(smem_u32(smem + F2Smem::kQ), 16, 1024);
      const uint64_t dK = make_sw128_desc(smem_u32(smem + F2Smem::kK), 16, 1024);
      const uint64_t dV = make_sw128_desc(smem_u32(smem + F2Smem::kV), kF2Box, 1024);
      auto issue_s = [&](int t, int i) {      // S_t = Q_t K_i^T
        const uint64_t q = dQ + static_cast<uint64_t>((t * kF2Tile) >> 4);
        const uint64_t k = dK + static_cast<uint64_t>(((i & 1) * kF2Tile) >> 4);
#pragma unroll
        for (int kk = 0; kk < kF2T / 16; ++kk) {
          const uint64_t off = static_cast<uint64_t>(((kk >> 2) * kF2Box + (kk & 3) * 32) >> 4);
          if (leader) umma_bf16_ss(tmem_base + t * 128, q + off, k + off, idesc_s, kk != 0);
        }
        if (leader) umma_commit(&s_full[t]);
      };
      auto issue_pv = [&](int t, int i) {     // O_t += P_t V_i, P from TMEM
        const uint64_t v = dV + static_cast<uint64_t>(((i & 1) * kF2Tile) >> 4);
#pragma unroll
        for (int kk = 0; kk < kF2T / 16; ++kk)
          if (leader)
            f2_umma_ts(tmem_base + 256 + t * 128, tmem_base + t * 128 + kk * 8,
                       v + static_cast<uint64_t>((kk * 2048) >> 4), idesc_o, (i | kk) != 0);
        if (leader) umma_commit(&o_done[t]);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      issue_s(0, 0);
      issue_s(1, 0);
      if (leader) umma_commit(&k_empty[0]);
      for (int i = 0; i < n_max; ++i) {
        const int st = i & 1;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          if (i >= nb[t]) continue;
          mbar_wait(&p_ready[t], i & 1);
          mbar_wait(&v_full[st], (i >> 1) & 1);
          tc_fence_after();
          issue_pv(t, i);
          if (t == 1 && leader) umma_commit(&v_empty[st]);
          if (i + 1 < nb[t]) {
            mbar_wait(&k_full[st ^ 1], ((i + 1) >> 1) & 1);
            tc_fence_after();
            issue_s(t, i + 1);
            if (t == 1 && leader) umma_commit(&k_empty[st ^ 1]);
          }
        }
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    const int r = ew * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(ew * 32) << 16;
    float m_ref[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};
    for (int i = 0; i < n_max; ++i) {
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (i >= nb[t]) continue;
        const int qi = m0 + t * kF2T + r;
        const uint32_t tS = tmem_base + t * 128 + lane_off;
        const uint32_t tO = tmem_base + 256 + t * 128 + lane_off;
        mbar_wait(&s_full[t], i & 1);
        tc_fence_after();
        // raw (unscaled) scores stay in registers; scale and running max are folded into one FFMA
        // in front of the exponential: p = ex2(s * scale_log2 - m)
        float s[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32(tS + c * 32, raw);
          tmem_ld_wait();
#pragma unroll
          for (int x = 0; x < 32; ++x) s[c * 32 + x] = __uint_as_float(raw[x]);
        }
        const int n0 = i * kF2T;
        const bool need_mask = (causal && (n0 + kF2T - 1 > m0 + t * kF2T + shift)) || (n0 + kF2T > sk);
        if (need_mask) {
          const int lim = causal ? min(qi + shift, sk - 1) : sk - 1;
#pragma unroll
          for (int x = 0; x < 128; ++x)
            if (n0 + x > lim) s[x] = -INFINITY;
        }
        float mb = -INFINITY;
#pragma unroll
        for (int x = 0; x < 128; ++x) mb = fmaxf(mb, s[x]);
        mb *= scale_log2;      // scale > 0: the max commutes with the scaling
        const bool grow = mb > m_ref[t] + 8.f;
        if (__any_sync(0xffffffffu, grow)) {
          const float m_new = grow ? mb : m_ref[t];
          const float alpha = (m_ref[t] == -INFINITY) ? 0.f : exp2f(m_ref[t] - m_new);
          m_ref[t] = m_new;
          l[t] *= alpha;
          if (i > 0) {
            mbar_wait(&o_done[t], (i - 1) & 1);     // PV_t(i-1) must have landed before O is scaled
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
              uint32_t o[32];
              tmem_ld_32x32(tO + c * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int x = 0; x < 32; ++x) o[x] = __float_as_uint(__uint_as_float(o[x]) * alpha);
              f2_tmem_st32(tO + c * 32, o);
            }
          }
        }
        // a row that has seen no key yet keeps m = -inf: use 0 there so -inf - m stays -inf
        const float neg_m = (m_ref[t] == -INFINITY) ? 0.f : -m_ref[t];
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t pk[32];
#pragma unroll
          for (int x = 0; x < 32; ++x) {
            const float p0 = f2_ex2(fmaf(s[c * 64 + 2 * x], scale_log2, neg_m));
            const float p1 = f2_ex2(fmaf(s[c * 64 + 2 * x + 1], scale_log2, neg_m));
            rs0 += p0;
            rs1 += p1;
            pk[x] = pack_bf16x2(p0, p1);
          }
          f2_tmem_st32(tS + c * 32, pk);          // P over the first 64 columns of S
        }
        const float rs = rs0 + rs1;
        l[t] += rs;
        f2_tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&p_ready[t]);
      }
    }
    // ------------------------------ epilogue ----------------------------------
    const int np = G * qpg;
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const int qi = m0 + t * kF2T + r;
      mbar_wait(&o_done[t], (nb[t] - 1) & 1);
      tc_fence_after();
      const float inv = l[t] > 0.f ? 1.f / l[t] : 0.f;
      if (qi < sq)
        lse[((size_t)b * np + head) * sq + qi] =
            l[t] > 0.f ? m_ref[t] * 0.69314718056f + __logf(l[t]) : -INFINITY;
      __nv_bfloat16* orow = out + ((size_t)qi * B + b) * ((size_t)np * kF2T) + (size_t)head * kF2T;
      const uint32_t tO = tmem_base + 256 + t * 128 + lane_off;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        tmem_ld_32x32(tO + c * 32, o);
        tmem_ld_wait();
        if (qi < sq) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(o[8 * q + 0]) * inv, __uint_as_float(o[8 * q + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(o[8 * q + 2]) * inv, __uint_as_float(o[8 * q + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(o[8 * q + 4]) * inv, __uint_as_float(o[8 * q + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(o[8 * q + 6]) * inv, __uint_as_float(o[8 * q + 7]) * inv);
            reinterpret_cast<uint4*>(orow + c * 32)[q] = w;
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

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_fn();

static int f2_make_tmap(CUtensorMap* map, const void* base, int rank, const long long* dims,
                        const long long* strides) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_UNSUPPORTED;
  cuuint64_t d[5], s[4];
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
  box[1] = kF2T;
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), d, s,
                  box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : B200_ERR_BAD_ARG;
}

}  // namespace b200

using namespace b200;

// Same contract as b200_attn_fwd (attn_fwd_tcgen05.cu).
B200_API int b200_attn_fwd_v2(const void* q, const void* k, const void* v, void* out, float* lse,
                              int sq, int sk, int B, int G, int qpg, int hn, long long q_ss,
                              long long q_sb, long long q_sg, long long q_sj, long long k_ss,
                              long long k_sb, long long k_sg, float softmax_scale, int causal,
                              cudaStream_t stream) {
  if (hn != kF2T) return B200_ERR_UNSUPPORTED;
  if (sq <= 0 || sk <= 0 || B <= 0 || G <= 0 || qpg <= 0) return B200_ERR_BAD_ARG;
  if (causal && sk < sq) return B200_ERR_UNSUPPORTED;
  CUtensorMap tq, tk, tv;
  long long qd[5] = {hn, sq, qpg, G, B}, qs[4] = {q_ss, q_sj, q_sg, q_sb};
  long long kd[4] = {hn, sk, G, B}, ks[3] = {k_ss, k_sg, k_sb};
  int rc = f2_make_tmap(&tq, q, 5, qd, qs);
  if (rc) return rc;
  rc = f2_make_tmap(&tk, k, 4, kd, ks);
  if (rc) return rc;
  rc = f2_make_tmap(&tv, v, 4, kd, ks);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         F2Smem::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  dim3 grid((sq + 2 * kF2T - 1) / (2 * kF2T), G * qpg, B);
  attn_fwd2_kernel<<<grid, kF2Threads, F2Smem::kTotal, stream>>>(
      tq, tk, tv, (__nv_bfloat16*)out, lse, sq, sk, B, G, qpg, softmax_scale * 1.4426950408889634f, causal);
  B200_CHECK_LAUNCH();
  return 0;
}
