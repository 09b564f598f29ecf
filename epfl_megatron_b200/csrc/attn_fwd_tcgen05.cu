// Causal / full GQA-native attention forward on tcgen05 + TMEM (K11 of SURVEY.md
// section 8). Replaces flash_attn_func at megatron/model/transformer.py:514-522 and the
// K/V head expansion + [s,b]<->[b,s] transposes in front of it (:448-455, :515-522).
//
// One CTA = 128 query rows of one (batch, head); KV is walked in blocks of 64.
//   * Q, K, V are fetched by TMA straight out of the fused projection output
//     [s, b, g, (q_per_g + 2), hn] through rank-5 / rank-4 tensor maps (any strides),
//     so GQA/MQA never materialises repeated K/V and nothing is transposed.
//   * S = Q K^T: tcgen05.mma (M=128, N=64, K=hn=128) into 64 TMEM columns.
//   * The 4 softmax warps own one query row per thread (no shuffles): tcgen05.ld S,
//     exp2 with a lazily updated reference max (O is only rescaled when the running max
//     grew by more than 2^8, FlashAttention-4 style), bf16 P is written to shared
//     memory in the 128B-swizzled K-major layout the next MMA reads.
//   * O += P V: tcgen05.mma (M=128, N=hn, K=64, B operand MN-major = V as stored)
//     accumulating in 128 TMEM columns for the whole KV walk.
//   * Two CTAs are resident per SM (80 KB smem, 256 TMEM columns each), so one CTA's
//     softmax overlaps the other's MMAs without an explicit ping-pong schedule.
// Outputs: O written as [s, b, np*hn] (the layout the dense projection consumes) and
// LSE [b, np, s] fp32 (natural log), the format FlashAttention's backward expects.
#include "common.cuh"

namespace b200 {

constexpr int kAttM = 128;   // query rows per CTA
constexpr int kAttN = 64;    // kv rows per block
constexpr int kAttD = 128;   // head dim
constexpr int kAttThreads = 256;

struct AttSmem {
  static constexpr int kQ = 0;                                // 2 x [128 x 128B] = 32 KB
  static constexpr int kK = kQ + kAttM * kAttD * 2;           // 2 x [64 x 128B]  = 16 KB
  static constexpr int kV = kK + kAttN * kAttD * 2;           // 2 x [64 x 128B]  = 16 KB
  static constexpr int kP = kV + kAttN * kAttD * 2;           // [128 x 128B]     = 16 KB
  static constexpr int kBar = kP + kAttM * kAttN * 2;
  static constexpr int kTotal = kBar + 128 + 1024;
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0),
      "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
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
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// q map: rank 5 {hn, s, j (q head inside group), g, b};  k/v maps: rank 4 {hn, s, g, b}
__global__ void __launch_bounds__(kAttThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, __nv_bfloat16* __restrict__ out,
                float* __restrict__ lse, int sq, int sk, int B, int G, int qpg, float scale_log2,
                int causal) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttSmem::kBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;
  uint64_t* k_empty = bars + 2;
  uint64_t* v_full = bars + 3;
  uint64_t* v_empty = bars + 4;
  uint64_t* s_full = bars + 5;
  uint64_t* p_ready = bars + 6;
  uint64_t* o_done = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (sq + kAttM - 1) / kAttM;
  const int mt = tiles_m - 1 - blockIdx.x;  // heaviest (most KV blocks) tiles first
  const int head = blockIdx.y, b = blockIdx.z;
  const int g = head / qpg, j = head - g * qpg;
  const int m0 = mt * kAttM;
  const int shift = sk - sq;  // query i sees keys <= i + shift when causal
  int kv_end = causal ? min(sk, m0 + kAttM + shift) : sk;
  const int nblk = (kv_end + kAttN - 1) / kAttN;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    mbar_init(k_full, 1);
    mbar_init(k_empty, 1);
    mbar_init(v_full, 1);
    mbar_init(v_empty, 1);
    mbar_init(s_full, 1);
    mbar_init(p_ready, 128);
    mbar_init(o_done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base;        // 64 columns
  const uint32_t tmem_O = tmem_base + 128;  // 128 columns

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, kAttM * kAttD * 2);
      tma_load_5d(smem + AttSmem::kQ, &tmQ, q_full, 0, m0, j, g, b);
      tma_load_5d(smem + AttSmem::kQ + kAttM * 128, &tmQ, q_full, 64, m0, j, g, b);
      for (int i = 0; i < nblk; ++i) {
        const uint32_t ph = i & 1;
        mbar_wait(k_empty, ph ^ 1);
        mbar_expect_tx(k_full, kAttN * kAttD * 2);
        tma_load_4d(smem + AttSmem::kK, &tmK, k_full, 0, i * kAttN, g, b);
        tma_load_4d(smem + AttSmem::kK + kAttN * 128, &tmK, k_full, 64, i * kAttN, g, b);
        mbar_wait(v_empty, ph ^ 1);
        mbar_expect_tx(v_full, kAttN * kAttD * 2);
        tma_load_4d(smem + AttSmem::kV, &tmV, v_full, 0, i * kAttN, g, b);
        tma_load_4d(smem + AttSmem::kV + kAttN * 128, &tmV, v_full, 64, i * kAttN, g, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kAttM, kAttN, 0, 0);
      constexpr uint32_t idesc_o = make_idesc_bf16(kAttM, kAttD, 0, 1);
      const uint32_t sQ = smem_u32(smem + AttSmem::kQ), sK = smem_u32(smem + AttSmem::kK);
      const uint32_t sV = smem_u32(smem + AttSmem::kV), sP = smem_u32(smem + AttSmem::kP);
      mbar_wait(q_full, 0);
      auto issue_s = [&](int i) {
        mbar_wait(k_full, i & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < kAttD / 16; ++k) {
          // hn 0..63 live in the first 128B-wide box, 64..127 in the second
          const uint32_t qa = sQ + (k >> 2) * (kAttM * 128) + (k & 3) * 32;
          const uint32_t ka = sK + (k >> 2) * (kAttN * 128) + (k & 3) * 32;
          umma_bf16_ss(tmem_S, make_sw128_desc(qa, 16, 1024), make_sw128_desc(ka, 16, 1024),
                       idesc_s, k != 0);
        }
        umma_commit(k_empty);
        umma_commit(s_full);
      };
      issue_s(0);
      for (int i = 0; i < nblk; ++i) {
        mbar_wait(p_ready, i & 1);
        mbar_wait(v_full, i & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < kAttN / 16; ++k) {
          const uint64_t pa = make_sw128_desc(sP + k * 32, 16, 1024);
          // V is [kv rows x hn]: MN-major B, two 64-wide hn chunks 8 KB apart, 16 kv rows
          // per MMA = 2 KB.
          const uint64_t vb = make_sw128_desc(sV + k * 2048, kAttN * 128, 1024);
          umma_bf16_ss(tmem_O, pa, vb, idesc_o, (i | k) != 0);
        }
        umma_commit(v_empty);
        if (i + 1 < nblk) issue_s(i + 1);
        else umma_commit(o_done);
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    const int r = ew * 32 + lane;  // row inside the tile == TMEM lane
    const int qi = m0 + r;
    const uint32_t lane_off = static_cast<uint32_t>(ew * 32) << 16;
    float m_ref = -INFINITY, l = 0.f;
    for (int i = 0; i < nblk; ++i) {
      mbar_wait(s_full, i & 1);
      tc_fence_after();
      uint32_t sr[64];
      tmem_ld_32x32(tmem_S + lane_off, sr);
      tmem_ld_32x32(tmem_S + lane_off + 32, sr + 32);
      tmem_ld_wait();
      const int n0 = i * kAttN;
      const int lim = causal ? min(qi + shift, sk - 1) : sk - 1;  // last visible key
      float mb = -INFINITY;
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        float s = __uint_as_float(sr[c]) * scale_log2;
        if (n0 + c > lim) s = -INFINITY;
        sr[c] = __float_as_uint(s);
        mb = fmaxf(mb, s);
      }
      // lazy rescale: keep the old reference unless the max grew by more than 2^8
      const bool grow = mb > m_ref + 8.f;
      const bool any_grow = __any_sync(0xffffffffu, grow);
      if (any_grow) {
        const float m_new = grow ? mb : m_ref;
        const float alpha = (m_ref == -INFINITY) ? 0.f : exp2f(m_ref - m_new);
        m_ref = m_new;
        l *= alpha;
        if (i > 0) {
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(tmem_O + lane_off + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int t = 0; t < 32; ++t) o[t] = __float_as_uint(__uint_as_float(o[t]) * alpha);
            tmem_st_32x32(tmem_O + lane_off + c * 32, o);
          }
          tmem_st_wait();
        }
      }
      float rs = 0.f;
      uint8_t* prow = smem + AttSmem::kP;
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        float p[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const float s = __uint_as_float(sr[ch * 8 + t]);
          p[t] = (m_ref == -INFINITY) ? 0.f : exp2f(s - m_ref);
          rs += p[t];
        }
        uint4 o;
        o.x = pack_bf16x2(p[0], p[1]);
        o.y = pack_bf16x2(p[2], p[3]);
        o.z = pack_bf16x2(p[4], p[5]);
        o.w = pack_bf16x2(p[6], p[7]);
        *reinterpret_cast<uint4*>(prow + sw128_offset(r, ch)) = o;
      }
      l += rs;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    mbar_wait(o_done, 0);
    tc_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    const int np = G * qpg;
    if (qi < sq) {
      lse[((size_t)b * np + head) * sq + qi] =
          l > 0.f ? m_ref * 0.69314718056f + __logf(l) : -INFINITY;
    }
    __nv_bfloat16* orow = out + ((size_t)qi * B + b) * ((size_t)np * kAttD) + (size_t)head * kAttD;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      tmem_ld_32x32(tmem_O + lane_off + c * 32, o);
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

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
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

// bf16 tensor map of rank `rank`; dims[0] is the contiguous head dim, strides are in
// ELEMENTS for dims 1..rank-1; box = {64, box_rows, 1, ...}; 128B swizzle.
static int make_tmap_nd(CUtensorMap* map, const void* base, int rank, const long long* dims,
                        const long long* strides, int box_rows) {
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
  box[1] = (cuuint32_t)box_rows;
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                  d, s, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : B200_ERR_BAD_ARG;
}
}  // namespace b200

using namespace b200;

// q: element (s, b, g, j, d) at q + s*q_ss + b*q_sb + g*q_sg + j*q_sj + d
// k/v: element (s, b, g, d) at k + s*k_ss + b*k_sb + g*k_sg + d   (same strides for v)
// out: [sq, B, G*qpg*128] bf16 contiguous; lse: [B, G*qpg, sq] fp32.
B200_API int b200_attn_fwd(const void* q, const void* k, const void* v, void* out, float* lse,
                           int sq, int sk, int B, int G, int qpg, int hn, long long q_ss,
                           long long q_sb, long long q_sg, long long q_sj, long long k_ss,
                           long long k_sb, long long k_sg, float softmax_scale, int causal,
                           cudaStream_t stream) {
  if (hn != kAttD) return B200_ERR_UNSUPPORTED;
  if (sq <= 0 || sk <= 0 || B <= 0 || G <= 0 || qpg <= 0) return B200_ERR_BAD_ARG;
  if (causal && sk < sq) return B200_ERR_UNSUPPORTED;
  CUtensorMap tq, tk, tv;
  long long qd[5] = {hn, sq, qpg, G, B}, qs[4] = {q_ss, q_sj, q_sg, q_sb};
  long long kd[4] = {hn, sk, G, B}, ks[3] = {k_ss, k_sg, k_sb};
  int rc = make_tmap_nd(&tq, q, 5, qd, qs, kAttM);
  if (rc) return rc;
  rc = make_tmap_nd(&tk, k, 4, kd, ks, kAttN);
  if (rc) return rc;
  rc = make_tmap_nd(&tv, v, 4, kd, ks, kAttN);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         AttSmem::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  dim3 grid((sq + kAttM - 1) / kAttM, G * qpg, B);
  attn_fwd_kernel<<<grid, kAttThreads, AttSmem::kTotal, stream>>>(
      tq, tk, tv, (__nv_bfloat16*)out, lse, sq, sk, B, G, qpg,
      softmax_scale * 1.4426950408889634f, causal);
  B200_CHECK_LAUNCH();
  return 0;
}
