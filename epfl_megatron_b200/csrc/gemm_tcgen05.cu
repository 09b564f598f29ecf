// Persistent warp-specialised bf16 GEMM on the 5th-gen tensor cores (K7 of SURVEY.md
// section 8): TMA (cp.async.bulk.tensor, 128B swizzle) -> 4-stage shared-memory ring ->
// tcgen05.mma (kind::f16, M=128, N=BN) issued by ONE thread -> fp32 accumulators in
// TMEM (two BN-column stages, so the epilogue of tile i overlaps the MMAs of tile
// i+1) -> tcgen05.ld -> registers -> global.
//
//   C[M,N] (op)= A_op[M,K] * B_op[N,K]^T,  fp32 accumulate
//
// Operand storage (the "major" of each operand is independent):
//   A K-major : A stored [M, K] row-major          A MN-major : A stored [K, M] row-major
//   B K-major : B stored [N, K] row-major          B MN-major : B stored [K, N] row-major
// which covers the three GEMMs of a linear layer without any transposed copies:
//   forward  Y  = X  W^T : A = X  (K-major),  B = W  (K-major)
//   dgrad    dX = dY W   : A = dY (K-major),  B = W  (MN-major)
//   wgrad    dW = dY^T X : A = dY (MN-major), B = X  (MN-major)
// Output modes: 0 = store bf16, 1 = store fp32, 2 = accumulate into fp32 (the
// `main_grad += dY^T X` fusion that the reference gets from
// fused_weight_gradient_dense.cu:128-152 through a cuBLAS beta=1 call).
//
// Replaces the cuBLAS calls at megatron/core/tensor_parallel/layers.py:240,267,307 on
// the fused paths.
#include "common.cuh"

namespace b200 {

constexpr int kBM = 128;
constexpr int kBK = 64;            // 64 bf16 = 128 bytes = one swizzle row
constexpr int kStages = 4;
constexpr int kGemmThreads = 256;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps4-7 epilogue
constexpr int kGroupM = 8;

template <int BN>
struct GemmSmem {
  static constexpr int kABytes = kBM * kBK * 2;   // 16 KB
  static constexpr int kBBytes = BN * kBK * 2;    // 16/32 KB
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarOffset = kStages * kStageBytes;
  static constexpr int kTotal = kBarOffset + 256 + 1024;  // + barriers + alignment slack
};

__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int& tm, int& tn) {
  const int group = kGroupM * tiles_n;
  const int gid = tile / group;
  const int first_m = gid * kGroupM;
  const int gsz = min(tiles_m - first_m, kGroupM);
  const int in_group = tile - gid * group;
  tm = first_m + in_group % gsz;
  tn = in_group / gsz;
}

// One thread's 32 consecutive fp32 accumulator columns of row `row` -> global memory.
__device__ __forceinline__ void store_acc_chunk(const uint32_t* r, void* __restrict__ Cout, int row,
                                                int col0, int M, int N, long long ldc,
                                                int out_mode) {
  if (row >= M || col0 >= N) return;
  const bool full = (col0 + 32 <= N);
  if (out_mode == 0) {
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(Cout) + (long long)row * ldc + col0;
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(r[8 * q + 0]), __uint_as_float(r[8 * q + 1]));
        o.y = pack_bf16x2(__uint_as_float(r[8 * q + 2]), __uint_as_float(r[8 * q + 3]));
        o.z = pack_bf16x2(__uint_as_float(r[8 * q + 4]), __uint_as_float(r[8 * q + 5]));
        o.w = pack_bf16x2(__uint_as_float(r[8 * q + 6]), __uint_as_float(r[8 * q + 7]));
        reinterpret_cast<uint4*>(dst)[q] = o;
      }
    } else {
      for (int j = 0; j < 32 && col0 + j < N; ++j)
        dst[j] = __float2bfloat16_rn(__uint_as_float(r[j]));
    }
  } else {
    float* dst = reinterpret_cast<float*>(Cout) + (long long)row * ldc + col0;
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 o = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                               __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
        if (out_mode == 2) {
          float4 p = reinterpret_cast<float4*>(dst)[q];
          o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
        }
        reinterpret_cast<float4*>(dst)[q] = o;
      }
    } else {
      for (int j = 0; j < 32 && col0 + j < N; ++j) {
        float v = __uint_as_float(r[j]);
        dst[j] = out_mode == 2 ? dst[j] + v : v;
      }
    }
  }
}

template <int BN, bool kAMN, bool kBMN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 void* __restrict__ Cout, int M, int N, int K, long long ldc, int out_mode) {
  using L = GemmSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = (M + kBM - 1) / kBM;
  const int tiles_n = (N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + kBK - 1) / kBK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 128);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<2 * BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int tm, tn;
        tile_coords(tile, tiles_m, tiles_n, tm, tn);
        const int m0 = tm * kBM, n0 = tn * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::kStageBytes;
          uint8_t* sb = sa + L::kABytes;
          mbar_expect_tx(&full_bar[stage], L::kStageBytes);
          if (!kAMN) {
            tma_load_2d(sa, &tmA, &full_bar[stage], kb * kBK, m0);
          } else {
#pragma unroll
            for (int j = 0; j < kBM / 64; ++j)
              tma_load_2d(sa + j * (kBK * 128), &tmA, &full_bar[stage], m0 + j * 64, kb * kBK);
          }
          if (!kBMN) {
            tma_load_2d(sb, &tmB, &full_bar[stage], kb * kBK, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (kBK * 128), &tmB, &full_bar[stage], n0 + j * 64, kb * kBK);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBM, BN, kAMN ? 1 : 0, kBMN ? 1 : 0);
      // K-major: 8-row groups are 1024 B apart (SBO); LBO unused inside one swizzle row.
      // MN-major: 64-element chunks along M/N are kBK*128 B apart (LBO), 8-row K groups
      // 1024 B apart (SBO).
      constexpr uint32_t lbo_a = kAMN ? kBK * 128 : 16, lbo_b = kBMN ? kBK * 128 : 16;
      constexpr uint32_t kstep_a = kAMN ? 16 * 128 : 32, kstep_b = kBMN ? 16 * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t sb = sa + L::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = make_sw128_desc(sa + k * kstep_a, lbo_a, 1024);
            const uint64_t db = make_sw128_desc(sb + k * kstep_b, lbo_b, 1024);
            umma_bf16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[as]);
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue ----------------------------------
    const int ew = warp - 4;  // TMEM lane quarter this warp may read
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int tm, tn;
      tile_coords(tile, tiles_m, tiles_n, tm, tn);
      const int row = tm * kBM + ew * 32 + lane;
      const int n0 = tn * BN;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        store_acc_chunk(r, Cout, row, n0 + c * 32, M, N, ldc, out_mode);
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<2 * BN>(tmem_base);
  }
}


// ---------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): a cluster of two CTAs on one TPC owns a 256 x 256
// output tile. Each CTA stages its own 128 rows of A and its own 128-row half of B
// (32 KB per stage instead of 48 KB, so 6 stages fit and the shared-memory read
// traffic per MMA halves), the leader CTA's elected thread issues ONE
// tcgen05.mma.cta_group::2 (M = 256, N = 256) per 16-element K step, and each CTA
// drains its own 128 x 256 fp32 accumulator half from its own TMEM.
//
// Barriers: full[stage] lives in the leader and counts the bytes of BOTH CTAs' TMA
// loads; empty[stage] / tmem_full[as] exist in both CTAs and are signalled by a
// multicast tcgen05.commit; tmem_empty[as] lives in the leader and is arrived on by one
// lane of each of the 8 epilogue warps of the pair (remote arrive for the peer).
// ---------------------------------------------------------------------------------
constexpr int kPairStages = 5;
constexpr int kPairBN = 256;       // instruction N; each CTA stages kPairBN / 2 rows of B
constexpr int kPairHalfN = kPairBN / 2;

struct PairSmem {
  static constexpr int kABytes = kBM * kBK * 2;          // 16 KB
  static constexpr int kBBytes = kPairHalfN * kBK * 2;   // 16 KB
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kOutBox = kBM * 128;              // [128 rows x 128 B]: 64 bf16 or 32 fp32 columns
  static constexpr int kOutOffset = kPairStages * kStageBytes;
  static constexpr int kBarOffset = kOutOffset + 2 * kOutBox;
  static constexpr int kTotal = kBarOffset + 256 + 1024;
};

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
      ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
      : "memory");
}

// kF8: 8-bit floating point operands (K-major only; a tile row is 128 one-byte elements instead of
// 64 bf16, every byte count of the pipeline stays the same), tcgen05.mma.kind::f8f6f4 with the
// runtime descriptor f8_idesc, and the epilogue multiplies by *alpha_a * *alpha_b (the two
// dequantisation factors, read from device memory so that no host round trip is needed).
template <bool kAMN, bool kBMN, bool kF8 = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmC, void* __restrict__ Cout, int M, int N, int K,
                      long long ldc, int out_mode, int tma_epilogue, uint32_t f8_idesc = 0,
                      const float* __restrict__ alpha_a = nullptr,
                      const float* __restrict__ alpha_b = nullptr) {
  static_assert(!kF8 || (!kAMN && !kBMN), "8-bit operands are staged K-major");
  constexpr int kKElems = kF8 ? 2 * kBK : kBK;      // elements per 128-byte tile row
  using L = PairSmem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + kPairStages;
  uint64_t* tmem_full = empty_bar + kPairStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int tiles_m = (M + 2 * kBM - 1) / (2 * kBM);
  const int tiles_n = (N + kPairBN - 1) / kPairBN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + kKElems - 1) / kKElems;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < kPairStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc_pair<2 * kPairBN>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int tm, tn;
        tile_coords(tile, tiles_m, tiles_n, tm, tn);
        const int m0 = tm * 2 * kBM + static_cast<int>(rank) * kBM;
        const int n0 = tn * kPairBN + static_cast<int>(rank) * kPairHalfN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::kStageBytes;
          uint8_t* sb = sa + L::kABytes;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * L::kStageBytes);
          if (!kAMN) {
            tma_load_2d_pair(sa, &tmA, &full_bar[stage], kb * kKElems, m0);
          } else {
#pragma unroll
            for (int j = 0; j < kBM / 64; ++j)
              tma_load_2d_pair(sa + j * (kBK * 128), &tmA, &full_bar[stage], m0 + j * 64, kb * kBK);
          }
          if (!kBMN) {
            tma_load_2d_pair(sb, &tmB, &full_bar[stage], kb * kKElems, n0);
          } else {
#pragma unroll
            for (int j = 0; j < kPairHalfN / 64; ++j)
              tma_load_2d_pair(sb + j * (kBK * 128), &tmB, &full_bar[stage], n0 + j * 64, kb * kBK);
          }
          if (++stage == kPairStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) --------------
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * kBM, kPairBN, kAMN ? 1 : 0, kBMN ? 1 : 0);
      constexpr uint32_t lbo_a = kAMN ? kBK * 128 : 16, lbo_b = kBMN ? kBK * 128 : 16;
      constexpr uint32_t kstep_a = kAMN ? 16 * 128 : 32, kstep_b = kBMN ? 16 * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kPairBN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t sb = sa + L::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = make_sw128_desc(sa + k * kstep_a, lbo_a, 1024);
            const uint64_t db = make_sw128_desc(sb + k * kstep_b, lbo_b, 1024);
            if (kF8) umma_f8_ss_pair(d_tmem, da, db, f8_idesc, (kb | k) != 0 ? 1u : 0u);   // K = 32 bytes
            else umma_bf16_ss_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);
          if (++stage == kPairStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(&tmem_full[as]);
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue (both CTAs) ----------------------
    // TMEM -> registers -> 128-byte-swizzled shared-memory box -> TMA: a plain tile store (bf16 / fp32)
    // or, for out_mode 2, cp.reduce.async.bulk .add.f32 so the `main_grad += dY^T X` accumulation
    // happens in L2 with full-line writes instead of a read-modify-write through the SM. Two boxes
    // alternate. C that TMA cannot address (odd strides / alignment) takes the direct-store path.
    const int ew = warp - 4;
    const int r = ew * 32 + lane;
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t leader_empty0 = mapa_u32(smem_u32(&tmem_empty[0]), 0);
    uint8_t* stage_out = smem + L::kOutOffset;
    const bool issuer = (ew == 0 && lane == 0);
    uint32_t box_count = 0;
    float alpha = 1.f;
    if (kF8) alpha = (alpha_a ? *alpha_a : 1.f) * (alpha_b ? *alpha_b : 1.f);
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int tm, tn;
      tile_coords(tile, tiles_m, tiles_n, tm, tn);
      const int row0 = tm * 2 * kBM + static_cast<int>(rank) * kBM;
      const int row = row0 + r;
      const int n0 = tn * kPairBN;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * kPairBN;
      if (!tma_epilogue) {
#pragma unroll 1
        for (int c = 0; c < kPairBN / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
          if (kF8) {
#pragma unroll
            for (int q = 0; q < 32; ++q) v[q] = __float_as_uint(__uint_as_float(v[q]) * alpha);
          }
          store_acc_chunk(v, Cout, row, n0 + c * 32, M, N, ldc, out_mode);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(leader_empty0 + as * 8);
      } else if (out_mode == 0) {
#pragma unroll 1
        for (int bx = 0; bx < kPairBN / 64; ++bx) {
          uint8_t* box = stage_out + (box_count & 1) * L::kOutBox;
          if (issuer) tma_store_wait_read<1>();      // the store two boxes ago has drained this buffer
          asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t v[32];
            tmem_ld_32x32(taddr + (2 * bx + h) * 32, v);
            tmem_ld_wait();
            if (kF8) {
#pragma unroll
              for (int q = 0; q < 32; ++q) v[q] = __float_as_uint(__uint_as_float(v[q]) * alpha);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 o;
              o.x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
              o.y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
              o.z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
              o.w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
              *reinterpret_cast<uint4*>(box + sw128_offset(r, h * 4 + q)) = o;
            }
          }
          if (bx == kPairBN / 64 - 1) {          // every accumulator column of this tile has been read
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(leader_empty0 + as * 8);
          }
          fence_proxy_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (issuer && row0 < M && n0 + bx * 64 < N) {
            tma_store_2d(&tmC, box, n0 + bx * 64, row0);
            tma_store_commit();
          }
          ++box_count;
        }
      } else {
#pragma unroll 1
        for (int bx = 0; bx < kPairBN / 32; ++bx) {
          uint8_t* box = stage_out + (box_count & 1) * L::kOutBox;
          if (issuer) tma_store_wait_read<1>();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          uint32_t v[32];
          tmem_ld_32x32(taddr + bx * 32, v);
          tmem_ld_wait();
          if (kF8) {
#pragma unroll
            for (int q = 0; q < 32; ++q) v[q] = __float_as_uint(__uint_as_float(v[q]) * alpha);
          }
#pragma unroll
          for (int q = 0; q < 8; ++q)
            *reinterpret_cast<uint4*>(box + sw128_offset(r, q)) =
                make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          if (bx == kPairBN / 32 - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(leader_empty0 + as * 8);
          }
          fence_proxy_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (issuer && row0 < M && n0 + bx * 32 < N) {
            if (out_mode == 2) tma_reduce_add_2d(&tmC, box, n0 + bx * 32, row0);
            else tma_store_2d(&tmC, box, n0 + bx * 32, row0);
            tma_store_commit();
          }
          ++box_count;
        }
      }
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
    if (issuer) tma_store_wait<0>();
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<2 * kPairBN>(tmem_base);
  }
}

}  // namespace b200

// ------------------------------- host side -----------------------------------
#include <mutex>

namespace b200 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled is a DRIVER entry point: unlike runtime calls it does not bind the
// device's primary context to the calling thread. A thread whose first CUDA action is one of
// our launches (an autograd worker starting a pipeline stage's backward with a GEMM) would get
// CUDA_ERROR_INVALID_CONTEXT from the encode, so bind the context once per thread first.
void ensure_thread_context() {
  static thread_local bool bound = false;
  if (!bound) {
    cudaFree(nullptr);
    bound = true;
  }
}

EncodeTiledFn get_encode_fn() {
  ensure_thread_context();
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D bf16 tensor [rows, cols] with row stride `ld` elements, box [box_rows, 64 cols],
// 128-byte swizzle, out-of-bounds elements read as zero.
int make_tmap_2d(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                 int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_UNSUPPORTED;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : B200_ERR_BAD_ARG;
}

// 2-D one-byte-element tensor [rows, cols] with row stride `ld` bytes, box [box_rows, 128 cols],
// 128-byte swizzle (FP8 operands; zero fill out of bounds = +0 in E4M3 and E5M2).
int make_tmap_2d_u8(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                    int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_UNSUPPORTED;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld};
  cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : B200_ERR_BAD_ARG;
}

// Output map for the TMA epilogue: [rows, cols] of bf16 (box 64 cols) or fp32 (box 32 cols), 128 rows
// per box, 128-byte swizzle. Returns B200_ERR_UNSUPPORTED when TMA cannot address this C.
int make_tmap_out(CUtensorMap* map, void* base, long long rows, long long cols, long long ld, bool fp32) {
  EncodeTiledFn fn = get_encode_fn();
  const int esz = fp32 ? 4 : 2;
  if (!fn || (reinterpret_cast<uintptr_t>(base) & 15) || (ld * esz) % 16) return B200_ERR_UNSUPPORTED;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esz), (cuuint32_t)kBM};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base,
                  dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : B200_ERR_UNSUPPORTED;
}

template <int BN, bool kAMN, bool kBMN>
int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, void* C, int M, int N, int K,
                long long ldc, int out_mode, int max_ctas, cudaStream_t stream) {
  auto kern = gemm_bf16_kernel<BN, kAMN, kBMN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         GemmSmem<BN>::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int tiles = ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN);
  int grid = tiles < kNumSMs ? tiles : kNumSMs;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  kern<<<grid, kGemmThreads, GemmSmem<BN>::kTotal, stream>>>(ta, tb, C, M, N, K, ldc, out_mode);
  B200_CHECK_LAUNCH();
  return 0;
}

template <bool kAMN, bool kBMN>
int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb, void* C, int M, int N, int K,
                     long long ldc, int out_mode, int max_ctas, cudaStream_t stream) {
  auto kern = gemm_bf16_pair_kernel<kAMN, kBMN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         PairSmem::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int tiles = ((M + 2 * kBM - 1) / (2 * kBM)) * ((N + kPairBN - 1) / kPairBN);
  int pairs = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
  if (max_ctas > 1 && pairs > max_ctas / 2) pairs = max_ctas / 2;
  CUtensorMap tc;
  const int tma_epilogue = make_tmap_out(&tc, C, M, N, ldc, out_mode != 0) == 0 ? 1 : 0;
  if (!tma_epilogue) tc = ta;   // unused; any valid map
  kern<<<2 * pairs, kGemmThreads, PairSmem::kTotal, stream>>>(ta, tb, tc, C, M, N, K, ldc, out_mode,
                                                              tma_epilogue, 0u, nullptr, nullptr);
  B200_CHECK_LAUNCH();
  return 0;
}

int launch_gemm_pair_f8(const CUtensorMap& ta, const CUtensorMap& tb, void* C, int M, int N, int K,
                        long long ldc, int out_mode, uint32_t idesc, const float* alpha_a,
                        const float* alpha_b, cudaStream_t stream) {
  auto kern = gemm_bf16_pair_kernel<false, false, true>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairSmem::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int tiles = ((M + 2 * kBM - 1) / (2 * kBM)) * ((N + kPairBN - 1) / kPairBN);
  const int pairs = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
  CUtensorMap tc;
  const int tma_epilogue = make_tmap_out(&tc, C, M, N, ldc, out_mode != 0) == 0 ? 1 : 0;
  if (!tma_epilogue) tc = ta;
  kern<<<2 * pairs, kGemmThreads, PairSmem::kTotal, stream>>>(ta, tb, tc, C, M, N, K, ldc, out_mode, tma_epilogue,
                                                              idesc, alpha_a, alpha_b);
  B200_CHECK_LAUNCH();
  return 0;
}

}  // namespace b200

using namespace b200;

// C[M, N] (bf16: out_mode 0, fp32 store: 1, fp32 accumulate: 2) = (*alpha_a * *alpha_b) * A[M, K] B[N, K]^T
// with 8-bit floating point operands, both K-major (row stride lda / ldb bytes, multiples of 16);
// a_e5m2 / b_e5m2: 0 = E4M3, 1 = E5M2. alpha_* are DEVICE pointers (the dequantisation factors 1 / scale)
// or NULL for 1. fp32 accumulation in tensor memory. K8 of the FP8 recipe the reference reaches through
// Transformer Engine (megatron/model/transformer.py:931-950); tcgen05.mma.kind::f8f6f4, CTA pairs.
B200_API int b200_gemm_fp8(const void* A, const void* B, void* C, int M, int N, int K, int a_e5m2, int b_e5m2,
                           long long lda, long long ldb, long long ldc, int out_mode, const float* alpha_a,
                           const float* alpha_b, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || out_mode < 0 || out_mode > 2) return B200_ERR_BAD_ARG;
  if ((lda % 16) || (ldb % 16) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return B200_ERR_UNSUPPORTED;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d_u8(&ta, A, M, K, lda, kBM);
  if (rc) return rc;
  rc = make_tmap_2d_u8(&tb, B, N, K, ldb, kPairHalfN);
  if (rc) return rc;
  const uint32_t idesc = make_idesc_f8(2 * kBM, kPairBN, a_e5m2 ? 1 : 0, b_e5m2 ? 1 : 0);
  return launch_gemm_pair_f8(ta, tb, C, M, N, K, ldc, out_mode, idesc, alpha_a, alpha_b, stream);
}

// variant: 0 = choose (CTA pairs when the problem has at least one 256 x 256 tile),
// 1 = force the single-CTA kernel, 2 = force the CTA-pair kernel.
static int gemm_dispatch(const void* A, const void* B, void* C, int M, int N, int K,
                         int a_mn_major, int b_mn_major, long long lda, long long ldb,
                         long long ldc, int out_mode, int max_ctas, int variant,
                         cudaStream_t stream);

B200_API int b200_gemm_bf16_variant(const void* A, const void* B, void* C, int M, int N, int K,
                                    int a_mn_major, int b_mn_major, long long lda, long long ldb,
                                    long long ldc, int out_mode, int max_ctas, int variant,
                                    cudaStream_t stream) {
  return gemm_dispatch(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, out_mode, max_ctas,
                       variant, stream);
}

// See the header comment for the operand conventions. lda/ldb/ldc are row strides in
// elements of the STORED matrices. max_ctas > 0 caps the persistent grid (used to leave
// SMs free for a concurrent communication kernel).
B200_API int b200_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                            int a_mn_major, int b_mn_major, long long lda, long long ldb,
                            long long ldc, int out_mode, int max_ctas, cudaStream_t stream) {
  return gemm_dispatch(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, out_mode, max_ctas,
                       0, stream);
}

static int gemm_dispatch(const void* A, const void* B, void* C, int M, int N, int K,
                         int a_mn_major, int b_mn_major, long long lda, long long ldb,
                         long long ldc, int out_mode, int max_ctas, int variant,
                         cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return B200_ERR_BAD_ARG;
  if ((lda % 8) || (ldb % 8) || out_mode < 0 || out_mode > 2) return B200_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return B200_ERR_UNSUPPORTED;
  CUtensorMap ta, tb;
  int rc;
  const bool pair = variant == 2 || (variant == 0 && M >= 2 * kBM && N >= kPairBN && max_ctas != 1);
  if (pair) {
    if (!a_mn_major) rc = make_tmap_2d(&ta, A, M, K, lda, kBM);
    else rc = make_tmap_2d(&ta, A, K, M, lda, kBK);
    if (rc) return rc;
    if (!b_mn_major) rc = make_tmap_2d(&tb, B, N, K, ldb, kPairHalfN);
    else rc = make_tmap_2d(&tb, B, K, N, ldb, kBK);
    if (rc) return rc;
    if (!a_mn_major && !b_mn_major)
      return launch_gemm_pair<false, false>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);
    if (!a_mn_major && b_mn_major)
      return launch_gemm_pair<false, true>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);
    if (a_mn_major && !b_mn_major)
      return launch_gemm_pair<true, false>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);
    return launch_gemm_pair<true, true>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);
  }
  const int BN = N > 128 ? 256 : 128;
  if (!a_mn_major) rc = make_tmap_2d(&ta, A, M, K, lda, kBM);
  else rc = make_tmap_2d(&ta, A, K, M, lda, kBK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_2d(&tb, B, N, K, ldb, BN);
  else rc = make_tmap_2d(&tb, B, K, N, ldb, kBK);
  if (rc) return rc;
#define B200_GEMM_DISPATCH(BN_)                                                                  \
  if (!a_mn_major && !b_mn_major)                                                                \
    return launch_gemm<BN_, false, false>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);  \
  if (!a_mn_major && b_mn_major)                                                                 \
    return launch_gemm<BN_, false, true>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);   \
  if (a_mn_major && !b_mn_major)                                                                 \
    return launch_gemm<BN_, true, false>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);   \
  return launch_gemm<BN_, true, true>(ta, tb, C, M, N, K, ldc, out_mode, max_ctas, stream);
  if (BN == 256) {
    B200_GEMM_DISPATCH(256)
  }
  B200_GEMM_DISPATCH(128)
#undef B200_GEMM_DISPATCH
}
