// Tensor-parallel GEMMs with the collective INSIDE the kernel (K8 / K9 of SURVEY.md section 8),
// over peer-mapped ("symmetric") memory of the GPUs of one NVSwitch domain.
//
//   mode AG  (all-gather -> GEMM): ColumnParallelLinear forward under sequence parallelism and
//     the dgrad of RowParallelLinear (megatron/core/tensor_parallel/layers.py:225-243 and
//     mappings.py:244-246 + the matmul behind them).  C[M, N] = gather_rows(X)[M, K] * B^T.
//     Every CTA donates two warps. They first copy the rank's own rows into its staging buffer
//     (where the peers pull them from; guarded by the peers' "consumed" flags of two calls ago)
//     and into their place in the gathered matrix, the last warp to finish raises the peers'
//     shard_ready flag; the own-chunk tiles meanwhile read the caller's tensor directly through
//     a second tensor map, so no separate staging kernel runs before the GEMM. Then the same warps
//     PULL the other ranks' row chunks out of their staging buffers (16-byte peer loads over
//     NVLink) into the local gathered matrix, 128-row block
//     by 128-row block; a counter per block is bumped when a warp's share has landed. The
//     TMA producers walk the output tiles starting with the rank's own chunk and only wait
//     on a block's counter right before its first A tile is loaded, so the transfer of chunk
//     r+1 runs under the MMAs of chunk r. The gathered matrix is an ordinary tensor the
//     caller keeps for the weight gradient (no second gather in backward).
//   mode RS  (GEMM -> reduce-scatter): RowParallelLinear forward and the dgrad of
//     ColumnParallelLinear (layers.py:691-692, :285-296).  out[M/tp, N] = sum_r (A_r B_r^T)[chunk].
//     Tiles whose rows another rank owns come first: the epilogue converts the accumulators to
//     bf16 and STORES the tile straight into the owner's slot for this source (TMA stores over
//     NVLink), then bumps the owner's arrival counter of that tile (release at system scope).
//     The rank's own tiles come last: by then the peers' partial tiles have long arrived, the
//     epilogue waits on the tile's counter, adds the tp-1 partials to the fp32 accumulators on
//     their way out of tensor memory and TMA-stores the FINAL tile -- the reduction is the
//     epilogue of the tiles that would have produced the last partial anyway; no partial of the
//     own chunk is ever written, no separate reduction pass reads the slots back.
//   mode AR  (GEMM -> all-reduce): RowParallelLinear forward and ColumnParallelLinear dgrad WITHOUT
//     sequence parallelism (layers.py:694, :278-283). Same as RS, but the owner stores each final
//     tile into the output buffer of EVERY rank (peer TMA stores) and counts it there; the kernel
//     ends once all tiles owned by other ranks have arrived.
//
// Protocol state lives in the symmetric workspace (csrc/comm.cu): monotonic u32 counters
// compared with wrap-around arithmetic, one epoch per call. Waits on other GPUs are bounded:
// after kWaitLimitCycles (about a minute, far above any host-side launch skew between ranks) the
// wait is recorded, printed and the kernel traps, so a lost peer or a protocol bug ends the job
// with a launch error instead of a hung device or a silently wrong step.
// The MMA pipeline is the CTA-pair kernel of gemm_tcgen05.cu (cta_group::2, 256 x 256 tiles).
#include "common.cuh"

namespace b200 {

int make_tmap_2d(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                 int box_rows);

namespace fc {

constexpr int kBM = 128, kBK = 64, kStages = 5, kBN = 256, kHalfN = 128, kThreads = 256, kGroupM = 8;
constexpr int kABytes = kBM * kBK * 2, kBBytes = kHalfN * kBK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kOutBoxBytes = kBM * 128;                       // [128 rows x 64 bf16], 128B swizzle
constexpr int kOutOffset = kStages * kStageBytes;             // 4 boxes = the CTA's 128 x 256 tile
constexpr int kBarOffset = kOutOffset + 4 * kOutBoxBytes, kSmemTotal = kBarOffset + 256 + 1024;
constexpr int kMaxTP = 8;
constexpr int kModeAG = 1, kModeRS = 2, kModeAR = 3;

struct CommArgs {
  // TMA store maps: [0] = C (AG mode) or, in RS mode, [o] = owner o's slot for source <me>
  alignas(64) CUtensorMap out_map[kMaxTP];
  int mode, tp, rank, rows_per_chunk;
  unsigned epoch;
  // ---- all-gather ----
  const uint4* peer_src[kMaxTP];     // rank r's staging shard [rows_per_chunk, K]
  uint4* gathered;                   // local [M, K]
  unsigned* row_ready;               // [M / 128] warps that delivered their share of a row block
  unsigned* row_ready_next;          // the other counter set: zeroed here for the next call
  const unsigned* shard_ready;       // [tp] epoch at which rank r's staging shard became valid
  unsigned* chunk_done;              // [tp] warps that finished pulling chunk r
  unsigned* chunk_done_next;
  int max_blocks;
  unsigned* peer_consumed[kMaxTP];   // rank r's "rank <me> has pulled my shard" flag
  long long row_bytes;
  const uint4* x_local;              // the caller's [rows_per_chunk, K] shard
  uint4* stage_local;                // my staging shard of this call
  const unsigned* consumed_local;    // [tp] epoch of the last pull each peer finished on stage_local's buffer pair
  unsigned consumed_wait;            // peers must have finished this epoch's pull before stage_local is rewritten
  int do_consumed_wait;
  unsigned* signal_ready[kMaxTP];    // rank r's shard_ready[<me>] word
  unsigned* stage_done;              // local monotonic counter of copier warps that finished staging
  unsigned stage_done_target;
  // ---- reduce-scatter / all-reduce ----
  unsigned* cnt_dst[kMaxTP];         // owner o's arrival counters [rows_per_chunk / 128][tiles_n]
  const __nv_bfloat16* my_parts;     // my slots [tp][rows_per_chunk][N]
  const unsigned* my_cnt;            // counters start at 0; a tile is complete at tp - 1
  alignas(64) CUtensorMap final_map[kMaxTP];   // RS: [rank] = out [rows_per_chunk, N]; AR: [r] = rank r's out [M, N]
  unsigned* ar_cnt_dst[kMaxTP];      // AR: rank r's "final tiles delivered" counter
  const unsigned* my_ar_cnt;
  unsigned ar_target;                // AR: final tiles (128 x 256) other ranks deliver to me
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add(unsigned* p, unsigned v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() {
  asm volatile("fence.proxy.async.global;" ::: "memory");
}
// wrap-safe "counter has reached target"
__device__ __forceinline__ bool reached(unsigned v, unsigned target) {
  return static_cast<int>(v - target) >= 0;
}
// First wait that timed out: {site, blockIdx, value seen, target, epoch, mode, rank, index}.
// A timed-out wait is fatal: the record below is kept for b200_comm_debug_read (readable while the
// context is alive, i.e. by other streams' checks before the error surfaces), one line is printed
// and the kernel traps like csrc/comm.cu's waits do. Continuing with a partially gathered or
// partially reduced tile would train on garbage.
__device__ unsigned g_comm_debug[8];
constexpr long long kWaitLimitCycles = 120000000000LL;   // ~60 s at 1.97 GHz, ~80 s under a power cap

template <bool kSys>
__device__ __forceinline__ void spin_until(const unsigned* p, unsigned target, unsigned site,
                                           unsigned index, unsigned epoch, int mode, int rank) {
  const long long t0 = clock64();
  unsigned v;
  while (!reached(v = (kSys ? ld_acquire_sys(p) : ld_acquire_gpu(p)), target)) {
    __nanosleep(64);
    if (clock64() - t0 > kWaitLimitCycles) {
      if (atomicCAS(&g_comm_debug[0], 0u, site) == 0u) {
        g_comm_debug[1] = blockIdx.x;
        g_comm_debug[2] = v;
        g_comm_debug[3] = target;
        g_comm_debug[4] = epoch;
        g_comm_debug[5] = (unsigned)mode;
        g_comm_debug[6] = (unsigned)rank;
        g_comm_debug[7] = index;
        __threadfence_system();
        printf("b200 fused TP collective: wait timed out (site %u, rank %d, cta %u, saw %u, wanted %u, "
               "epoch %u, mode %d, index %u)\n", site, rank, blockIdx.x, v, target, epoch, mode, index);
      }
      asm volatile("trap;");
    }
  }
}

// All-gather mode, warps 2..7 of every CTA at kernel start (the epilogue warps have nothing to do
// until the first tile's MMAs are through): copy the rank's own rows into its staging buffer, where
// the peers pull them from, and into their place in the gathered matrix. The buffer's previous
// content (two calls ago) must have been pulled by every peer first; the last warp to finish raises
// the peers' shard_ready flags.
constexpr int kStageWarps = 6;
__device__ __forceinline__ void stage_own_rows(const CommArgs& ca, int w, int lane) {
  const int nw = kStageWarps * gridDim.x;
  if (ca.do_consumed_wait) {
    if (lane < ca.tp && lane != ca.rank)
      spin_until<true>(&ca.consumed_local[lane], ca.consumed_wait, 5u, lane, ca.epoch, ca.mode, ca.rank);
    __syncwarp();
  }
  const long long nseg = (long long)ca.rows_per_chunk * (ca.row_bytes / 512);   // 512 B = one warp access
  const uint4* xs = ca.x_local + lane;
  uint4* sd = ca.stage_local + lane;
  uint4* gd = ca.gathered + (long long)ca.rank * ca.rows_per_chunk * (ca.row_bytes / 16) + lane;
  constexpr int U0 = 8;
  for (long long sg = w; sg < nseg; sg += (long long)nw * U0) {
    uint4 v[U0];
#pragma unroll
    for (int u = 0; u < U0; ++u) {
      const long long g = sg + (long long)u * nw;
      if (g < nseg) v[u] = xs[g * 32];
    }
#pragma unroll
    for (int u = 0; u < U0; ++u) {
      const long long g = sg + (long long)u * nw;
      if (g < nseg) {
        sd[g * 32] = v[u];
        gd[g * 32] = v[u];
      }
    }
  }
  __threadfence_system();
  __syncwarp();
  if (lane == 0) {
    const unsigned old = atomicAdd(ca.stage_done, 1u);
    if (old + 1u == ca.stage_done_target) {
      __threadfence_system();
      for (int q = 1; q < ca.tp; ++q) st_release_sys(ca.signal_ready[(ca.rank + q) % ca.tp], ca.epoch);
    }
  }
}

__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int& tm, int& tn) {
  const int group = kGroupM * tiles_n;
  const int gid = tile / group;
  const int first_m = gid * kGroupM;
  const int gsz = min(tiles_m - first_m, kGroupM);
  const int in_group = tile - gid * group;
  tm = first_m + in_group % gsz;
  tn = in_group / gsz;
}

// Reduce modes at tp 2: position in a CTA pair's work sequence -> tile index. The plain order is
// "all tiles of the peer's rows, then all own tiles": every partial tile crosses NVLink in the first
// half of the kernel, which is then bound by the link (~64 KB per CTA and tile), while the link idles
// in the second half. Here the first `lag` positions are remote tiles, then remote and own tiles
// alternate, and the last `lag` positions are own tiles: the transfers spread over the kernel, and
// an own tile is still reached at least `lag` positions after the peer produced its partial
// (both ranks walk the same sequence). half = number of remote (= own) tiles.
__host__ __device__ __forceinline__ int interleaved_tile(int i, int half, int lag) {
  if (i < lag) return i;
  const int mid = 2 * (half - lag);
  if (i < lag + mid) {
    const int j = i - lag;
    return (j & 1) ? half + (j >> 1) : lag + (j >> 1);
  }
  return half + (half - lag) + (i - lag - mid);
}

template <bool kBMN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_comm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmAloc,
                 const __grid_constant__ CUtensorMap tmB, __nv_bfloat16* __restrict__ Cout, int M, int N,
                 int K, long long ldc, const __grid_constant__ CommArgs ca) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kBarOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int tiles_m = (M + 2 * kBM - 1) / (2 * kBM);
  const int tiles_n = (N + kBN - 1) / kBN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + kBK - 1) / kBK;
  const int tiles_per_chunk = ca.rows_per_chunk / (2 * kBM);
  // AG: start with the rank's own rows; RS: finish with them
  const int tm_rot = (ca.mode == kModeAG ? ca.rank : (ca.rank + 1) % ca.tp) * tiles_per_chunk;
  // tp 2 reduce modes with chunk-aligned tile groups: interleave remote and own tiles (see above)
  const bool interleave = ca.mode != kModeAG && ca.tp == 2 && (tiles_per_chunk % kGroupM) == 0;
  const int il_half = num_tiles >> 1;
  const int il_lag = min(num_pairs, il_half);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmAloc);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc_pair<2 * kBN>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const unsigned ready_target = 2 * gridDim.x;   // two copier warps per CTA, counters start at 0
      for (int seq = pair; seq < num_tiles; seq += num_pairs) {
        const int tile = interleave ? interleaved_tile(seq, il_half, il_lag) : seq;
        int tm, tn;
        tile_coords(tile, tiles_m, tiles_n, tm, tn);
        tm = (tm + tm_rot) % tiles_m;
        const int m0 = tm * 2 * kBM + static_cast<int>(crank) * kBM;
        const int n0 = tn * kBN + static_cast<int>(crank) * kHalfN;
        // own rows (all-gather mode) are read from the caller's shard, not from the gathered matrix
        const bool local_a = ca.mode == kModeAG && m0 / ca.rows_per_chunk == ca.rank;
        const CUtensorMap* ma = local_a ? &tmAloc : &tmA;
        const int ma_row = local_a ? m0 - ca.rank * ca.rows_per_chunk : m0;
        if (ca.mode == kModeAG && m0 < M && !local_a) {
          // rows pulled from another GPU by the copier warps: wait for this 128-row block
          spin_until<false>(&ca.row_ready[m0 / kBM], ready_target, 1u, m0 / kBM, ca.epoch, ca.mode, ca.rank);
          fence_proxy_async_global();
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          if (crank == 0) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);
          tma_load_2d_pair(sa, ma, &full_bar[stage], kb * kBK, ma_row);
          if (!kBMN) {
            tma_load_2d_pair(sb, &tmB, &full_bar[stage], kb * kBK, n0);
          } else {
#pragma unroll
            for (int j = 0; j < kHalfN / 64; ++j)
              tma_load_2d_pair(sb + j * (kBK * 128), &tmB, &full_bar[stage], n0 + j * 64, kb * kBK);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) --------------
    if (lane == 0 && crank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * kBM, kBN, 0, kBMN ? 1 : 0);
      constexpr uint32_t lbo_b = kBMN ? kBK * 128 : 16;
      constexpr uint32_t kstep_b = kBMN ? 16 * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kBN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint32_t sb = sa + kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = make_sw128_desc(sa + k * 32, 16, 1024);
            const uint64_t db = make_sw128_desc(sb + k * kstep_b, lbo_b, 1024);
            umma_bf16_ss_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);
          if (++stage == kStages) {
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
  } else if (warp == 2 || warp == 3) {
    // ------------------------------ all-gather copiers ------------------------
    if (ca.mode == kModeAG) {
      const int nw = 2 * gridDim.x, w = 2 * blockIdx.x + (warp - 2);
      if (w == 0) {   // nobody reads the other counter set during this call: clear it for the next
        for (int i = lane; i < ca.max_blocks; i += 32) ca.row_ready_next[i] = 0u;
        if (lane < kMaxTP) ca.chunk_done_next[lane] = 0u;
      }
      const long long segs_per_block = (long long)kBM * (ca.row_bytes / 512);  // 512 B = one warp access
      const int nrb = ca.rows_per_chunk / kBM;
      stage_own_rows(ca, blockIdx.x * kStageWarps + (warp - 2), lane);
      // Flat, software-pipelined pull: a unit = one pass of all copier warps over (part of) a 128-row
      // block of a remote chunk, chunks in the order the tiles need them. The loads of unit u+1 are in
      // flight (second register buffer) while unit u is stored, fenced and counted, so the NVLink
      // round trip is paid once per warp instead of once per block: at tp 8 the 14 blocks of a
      // 2048-row activation no longer queue behind each other.
      constexpr int U = 16;         // 16 x 512 B per warp and buffer: 2 x 2.4 MB in flight over the grid
      const long long per_pass = (long long)nw * U;
      const int sub_per_block = (int)((segs_per_block + per_pass - 1) / per_pass);
      const int units = (ca.tp - 1) * nrb * sub_per_block;
      const long long chunk_vec = (long long)ca.rows_per_chunk * (ca.row_bytes / 16);
      auto issue = [&](int unit, uint4* v) {
        const int blk = unit / sub_per_block, sub = unit - blk * sub_per_block;
        const int step = blk / nrb + 1, rb = blk - (step - 1) * nrb;
        const int src = (ca.rank + step) % ca.tp;
        if (sub == 0 && rb == 0) {
          if (lane == 0) spin_until<true>(&ca.shard_ready[src], ca.epoch, 2u, src, ca.epoch, ca.mode, ca.rank);
          __syncwarp();
        }
        const uint4* sp = ca.peer_src[src] + lane + (long long)rb * segs_per_block * 32;
        const long long g0 = (long long)sub * per_pass + w;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long g = g0 + (long long)u * nw;
          if (g < segs_per_block) v[u] = __ldcv(sp + g * 32);
        }
      };
      auto retire = [&](int unit, const uint4* v) {
        const int blk = unit / sub_per_block, sub = unit - blk * sub_per_block;
        const int step = blk / nrb + 1, rb = blk - (step - 1) * nrb;
        const int src = (ca.rank + step) % ca.tp;
        uint4* dp = ca.gathered + (long long)src * chunk_vec + lane + (long long)rb * segs_per_block * 32;
        const long long g0 = (long long)sub * per_pass + w;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long g = g0 + (long long)u * nw;
          if (g < segs_per_block) dp[g * 32] = v[u];
        }
        if (sub == sub_per_block - 1) {
          fence_proxy_async_global();   // the rows are read back by TMA (async proxy)
          __syncwarp();
          if (lane == 0) {
            red_release_gpu_add(&ca.row_ready[src * nrb + rb], 1u);
            if (rb == nrb - 1) {
              __threadfence();
              const unsigned old = atomicAdd(&ca.chunk_done[src], 1u);
              if (old + 1u == (unsigned)nw) st_release_sys(ca.peer_consumed[src], ca.epoch);
            }
          }
        }
      };
      uint4 va[U], vb[U];
      if (units > 0) issue(0, va);
      for (int unit = 0; unit < units; unit += 2) {
        if (unit + 1 < units) issue(unit + 1, vb);
        retire(unit, va);
        if (unit + 2 < units) issue(unit + 2, va);
        if (unit + 1 < units) retire(unit + 1, vb);
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue (both CTAs) ----------------------
    // TMEM -> registers -> bf16 -> swizzled shared-memory boxes -> TMA store (to the local C or
    // straight into the owning rank's partial-sum slot over NVLink: full 128-byte lines).
    const int ew = warp - 4;
    const int r = ew * 32 + lane;     // row inside the CTA's 128-row half tile == TMEM lane
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t leader_empty0 = mapa_u32(smem_u32(&tmem_empty[0]), 0);
    uint8_t* stage_out = smem + kOutOffset;
    const bool issuer = (ew == 0 && lane == 0);
    const bool reduce_mode = ca.mode != kModeAG;
    if (!reduce_mode) stage_own_rows(ca, blockIdx.x * kStageWarps + (warp - 2), lane);
    const size_t slot = (size_t)ca.rows_per_chunk * N;
    for (int seq = pair; seq < num_tiles; seq += num_pairs) {
      const int tile = interleave ? interleaved_tile(seq, il_half, il_lag) : seq;
      int tm, tn;
      tile_coords(tile, tiles_m, tiles_n, tm, tn);
      tm = (tm + tm_rot) % tiles_m;
      const int row0 = tm * 2 * kBM + static_cast<int>(crank) * kBM;
      const int n0 = tn * kBN;
      int owner = 0, drow = row0;
      if (reduce_mode) {
        owner = row0 / ca.rows_per_chunk;
        drow = row0 - owner * ca.rows_per_chunk;
      }
      const bool own = reduce_mode && owner == ca.rank;
      const int cidx = (drow / kBM) * tiles_n + tn;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      if (own) {
        // every peer's partial of this 128 x 256 tile has landed in my slots
        if (lane == 0)
          spin_until<true>(&ca.my_cnt[cidx], (unsigned)(ca.tp - 1), 3u, cidx, ca.epoch, ca.mode, ca.rank);
        __syncwarp();
      }
      if (issuer) tma_store_wait_read<0>();          // the previous tile's boxes have been read
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * kBN;
      const __nv_bfloat16* prow = ca.my_parts + (size_t)(drow + r) * N + n0;
#pragma unroll 1
      for (int c = 0; c < kBN / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c * 32, v);
        if (own) {
          // fp32 sum of the peers' bf16 partials (rank order), added to the accumulators
          float pacc[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) pacc[k] = 0.f;
          // four peers' loads are issued together (one L2 / HBM round trip per batch, not per peer)
          for (int pb = 0; pb < ca.tp; pb += 4) {
            uint4 q[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int p = pb + i;
              const bool on = p < ca.tp && p != ca.rank;
              const uint4* src = reinterpret_cast<const uint4*>(prow + (size_t)(on ? p : 0) * slot + c * 32);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                q[i][j] = (on && n0 + c * 32 + j * 8 < N) ? __ldcv(src + j) : make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                float f[8];
                unpack8(q[i][j], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) pacc[j * 8 + k] += f[k];
              }
            }
          }
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 32; ++k) v[k] = __float_as_uint(__uint_as_float(v[k]) + pacc[k]);
        } else {
          tmem_ld_wait();
        }
        uint8_t* box = stage_out + (c >> 1) * kOutBoxBytes;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
          o.y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
          o.z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
          o.w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
          *reinterpret_cast<uint4*>(box + sw128_offset(r, (c & 1) * 4 + q)) = o;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(leader_empty0 + as * 8);   // TMEM stage free for the MMAs
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (issuer && row0 < M) {
        if (!reduce_mode) {
#pragma unroll
          for (int bx = 0; bx < 4; ++bx)
            if (n0 + bx * 64 < N) tma_store_2d(&ca.out_map[0], stage_out + bx * kOutBoxBytes, n0 + bx * 64, row0);
          tma_store_commit();
        } else if (!own) {
          // partial tile -> the owner's slot for this source, then tell the owner
#pragma unroll
          for (int bx = 0; bx < 4; ++bx)
            if (n0 + bx * 64 < N) tma_store_2d(&ca.out_map[owner], stage_out + bx * kOutBoxBytes, n0 + bx * 64, drow);
          tma_store_commit();
          tma_store_wait<0>();              // the partial rows have been written ...
          fence_proxy_async_global();
          __threadfence_system();           // ... and are visible before the owner counts them
          red_release_sys_add(ca.cnt_dst[owner] + cidx, 1u);
        } else if (ca.mode == kModeRS) {
#pragma unroll
          for (int bx = 0; bx < 4; ++bx)
            if (n0 + bx * 64 < N)
              tma_store_2d(&ca.final_map[ca.rank], stage_out + bx * kOutBoxBytes, n0 + bx * 64, drow);
          tma_store_commit();
        } else {
          // all-reduce: the final tile goes to every rank's output (mine first)
          for (int q = 0; q < ca.tp; ++q) {
            const int dst = (ca.rank + q) % ca.tp;
#pragma unroll
            for (int bx = 0; bx < 4; ++bx)
              if (n0 + bx * 64 < N)
                tma_store_2d(&ca.final_map[dst], stage_out + bx * kOutBoxBytes, n0 + bx * 64, row0);
          }
          tma_store_commit();
          tma_store_wait<0>();
          fence_proxy_async_global();
          __threadfence_system();
          for (int q = 1; q < ca.tp; ++q) red_release_sys_add(ca.ar_cnt_dst[(ca.rank + q) % ca.tp], 1u);
        }
      }
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
    if (issuer) tma_store_wait<0>();
  }

  if (ca.mode == kModeAR && threadIdx.x == 0) {
    // the output is complete on this rank once every tile owned by another rank has arrived
    spin_until<true>(ca.my_ar_cnt, ca.ar_target, 4u, 0u, ca.epoch, ca.mode, ca.rank);
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<2 * kBN>(tmem_base);
  }
}

// bf16 [rows, cols] row stride ld: store box [128 rows x 64 cols], 128-byte swizzle
static int make_out_map(CUtensorMap* map, void* base, long long rows, long long cols, long long ld) {
  return make_tmap_2d(map, base, rows, cols, ld, kBM);
}

template <bool kBMN>
int launch(const CUtensorMap& ta, const CUtensorMap& taloc, const CUtensorMap& tb, void* C, int M, int N, int K,
           long long ldc, const CommArgs& ca, cudaStream_t stream) {
  auto kern = gemm_comm_kernel<kBMN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  // every CTA takes part in the collective, so the grid is always the full machine
  kern<<<kNumSMs, kThreads, kSmemTotal, stream>>>(ta, taloc, tb, (__nv_bfloat16*)C, M, N, K, ldc, ca);
  B200_CHECK_LAUNCH();
  return 0;
}

}  // namespace fc
}  // namespace b200

using namespace b200;
using b200::fc::CommArgs;

// Workspace layout (all pointers are device addresses valid on THIS GPU; peer_* entries point
// into the other ranks' workspaces through the IPC mappings of csrc/comm.cu):
//   flags (u32): row_ready[2][max_blocks] | shard_ready[8] | chunk_done[2][8] | consumed[8] |
//                (legacy rs_cnt[2][max_blocks]) | tile_cnt[2][max_blocks * max_tn] | ar_cnt[2] (+ pad)
struct B200CommDesc {
  int tp, rank;
  void* peer_stage[8];        // rank r's staging shard
  unsigned* peer_flags[8];    // rank r's flag block
  int max_blocks;
  int max_tn;                 // widest reduce-scattered output in 256-column tiles
};

static inline unsigned* f_row_ready(unsigned* f, int mb, int set) { return f + set * mb; }
static inline unsigned* f_shard_ready(unsigned* f, int mb) { return f + 2 * mb; }
static inline unsigned* f_chunk_done(unsigned* f, int mb, int set) { return f + 2 * mb + 8 + set * 8; }
static inline unsigned* f_consumed(unsigned* f, int mb) { return f + 2 * mb + 24; }
static inline unsigned* f_tile_cnt(unsigned* f, int mb, int mtn, int set) {
  return f + 4 * mb + 32 + set * mb * mtn;
}
static inline unsigned* f_ar_cnt(unsigned* f, int mb, int mtn, int set) {
  return f + 4 * mb + 32 + 2 * mb * mtn + set;
}

// Copies the 8-word record of the first timed-out wait on this GPU (all zero = none) and clears it.
B200_API int b200_comm_debug_read(unsigned* out8) {
  cudaError_t e = cudaMemcpyFromSymbol(out8, fc::g_comm_debug, 8 * sizeof(unsigned));
  if (e != cudaSuccess) return (int)e;
  unsigned z[8] = {};
  return (int)cudaMemcpyToSymbol(fc::g_comm_debug, z, sizeof(z));
}

// Host view of the tile order of the tp 2 reduce modes (tests): position i of a CTA pair's sequence ->
// tile index, for `half` remote + `half` own tiles and a lag of `lag` positions.
B200_API int b200_comm_interleaved_tile(int i, int half, int lag) { return fc::interleaved_tile(i, half, lag); }

B200_API int b200_comm_flag_words(int max_blocks, int max_tn) {
  return 4 * max_blocks + 32 + 2 * max_blocks * max_tn + 16;
}
// word offsets of the flag blocks the host side signals / waits on
B200_API int b200_comm_shard_ready_word(int max_blocks) { return 2 * max_blocks; }
B200_API int b200_comm_consumed_word(int max_blocks) { return 2 * max_blocks + 24; }

// C[M, N] = gathered[M, K] * B^T where gathered rows [r*M/tp, (r+1)*M/tp) are rank r's x_shard
// [M/tp, K]: the kernel stages the own shard for the peers, fills `gathered` (own rows copied, the
// others pulled over NVLink) and multiplies, own rows first. epoch: 1, 2, ... per call on this
// workspace, identical on all ranks; `d` must be the descriptor of staging buffer (epoch & 1);
// stage_done: a local u32 counter used only by this entry point (zero at workspace creation),
// stage_call: number of calls of this entry point on the workspace so far, this one included.
B200_API int b200_gemm_allgather(const B200CommDesc* d, const void* x_shard, void* gathered, const void* B,
                                 void* C, int M, int N, int K, int b_mn_major, long long ldb,
                                 long long ldc, unsigned epoch, unsigned* stage_done,
                                 unsigned stage_call, cudaStream_t stream) {
  if (!d || d->tp < 2 || d->tp > 8 || M % d->tp || !x_shard || !stage_done) return B200_ERR_BAD_ARG;
  const int rpc = M / d->tp;
  if (rpc % 256 || (K * 2) % 512 || M / 128 > d->max_blocks) return B200_ERR_UNSUPPORTED;
  CUtensorMap ta, taloc, tb;
  int rc = make_tmap_2d(&ta, gathered, M, K, K, fc::kBM);
  if (rc) return rc;
  rc = make_tmap_2d(&taloc, x_shard, rpc, K, K, fc::kBM);
  if (rc) return rc;
  rc = b_mn_major ? make_tmap_2d(&tb, B, K, N, ldb, fc::kBK) : make_tmap_2d(&tb, B, N, K, ldb, fc::kHalfN);
  if (rc) return rc;
  CommArgs ca = {};
  ca.mode = fc::kModeAG;
  ca.tp = d->tp;
  ca.rank = d->rank;
  ca.rows_per_chunk = rpc;
  ca.epoch = epoch;
  unsigned* mine = d->peer_flags[d->rank];
  for (int r = 0; r < d->tp; ++r) {
    ca.peer_src[r] = (const uint4*)d->peer_stage[r];
    ca.peer_consumed[r] = f_consumed(d->peer_flags[r], d->max_blocks) + d->rank;
    ca.signal_ready[r] = f_shard_ready(d->peer_flags[r], d->max_blocks) + d->rank;
  }
  ca.gathered = (uint4*)gathered;
  const int set = (int)(epoch & 1u);   // per-call counters alternate; each call clears the other set
  ca.row_ready = f_row_ready(mine, d->max_blocks, set);
  ca.row_ready_next = f_row_ready(mine, d->max_blocks, set ^ 1);
  ca.shard_ready = f_shard_ready(mine, d->max_blocks);
  ca.chunk_done = f_chunk_done(mine, d->max_blocks, set);
  ca.chunk_done_next = f_chunk_done(mine, d->max_blocks, set ^ 1);
  ca.max_blocks = d->max_blocks;
  ca.row_bytes = (long long)K * 2;
  ca.x_local = (const uint4*)x_shard;
  ca.stage_local = (uint4*)d->peer_stage[d->rank];
  ca.consumed_local = f_consumed(mine, d->max_blocks);
  ca.consumed_wait = epoch - 2u;       // the previous user of this staging buffer
  ca.do_consumed_wait = epoch > 2u;
  ca.stage_done = stage_done;
  // warps 2..7 of every CTA add one per call of THIS entry point (stage_call = 1, 2, ...; the epoch
  // sequence is shared with other users of the staging buffers and may skip numbers here)
  ca.stage_done_target = stage_call * (unsigned)(fc::kStageWarps * kNumSMs);
  rc = fc::make_out_map(&ca.out_map[0], C, M, N, ldc);
  if (rc) return rc;
  return b_mn_major ? fc::launch<true>(ta, taloc, tb, C, M, N, K, ldc, ca, stream)
                    : fc::launch<false>(ta, taloc, tb, C, M, N, K, ldc, ca, stream);
}

// Shared set-up of the reduce-scatter / all-reduce GEMMs. peer_parts[r] = rank r's partial-sum
// slots of `set`; the workspace has two sets of slots and counters and callers alternate `set`
// call by call, which is enough to keep a fast rank from touching state a slow rank still reads
// (a rank can finish call e+1 only after every peer has delivered its partials of call e+1, i.e.
// after every peer has finished call e). The owner zeroes the set's counters after its kernel.
static int reduce_gemm(const B200CommDesc* d, void* const* peer_parts, void* const* peer_out, int set,
                       const void* A, const void* B, void* out, int M, int N, int K, int b_mn_major,
                       long long lda, long long ldb, cudaStream_t stream) {
  if (!d || d->tp < 2 || d->tp > 8 || M % d->tp || set < 0 || set > 1) return B200_ERR_BAD_ARG;
  const int rpc = M / d->tp;
  const int tiles_n = (N + fc::kBN - 1) / fc::kBN;
  if (rpc % 256 || N % 8 || M / 128 > d->max_blocks || tiles_n > d->max_tn || (lda % 8) || (ldb % 8))
    return B200_ERR_UNSUPPORTED;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d(&ta, A, M, K, lda, fc::kBM);
  if (rc) return rc;
  rc = b_mn_major ? make_tmap_2d(&tb, B, K, N, ldb, fc::kBK) : make_tmap_2d(&tb, B, N, K, ldb, fc::kHalfN);
  if (rc) return rc;
  CommArgs ca = {};
  ca.mode = peer_out ? fc::kModeAR : fc::kModeRS;
  ca.tp = d->tp;
  ca.rank = d->rank;
  ca.rows_per_chunk = rpc;
  const size_t slot = (size_t)rpc * N;
  for (int o = 0; o < d->tp; ++o) {
    rc = fc::make_out_map(&ca.out_map[o], (__nv_bfloat16*)peer_parts[o] + slot * d->rank, rpc, N, N);
    if (rc) return rc;
    ca.cnt_dst[o] = f_tile_cnt(d->peer_flags[o], d->max_blocks, d->max_tn, set);
    if (peer_out) {
      rc = fc::make_out_map(&ca.final_map[o], peer_out[o], M, N, N);
      if (rc) return rc;
      ca.ar_cnt_dst[o] = f_ar_cnt(d->peer_flags[o], d->max_blocks, d->max_tn, set);
    }
  }
  ca.my_parts = (const __nv_bfloat16*)peer_parts[d->rank];
  unsigned* my_cnt = f_tile_cnt(d->peer_flags[d->rank], d->max_blocks, d->max_tn, set);
  unsigned* my_ar = f_ar_cnt(d->peer_flags[d->rank], d->max_blocks, d->max_tn, set);
  ca.my_cnt = my_cnt;
  ca.my_ar_cnt = my_ar;
  ca.ar_target = (unsigned)(((M - rpc) / fc::kBM) * tiles_n);
  if (!peer_out) {
    rc = fc::make_out_map(&ca.final_map[d->rank], out, rpc, N, N);
    if (rc) return rc;
  }
  rc = b_mn_major ? fc::launch<true>(ta, ta, tb, nullptr, M, N, K, N, ca, stream)
                  : fc::launch<false>(ta, ta, tb, nullptr, M, N, K, N, ca, stream);
  if (rc) return rc;
  // peers touch this set again two calls from now at the earliest (see above)
  cudaError_t e = cudaMemsetAsync(my_cnt, 0, sizeof(unsigned) * (rpc / fc::kBM) * tiles_n, stream);
  if (e == cudaSuccess && peer_out) e = cudaMemsetAsync(my_ar, 0, sizeof(unsigned), stream);
  return (int)e;
}

// out[M/tp, N] = rows chunk <rank> of sum over ranks of A_r[M, K] * B_r^T.
B200_API int b200_gemm_reduce_scatter(const B200CommDesc* d, void* const* peer_parts, int set,
                                      const void* A, const void* B, void* out, int M, int N, int K,
                                      int b_mn_major, long long lda, long long ldb,
                                      cudaStream_t stream) {
  return reduce_gemm(d, peer_parts, nullptr, set, A, B, out, M, N, K, b_mn_major, lda, ldb, stream);
}

// Every rank's peer_out[r] (rank r's [M, N] output buffer of `set`, peer-mapped) receives
// sum over ranks of A_r[M, K] * B_r^T: the reduce-scatter above with the owner of a row chunk
// storing its final tiles into all tp outputs.
B200_API int b200_gemm_allreduce(const B200CommDesc* d, void* const* peer_parts, void* const* peer_out,
                                 int set, const void* A, const void* B, int M, int N, int K,
                                 int b_mn_major, long long lda, long long ldb, cudaStream_t stream) {
  if (!peer_out) return B200_ERR_BAD_ARG;
  return reduce_gemm(d, peer_parts, peer_out, set, A, B, nullptr, M, N, K, b_mn_major, lda, ldb, stream);
}
