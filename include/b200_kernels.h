/* C ABI of libb200kernels.so (sm_100a only).
 *
 * Conventions
 *   - every entry point returns 0 on success, a positive cudaError_t when a launch or a
 *     runtime call fails, or a negative argument error:
 *         -1 bad argument, -2 unsupported shape / alignment, -3 size beyond a stated maximum;
 *   - all tensors are device pointers; bf16 tensors are `void*`, fp32 tensors `float*`;
 *   - the last argument is the CUDA stream the work is enqueued on; nothing synchronises;
 *   - unless stated otherwise the library allocates nothing: scratch is passed in.
 * Each comment names the reference code whose job the kernel does (paths relative to the
 * reference tree).
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- norms (csrc/norm.cu) ------------------------------------------------------------ */
/* rows of the [grid, H] fp32 scratch the backward kernels need for M rows */
int b200_norm_bwd_grid(int M);
/* y = rmsnorm(x [+ residual]) * w; h_out = x + residual when residual != NULL.
 * Reference: megatron/model/fused_layer_norm.py:125-139 (eager RMSNorm). */
int b200_rmsnorm_fwd(const void* x, const void* residual, const void* w, void* y, void* h_out,
                     float* rstd, int M, int H, float eps, cudaStream_t stream);
/* dx (+= dres), dw (+= when accumulate). scratch: fp32 [b200_norm_bwd_grid(M), H]. */
int b200_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                     const void* dres, void* dx, float* dw, float* scratch, int M, int H,
                     int accumulate, cudaStream_t stream);
/* RMSNorm backward for callers that cannot allocate device memory (a launch sequence under
 * recording, a host with its memory budget already handed out, a plug-in that only receives input /
 * output pointers). No counterpart in the reference. Same results as b200_rmsnorm_bwd, but the
 * [rows, H] fp32 partial-gradient scratch is a 148 x 8192 array built into the library: it occupies
 * 4.6 MiB on every GPU that loads the library until the process exits. H > 8192
 * (b200_rmsnorm_bwd_noalloc_max_h()) returns -3. The entry point calls no allocator. All calls on a
 * GPU share the one array: it needs no clearing between calls, but calls that overlap on the GPU
 * (different streams / threads) race -- keep them on one stream. The first call on each GPU looks the
 * array's device address up (cudaGetSymbolAddress, that GPU current); make it before any stream capture. */
int b200_rmsnorm_bwd_noalloc_max_h(void);
int b200_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* rstd,
                             const void* dres, void* dx, float* dw, int M, int H, int accumulate,
                             cudaStream_t stream);
/* Reference: megatron/fused_kernels/layer_norm_cuda_kernel.cu (cuApplyLayerNorm, cuComputeGradInput,
 * cuComputePartGradGammaBeta). scratch: fp32 [b200_norm_bwd_grid(M), 2, H]. */
int b200_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean,
                       float* rstd, int M, int H, float eps, cudaStream_t stream);
int b200_layernorm_bwd(const void* dy, const void* x, const void* w, const float* mean,
                       const float* rstd, void* dx, float* dgamma, float* dbeta, float* scratch,
                       int M, int H, int accumulate, cudaStream_t stream);

/* ---- elementwise (csrc/elementwise.cu) ------------------------------------------------ */
/* In-place RoPE on x [S*B tokens, G, nslot, HN]; the first nrot slots of each group are rotated.
 * Reference: megatron/model/positional_embeddings.py:24-51. */
int b200_rope(void* x, const float* cos_sin, const long long* pos_ids, int S, int B, int G,
              int nslot, int nrot, int HN, int pos_offset, int max_pos, int inverse,
              cudaStream_t stream);
/* kind: 0 liglu, 1 geglu, 2 reglu, 3 swiglu. Reference: megatron/model/glu_activations.py:11-54. */
int b200_glu_fwd(const void* in, void* out, long long M, int F, int kind, cudaStream_t stream);
int b200_glu_bwd(const void* dy, const void* in, void* din, long long M, int F, int kind,
                 cudaStream_t stream);
/* Reference: megatron/model/fused_bias_gelu.py:14-43. */
int b200_bias_gelu_fwd(const void* x, const void* bias, void* y, long long M, int H,
                       int tanh_approx, cudaStream_t stream);
int b200_bias_gelu_bwd(const void* dy, const void* x, const void* bias, void* dx, long long M,
                       int H, int tanh_approx, cudaStream_t stream);
/* out = residual + dropout(x + bias); Philox mask from (seed, offset), regenerated in backward.
 * Reference: megatron/model/transformer.py:532-578. */
int b200_dropout_add_fwd(const void* x, const void* bias, const void* residual, void* out,
                         long long M, int H, float p, unsigned long long seed,
                         unsigned long long offset, cudaStream_t stream);
int b200_dropout_bwd(const void* dy, void* dx, long long M, int H, float p,
                     unsigned long long seed, unsigned long long offset, cudaStream_t stream);
/* Reference: megatron/model/distributed.py:174-184 (main_grad.add_) and the fp32<->bf16 copies of
 * megatron/optimizer/optimizer.py:48-51. */
int b200_accum_bf16_to_f32(const void* src, float* dst, long long n, float scale, int accumulate,
                           cudaStream_t stream);
int b200_cast_f32_to_bf16(const float* src, void* dst, long long n, cudaStream_t stream);

/* ---- softmax (csrc/softmax.cu) -------------------------------------------------------- */
/* mode 0 plain, 1 causal, 2 masked (uint8 mask [mask_batch,1,sq,sk]). Reference:
 * megatron/fused_kernels/scaled_{upper_triang_,}masked_softmax.h, scaled_softmax_cuda.cu. */
int b200_softmax_fwd(const void* x, const uint8_t* mask, void* y, int b, int np, int sq, int sk,
                     int mask_batch, float scale, int mode, cudaStream_t stream);
int b200_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int sk, float scale,
                     cudaStream_t stream);

/* ---- vocab-parallel cross entropy (csrc/xent.cu) --------------------------------------- */
/* Reference: megatron/core/tensor_parallel/cross_entropy.py:16-126. */
int b200_xent_stats(const void* logits, const long long* labels, float* stats, int T, int V,
                    long long vocab_start, cudaStream_t stream);
int b200_xent_combine(const float* all_stats, float* loss, float* lse, int T, int tp,
                      cudaStream_t stream);
/* Exchange inside the kernels (tensor-parallel group on one NVSwitch domain): the statistics pass stores its
 * triples into every rank's peer-mapped table and counts itself in there; the combine waits for `target`
 * arrivals (running total: tp * b200_xent_push_grid(T) per call on that counter). */
int b200_xent_push_grid(int T);
int b200_xent_stats_push(const void* logits, const long long* labels, float* const* peer_stats,
                         unsigned* const* peer_cnt, int tp, int rank, int T, int V, long long vocab_start,
                         cudaStream_t stream);
int b200_xent_combine_wait(const float* all_stats, const unsigned* cnt, unsigned target, float* loss, float* lse,
                           int T, int tp, cudaStream_t stream);
int b200_xent_bwd(void* logits, const long long* labels, const float* lse, const float* dloss,
                  int T, int V, long long vocab_start, cudaStream_t stream);

/* ---- optimizer (csrc/adam.cu) ---------------------------------------------------------- */
/* Reference: apex FusedAdam / amp_C as used by megatron/optimizer/__init__.py:75-86 and
 * megatron/optimizer/clip_grads.py:75-105. coef/skip are DEVICE scalars (may be NULL). */
int b200_adam_step(float* p, const float* g, float* m, float* v, void* p16, long long n, float lr,
                   float beta1, float beta2, float eps, float wd, int step, int adamw,
                   int bias_correction, const float* coef, const float* skip, cudaStream_t stream);
int b200_sgd_step(float* p, const float* g, float* mom, void* p16, long long n, float lr,
                  float momentum, float wd, int first_step, const float* coef, const float* skip,
                  cudaStream_t stream);
int b200_sumsq_scratch_elems(void);
/* out[0] (+)= sum(x^2), out[1] = non-finite flag; scratch: fp32 [b200_sumsq_scratch_elems()]. */
int b200_sumsq(const float* x, long long n, float* out, float* scratch, int accumulate,
               cudaStream_t stream);
int b200_scale(float* x, long long n, const float* coef, float host_coef, cudaStream_t stream);

/* ---- tcgen05 GEMM (csrc/gemm_tcgen05.cu) ----------------------------------------------- */
/* C[M,N] (op)= A_op[M,K] * B_op[N,K]^T, bf16 in, fp32 accumulate in TMEM.
 * a_mn_major / b_mn_major: operand stored with its M / N dimension contiguous (i.e. [K, M] / [K, N]).
 * out_mode: 0 store bf16, 1 store fp32, 2 accumulate into fp32 (wgrad into main_grad).
 * Reference: torch.matmul call sites megatron/core/tensor_parallel/layers.py:240,267,307 and
 * megatron/fused_kernels/fused_weight_gradient_dense.cu:128-152. */
int b200_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                   int b_mn_major, long long lda, long long ldb, long long ldc, int out_mode,
                   int max_ctas, cudaStream_t stream);
/* Same GEMM with the kernel chosen by the caller: variant 0 = automatic (CTA pairs, i.e.
 * tcgen05.mma.cta_group::2 over 256 x 256 tiles, whenever M >= 256 and N >= 256), 1 = single-CTA
 * 128 x 256 tiles, 2 = CTA pairs. Same reference call sites as b200_gemm_bf16. */
int b200_gemm_bf16_variant(const void* A, const void* B, void* C, int M, int N, int K,
                           int a_mn_major, int b_mn_major, long long lda, long long ldb,
                           long long ldc, int out_mode, int max_ctas, int variant,
                           cudaStream_t stream);

/* ---- peer memory + fused tensor-parallel GEMMs (csrc/comm.cu, csrc/gemm_comm_tcgen05.cu) ----
 * The reference moves every tensor-parallel activation through NCCL calls around cuBLAS GEMMs
 * (megatron/core/tensor_parallel/layers.py:225-243, :691-692; mappings.py:89-124); here the gather /
 * scatter happens inside the GEMM kernel over CUDA-IPC mapped peer memory. */
int b200_symm_alloc(long long bytes, void** out);            /* cudaMalloc + zero */
int b200_symm_free(void* p);
int b200_symm_export(void* p, void* handle64);               /* cudaIpcMemHandle_t, 64 bytes */
int b200_symm_import(const void* handle64, void** out);      /* map a peer's region */
int b200_symm_unmap(void* p);
int b200_symm_signal(unsigned* const* addrs, int n, unsigned value, cudaStream_t stream);
int b200_symm_wait(const unsigned* flags, int n, int skip, unsigned value, cudaStream_t stream);
int b200_copy2(const void* src, void* dst0, void* dst1, long long bytes, cudaStream_t stream);
int b200_stage_shard(const void* src, void* stage, void* own_rows, long long bytes, const unsigned* consumed,
                     int tp, int rank, unsigned wait_value, int do_wait, unsigned* done_counter,
                     unsigned call_index, unsigned* const* signal_addrs, int n_signal, unsigned epoch,
                     cudaStream_t stream);
int b200_comm_interleaved_tile(int i, int half, int lag);  /* tile order of the tp 2 reduce GEMMs (host view) */
int b200_comm_flag_words(int max_blocks, int max_tn);     /* max_tn: widest reduced output / 256 columns */
int b200_comm_shard_ready_word(int max_blocks);
int b200_comm_consumed_word(int max_blocks);
int b200_comm_debug_read(unsigned* out8);
typedef struct { int tp, rank; void* peer_stage[8]; unsigned* peer_flags[8]; int max_blocks; int max_tn; } B200CommDesc;
int b200_gemm_allgather(const B200CommDesc* d, const void* x_shard, void* gathered, const void* B, void* C,
                        int M, int N, int K, int b_mn_major, long long ldb, long long ldc, unsigned epoch,
                        unsigned* stage_done, unsigned stage_call, cudaStream_t stream);
/* GEMM -> reduce-scatter (megatron/core/tensor_parallel/layers.py:691-692, :285-296): partial tiles of rows
 * another rank owns are TMA-stored into that rank's slots; the owner adds them in the epilogue of its own
 * tiles. b200_gemm_allreduce (layers.py:694, :278-283, no sequence parallelism): the owner additionally
 * stores every final tile into peer_out[r] ([M, N], peer-mapped) of all ranks. `set` alternates 0/1 per call. */
int b200_gemm_reduce_scatter(const B200CommDesc* d, void* const* peer_parts, int set, const void* A,
                             const void* B, void* out, int M, int N, int K, int b_mn_major, long long lda,
                             long long ldb, cudaStream_t stream);
int b200_gemm_allreduce(const B200CommDesc* d, void* const* peer_parts, void* const* peer_out, int set,
                        const void* A, const void* B, int M, int N, int K, int b_mn_major, long long lda,
                        long long ldb, cudaStream_t stream);

/* ---- FP8 (csrc/fp8.cu, csrc/gemm_tcgen05.cu) ---------------------------------------------------------------
 * The reference reaches FP8 through Transformer Engine only (megatron/model/transformer.py:931-950).
 * b200_fp8_quantize: y[rows, cols] = fp8(x * *scale) (+ yT[cols, rows], + *amax = max(*amax, max|x|)) in one pass
 *   over bf16 x; rows % 16 == 0, cols % 16 == 0; e5m2: 0 = E4M3, 1 = E5M2; saturating.
 * b200_gemm_fp8: C[M, N] = (*alpha_a * *alpha_b) * A[M, K] B[N, K]^T, both operands one byte per element and
 *   K-major (strides in bytes, multiples of 16), fp32 accumulation (tcgen05.mma.kind::f8f6f4, CTA pairs);
 *   out_mode 0 bf16, 1 fp32, 2 fp32 accumulate; alpha_* device pointers or NULL. */
int b200_fp8_quantize(const void* x, void* y, void* yT, int rows, int cols, long long ldx, const float* scale,
                      float* amax, int e5m2, cudaStream_t stream);
int b200_gemm_fp8(const void* A, const void* B, void* C, int M, int N, int K, int a_e5m2, int b_e5m2, long long lda,
                  long long ldb, long long ldc, int out_mode, const float* alpha_a, const float* alpha_b,
                  cudaStream_t stream);

/* ---- vocab-parallel embedding over peer memory (csrc/embed.cu) -------------------------------------------
 * Replaces megatron/core/tensor_parallel/layers.py:187-210 (masked local lookup + all-reduce) and its backward.
 * b200_ipc_export_range: CUDA IPC handle of the allocation containing `ptr` + the offset of `ptr` inside it.
 * b200_embed_fwd: out[(si - s_off) * B + bi] = tables[id / Vp][id % Vp], id = ids[bi * S + si] (ids [B, S] int64,
 *   tables[r] = rank r's [Vp, H] bf16 shard, peer-mapped); out is [s_out * B, H] in [s, b, h] order.
 * b200_embed_bwd: main_grad[id - lo] (fp32 [Vp, H]) += output-gradient row of every token with lo <= id < lo + Vp;
 *   srcs[c] = rows of sequence chunk c (local tensor or a peer's staging buffer), vector atomics. */
int b200_ipc_export_range(const void* ptr, void* handle64, long long* offset_out);
int b200_embed_fwd(const long long* ids, const void* const* tables, void* out, int S, int B, int s_off, int s_out,
                   int H, int Vp, int tp, cudaStream_t stream);
int b200_embed_bwd(const long long* ids, const void* const* srcs, int n_chunks, float* main_grad, int S, int B,
                   int H, long long lo, int Vp, cudaStream_t stream);

/* ---- data-parallel reduce-scatter / Adam with parameter all-gather over peer memory (csrc/dp_comm.cu) --
 * Replaces megatron/optimizer/distrib_optimizer.py:553-567 and :592-608. Buffers come from b200_symm_alloc and
 * are mapped by the peers (b200_symm_export / b200_symm_import).
 * b200_dp_reduce: in place, grads[rank][i] = scale * sum_r grads[r][i] for n floats (n % 4 == 0); grads[r] is
 *   rank r's address of the slice (peer-mapped or local staging); waits until the local flag words ready[r]
 *   (r != rank) have reached `epoch`; `blocks` = grid size.
 * b200_adam_step_push: b200_adam_step on n elements (n % 8 == 0) whose bf16 result is stored to n_dst
 *   parameter buffers (own + peers).
 * b200_dp_wait: bounded stream-ordered wait for flags[i] >= value, i < n, i != skip; a time-out is recorded
 *   (site, value seen, target, index, block) and returned once by b200_dp_debug_read (all zero: none).
 * b200_copy_async: copy-engine transfer, src may be peer-mapped. */
int b200_dp_reduce(const void* const* grads, int dp, int rank, long long n, float scale, const unsigned* ready,
                   unsigned epoch, int blocks, cudaStream_t stream);
int b200_adam_step_push(float* p, const float* g, float* m, float* v, void* const* p16_dst, int n_dst,
                        long long n, float lr, float beta1, float beta2, float eps, float wd, int step,
                        int adamw, int bias_correction, const float* coef, const float* skip,
                        cudaStream_t stream);
int b200_dp_wait(const unsigned* flags, int n, int skip, unsigned value, unsigned site, cudaStream_t stream);
int b200_copy_async(void* dst, const void* src, long long bytes, cudaStream_t stream);
int b200_dp_debug_read(unsigned* out8);

/* ---- tcgen05 attention (csrc/attn_fwd_tcgen05.cu) --------------------------------------- */
/* Causal / full GQA attention forward reading q, k, v through arbitrary element strides
 * (the fused QKV projection layout [s, b, g, q_per_g+2, hn] needs no copies). hn must be 128.
 * out [sq, B, G*qpg*hn] bf16, lse [B, G*qpg, sq] fp32 (natural log).
 * Reference: flash_attn_func call at megatron/model/transformer.py:514-522 plus the K/V expansion and
 * transposes at :448-455,515-522. */
int b200_attn_fwd(const void* q, const void* k, const void* v, void* out, float* lse, int sq, int sk,
                  int B, int G, int qpg, int hn, long long q_ss, long long q_sb, long long q_sg,
                  long long q_sj, long long k_ss, long long k_sb, long long k_sg,
                  float softmax_scale, int causal, cudaStream_t stream);

/* Second-generation forward (csrc/attn_fwd_v2_tcgen05.cu): two 128-row query tiles per CTA ping-pong
 * through the tensor core, P stays in TMEM, KV blocks of 128. Same arguments and reference call site. */
int b200_attn_fwd_v2(const void* q, const void* k, const void* v, void* out, float* lse, int sq, int sk,
                     int B, int G, int qpg, int hn, long long q_ss, long long q_sb, long long q_sg,
                     long long q_sj, long long k_ss, long long k_sb, long long k_sg,
                     float softmax_scale, int causal, cudaStream_t stream);

/* Attention backward (csrc/attn_bwd_tcgen05.cu): dq/dk/dv written with their own element strides,
 * e.g. straight into the gradient of the fused QKV tensor; the GQA reduction of dk/dv over the
 * query heads of a group happens in TMEM (plus fp32 partial sums in scratch when few KV heads would
 * leave SMs idle). scratch: fp32 [b200_attn_bwd_scratch_elems(sq, sk, B, G, qpg)].
 * out, dout: [sq, B, G*qpg*hn] bf16 contiguous; lse from b200_attn_fwd. hn must be 128.
 * Reference: the FlashAttention-2 backward behind flash_attn_func, megatron/model/transformer.py:514-522. */
long long b200_attn_bwd_scratch_elems(int sq, int sk, int B, int G, int qpg);
int b200_attn_bwd(const void* q, const void* k, const void* v, const void* out, const void* dout,
                  const float* lse, void* dq, void* dk, void* dv, float* scratch, int sq, int sk,
                  int B, int G, int qpg, int hn, long long q_ss, long long q_sb, long long q_sg,
                  long long q_sj, long long k_ss, long long k_sb, long long k_sg, long long dq_ss,
                  long long dq_sb, long long dq_sg, long long dq_sj, long long dk_ss, long long dk_sb,
                  long long dk_sg, float softmax_scale, int causal, cudaStream_t stream);
/* Second-generation backward kernels (64-row streamed blocks, S / dP double-buffered in tensor memory): same
 * contract, scratch and result; selected with B200_ATTN_BWD_ENTRY=b200_attn_bwd_v2 for A/B runs. */
int b200_attn_bwd_v2(const void* q, const void* k, const void* v, const void* out, const void* dout,
                     const float* lse, void* dq, void* dk, void* dv, float* scratch, int sq, int sk,
                     int B, int G, int qpg, int hn, long long q_ss, long long q_sb, long long q_sg,
                     long long q_sj, long long k_ss, long long k_sb, long long k_sg, long long dq_ss,
                     long long dq_sb, long long dq_sg, long long dq_sj, long long dk_ss, long long dk_sb,
                     long long dk_sg, float softmax_scale, int causal, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
