/* dsb200.h — C ABI of the deeperspeed_b200 native kernel library (libdsb200.so, sm_100a).
 *
 * Every entry point takes raw device pointers plus an opaque `stream` (a cudaStream_t cast to void*; NULL is the
 * legacy default stream) and returns 0 on success or a negative DSB_ERR_* / -cudaError code. Nothing here
 * synchronises the device. Python binds these through ctypes (deeperspeed_b200/ops/native.py).
 *
 * Citations name the reference implementation (MLRichter/DeeperSpeed) whose behaviour each entry point covers.
 */
#ifndef DSB200_H
#define DSB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSB_OK 0
#define DSB_ERR_ALIGN (-2001)  /* pointer / leading dimension not 16-byte aligned */
#define DSB_ERR_ARG (-2002)    /* inconsistent arguments */
#define DSB_ERR_SIZE (-2003)   /* problem size outside what the entry point supports */
#define DSB_ERR_DTYPE (-2004)

#define DSB_DTYPE_F16 0
#define DSB_DTYPE_BF16 1
#define DSB_DTYPE_F32 2

/* ---- tcgen05 GEMM with fused epilogues (csrc/gemm/gemm_tcgen05.cu) --------------------------------------------
 * D[b] = epi(alpha * A[b] * B[b]^T), A: M x K, B: N x K. a_mn_major / b_mn_major = 1 means the operand is stored
 * [K, rows] row-major (i.e. "transposed" in BLAS terms). Replaces cublasGemmEx/cublasGemmStridedBatchedEx
 * (reference csrc/transformer/cublas_wrappers.cu:16-164) plus the SIMT epilogues that followed them. */
#define DSB_EPI_NONE 0
#define DSB_EPI_BIAS 1          /* + bias[n] */
#define DSB_EPI_BIAS_GELU 2     /* gelu(acc + bias); optional aux_out receives the pre-activation */
#define DSB_EPI_BIAS_RESIDUAL 3 /* acc (+ bias) + aux[m, n]; aux may alias D (accumulate) */
#define DSB_EPI_DGELU 4         /* acc * gelu'(aux[m, n]) */
int dsb_gemm(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb, long ldd,
             long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major, int dtype, int epilogue,
             float alpha, const void* bias, const void* aux, long ldaux, long stride_aux, void* aux_out,
             long ldaux_out, long stride_aux_out, void* stream);

/* ---- optimizers (csrc/optim) -------------------------------------------------------------------------------- */
/* Flat-arena Adam/AdamW: fp32 master/m/v, grads in g_dtype, optional low-precision copy of the updated params.
 * inv_scale_dev (optional, device float) multiplies the gradient (1 / combined loss scale); skip_dev (optional,
 * device int) != 0 turns the launch into a no-op (overflow skip without a host sync).
 * Reference: csrc/adam/multi_tensor_adam.cu:28-121. */
int dsb_adam_flat(float* p, float* m, float* v, const void* g, int g_dtype, void* p_lp, int lp_dtype, long n,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int adamw_mode,
                  int bias_correction, int step, float inv_scale, const float* inv_scale_dev, const int* skip_dev,
                  void* stream);
/* Multi-tensor Adam over a host table of device pointers (params updated in their own dtype, fp32 math).
 * Reference: csrc/adam/multi_tensor_apply.cuh:43-127. */
int dsb_adam_multi(int ntensors, void* const* p, void* const* g, void* const* m, void* const* v, const long* numel,
                   int p_dtype, int g_dtype, int state_dtype, float lr, float beta1, float beta2, float eps,
                   float weight_decay, int adamw_mode, int bias_correction, int step, float inv_scale, void* stream);
/* LAMB for one tensor; `scratch` needs dsb_lamb_scratch_bytes(n) bytes. coeff_out (device float) gets the trust
 * ratio. Reference: csrc/lamb/fused_lamb_cuda_kernel.cu:186-310. */
long dsb_lamb_scratch_bytes(long n);
int dsb_lamb_step(void* p, int p_dtype, void* p_copy, int copy_dtype, const void* g, int g_dtype, float* m, float* v,
                  long n, float lr, float beta1, float beta2, float eps, float weight_decay, float max_coeff,
                  float min_coeff, int eps_inside_sqrt, int bias_correction, int step, float grad_scale,
                  float* coeff_out, void* scratch, void* stream);
/* dtype cast (fp32 -> fp16/bf16 param write-back; reference csrc/adam/custom_cuda_kernel.cu:5-20). */
int dsb_cast(const void* src, int src_dtype, void* dst, int dst_dtype, long n, void* stream);
/* sum of squares + inf/nan flag of a flat buffer, no host sync. `scratch` needs dsb_sumsq_scratch_bytes() bytes.
 * out[0] += sumsq (or = if !accumulate), flag_out |= non-finite seen. Replaces the per-tensor host-synchronising
 * overflow checks (reference deepspeed/runtime/utils.py:157-175, zero/stage2.py:1600-1618). */
long dsb_sumsq_scratch_bytes(void);
int dsb_sumsq(const void* x, int dtype, long n, float* out, int* flag_out, int accumulate, void* scratch,
              void* stream);
/* x *= scale (device or host scalar), in place. */
int dsb_scale(void* x, int dtype, long n, float scale, const float* scale_dev, void* stream);

/* ---- transformer building blocks (csrc/transformer) --------------------------------------------------------- */
/* y = LN(x (+ residual)) * gamma + beta; optionally writes sum_out = x + residual, mean/rstd for backward.
 * Reference: csrc/transformer/normalize_kernels.cu:16-600. */
int dsb_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                      void* sum_out, float* mean, float* rstd, long rows, int hidden, float eps, int dtype,
                      void* stream);
/* dx = LN'(dy) (+ dres); dgamma/dbeta via per-CTA partials in `scratch` (dsb_layernorm_bwd_scratch_bytes).
 * Reference: csrc/transformer/normalize_kernels.cu:607-2103. */
long dsb_layernorm_bwd_scratch_bytes(int hidden);
int dsb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const void* gamma,
                      const void* dres, void* dx, void* dgamma, void* dbeta, long rows, int hidden, int dtype,
                      void* scratch, void* stream);
/* "Invertible" LayerNorm backward: the normalised input is recovered from the OUTPUT y as (y - beta) / gamma, so the
 * forward pass does not keep its input (DeepSpeedTransformerConfig.normalize_invertible). gamma entries equal to zero
 * contribute a zero x-hat. Reference: csrc/transformer/normalize_kernels.cu:602-672,1056-1348. */
int dsb_layernorm_bwd_inv(const void* dy, const void* y, const float* rstd, const void* gamma, const void* beta,
                          const void* dres, void* dx, void* dgamma, void* dbeta, long rows, int hidden, int dtype,
                          int accumulate, void* scratch, void* stream);
/* Same; accumulate != 0 adds the gamma/beta gradients into dgamma/dbeta (the parameters' .grad storage). */
int dsb_layernorm_bwd_acc(const void* dy, const void* x, const float* mean, const float* rstd, const void* gamma,
                          const void* dres, void* dx, void* dgamma, void* dbeta, long rows, int hidden, int dtype,
                          int accumulate, void* scratch, void* stream);
/* y = softmax(x * scale + mask) over the last dim. x: [batch, heads, q_len, k_len]; mask: additive [batch, k_len]
 * or NULL; causal != 0 masks k > q + (k_len - q_len). In place allowed.
 * Reference: csrc/transformer/softmax_kernels.cu:8-431. */
int dsb_softmax_fwd(const void* x, const void* mask, void* y, int batch, int heads, int q_len, int k_len,
                    float scale, int causal, int dtype, void* stream);
/* dx = y * (dy - sum(dy * y)) * scale. Reference: softmax_kernels.cu:433-577. */
int dsb_softmax_bwd(const void* dy, const void* y, void* dx, long rows, int k_len, float scale, int dtype,
                    void* stream);
/* Flash-attention forward on tcgen05 (head dim 128, S a multiple of 128): out = softmax(scale * q k^T (+ causal
 * mask)) v per (batch, head), never materialising the [S, S] probabilities. q / k / v / out are [B, S, H, 128] with
 * element strides (sb, ss, sh, 1) each (multiples of 8), so slices of a fused QKV activation are read in place.
 * `lse` (optional, [B, H, S] fp32): natural-log sum of exponentials of the scaled scores, for a backward pass.
 * Reference: the score / softmax / context chain of csrc/transformer/ds_transformer_cuda.cpp:186-214. */
int dsb_flash_attn_fwd(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S, int H, int D,
                       long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh, long v_sb, long v_ss,
                       long v_sh, long o_sb, long o_ss, long o_sh, float scale, int causal, int dtype, void* stream);
/* Forward kernel: 2 = two query tiles per CTA with P kept in tensor memory (default, needs S % 256 == 0, otherwise the
 * 64-key kernel runs); 64 = 64-key blocks, two CTAs per SM; 128 = 128-key blocks, one CTA per SM. Returns the previous
 * setting; other values only query. dsb_flash_attn_bwd_variant: 2 (default) / 1 selects which backward the Python
 * wrapper calls (dsb_flash_attn_bwd2 / dsb_flash_attn_bwd). */
int dsb_flash_attn_fwd_variant(int keys_per_step);
int dsb_flash_attn_bwd_variant(int generation);
/* Debug aid: when `p` is not NULL, CTA (0, 0) of the following attention launches adds its clock64 phase totals to
 * the 16 int64 at `p` ([0..4] softmax warp, [8..12] MMA thread, [15] blocks walked); NULL switches it off. */
void dsb_flash_attn_debug_timers(long long* p);
/* Backward of dsb_flash_attn_fwd: one CTA per (key block, batch, head) keeps K / V resident, walks the query blocks
 * and accumulates dV, dK in TMEM; dQ contributions are added to `dq_acc` ([B, S, H, 128] fp32, contiguous, zeroed by
 * the caller) with vector reductions. `strides`: 7 x (sb, ss, sh) for dO, q, k, v, out, dk, dv. `delta`: [B, H, S]
 * fp32 scratch (rowsum(dO * O), computed here). */
int dsb_flash_attn_bwd(const void* dO, const void* q, const void* k, const void* v, const void* out, const float* lse,
                       float* delta, float* dq_acc, void* dk, void* dv, int B, int S, int H, int D,
                       const long* strides, float scale, int causal, int dtype, void* stream);
/* Second-generation backward (default): everything is computed transposed (lane = key row) so that P^T and dS^T stay
 * in tensor memory as UMMA A operands; dq / dk / dv come back in the input dtype. `strides`: 8 x (sb, ss, sh) for dO, q,
 * k, v, out, dq, dk, dv. `scratch`: (2 + 128) * B * H * S floats (cleared here; nothing is allocated). rot_dim > 0:
 * q / k were rotated (dsb_rotary_oop) before the forward; dq / dk are returned for the unrotated tensors.
 * Reference: the backward chain of csrc/transformer/ds_transformer_cuda.cpp:340-420. */
int dsb_flash_attn_bwd2(const void* dO, const void* q, const void* k, const void* v, const void* out, const float* lse,
                        float* scratch, void* dq, void* dk, void* dv, int B, int S, int H, int D, const long* strides,
                        float scale, int causal, int dtype, const float* rot_cos, const float* rot_sin, int rot_dim,
                        void* stream);
/* Block-sparse / masked variants (S a multiple of 256): `sched` is a device int32 schedule of the visible 64-key blocks
 * per (head, 256-row query block) with one 32-bit visibility word per 128 x 64 tile (16 x 16 granularity); `kpm` an
 * additive key mask [B, S] in log2 units. The backward takes the transposed schedule (per 128-key block, the 64-row
 * query blocks that see it). Record layouts: csrc/transformer/flash_attn.cu. The kernels skip every tile that is not
 * listed, so time scales with the layout's density.
 * Reference: deepspeed/ops/sparse_attention/{matmul,softmax}.py + trsrc/*.tr (SDD + softmax + DSD as separate kernels). */
int dsb_flash_attn_fwd_ex(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S, int H, int D,
                          const long* strides, float scale, int causal, int dtype, const int* sched, int sched_hstride,
                          int sched_pstride, const float* kpm, void* stream);
int dsb_flash_attn_bwd2_ex(const void* dO, const void* q, const void* k, const void* v, const void* out,
                           const float* lse, float* scratch, void* dq, void* dk, void* dv, int B, int S, int H, int D,
                           const long* strides, float scale, int causal, int dtype, const int* sched, int sched_hstride,
                           int sched_pstride, const float* kpm, void* stream);
/* The same with attention dropout inside the kernels: probabilities are dropped with probability dropout_p (quantised
 * to 1 / 65536) and the kept ones scaled by 1 / (1 - p); the keep bits are a hash of (dropout_seed, batch, head, query,
 * key), so the backward call reproduces the forward's mask from the same seed - no mask tensor exists. */
int dsb_flash_attn_fwd_drop(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S, int H,
                            int D, const long* strides, float scale, int causal, int dtype, const int* sched,
                            int sched_hstride, int sched_pstride, const float* kpm, float dropout_p,
                            unsigned dropout_seed, void* stream);
int dsb_flash_attn_bwd2_drop(const void* dO, const void* q, const void* k, const void* v, const void* out,
                             const float* lse, float* scratch, void* dq, void* dk, void* dv, int B, int S, int H, int D,
                             const long* strides, float scale, int causal, int dtype, const int* sched,
                             int sched_hstride, int sched_pstride, const float* kpm, float dropout_p,
                             unsigned dropout_seed, void* stream);
/* Experiment switches of the second-generation kernels (0 = production); see csrc/transformer/flash_attn_v2.cu. */
void dsb_flash_attn_debug_mode(int mode);
/* Embedding lookup out[t] = table[ids[t]] (ids int64, dim % 8 == 0) and its gradient dtable[ids[t]] += dout[t], added
 * with packed 16-bit atomics straight into the caller's (pre-existing, e.g. arena) gradient buffer: no dense temporary,
 * no sort, no separate accumulate pass. */
int dsb_embedding_fwd(const void* table, const long* ids, void* out, long ntok, int dim, long vocab, int dtype,
                      void* stream);
int dsb_embedding_bwd_acc(const void* dout, const long* ids, void* dtable, long ntok, int dim, long vocab, int dtype,
                          void* stream);
/* y = gelu(x + bias); dx = dy * gelu'(x + bias). Reference: csrc/transformer/gelu_kernels.cu:38-330. */
int dsb_bias_gelu_fwd(const void* x, const void* bias, void* y, long rows, int cols, int dtype, void* stream);
int dsb_bias_gelu_bwd(const void* dy, const void* x, const void* bias, void* dx, long rows, int cols, int dtype,
                      void* stream);
/* out = dropout(x + bias) + residual with a Philox mask (uint8 per element); ratio = drop probability.
 * Reference: csrc/transformer/dropout_kernels.cu:5-721. */
int dsb_dropout_fwd(const void* x, const void* bias, const void* residual, void* out, unsigned char* mask, long rows,
                    int cols, float ratio, unsigned long long seed, unsigned long long offset, int dtype,
                    void* stream);
int dsb_dropout_bwd(const void* dy, const unsigned char* mask, void* dx, long n, float ratio, int dtype,
                    void* stream);
/* [B, S, 3, H, D] (+ bias[3*H*D]) -> q, k, v each [B, H, S, D]; and [B, H, S, D] -> [B, S, H*D].
 * Reference: csrc/transformer/transform_kernels.cu:56-575. */
int dsb_qkv_bias_split_transpose(const void* qkv, const void* bias, void* q, void* k, void* v, int B, int S, int H,
                                 int D, int dtype, void* stream);
int dsb_transpose_0213(const void* x, void* y, int B, int d1, int d2, int D, int dtype, void* stream);
/* out[cols] = sum over rows of x[rows, cols] (bias gradient); `scratch` = cols floats of fp32 accumulators.
 * Reference: general_kernels.cu:6-48. */
int dsb_column_sum(const void* x, void* out, long rows, int cols, int dtype, int accumulate, float* scratch,
                   void* stream);
/* dsb_column_sum without a scratch argument, for callers that may not allocate device memory (code running where
 * the CUDA allocation API is off limits, or in a process whose HBM is already owned by another allocator). The
 * accumulators sit in a fixed device array inside the library: cols must be <= DSB_COLSUM_NOALLOC_MAX_COLS, larger
 * requests return DSB_ERR_SIZE. The entry point itself allocates and frees nothing. Its first call on each device
 * resolves the array's address (do that call outside stream capture); the array is shared by all calls on a GPU,
 * is re-zeroed at the start of every call, and calls that can overlap on one GPU race and must be serialised.
 * This entry point has no counterpart in the reference. */
#define DSB_COLSUM_NOALLOC_MAX_COLS 65536
int dsb_column_sum_noalloc(const void* x, void* out, long rows, int cols, int dtype, int accumulate, void* stream);
/* out = a + b (+ c (+ d)). Reference: general_kernels.cu:89-323. */
int dsb_fused_add(const void* a, const void* b, const void* c, const void* d, void* out, long n, int dtype,
                  void* stream);
/* Token-level softmax cross entropy: loss[t] = lse(logits[t]) - logits[t, target[t]] (0 for ignore_index);
 * backward overwrites logits with (softmax - onehot) * grad_scale. */
int dsb_cross_entropy_fwd(const void* logits, const long* targets, float* loss, float* lse, long tokens, int vocab,
                          long ld, long ignore_index, int dtype, void* stream);
int dsb_cross_entropy_bwd(void* logits, const long* targets, const float* lse, const float* grad_scale_dev,
                          float grad_scale, long tokens, int vocab, long ld, long ignore_index, int dtype,
                          void* stream);
/* Rotary embedding applied in place to q and k laid out [tokens_b, seq, heads, dim] (first rot_dim dims rotate,
 * GPT-NeoX half-split convention); inverse != 0 applies the transpose rotation (backward). */
int dsb_rotary(void* q, void* k, int B, int S, int H, int D, int rot_dim, long stride_b, long stride_s, long stride_h,
               float base, int inverse, int dtype, void* stream);

/* Out-of-place rotary through a cos/sin table [S, rot_dim/2] (dsb_rotary_table): q/k read with one set of
 * (batch, seq, head) strides and written with another, e.g. out of the fused QKV activation / into the dQKV buffer.
 * The _v form also copies a third tensor (v / dv) verbatim in the same launch. */
int dsb_rotary_table(float* cos_out, float* sin_out, int S, int rot_dim, float base, void* stream);
int dsb_rotary_oop(const void* q_in, const void* k_in, void* q_out, void* k_out, int B, int S, int H, int D,
                   int rot_dim, long in_sb, long in_ss, long in_sh, long out_sb, long out_ss, long out_sh,
                   const float* cos_tab, const float* sin_tab, int inverse, int dtype, void* stream);
int dsb_rotary_oop_v(const void* q_in, const void* k_in, void* q_out, void* k_out, int B, int S, int H, int D,
                     int rot_dim, long in_sb, long in_ss, long in_sh, long out_sb, long out_ss, long out_sh,
                     const float* cos_tab, const float* sin_tab, int inverse, int dtype, const void* v_in, void* v_out,
                     long v_sb, long v_ss, long v_sh, void* stream);

/* ---- GEMM variants (csrc/gemm) ----------------------------------------------------------------------------- */
/* CTA-pair (cta_group::2) kernel, same contract as dsb_gemm. _ex: split_mode -1 = choose the tail-wave K split,
 * 1 = whole tiles, 2..4 = that split. force_* are test / tuning hooks for the plain entry point. */
int dsb_gemm_pair(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb, long ldd,
                  long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major, int dtype, int epilogue,
                  float alpha, const void* bias, const void* aux, long ldaux, long stride_aux, void* aux_out,
                  long ldaux_out, long stride_aux_out, void* stream);
int dsb_gemm_pair_ex(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb,
                     long ldd, long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major, int dtype,
                     int epilogue, float alpha, const void* bias, const void* aux, long ldaux, long stride_aux,
                     void* aux_out, long ldaux_out, long stride_aux_out, int split_mode, void* stream);
int dsb_gemm_pair_force_bn(int bn);
int dsb_gemm_pair_force_split(int split);
/* Dispatcher (tuned table, else static rule) and the autotuner that fills the table: variant 0 = single-CTA,
 * 1..4 = pair with that split, 5 = pair choosing its split. Reference: csrc/includes/gemm_test.h:27-293.
 * dsb_gemm_tune ALLOCATES device memory (cudaMalloc / cudaFree), records events and synchronises `stream`: call it at
 * set-up time only, never while a stream is being captured into a CUDA graph. dsb_gemm_auto is capture-safe; it may be
 * called from any thread while another thread tunes (table under a mutex, empty check on an atomic flag). */
int dsb_gemm_auto(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb, long ldd,
                  long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major, int dtype, int epilogue,
                  float alpha, const void* bias, const void* aux, long ldaux, long stride_aux, void* aux_out,
                  long ldaux_out, long stride_aux_out, void* stream);
int dsb_gemm_tune(int M, int N, int K, int batch, int a_mn, int b_mn, int dtype, int iters, float* ms_out,
                  void* stream);
int dsb_gemm_tuned_variant(int M, int N, int K, int batch, int a_mn, int b_mn, int dtype);
void dsb_gemm_tune_clear(void);
/* Fused all-gather + GEMM: B [N, K] is row-sharded over `nranks` peers (shard_ptrs = peer-mapped device pointers,
 * N / nranks rows each, a multiple of 256); TMA loads each B tile straight from its owner's memory over NVLink, so
 * no gathered copy of the weight ever exists (ZeRO-3 linear). No reference counterpart. */
int dsb_gemm_gather_b(const void* A, void* const* shard_ptrs, int nranks, void* D, int M, int N, int K, long lda,
                      long ldb, long ldd, int dtype, int epilogue, float alpha, const void* bias, const void* aux,
                      long ldaux, void* stream);

/* Overlapped all-gather + GEMM (the ZeRO-3 linear the framework uses): dense shards of N / nranks rows; two warps of
 * every CTA copy the remote shards into `gathered` [N, K] panel by panel (N / 256 flags in `panel_flags`, zeroed by
 * the call) while the tensor pipe consumes the panels already present, own shard first. One launch; every weight
 * byte crosses NVLink once. No reference counterpart (reference: all_gather then F.linear,
 * runtime/zero/partition_parameters.py:746-855). */
int dsb_gemm_ag_overlap(const void* A, void* const* shard_ptrs, int nranks, int me, void* gathered, int* panel_flags,
                        void* D, int M, int N, int K, long lda, long ldd, int dtype, int epilogue, float alpha,
                        const void* bias, const void* aux, long ldaux, void* stream);

/* ---- fused compute + collective kernels over peer memory (csrc/comm) ---------------------------------------- */
/* Symmetric-heap plumbing: plain cudaMalloc + CUDA IPC handles (64 bytes) exchanged by the Python side. */
int dsb_symm_alloc(long bytes, void** out);
int dsb_symm_free(void* p);
int dsb_ipc_get_handle(void* devptr, void* handle64);
int dsb_ipc_open(const void* handle64, void** out);
int dsb_ipc_close(void* p);
/* Same overlap for the dgrad of a row-sharded weight W [K, N]: D [M, N] = alpha * A [M, K] * gather(W). The shards
 * split the reduction dimension; every tile's k loop starts at the own shard's range and follows the copiers' panel
 * order over the remote ranges. rows per shard a multiple of 256; `panel_flags`: K / 256 ints.
 * Replaces `_allgather_param` before the backward of a linear (reference zero/stage3.py:1471-1559). */
int dsb_gemm_ag_overlap_k(const void* A, void* const* shard_ptrs, int nranks, int me, void* gathered, int* panel_flags,
                          void* D, int M, int N, int K, long lda, long ldd, int dtype, float alpha, void* stream);
/* GEMM + reduce-scatter in one cooperative kernel (ZeRO-3 weight gradient): every rank computes
 * D = alpha * op(A) op(B)^T ([M, N], rows owned by nranks ranks, M / nranks a multiple of 128); tiles are TMA-stored
 * into slot `me` of the owner's staging buffer (stage_ptrs[r]: [nranks, M / nranks, N], peer-mapped) as they are
 * produced, and after a cross-GPU barrier (flag blocks at flag_bases[r]) each rank sums its nranks slots into
 * out [M / nranks, N] times `scale`. `sync2`: two ints of local scratch. Callers alternate two staging buffers.
 * Replaces the wgrad GEMM + dist.reduce_scatter of reference zero/partition_parameters.py:888-926. */
int dsb_gemm_reduce_scatter(const void* A, const void* B, void* out, void* const* stage_ptrs, void* const* flag_bases,
                            int nranks, int me, int* sync2, int M, int N, int K, long lda, long ldb, int a_mn_major,
                            int b_mn_major, int dtype, float alpha, float scale, unsigned int epoch, void* stream);
/* ZeRO-1/2 step as two cooperative kernels. rs_norm: every rank pulls its slice of all peers' gradient arenas,
 * reduces in fp32, unscales, accumulates sum-of-squares + non-finite flag. adam_ag: Adam on the owned slice and
 * stores the new low-precision parameters into every peer's parameter arena. Replaces reduce_scatter + cast +
 * optimizer + all_gather (reference stage1.py:629-773). */
long dsb_fused_zero_scratch_bytes(void);
int dsb_fused_rs_norm(void* const* grad_ptrs, void* const* flag_ptrs, int R, int me, long slice_off, long n,
                      int dtype, float scale, float* red32, float* sumsq_out, int* flag_out, void* scratch,
                      unsigned int epoch, void* stream);
int dsb_fused_adam_ag(const float* red32, float* p32, float* m, float* v, void* const* param_ptrs,
                      void* const* flag_ptrs, int R, int me, long slice_off, long n, int dtype, float lr, float beta1,
                      float beta2, float eps, float weight_decay, int adamw_mode, int bias_correction, int step,
                      const float* inv_scale_dev, const int* skip_dev, unsigned int epoch, void* stream);
/* The whole ZeRO-1/2 step of one gradient bucket as one cooperative kernel (csrc/comm/fused_zero_step.cu). `phases`:
 * 1 = reduce this rank's slice over all ranks (+ local norm terms into sumsq2), 2 = Adam + cast + all-gather + zero the
 * local gradient bucket, 4 = trade the norm terms with the peers inside the kernel and derive inv_scale / skip /
 * host_out2 = {norm, overflow} on the device (only together with 2). mc_grads / mc_params: multicast addresses of the
 * gradient / parameter arenas (NVLS: multimem.ld_reduce / multimem.st), NULL = unicast peer loads / stores.
 * slice_off, bucket_off, bucket_n, n in elements (multiples of 8); scratch: dsb_fused_zero_scratch_bytes(). */
int dsb_fused_zero_step(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs, const void* mc_grads,
                        void* mc_params, int R, int me, int phases, int zero_grads, long slice_off, long bucket_off,
                        long bucket_n, long n, int dtype, void* red, float* p32, float* m, float* v, float lr,
                        float beta1, float beta2, float eps, float weight_decay, int adamw_mode, int bias_correction,
                        int step, float loss_scale, float clip, float* sumsq2, float* inv_scale, int* skip,
                        float* host_out2, void* scratch, unsigned int epoch, void* stream);
/* ... with at most max_ctas CTAs (0 = fill the GPU): for a reduce phase overlapped with backward on a side stream. */
int dsb_fused_zero_step_grid(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs,
                             const void* mc_grads, void* mc_params, int R, int me, int phases, int zero_grads,
                             long slice_off, long bucket_off, long bucket_n, long n, int dtype, void* red, float* p32,
                             float* m, float* v, float lr, float beta1, float beta2, float eps, float weight_decay,
                             int adamw_mode, int bias_correction, int step, float loss_scale, float clip,
                             float* sumsq2, float* inv_scale, int* skip, float* host_out2, void* scratch,
                             unsigned int epoch, int max_ctas, void* stream);
/* LAMB counterpart of dsb_fused_adam_ag, one launch per parameter group over all `n / sub` bucket slices this rank
 * owns. `upd32` holds the reduced fp32 gradients on entry (it is overwritten with the update). The owned range is
 * described by `nseg` segments: seg_start[i] .. seg_start[i+1] (local element offsets, multiples of 8,
 * seg_start[nseg] == n) belongs to tensor seg_tid[i] < ntensors. `norm_ptrs[r]` is rank r's table of 2 * ntensors
 * floats in the symmetric heap; per-tensor |w|^2, |u|^2 are summed over the ranks through it, so the trust ratio is
 * per tensor even when a tensor is spread over several ranks. `coeff_out` (ntensors floats, may be NULL) receives
 * the ratios of the tensors this rank touches. Reference math: csrc/lamb/fused_lamb_cuda_kernel.cu:186-310. */
int dsb_fused_lamb_ag(float* upd32, float* p32, float* m, float* v, void* const* param_ptrs, void* const* norm_ptrs,
                      void* const* flag_ptrs, int R, int me, long sub, long n, const long* seg_start,
                      const int* seg_tid, int nseg, int ntensors, int dtype, float lr, float beta1, float beta2,
                      float eps, float weight_decay, float max_coeff, float min_coeff, int eps_inside_sqrt,
                      int bias_correction, int step, const float* inv_scale_dev, const int* skip_dev, float* coeff_out,
                      unsigned int epoch, void* stream);
/* 1-bit (sign + scale, error feedback) allreduce as ONE cooperative kernel: compress, exchange packed signs through
 * the peers' heaps, server-side average + recompress, gather. Reference: runtime/comm/nccl.py:47-186. */
long dsb_onebit_heap_bytes(long n, int W);
int dsb_onebit_allreduce(float* m, float* e_w, float* e_s, long n, int W, int me, void* const* heap_bases,
                         long region_off, void* const* flag_bases, void* scratch, unsigned int epoch, void* stream);
/* The compression-stage step of 1-bit Adam in the same kernel: m <- beta1 m + (1 - beta1) g on the way in, the
 * averaged momentum is decompressed straight into p <- p - lr (m / (sqrt(v) + eps) + wd p) on the way out.
 * m, g, p, v: n_real fp32 elements (multiple of 4); e_w: n, e_s: n / W elements (n = n_real padded to 32 W). */
int dsb_onebit_adam_step(float* m, const float* g, float* p, const float* v, float* e_w, float* e_s, long n,
                         long n_real, int W, int me, void* const* heap_bases, long region_off,
                         void* const* flag_bases, void* scratch, unsigned int epoch, float beta1, float lr, float eps,
                         float weight_decay, void* stream);

/* ---- block-sparse attention (csrc/sparse_attn) -------------------------------------------------------------- */
/* Host LUT builders: layout [H, nb, nb] (0/1) -> block list (h, i, j), row CSR, column CSC; super-block packing. */
long dsb_bs_count_nnz(const int64_t* layout, int H, int nb);
int dsb_bs_build_luts(const int64_t* layout, int H, int nb, int32_t* blocks, int32_t* row_ptr, int32_t* row_idx,
                      int32_t* col_ptr, int32_t* col_idx);
int dsb_sdd_segment(const int64_t* layout, int H, int nb, int max_width, int32_t* out_id, int32_t* out_width,
                    int64_t* counts);
/* C[nnz blocks] = alpha * A B^T on the non-zero blocks; O = P_sparse X (or P^T X); in-place block-sparse softmax.
 * Reference: ops/sparse_attention/trsrc/{matmul,softmax_fwd,softmax_bwd}.tr. */
int dsb_bs_sdd(const void* A, const void* B, void* C, const int* blocks, int batch, int H, int S, int D, int block,
               long nnz, float alpha, int dtype, void* stream);
int dsb_bs_dsd(const void* P, const void* X, void* O, const int* ptr, const int* idx, int batch, int H, int S, int D,
               int block, long nnz, int transposed, int dtype, void* stream);
int dsb_bs_softmax_fwd(void* X, const int* ptr, const int* idx, int batch, int H, int S, int block, long nnz,
                       float scale, const void* rpe, const void* kp_mask, int kp_mul, const void* attn_mask,
                       int attn_mul, int dtype, void* stream);
int dsb_bs_softmax_bwd(void* dY, const void* Y, const int* ptr, const int* idx, int batch, int H, int S, int block,
                       long nnz, float scale, int dtype, void* stream);

/* ---- host library (libdsb200_host.so, csrc/host) ------------------------------------------------------------ */
/* CPU Adam: AVX-512 / AVX2 / scalar chosen at run time (simd_level 2 / 1 / 0), OpenMP over tiles; optional
 * low-precision copy of the updated parameters to a host buffer or, through pinned double buffering, to the GPU.
 * Reference: csrc/adam/cpu_adam.cpp:1-684. */
int dsb_cpu_adam_simd_level(void);
int dsb_cpu_adam_create(int id, float lr, float beta1, float beta2, float eps, float weight_decay, int adamw_mode);
int dsb_cpu_adam_destroy(int id);
int dsb_cpu_adam_step(int id, long step, float lr, float beta1, float beta2, float eps, float weight_decay,
                      int bias_correction, float* p, const float* g, float* m, float* v, long n, void* dev_lp,
                      void* host_lp, int lp_bf16);
int dsb_cpu_cast_lp(const float* src, void* dst, long n, int lp_bf16);
/* Async file I/O: kernel AIO (io_submit / io_getevents by raw syscall) with O_DIRECT where alignment allows, thread
 * pool per handle. Reference: csrc/aio/py_lib/deepspeed_py_aio_handle.cpp. */
void* dsb_aio_handle_create(long block_size, int queue_depth, int single_submit, int overlap_events, int num_threads);
void dsb_aio_handle_destroy(void* handle);
int dsb_aio_submit(void* handle, int is_read, void* buf, long nbytes, const char* path);
long dsb_aio_wait(void* handle);
long dsb_aio_pending(void* handle);
long dsb_aio_sync_rw(int is_read, void* buf, long nbytes, const char* path, long block_size, int queue_depth,
                     int single_submit, int overlap_events);
void* dsb_aio_alloc_aligned(long nbytes, int lock_pages);
void dsb_aio_free_aligned(void* p, long nbytes, int locked);
int dsb_parallel_memcpy(void* dst, const void* src, long nbytes);
/* flatten / unflatten of n host or device-agnostic byte ranges. Reference: csrc/utils/flatten_unflatten.cpp. */
int dsb_flatten_bytes(const void* const* ptrs, const long* bytes, int n, void* dst);
int dsb_unflatten_bytes(const void* src, void* const* ptrs, const long* bytes, int n);

#ifdef __cplusplus
}
#endif
#endif /* DSB200_H */
