/* dlti_b200 — C ABI of the native sm_100a kernel library (libdlti_b200.so).
 *
 * Every entry point:
 *   - takes raw device pointers plus a cudaStream_t and is asynchronous on that stream;
 *   - returns 0 on success, a positive cudaError_t, or a negative DLTI_ERR_* code
 *     (dlti_error_string() / dlti_last_error() describe it);
 *   - allocates no device memory unless its comment says so (workspaces are passed in).
 *
 * dtype codes: 0 = bf16, 1 = fp16, 2 = fp32.
 *
 * The reference repository (rokulkarni15/distributed-llm-training-inference) has no native
 * code; each declaration cites the library call it replaces in the reference's stack
 * (SURVEY.md section 2.4 / 2.7 numbering).
 */
#ifndef DLTI_B200_H_
#define DLTI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* dlti_stream_t;

#define DLTI_BF16 0
#define DLTI_FP16 1
#define DLTI_FP32 2

/* ---- runtime ---------------------------------------------------------------------------- */
int dlti_version(void);
int dlti_sm_count(void);
/* Programmatic dependent launch on the decode path (skinny GEMM, dlti_acc_*, paged decode). Level 1 (default):
 * those kernels start while their predecessor in the stream drains and order all global accesses behind a
 * dependency wait. Level 2: dlti_gemm_skinny additionally requests its first weight tiles before that wait —
 * only for weights that no kernel still in flight writes (the serving engine sets this). Level 0 = plain stream
 * ordering. Environment: DLTI_PDL=0|1|2. */
void dlti_set_pdl(int level);
int dlti_get_pdl(void);
long long dlti_launch_count(void);      /* kernels launched by this library since reset */
void dlti_reset_launch_count(void);
const char* dlti_last_error(void);
const char* dlti_error_string(int code);

/* ---- K2/K3: tcgen05 GEMM with K-extension (fused LoRA) ------------------------------------
 * C[M,N] = alpha * ( op(A)·op(B) + A2·op(B2) ) + R          (out bf16/fp16, optional residual R)
 * C[M,N] (+)= alpha * ( ... )                               (out fp32, accumulate = 0/1)
 *   a_mn = 0: A is [M,K] row-major (pitch lda).   a_mn = 1: A is stored [K,M] row-major.
 *   b_mn = 0: B is [N,K] row-major (x·W^T).       b_mn = 1: B is stored [K,N] row-major.
 *   A2/B2 (optional, K2 % 64 == 0) use the same majorness as A/B and are accumulated into the
 *   same tensor-memory tile after the main K loop (LoRA: A2 = dropout(x)·A^T, B2 = (alpha/r)·B).
 *   split_k > 1 requires fp32 output with accumulate = 1 (atomic adds).
 *   block_n: 0 = auto, or 64/128/256. max_ctas: 0 = all SMs.
 * Replaces cuBLAS GEMM under torch.nn.functional.linear + PEFT lora.Linear
 * (reference training/train_baseline.py:122-140). */
int dlti_gemm(const void* A, const void* B, void* C, const void* R, int M, int N, int K,
              long long lda, long long ldb, long long ldc, long long ldr, int a_mn, int b_mn,
              const void* A2, const void* B2, int K2, long long lda2, long long ldb2,
              int in_dtype, int out_dtype, float alpha, int accumulate, int split_k, int block_n,
              int max_ctas, dlti_stream_t stream);

/* ---- K4: RMSNorm ---------------------------------------------------------------------------
 * fwd: y = w * T(x * rsqrt(mean(x^2) + eps)); optional residual_inout (+= x, then normalised: the
 * serving path's fused add+norm), optional ydrop = dropout(y) (LoRA-branch input, Philox mask
 * keyed by (seed, stream_id, element index)), rstd (fp32 per row) for the backward.
 * bwd: dx (norm weights are frozen: no dw); dres (optional) is added to dx.
 * Replaces HF LlamaRMSNorm's ~5 ATen kernels / vLLM rms_norm + fused_add_rms_norm. */
int dlti_rmsnorm_fwd(const void* x, void* residual_inout, const void* w, void* y, void* ydrop,
                     float* rstd, long long rows, int d, float eps, float drop_p,
                     unsigned long long seed, unsigned long long stream_id, int dtype,
                     dlti_stream_t stream);
int dlti_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                     const void* dres, void* dx, long long rows, int d, int dtype,
                     dlti_stream_t stream);

/* ---- K5: RoPE (rotate-half), in place on the first n_rot_heads heads of each [tokens, row_stride]
 * row (merged qkv: q heads then k heads). cos/sin: fp32 [max_pos, hd/2]. backward = inverse
 * rotation. Replaces HF apply_rotary_pos_emb / vLLM rotary_embedding. */
int dlti_rope(void* buf, const int* pos, const float* cos_t, const float* sin_t, long long tokens,
              int n_rot_heads, int hd, long long row_stride, int backward, int dtype,
              dlti_stream_t stream);

/* ---- K7: SwiGLU on gu = [tokens, 2*ff] (gate | up). Replaces silu + mul / vLLM silu_and_mul
 * (SURVEY O7, S8). */
int dlti_swiglu_fwd(const void* gu, void* h, long long tokens, int ff, long long gu_stride,
                    long long h_stride, int dtype, dlti_stream_t stream);
int dlti_swiglu_bwd(const void* dh, const void* gu, void* dgu, long long tokens, int ff,
                    long long dh_stride, long long gu_stride, int dtype, dlti_stream_t stream);

/* ---- dropout / a + dropout(b): counter-based masks (mask index = row*cols + col), reproducible in
 * layer recompute. Replaces torch.nn.Dropout inside PEFT lora.Linear (lora_dropout = 0.05). */
int dlti_dropout(const void* x, void* y, long long rows, int cols, long long x_stride,
                 long long y_stride, float p, unsigned long long seed, unsigned long long stream_id,
                 int dtype, dlti_stream_t stream);
int dlti_add_dropout(const void* a, const void* b, void* y, long long rows, int cols, float p,
                     unsigned long long seed, unsigned long long stream_id, int dtype,
                     dlti_stream_t stream);

/* ---- K15: embedding gather; ids outside [vocab_start, vocab_start + vocab_rows) give zero rows
 * (vocab-parallel shard contribution). Replaces ATen embedding. */
int dlti_embedding(const long long* ids, const void* table, void* out, long long tokens, int d,
                   long long vocab_start, long long vocab_rows, int dtype, dlti_stream_t stream);

/* ---- K8 (loss): softmax cross-entropy over 16-bit logits, fp32 math. loss_rows[t] = lse - logit
 * [label] (0 when label == ignore_index); dlogits (may alias logits) = (softmax - onehot) * *gscale.
 * Replaces ATen log_softmax + nll_loss (+ backward). */
int dlti_cross_entropy(const void* logits, void* dlogits, const long long* labels, float* loss_rows,
                       const float* gscale, long long rows, int V, long long ld, long long ldd,
                       long long ignore_index, int dtype, dlti_stream_t stream);

/* ---- K9 (single-GPU part): out[0] (+)= sum(g^2) over a flat fp32 vector (gradient norm; a
 * non-finite element makes it non-finite = fp16 overflow). `partials` is caller-provided SCRATCH of
 * dlti_sumsq_scratch_floats() floats. Replaces clip_grad_norm_ + GradScaler's finite check
 * (SURVEY N7, O12, C6, C7). */
int dlti_sumsq_scratch_floats(void);
int dlti_sumsq(const float* g, long long n, float* partials, float* out, int accumulate,
               dlti_stream_t stream);

/* Same reduction as dlti_sumsq for callers that cannot hand the library a device workspace: a
 * process whose own pool (e.g. a KV-cache arena) already holds all free HBM, a loop that must not
 * pay the device-wide synchronisation of cudaMalloc/cudaFree, or a plain C host with no tensor
 * allocator. The partial sums live in a 32 KB array built into the library (one copy per GPU that
 * loads it, until exit), so n is limited to dlti_sumsq_noalloc_max_elems() = 134,217,728 floats;
 * larger n returns DLTI_ERR_TOO_LARGE (-4) with a message in dlti_last_error(). The first call on
 * each device looks up the array's device address (cudaGetSymbolAddress) — make it before any
 * stream capture. Calls share the array: calls whose kernels overlap on the GPU race; keep them
 * ordered. No cudaMalloc/cudaFree is made. This entry point has no counterpart in the reference. */
long long dlti_sumsq_noalloc_max_elems(void);
int dlti_sumsq_noalloc(const float* g, long long n, float* out, int accumulate,
                       dlti_stream_t stream);

/* Fused unscale + global-norm clip + AdamW (decoupled weight decay) on a flat fp32 partition with
 * optional 16-bit copy-out; skipped (and *skipped_out = 1) when *gnorm_sq is not finite.
 * Replaces torch.optim.AdamW / DeepSpeed FusedAdam multi_tensor_adam.cu. */
int dlti_adamw(float* p, float* m, float* v, float* g, void* p16, int p16_dtype, long long n,
               float lr, float b1, float b2, float eps, float wd, int step, const float* gnorm_sq,
               float inv_scale, float max_norm, int* skipped_out, int zero_grad,
               dlti_stream_t stream);

/* Batched 2-D fp32 -> 16-bit casts with per-job scale; jobs_dev = device array of 48-byte records
 * {src*, dst*, rows, cols, src_pitch, dst_pitch, scale, pad}. Refreshes the packed LoRA buffers. */
int dlti_cast_job_bytes(void);
int dlti_multi_cast(const void* jobs_dev, int njobs, int dst_dtype, dlti_stream_t stream);

/* ---- K6: causal varlen flash attention, head_dim 128, GQA. qkv = [T, (H+2KV)*128] (q|k|v per
 * token row, RoPE applied); out = [T, H*128]; lse = [H, T] fp32 (NULL allowed for inference).
 * Backward writes dqkv = [T, (H+2KV)*128]; D_scratch = [H, T] fp32 workspace.
 * Replaces PyTorch SDPA / flash-attn. */
int dlti_attn_fwd(const void* qkv, void* out, float* lse, const int* cu_seqlens, int batch,
                  int max_seqlen, long long T_total, int H, int KV, int hd, float scale, int dtype,
                  dlti_stream_t stream);
int dlti_attn_bwd(const void* dout, const void* qkv, const void* out, const float* lse,
                  float* D_scratch, void* dqkv, const int* cu_seqlens, int batch, int max_seqlen,
                  long long T_total, int H, int KV, int hd, float scale, int dtype,
                  dlti_stream_t stream);

/* ---- K2 (CTA-pair form): same contract as dlti_gemm with A K-major and no split-K, computed by
 * tcgen05 cta_group::2 pairs (256x256 tile per pair, each CTA stages half of the B tile). A ragged
 * last wave is removed by stream-K: the last waves' tiles are cut into equal (tile, k-block) ranges and
 * the owner of a split tile adds its helper's fp32 partial from a per-stream workspace. */
int dlti_gemm_2cta(const void* A, const void* B, void* C, const void* R, int M, int N, int K,
                   long long lda, long long ldb, long long ldc, long long ldr, int b_mn, const void* A2,
                   const void* B2, int K2, long long lda2, long long ldb2, int in_dtype, int out_dtype,
                   float alpha, int accumulate, int stream_k /* -1 auto, 0 off, 1 force */,
                   dlti_stream_t stream);

/* K1 on the CTA-pair kernel (ZeRO-3 all-gather fused into the GEMM; replaces the reference's DeepSpeed
 * stage-3 parameter all-gather, /root/reference/configs/ds_config_zero3.json:16-37). Attaches to the NEXT
 * dlti_gemm_2cta / _rope_2cta / _swiglu_2cta / _swiglu_bwd_2cta launch made from this host thread:
 *  - a gather job (j_count > 0): one extra warp per CTA pulls the 64-row units (rot_unit + j) % n_units,
 *    j in [j_begin, j_begin + j_count), of a row-sharded matrix [world * rows_per_rank, row_bytes] from their
 *    owner ranks (peer loads; own rows locally) into `staging` and publish per-unit counters: unit u is
 *    complete once counters[u] >= epoch * 16. The job may feed this launch's own B operand or, normally, a
 *    matrix that a later launch consumes (the next layer's weights ride on this layer's GEMMs);
 *  - a wait spec (wait_counters != NULL): the launch's B operand is such a staging buffer; the TMA producer
 *    acquires the counters of the units each tile needs (fast path when everything has landed already).
 * dlti_gather_job_run executes an attached job as a kernel of its own. */
int dlti_gemm2_attach(const void* const* peer_shards, void* staging, unsigned int* counters, long long row_bytes,
                      int rows_per_rank, int world, int rot_unit, int j_begin, int j_count, unsigned int epoch,
                      const unsigned int* wait_counters, int wait_units, unsigned int wait_epoch, int* err_flag);
int dlti_gather_job_run(int ctas, dlti_stream_t stream);
/* K11 (all-gather -> column-parallel GEMM; the reference's surface is vLLM tensor parallelism,
 * /root/reference/README.md:10): after dlti_gemm2_attach, declare that the job / wait spec concern the A operand
 * (activation rows pulled from their owner ranks by the launch's gather warp; the A producer waits per 64-row
 * unit) and that the M blocks are visited starting at block m_rot (this rank's own rows, which land first). */
int dlti_gemm2_attach_a_side(int m_rot);
/* K10 for prefill-sized steps (row-parallel GEMM -> reduce-scatter): the next CTA-pair launch stores output row r
 * into the receive slot of rank r / rows_per_owner (peer stores from the epilogue, tile by tile); the owner then
 * sums its slots in rank order (dlti_tp_own_reduce_norm). */
int dlti_gemm2_attach_scatter(const void* const* slot_peers, int world, int rank, int rows_per_owner);

/* Llama gate_up projection + SwiGLU in one kernel: h[M, ff] = silu(A·Wg^T) * (A·Wu^T), W = [gate; up]
 * stored [2*ff, K]. CTA rank 0 of a pair stages gate rows, rank 1 the matching up rows; the [M, 2*ff]
 * intermediate is never written. Rounding points match dlti_swiglu_fwd on a stored gate|up tensor. */
int dlti_gemm_swiglu_2cta(const void* A, const void* W, void* Hout, int M, int ff, int K, long long lda,
                          long long ldw, long long ldh, int dtype, dlti_stream_t stream);

/* Fused QKV projection (+ LoRA K-extension) with the rotary embedding applied in the epilogue to the first
 * rope_cols output columns (heads of 128, rotate-half; cos/sin fp32 [max_pos, 64]; pos[M] int32). */
int dlti_gemm_rope_2cta(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                        long long ldb, long long ldc, const void* A2, const void* B2, int K2, long long lda2,
                        long long ldb2, int dtype, float alpha, const int* pos, const float* cos_t,
                        const float* sin_t, int rope_cols, dlti_stream_t stream);

/* Input-gradient GEMM of the down projection with the SwiGLU backward in its epilogue:
 * dGU[M, 2*ff] = swiglu_bwd(dY[M, K] · Wdown[K, ff], GU[M, 2*ff]); dh is never written. */
int dlti_gemm_swiglu_bwd_2cta(const void* dY, const void* Wdown, const void* GU, void* dGU, int M, int ff, int K,
                              long long ldy, long long ldw, long long ldgu, long long lddgu, int dtype,
                              dlti_stream_t stream);

/* ---- Decode-sized GEMM (M <= 256): swap-AB (the weight is the 128-row tensor-core operand), K split
 * over CTAs, fp32 partials reduced with atomics into acc32[M,N] (must be zero on entry);
 * dlti_skinny_finish converts acc32 -> out and re-zeroes it. split_k: 0 = stream-K (the (tile, K block)
 * space cut into equal ranges per CTA; default), n > 0 = n tile-major splits, < 0 = automatic tile-major
 * split. Replaces the cuBLAS GEMV/GEMM calls of the vLLM decode path. */
int dlti_gemm_skinny(const void* X, const void* W, float* acc32, int M, int N, int K, long long ldx,
                     long long ldw, int in_dtype, int split_k, dlti_stream_t stream);
int dlti_skinny_finish(float* acc32, void* out, long long rows, int cols, long long out_stride, int dtype,
                       dlti_stream_t stream);

/* Consumers of acc32 that replace dlti_skinny_finish + the next elementwise op of a decode step (each value
 * is rounded to `dtype` first, so results equal finish -> op bit for bit; acc32 is zero on return):
 *   residual += acc ; y = RMSNorm(residual) * w          (o_proj / down_proj; vLLM fused_add_rms_norm, S7/S9)
 *   h = silu(acc[:, :ff]) * acc[:, ff:]                  (gate_up; vLLM silu_and_mul, S8)
 *   qkv = rope(acc) + k/v rows into the paged cache      (qkv_proj; vLLM rotary_embedding + reshape_and_cache, S4) */
int dlti_acc_add_rmsnorm(float* acc32, void* residual_inout, const void* w, void* y, long long rows, int d,
                         float eps, int dtype, dlti_stream_t stream);
int dlti_acc_swiglu(float* acc32, void* h, long long rows, int ff, long long h_stride, int dtype,
                    dlti_stream_t stream);
int dlti_acc_rope_cache(float* acc32, void* qkv, const int* pos, const float* cos_t, const float* sin_t,
                        const long long* slot_mapping, void* k_cache, void* v_cache, long long tokens, int H,
                        int KV, int hd, long long row_stride, int dtype, dlti_stream_t stream);

/* ---- Peer memory plumbing (symmetric arena fallbacks): CUDA IPC handle export/open/close and the
 * base/size of the allocation that contains a pointer. */
int dlti_ipc_get_handle(const void* dev_ptr, void* handle_out_64);
int dlti_ipc_open(const void* handle_64, void** out_ptr);
int dlti_ipc_close(void* ptr);
int dlti_mem_base(const void* dev_ptr, void** base_out, unsigned long long* size_out);

/* ---- K1/K9/K10/K11: fused compute + collective kernels over NVLink peer memory. All pointer tables
 * are DEVICE arrays of `world` peer pointers (index = rank); `err_flag` receives a non-zero code when a
 * bounded wait expires (the kernels never spin forever). Replace NCCL all-gather / reduce-scatter /
 * all-reduce in DeepSpeed ZeRO and vLLM TP. */
int dlti_comm_max_ranks(void);
/* all ranks arrive on `channel` with a host-chosen epoch (eager) or a device-side counter (graph-safe) */
int dlti_peer_barrier(const void* const* signal_pads, int rank, int world, int channel,
                      unsigned int epoch, int* err_flag, dlti_stream_t stream);
int dlti_peer_barrier_auto(const void* const* signal_pads, int rank, int world, int channel,
                           unsigned int* device_counter, int* err_flag, dlti_stream_t stream);
/* K9, overlapped form (replaces DeepSpeed's bucketed gradient reduce-scatter with overlap_comm, the global-norm
 * all-reduce, the partitioned optimizer step and the parameter all-gather:
 * /root/reference/configs/ds_config_zero3.json:29-31, ds_config_zero1.json:38-40).
 * dlti_grad_ready_signal: "my gradients of range `slot` are final for step `epoch`" to every rank.
 * dlti_rs_wait_accumulate: once all ranks have signalled, part[plo..] += sum_r grads[r][glo..] (fixed rank order).
 * dlti_k9_step: exchange of {sum of squares, loss sum}, unscale/clip/AdamW on [lo, lo+n) with the new fp32 values
 * stored into every rank's master, then a completion flag to every peer; dlti_peer_wait waits for those flags. */
int dlti_grad_ready_signal(const void* const* flags, int rank, int world, int slot, unsigned int epoch,
                           dlti_stream_t stream);
int dlti_rs_wait_accumulate(const void* const* grads, const unsigned int* my_flags, float* part, long long glo,
                            long long plo, long long n, int world, int slot, unsigned int epoch, int ctas,
                            int* err_flag, dlti_stream_t stream);
int dlti_k9_step(const void* const* masters, const void* const* boards, const void* const* signal_pads, float* m,
                 float* v, float* g, long long lo, long long n, int rank, int world, float lr, float b1, float b2,
                 float eps, float wd, int step, const float* local2, float* out2, float inv_scale, float max_norm,
                 int* skipped_out, unsigned int* done_scratch, int ch_norm, int ch_push, unsigned int epoch,
                 int* err_flag, dlti_stream_t stream);
int dlti_peer_wait(const unsigned int* my_pads, int world, int channel, unsigned int epoch, int* err_flag,
                   dlti_stream_t stream);

/* Owner side of K10 for prefill-sized steps: sum this rank's receive slots [world * per, d] in rank order, add the
 * (row-sharded) residual, RMSNorm, write the normalised rows to this rank's symmetric [per, d] buffer. */
int dlti_tp_own_reduce_norm(const void* slots, int world, long long per, long long row_lo, long long rows,
                            void* residual, const void* w, void* xn_own, int d, float eps, int dtype,
                            dlti_stream_t stream);

/* K15: embedding lookup over a vocabulary-sharded table in peer memory: out[t] = row ids[t] fetched from its
 * owner rank (ids outside [0, vocab) give zero rows). Replaces all-gather + lookup of the reference's ZeRO-3
 * embedding (HF LlamaModel.embed_tokens under DeepSpeed stage 3). */
int dlti_embedding_peer(const long long* ids, const void* const* shards, void* out, long long tokens, int d,
                        long long rows_per_rank, long long vocab, int world, int dtype, dlti_stream_t stream);
/* pull-based all-gather of equally sized row shards into `out` (K1 stand-alone form) */
int dlti_allgather_rows(const void* const* shards, void* out, long long shard_bytes, int rank, int world,
                        int blocks_per_peer, dlti_stream_t stream);
/* K1: C = A · op(gathered B): B is row-sharded over `world` ranks; gather warps inside the GEMM pull
 * the shards into B_staging (own rows are read in place) and publish per-unit counters (= epoch*16)
 * that the TMA producer waits on, so the transfer overlaps the math tile by tile. */
int dlti_gemm_gather(const void* A, void* B_staging, void* C, const void* R, int M, int N, int K,
                     long long lda, long long ldb, long long ldc, long long ldr, int b_mn, const void* A2,
                     const void* B2, int K2, long long lda2, long long ldb2, int in_dtype, int out_dtype,
                     float alpha, int block_n, int max_ctas, const void* const* peer_shards,
                     unsigned int* counters, unsigned int epoch, int rank, int world, int rows_per_rank,
                     int* err_flag, dlti_stream_t stream);
/* K10: row-parallel GEMM whose epilogue stores the partial tile into slot `rank` of every peer's
 * receive buffer [world, slot_stride] (followed by a barrier and dlti_tp_reduce_add_norm). */
int dlti_gemm_push(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb,
                   long long ldc, int in_dtype, int block_n, const void* const* recv_peers, int rank,
                   int world, long long slot_stride, dlti_stream_t stream);
/* K11: residual += sum over ranks (fixed order) of the received partials; y = rmsnorm(residual) * w */
int dlti_tp_reduce_add_norm(const void* recv, long long slot_stride, int world, void* residual,
                            const void* w, void* y, long long rows, int d, float eps, int dtype,
                            dlti_stream_t stream);
/* K11 with the rank barrier inside the kernel (CTA 0 signals the peers, every CTA waits for all of them; the
 * epoch is a device counter bumped by the last CTA, so the launch replays from a CUDA graph): replaces
 * dlti_peer_barrier_auto + dlti_tp_reduce_add_norm on the decode path. done_scratch: zero-initialised u32. */
int dlti_tp_barrier_reduce_add_norm(const void* recv, long long slot_stride, int world, void* residual,
                                    const void* w, void* y, long long rows, int d, float eps, int dtype,
                                    const void* const* signal_pads, int rank, int channel,
                                    unsigned int* device_counter, unsigned int* done_scratch, int* err_flag,
                                    dlti_stream_t stream);
/* Prefill-sized K10/K11: reduce-scatter + all-gather over peer memory. Each rank's partial [T, d] sits in its own
 * symmetric buffer; this call handles rows [row_lo, row_lo + rows) (the caller's share): pull the rows of all
 * ranks' partials (rank order), add the residual rows (row-sharded on this path), RMSNorm, push the normalised
 * rows into every rank's activation buffer. Needs a rank barrier before (partials complete) and after. */
int dlti_tp_pull_reduce_norm_push(const void* const* partial_peers, const void* const* xn_peers, int world,
                                  long long row_lo, long long rows, void* residual, const void* w, int d, float eps,
                                  int dtype, dlti_stream_t stream);
/* K9: reduce-scatter by pull: part[0:n] += sum over ranks of grads[r][lo:lo+n] (fixed rank order) */
int dlti_rs_accumulate(const void* const* grads, float* part, long long lo, long long n, int world,
                       dlti_stream_t stream);
/* K9: all-reduce of one fp32 scalar (the grad-norm partial) through per-rank boards + signal pads */
int dlti_sumsq_exchange(const float* local, const void* const* boards, const void* const* signal_pads,
                        int rank, int world, int channel, unsigned int epoch, float* out, int* err_flag,
                        dlti_stream_t stream);
/* K9: AdamW on this rank's fp32 partition whose updated values are pushed into every rank's fp32
 * master copy (the all-gather of ZeRO-1/2 folded into the optimizer kernel). */
int dlti_adamw_push(const void* const* masters, float* m, float* v, float* g, long long lo, long long n,
                    int rank, int world, float lr, float b1, float b2, float eps, float wd, int step,
                    const float* gnorm_sq, float inv_scale, float max_norm, int* skipped_out,
                    dlti_stream_t stream);

/* ---- Serving kernels (replace vLLM's rotary_embedding + reshape_and_cache,
 * paged_attention_v2 and the sampler). */
/* RoPE on q,k in place + write of the rotated k and v rows into the paged cache slots (one kernel) */
int dlti_rope_cache(void* qkv, const int* pos, const float* cos_t, const float* sin_t,
                    const long long* slot_mapping, void* k_cache, void* v_cache, long long tokens, int H,
                    int KV, int hd, long long row_stride, int dtype, dlti_stream_t stream);
/* split-KV paged decode attention; workspace holds per-split (m, l, acc) in fp32 */
int dlti_paged_decode_workspace_floats(int batch, int H, int splits);
int dlti_paged_decode(const void* q, long long q_stride, const void* k_cache, const void* v_cache,
                      const int* block_tables, int max_blocks, const int* ctx_lens, int batch,
                      int block_size, int H, int KV, int hd, float scale, int splits, float* workspace,
                      void* out, long long out_stride, int dtype, dlti_stream_t stream);
/* temperature / top-k / top-p / greedy sampling, one CTA per sequence, counter-based RNG keyed by
 * (seeds[b], step); out_logprob may be NULL */
int dlti_sample(const void* logits, long long ld, int batch, int V, const float* temperature,
                const int* top_k, const float* top_p, const unsigned long long* seeds,
                unsigned long long step, long long* out, float* out_logprob, int dtype,
                dlti_stream_t stream);

#ifdef DLTI_ATTN_TRACE
/* Developer builds only (-DDLTI_ATTN_TRACE): clock64() timeline of CTA (0,0,0) of the last backward
 * kernels, [dQ | dK/dV][role: softmax, mma, producer][iteration < 64][event < 4]. */
int dlti_attn_trace_read(long long* host_out);
#endif

/* ---- libdlti_cpu.so: shared-memory control ring for the tensor-parallel workers of one node (no
 * counterpart in the reference, which ships no serving code; vLLM's shm_broadcast queue plays this role there).
 * One writer (reader < 0 creates the ring), up to 16 readers, every reader sees every message in order. */
void* dlti_shm_ring_open(const char* name, int reader, int n_readers, int n_slots, int slot_bytes);
int dlti_shm_ring_write(void* ring, const void* data, long long nbytes, int timeout_ms);
long long dlti_shm_ring_read(void* ring, void* out, long long capacity, int timeout_ms);
void dlti_shm_ring_close(void* ring);

#ifdef __cplusplus
}
#endif
#endif /* DLTI_B200_H_ */
