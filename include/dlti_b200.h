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

#ifdef __cplusplus
}
#endif
#endif /* DLTI_B200_H_ */
