/* hyperion_b200.h -- C ABI of libhyperion_b200.so (sm_100a only).
 *
 * Conventions
 *  - every entry point returns 0 on success; otherwise a cudaError_t (> 0) or
 *    -1 for an argument error.  hyp_last_error() gives the text (thread-local).
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *  - dtype codes: 0 = fp32, 1 = fp16, 2 = bf16.
 *  - no entry point allocates device memory unless its comment says so; scratch
 *    is passed in by the caller, sized by the matching *_scratch_bytes().
 *  - pointers are device pointers unless stated otherwise.
 */
#ifndef HYPERION_B200_H_
#define HYPERION_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

enum { HYP_F32 = 0, HYP_F16 = 1, HYP_BF16 = 2 };
enum { HYP_ACT_NONE = 0, HYP_ACT_RELU = 1, HYP_ACT_GELU_ERF = 2, HYP_ACT_GELU_TANH = 3, HYP_ACT_SILU = 4 };

const char* hyp_last_error(void);
int hyp_version(void);
int hyp_device_supported(void); /* 1 iff the current device is compute capability 10.x */

/* ---- gemm_tcgen05 (csrc/gemm_tcgen05.cu) -----------------------------------
 * D[M,N] = epilogue(alpha * A . B^T), fp32 accumulation in tensor memory.
 *   a_mn_major = 0: A is row-major [M,K] with leading dim lda
 *   a_mn_major = 1: A is stored transposed, row-major [K,M] with leading dim lda
 *   b_mn_major = 0: B is row-major [N,K] (the nn.Linear weight layout), ld ldb
 *   b_mn_major = 1: B is row-major [K,N], ld ldb
 * Operand pointers must be 16-byte aligned and leading dims a multiple of 16 bytes.
 * in_dtype fp32 runs on the tensor cores as tf32.
 *   mode 0: D = act(alpha*acc + bias[n]) + aux[m,n] + beta*D ; preact_out (optional, same
 *           ld/dtype as D) receives alpha*acc + bias[n] for the activation backward
 *   mode 1: D = alpha*acc * act'(aux[m,n]) + beta*D        (activation backward)
 * splits > 1: split-K; D must be fp32 and zero-initialised (or hold a value to
 * accumulate into), plain epilogue only.
 * Reference: torch.matmul (Phase 1/01_hardware_exploration.ipynb c1 L92), every
 * nn.Linear in 02_development/distributed_utils.py:75-88. */
int hyp_gemm(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb, long long ldd,
             int in_dtype, int out_dtype, int a_mn_major, int b_mn_major, const void* bias, int bias_dtype,
             const void* aux, long long ld_aux, int aux_dtype, void* preact_out, int act, int mode, float alpha,
             float beta, int splits, void* stream);

/* ---- streaming kernels (csrc/elementwise.cu) --------------------------------
 * Reference: z = x + y bandwidth sweep, Phase 1/01_hardware_exploration.ipynb c1 L148. */
int hyp_stream_add(const void* x, const void* y, void* z, long long n, int dtype, void* stream);
int hyp_stream_copy(const void* x, void* z, long long nbytes, void* stream);
/* out[c] += sum_r x[r,c]; out is fp32 and must be initialised by the caller. */
int hyp_colsum(const void* x, float* out, int M, int N, long long ld, int dtype, void* stream);
int hyp_swiglu_fwd(const void* gate, const void* up, void* out, long long n, int dtype, void* stream);
int hyp_swiglu_bwd(const void* gate, const void* up, const void* dout, void* dgate, void* dup, long long n, int dtype,
                   void* stream);
int hyp_cast(const void* x, void* y, long long n, int in_dtype, int out_dtype, float scale, void* stream);

/* ---- fused_norm (csrc/norm.cu) ----------------------------------------------
 * z = x (+ residual); y = LayerNorm(z) or RMSNorm(z) (rms != 0). sum_out (optional)
 * receives z; mean/rstd (fp32 [rows]) are saved for the backward. d % 8 == 0, d <= 8192.
 * Reference: nn.LayerNorm in nn.TransformerEncoderLayer (distributed_utils.py:79),
 * LlamaRMSNorm via distributed_utils.py:465. */
int hyp_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y, void* sum_out,
                 float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype, void* stream);
long long hyp_norm_bwd_scratch_bytes(int rows, int d);
/* dz, and fp32 dgamma / dbeta (either may be NULL); `scratch` is caller-owned
 * global memory of at least hyp_norm_bwd_scratch_bytes(rows, d) bytes. */
int hyp_norm_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd, void* dz,
                 float* dgamma, float* dbeta, void* scratch, long long scratch_bytes, int rows, int d, int rms,
                 int dtype, int accumulate, void* stream);

/* hyp_norm_bwd without caller scratch.  It has no counterpart in the reference: it exists for callers that
 * cannot allocate device memory at the point where they call us -- a plugin inside another runtime's callback,
 * a fixed-budget real-time loop, a C wrapper with no allocator.  The partial sums go to a 16 MiB array built
 * into the library (held on every GPU that loads it, until exit); shapes needing more
 * (hyp_norm_bwd_scratch_bytes(rows, d) > hyp_norm_bwd_noalloc_max_scratch_bytes()) are refused with an error.
 * The first call on a device resolves the array's device address -- make it before any stream capture.  All calls
 * share the array: calls that overlap on the GPU race; serialise them.  Never allocates or frees. */
long long hyp_norm_bwd_noalloc_max_scratch_bytes(void);
int hyp_norm_bwd_noalloc(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                         void* dz, float* dgamma, float* dbeta, int rows, int d, int rms, int dtype, int accumulate,
                         void* stream);

/* ---- fused_adamw (csrc/optim.cu) --------------------------------------------
 * One pass over a flat fp32 parameter buffer; optionally writes a 16-bit compute
 * copy.  sumsq / inv_scale / found_inf are optional DEVICE scalars (global-norm
 * clipping, AMP unscale, skip-on-overflow) so the step needs no host sync.  step_ptr (optional device fp32)
 * overrides `step` so a captured CUDA graph can be replayed with an advancing step count.
 * Reference: torch.optim.AdamW.step (distributed_utils.py:161, 231, 334, 503),
 * GradScaler.step (:179), clip_grad_norm_ (:351, 522). */
int hyp_adamw(float* p, const void* g, float* m, float* v, void* p_copy, long long n, int grad_dtype, int copy_dtype,
              float lr, float beta1, float beta2, float eps, float weight_decay, int step, float max_norm,
              const float* sumsq, const float* inv_scale, const float* found_inf, const float* step_ptr, void* stream);
/* out2[0] += sum(x^2); out2[1] = 1 if a non-finite value was seen. */
int hyp_sumsq(const void* x, long long n, int dtype, float* out2, void* stream);

/* ---- fused cross-entropy (csrc/cross_entropy.cu) -------------------------------
 * Per row: lse, NLL (accumulated into loss_sum_count[0], count into [1]) and, when
 * write_grad != 0, logits are overwritten by (softmax - onehot) * gscale * (*gscale_ptr).
 * Reference: nn.CrossEntropyLoss(ignore_index=pad) (distributed_utils.py:162, 177). */
int hyp_ce_fwd_bwd(void* logits, long long ld, const long long* targets, int rows, int V, long long ignore_index,
                   float* loss_sum_count, float* row_lse, const float* gscale_ptr, float gscale, int write_grad,
                   int dtype, void* stream);
int hyp_ce_finalize(const float* sum_count, float* loss_out, float* inv_count_out, void* stream);

/* ---- embedding (csrc/embedding.cu) ------------------------------------------
 * Reference: nn.Embedding (distributed_utils.py:78). dtable is fp32, accumulated into. */
int hyp_embedding_fwd(const void* table, const long long* idx, void* out, long long n_tok, int d, long long vocab,
                      int dtype, void* stream);
int hyp_embedding_bwd(const void* dout, const long long* idx, float* dtable, long long n_tok, int d, long long vocab,
                      long long padding_idx, int dtype, void* stream);

/* ---- rope (csrc/rope.cu) -------------------------------------------------------
 * In-place rotate-half RoPE on x[b,s,h,:] = x + b*stride_b + s*stride_s + h*stride_h.
 * positions: optional int64 [B,S]; inverse != 0 applies the transpose (backward).
 * Reference: HF LlamaAttention rotary embedding via distributed_utils.py:465-468. */
int hyp_rope(void* x, int B, int S, int H, int D, long long stride_b, long long stride_s, long long stride_h,
             const long long* positions, float theta, int inverse, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HYPERION_B200_H_ */
