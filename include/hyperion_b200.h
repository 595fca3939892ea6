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
/* mode 1 (D = alpha*acc * act'(aux), the activation-backward epilogue) has no pre-activation output: there `preact_out`,
 * when non-NULL, is an fp32 [N] accumulator to which the epilogue ADDS the column sums of the rounded 16-bit output (the bias
 * gradient of the layer below), N % 32 == 0; the caller zeroes it. */
int hyp_gemm(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb, long long ldd,
             int in_dtype, int out_dtype, int a_mn_major, int b_mn_major, const void* bias, int bias_dtype,
             const void* aux, long long ld_aux, int aux_dtype, void* preact_out, int act, int mode, float alpha,
             float beta, int splits, void* stream);
/* D = epilogue(alpha * (A B^T + A2 B2^T)): the reduction runs over a second, K-major operand pair as well (K2 % 64 == 0,
 * zero-pad).  One launch for a LoRA projection (A2 = s * x A_l^T, B2 = B_l).  16-bit, M >= 256, N > 128; bias / residual
 * epilogue only.  Reference: peft LoRA forward base(x) + B(A(dropout(x))) * alpha/r (distributed_utils.py:469-474). */
int hyp_gemm_dualk(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb, long long ldd,
                   int in_dtype, int out_dtype, int a_mn_major, int b_mn_major, const void* A2, const void* B2, int K2,
                   long long lda2, long long ldb2, const void* bias, int bias_dtype, const void* aux, long long ld_aux,
                   int aux_dtype, float alpha, void* stream);

/* ---- streaming kernels (csrc/elementwise.cu) --------------------------------
 * Reference: z = x + y bandwidth sweep, Phase 1/01_hardware_exploration.ipynb c1 L148. */
int hyp_stream_add(const void* x, const void* y, void* z, long long n, int dtype, void* stream);
int hyp_stream_copy(const void* x, void* z, long long nbytes, void* stream);
/* y = x * keep * 1/(1-p_eff) over a row-major [rows, cols] tensor, Philox mask as described at hyp_norm_fwd_drop; applying
 * it to a gradient with the same (rng_state, salt, p) is the backward.  y == x is allowed.
 * Reference: nn.Dropout inside nn.TransformerEncoderLayer (distributed_utils.py:79), lora_dropout (:469-473). */
int hyp_dropout(const void* x, void* y, long long rows, int cols, int dtype, const unsigned long long* rng_state, int salt,
                float dropout_p, void* stream);
/* out[c] += sum_r x[r,c]; out is fp32 and must be initialised by the caller. */
int hyp_colsum(const void* x, float* out, int M, int N, long long ld, int dtype, void* stream);
/* Column sums of a 16-bit matrix rounded to the same dtype in one launch.  acc_ws (>= N fp32) and counter (one
 * word) must be zero on entry and are zero again on return (the last block converts and clears). */
int hyp_colsum_cast(const void* x, void* out16, int M, int N, long long ld, int dtype, float* acc_ws, unsigned* counter,
                    void* stream);
int hyp_swiglu_fwd(const void* gate, const void* up, void* out, long long n, int dtype, void* stream);
int hyp_swiglu_bwd(const void* gate, const void* up, const void* dout, void* dgate, void* dup, long long n, int dtype,
                   void* stream);
int hyp_cast(const void* x, void* y, long long n, int in_dtype, int out_dtype, float scale, void* stream);
/* hyp_cast with one more factor read from device memory (fp32 scalar, may be NULL). */
int hyp_cast_dscale(const void* x, void* y, long long n, int in_dtype, int out_dtype, float scale, const float* scale_ptr,
                    void* stream);

/* ---- fused_norm (csrc/norm.cu) ----------------------------------------------
 * z = x (+ residual); y = LayerNorm(z) or RMSNorm(z) (rms != 0). sum_out (optional)
 * receives z; mean/rstd (fp32 [rows]) are saved for the backward. d % 8 == 0, d <= 8192.
 * Reference: nn.LayerNorm in nn.TransformerEncoderLayer (distributed_utils.py:79),
 * LlamaRMSNorm via distributed_utils.py:465. */
int hyp_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y, void* sum_out,
                 float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype, void* stream);
/* Same with z = dropout(x, p) + residual: the residual-branch dropout (nn.TransformerEncoderLayer dropout1 / dropout2,
 * distributed_utils.py:79) folded into the add + norm pass.  rng_state: device uint64[2] = (seed, step); salt tells the
 * dropout calls of one step apart; the mask is a Philox4x32-10 function of (seed, step, salt, row, col) with 8 bits per
 * element (effective p = round(256 p) / 256, survivors scaled by 1 / (1 - p_eff)).  Gradient of x: hyp_dropout(dz) with
 * the same three values; gradient of residual: dz. */
int hyp_norm_fwd_drop(const void* x, const void* residual, const void* gamma, const void* beta, void* y, void* sum_out,
                      float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype,
                      const unsigned long long* rng_state, int salt, float dropout_p, void* stream);
long long hyp_norm_bwd_scratch_bytes(int rows, int d);
/* dz, and fp32 dgamma / dbeta (either may be NULL); `scratch` is caller-owned
 * global memory of at least hyp_norm_bwd_scratch_bytes(rows, d) bytes. */
int hyp_norm_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd, void* dz,
                 float* dgamma, float* dbeta, void* scratch, long long scratch_bytes, int rows, int d, int rms,
                 int dtype, int accumulate, void* stream);

/* hyp_norm_bwd as the framework's ops use it: `dres` (may be NULL) is the gradient reaching the norm's input through
 * the skip connection and is added to dz in the same pass; dgamma / dbeta are written as `param_dtype` (0 fp32,
 * 1 fp16, 2 bf16) so they can be stored straight into a 16-bit flat gradient buffer; `dz_colsum` (may be NULL, d <= 2048)
 * receives the column sums of dz (the bias gradient of the layer feeding this norm); and there is no second kernel --
 * partial sums are added (fp32 atomics) into the accumulator in `ws`, and the last block rounds it out.  `ws`
 * (hyp_norm_bwd_fused_ws_bytes(d) bytes, 16-byte aligned) must be zero on entry and is zero again on return.
 * No reference counterpart (autograd adds the two branches with a separate elementwise kernel there). */
long long hyp_norm_bwd_fused_ws_bytes(int d);
int hyp_norm_bwd_fused(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                       const void* dres, void* dz, void* dgamma, void* dbeta, void* dz_colsum, int param_dtype, void* ws,
                       long long ws_bytes, int rows, int d, int rms, int dtype, void* stream);

/* hyp_norm_bwd without caller scratch.  It has no counterpart in the reference: it exists for callers that
 * cannot allocate device memory at the point where they call us -- a plugin inside another runtime's callback,
 * a fixed-budget real-time loop, a C wrapper with no allocator.  The partial sums go to a 8 MiB array built
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
/* Loss-scaler bookkeeping on the device (GradScaler of the reference's DDP trainers, distributed_utils.py:163, 178-180):
 * note = raise found_inf when the (sum of squares, non-finite flag) pair reports an overflow; update = back off on overflow,
 * grow after growth_interval clean steps, refresh inv_scale, clear found_inf.  All arguments are device fp32 scalars. */
int hyp_scaler_note(float* found_inf, const float* sumsq_flag, void* stream);
int hyp_scaler_update(float* scale, float* inv_scale, float* growth_tracker, float* found_inf, float growth_factor,
                      float backoff_factor, float growth_interval, void* stream);

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

/* ---- flash_attn_sm100 (csrc/attention.cu) -------------------------------------------
 * softmax(Q K^T * scale + mask) V on tcgen05 with TMEM accumulators; head dim 64 or 128, fp16 / bf16.
 * q/k/v/o (and d_o, dk, dv) are indexed [b, h, s, d] through ELEMENT strides (unit stride in d; strides multiples of 8),
 * read in place by 4-D TMA descriptors.  key_padding_mask: uint8 [B, Sk], 1 = attend, or NULL.  lse: fp32 [B, H, Sq].
 * Backward: dq_acc is fp32 [B, H, Sq, D] scratch (cleared by the entry point, then accumulated with vector reductions); delta is fp32
 * [B, H, Sq] scratch; dk / dv are written in full.
 * Reference: nn.MultiheadAttention in nn.TransformerEncoderLayer (distributed_utils.py:79), HF Llama SDPA (:465). */
int hyp_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Sk, int D,
                       long long q_sb, long long q_sh, long long q_ss, long long k_sb, long long k_sh, long long k_ss,
                       long long v_sb, long long v_sh, long long v_ss, long long o_sb, long long o_sh, long long o_ss,
                       const unsigned char* key_padding_mask, int causal, float scale, int dtype, void* stream);
int hyp_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                       float* dq_acc, void* dk, void* dv, float* delta, const unsigned char* key_padding_mask, int B, int H,
                       int Sq, int Sk, int D, long long q_sb, long long q_sh, long long q_ss, long long k_sb, long long k_sh,
                       long long k_ss, long long v_sb, long long v_sh, long long v_ss, long long o_sb, long long o_sh,
                       long long o_ss, long long g_sb, long long g_sh, long long g_ss, long long dkv_sb, long long dkv_sh,
                       long long dkv_ss, int causal, float scale, int dtype, void* stream);
/* The same two entry points with dropout on the attention probabilities (attention dropout 0.1 of
 * nn.TransformerEncoderLayer, distributed_utils.py:79).  The forward drops P after the softmax normalisation; the backward
 * regenerates the mask from (rng_state, salt, dropout_p) -- pass the forward's values, and run it before the step word of
 * rng_state is advanced.  Mask element = Philox(seed, step, salt, b * H + h, query, key), see hyp_norm_fwd_drop. */
int hyp_flash_attn_fwd_drop(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Sk,
                            int D, long long q_sb, long long q_sh, long long q_ss, long long k_sb, long long k_sh,
                            long long k_ss, long long v_sb, long long v_sh, long long v_ss, long long o_sb, long long o_sh,
                            long long o_ss, const unsigned char* key_padding_mask, int causal, float scale, int dtype,
                            const unsigned long long* rng_state, int salt, float dropout_p, void* stream);
int hyp_flash_attn_bwd_drop(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                            float* dq_acc, void* dk, void* dv, float* delta, const unsigned char* key_padding_mask, int B,
                            int H, int Sq, int Sk, int D, long long q_sb, long long q_sh, long long q_ss, long long k_sb,
                            long long k_sh, long long k_ss, long long v_sb, long long v_sh, long long v_ss, long long o_sb,
                            long long o_sh, long long o_ss, long long g_sb, long long g_sh, long long g_ss, long long dkv_sb,
                            long long dkv_sh, long long dkv_ss, int causal, float scale, int dtype,
                            const unsigned long long* rng_state, int salt, float dropout_p, void* stream);
/* Rounds the fp32 dQ accumulator ([B,H,S,D] contiguous) into a 16-bit tensor addressed [b,h,s,d] through element
 * strides (e.g. the q slice of the packed [B,S,3,H,D] gradient buffer). */
int hyp_flash_dq_convert(const float* dq_acc, void* dq, int B, int H, int S, int D, long long sb, long long sh,
                         long long ss, int dtype, void* stream);

/* ---- NVLink communication kernels (csrc/comm.cu) ------------------------------------
 * peer_bufs / peer_flags are HOST arrays of W device addresses: this rank's mapping of every rank's symmetric data
 * buffer / flag region (hyp_comm_flag_words() uint32 words, zero-initialised).  mc_ptr: multicast (NVLS) address of
 * the data buffer or 0.  The kernels spin on flags written by peer GPUs: every rank must launch the same sequence, on
 * one stream, with at most hyp_comm_max_blocks() co-resident blocks.  The barrier epoch of a launch is
 * epoch + (epoch_base ? *epoch_base : 0); it must be equal on all ranks and grow by >= 4 per launch.
 * Reference: the implicit DDP all-reduce / FSDP reduce-scatter + all-gather of distributed_utils.py:159, 328-332. */
int hyp_comm_flag_words(void);
int hyp_comm_inf_word(void);   /* index (as float) of the word raised to 1.0 on every rank when a non-finite value is reduced */
int hyp_comm_max_blocks(void);
/* K-C1: in-place average (x extra_scale, x *inv_scale) of n elements at off_bytes, all ranks. */
int hyp_comm_allreduce(const long long* peer_bufs, const long long* peer_flags, long long mc_ptr, int rank, int W,
                       long long off_bytes, long long n, int dtype, unsigned epoch, const unsigned* epoch_base,
                       const float* inv_scale, float extra_scale, int max_blocks, void* stream);
/* K-C2: rank r reduces slice r (per_elems elements) of the region; fused_adam != 0 applies AdamW to the local fp32
 * shard (master, m, v; optional 16-bit copy p16) in the same pass, else the fp32 slice goes to gshard and its sum of
 * squares is added to sumsq[0]. */
int hyp_comm_reduce_scatter(const long long* peer_bufs, const long long* peer_flags, long long mc_ptr, int rank, int W,
                            long long off_bytes, long long per_elems, int dtype, unsigned epoch, const unsigned* epoch_base,
                            float scale, float* gshard, float* sumsq, int fused_adam, float* master, float* m, float* v,
                            void* p16, int p16_dtype, float lr, float beta1, float beta2, float eps, float weight_decay,
                            int step, const float* step_ptr, int max_blocks, void* stream);
/* K-C3 (gather only): dst[q*per_bytes ...] = rank q's bytes [shard_off_bytes, +per_bytes) for every q. */
int hyp_comm_allgather(const long long* peer_bufs, const long long* peer_flags, int rank, int W, long long shard_off_bytes,
                       long long per_bytes, void* dst, unsigned epoch, const unsigned* epoch_base, unsigned* chunk_flags,
                       int max_blocks, void* stream);
/* K-C3 fused: D = epilogue(A . B^T) where B (and bias) live in the flat unit buffer `unit_dst` (W slices of per_bytes);
 * gather warps inside the GEMM kernel pull slice q from peer_src[q] while tiles whose weights have landed are computed.
 * arrived: device uint32[W], zero at creation; a launch completes when arrived[q] == target + (target_base ? *target_base : 0). */
int hyp_gemm_ag(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb, long long ldd,
                int in_dtype, int out_dtype, const void* bias, int bias_dtype, const void* aux, long long ld_aux,
                int aux_dtype, void* preact_out, int act, float alpha, const long long* peer_src,
                const long long* peer_flags, int rank, int W, unsigned epoch, const unsigned* epoch_base, void* unit_dst,
                long long per_bytes, unsigned* arrived, unsigned target, const unsigned* target_base, void* stream);
/* Plain CUDA IPC provider of peer memory (set-up time only; hyp_ipc_alloc calls cudaMalloc). */
int hyp_ipc_alloc(long long bytes, long long* out_ptr, unsigned char* out_handle64);
int hyp_ipc_open(const unsigned char* handle64, long long* out_ptr);
int hyp_ipc_close(long long ptr);
int hyp_ipc_free(long long ptr);

/* ---- conv_igemm_tcgen05 (csrc/conv_igemm.cu) -----------------------------------------
 * NHWC convolution forward as an implicit GEMM: x [N,H,W,C], w [K,R,S,C], y [N,Ho,Wo,K], all 16-bit (dtype 1/2),
 * C % 64 == 0, K % 32 == 0, stride 1 or 2, symmetric zero padding.  Optional epilogue: y = conv * scale[k] + shift[k]
 * (fp32 [K]; a folded eval-mode BatchNorm or a bias), + residual (shaped like y), ReLU.
 * Reference: torchvision resnet18/50 Conv2d -> BatchNorm2d -> ReLU (distributed_utils.py:229, compilation_optimization.py:89). */
int hyp_conv2d_fwd(const void* x, const void* w, void* y, int N, int H, int W, int C, int K, int R, int S, int stride,
                   int pad, const float* scale, const float* shift, const void* residual, int relu, int dtype,
                   void* stream);

/* Weight gradient of the same convolution: dw [K,R,S,C] fp32 (ZEROED by the caller) += sum over output pixels of
 * dy[pix, k] * x[pix shifted by the tap, c]; x [N,H,W,C], dy [N,Ho,Wo,K] 16-bit NHWC; C % 64 == 0, K % 64 == 0. */
int hyp_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int K, int R, int S, int stride,
                     int pad, int dtype, void* stream);

/* ---- bn_act (csrc/bn_act.cu) ------------------------------------------------------
 * Training-mode BatchNorm2d + residual add + ReLU on NHWC-contiguous [rows = N*H*W, C] 16-bit tensors.
 * sums / dsums: fp32 [2*C] scratch ZEROED by the caller (batch statistics / dbeta,dgamma on return).
 * Reference: torchvision ResNet Conv2d->BatchNorm2d->ReLU(+residual) (distributed_utils.py:229). */
int hyp_bn_act_fwd(const void* x, const void* res, const float* gamma, const float* beta, void* y, float* sums, float* mean,
                   float* rstd, float* running_mean, float* running_var, long long rows, int C, float eps, float momentum,
                   int relu, int dtype, void* stream);
int hyp_bn_act_bwd(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, const float* gamma,
                   float* dsums, void* dx, void* dres, long long rows, int C, int relu, int dtype, void* stream);
/* The same pair with gamma / beta in the activations' 16-bit dtype (no per-layer cast launches in a mixed-precision module). */
int hyp_bn_act_fwd_p16(const void* x, const void* res, const void* gamma, const void* beta, void* y, float* sums, float* mean,
                       float* rstd, float* running_mean, float* running_var, long long rows, int C, float eps, float momentum,
                       int relu, int dtype, void* stream);
int hyp_bn_act_bwd_p16(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, const void* gamma,
                       float* dsums, void* dx, void* dres, long long rows, int C, int relu, int dtype, void* dgamma16,
                       void* dbeta16, void* stream);   /* dgamma16 / dbeta16: optional 16-bit [C] outputs (both or neither) */

#ifdef __cplusplus
}
#endif
#endif /* HYPERION_B200_H_ */
