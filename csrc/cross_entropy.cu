// fused cross-entropy over a (chunk of) logits: one block per row computes the
// log-sum-exp online, the NLL for the target, and -- in the same launch --
// overwrites the logits with d(loss)/d(logits) = (softmax - onehot) * gscale,
// so the forward and the backward of log_softmax + nll_loss are a single pass
// plus one re-read that hits L2.  ignore_index rows contribute zero loss and a
// zero gradient.  The row losses and the count of non-ignored rows are
// accumulated into a device fp32[2] (sum, count); no host sync.
//
// Used by ops.linear_cross_entropy: the vocabulary projection writes 16-bit
// logits through gemm_tcgen05 (one slice by default; HYPERION_CE_L2_ROWS=n
// slices them by rows), this kernel turns them into the loss and, in place,
// into d(loss)/d(logits), and the dgrad / wgrad GEMMs read that -- loss and all
// gradients of the head in one op; no fp32 logits, no separate softmax or
// backward pass (DESIGN.md has the numbers against the recompute variant).
// The default kernel keeps a row in registers (1024 threads x 7 vectors);
// rows equal to ignore_index store zeros without being read.
// Reference call sites: nn.Linear(emb, vocab) + nn.CrossEntropyLoss(
// ignore_index=pad) at 02_development/distributed_utils.py:81, 162, 177, 335, 348.
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {

template <typename T>
__global__ void __launch_bounds__(512)
ce_fwd_bwd_kernel(T* __restrict__ logits, long long ld, const long long* __restrict__ targets, int rows, int V,
                  long long ignore_index, float* __restrict__ loss_sum_count, float* __restrict__ row_lse,
                  const float* __restrict__ gscale_ptr, float gscale, int write_grad) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  T* lrow = logits + static_cast<long long>(row) * ld;
  const long long tgt = targets[row];
  const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= V;

  // pass 1: online max / sum-exp (16-byte vector loads when the row is aligned)
  float mx = -INFINITY, se = 0.f;
  constexpr int VE = 16 / sizeof(T);
  const bool vec = (sizeof(T) == 2) && ((ld * sizeof(T)) % 16 == 0) && (reinterpret_cast<uintptr_t>(logits) % 16 == 0);
  const int Vv = vec ? (V / VE) * VE : 0;
  if (vec) {
    for (int c = threadIdx.x * VE; c < Vv; c += blockDim.x * VE) {
      float f[8];
      unpack8<T>(*reinterpret_cast<const uint4*>(lrow + c), f);
      float cm = f[0];
#pragma unroll
      for (int i = 1; i < 8; ++i) cm = fmaxf(cm, f[i]);
      if (cm > mx) { se *= __expf(mx - cm); mx = cm; }
#pragma unroll
      for (int i = 0; i < 8; ++i) se += __expf(f[i] - mx);
    }
  }
  for (int c = Vv + threadIdx.x; c < V; c += blockDim.x) {
    const float x = to_f32<T>(lrow[c]);
    if (x > mx) { se = se * __expf(mx - x) + 1.f; mx = x; }
    else se += __expf(x - mx);
  }
  const float bmx = block_max(mx, red);
  se = (mx == -INFINITY) ? 0.f : se * __expf(mx - bmx);
  se = block_sum(se, red);
  const float lse = bmx + __logf(se);
  if (threadIdx.x == 0) {
    if (row_lse) row_lse[row] = lse;
    if (!ignored) {
      atomicAdd(loss_sum_count, lse - to_f32<T>(lrow[tgt]));
      atomicAdd(loss_sum_count + 1, 1.f);
    }
  }
  if (!write_grad) return;
  __syncthreads();  // the target logit was read above; now overwrite the row
  const float gs = ignored ? 0.f : (gscale_ptr ? *gscale_ptr * gscale : gscale);
  if (vec) {
    for (int c = threadIdx.x * VE; c < Vv; c += blockDim.x * VE) {
      float f[8];
      unpack8<T>(*reinterpret_cast<const uint4*>(lrow + c), f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float gr = __expf(f[i] - lse);
        if (c + i == tgt) gr -= 1.f;
        f[i] = gr * gs;
      }
      *reinterpret_cast<uint4*>(lrow + c) = pack8<T>(f);
    }
  }
  for (int c = Vv + threadIdx.x; c < V; c += blockDim.x) {
    const float x = to_f32<T>(lrow[c]);
    float gr = __expf(x - lse);
    if (c == tgt) gr -= 1.f;
    lrow[c] = from_f32<T>(gr * gs);
  }
}

// Same computation for 16-bit rows of up to 512 * 8 * RV elements with the whole row held in registers (RV 16-byte
// vectors per thread): the logits are read ONCE, the gradient is produced from the registers.  A GPT-2 row (50257
// bf16) is 13 vectors per thread at 512 threads (108 registers: one 16-warp block per SM) or 7 vectors at 1024 threads
// (<= 64 registers: 32 warps per SM, twice the loads in flight) -- the latter is used whenever the row fits.
template <typename T, int RV, int BLOCK = 512>
__global__ void __launch_bounds__(BLOCK)
ce_fwd_bwd_reg_kernel(T* __restrict__ logits, long long ld, const long long* __restrict__ targets, int rows, int V,
                      long long ignore_index, float* __restrict__ loss_sum_count, float* __restrict__ row_lse,
                      const float* __restrict__ gscale_ptr, float gscale, int write_grad) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  T* lrow = logits + static_cast<long long>(row) * ld;
  const long long tgt = targets[row];
  const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= V;
  const int Vv = (V / 8) * 8;
  if (ignored && write_grad && row_lse == nullptr) {
    // a padding position (ignore_index) adds nothing to the loss and its gradient row is zero: store the zeros without
    // reading the logits (about 40 % of the rows of WikiText-2-shaped batches)
#pragma unroll
    for (int k = 0; k < RV; ++k) {
      const int c = (k * BLOCK + threadIdx.x) * 8;
      if (c < Vv) *reinterpret_cast<uint4*>(lrow + c) = make_uint4(0u, 0u, 0u, 0u);
    }
    if (threadIdx.x < V - Vv) lrow[Vv + threadIdx.x] = from_f32<T>(0.f);
    return;
  }
  uint4 raw[RV];
#pragma unroll
  for (int k = 0; k < RV; ++k) {
    const int c = (k * BLOCK + threadIdx.x) * 8;
    raw[k] = make_uint4(0u, 0u, 0u, 0u);
    if (c < Vv) raw[k] = *reinterpret_cast<const uint4*>(lrow + c);
  }
  float mx = -INFINITY;
#pragma unroll
  for (int k = 0; k < RV; ++k) {
    const int c = (k * BLOCK + threadIdx.x) * 8;
    if (c < Vv) {
      float f[8];
      unpack8<T>(raw[k], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) mx = fmaxf(mx, f[i]);
    }
  }
  float tail = 0.f;  // up to 7 trailing elements: one per thread of the first warp
  const bool has_tail = threadIdx.x < V - Vv;
  if (has_tail) { tail = to_f32<T>(lrow[Vv + threadIdx.x]); mx = fmaxf(mx, tail); }
  const float bmx = block_max(mx, red);
  const float nb = -bmx * 1.4426950408889634f;
  float se = 0.f;
#pragma unroll
  for (int k = 0; k < RV; ++k) {
    const int c = (k * BLOCK + threadIdx.x) * 8;
    if (c < Vv) {
      float f[8];
      unpack8<T>(raw[k], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) se += ex2_fast(fmaf(f[i], 1.4426950408889634f, nb));
    }
  }
  if (has_tail) se += ex2_fast(fmaf(tail, 1.4426950408889634f, nb));
  se = block_sum(se, red);
  const float lse = bmx + __logf(se);
  if (threadIdx.x == 0) {
    if (row_lse) row_lse[row] = lse;
    if (!ignored) {
      atomicAdd(loss_sum_count, lse - to_f32<T>(lrow[tgt]));
      atomicAdd(loss_sum_count + 1, 1.f);
    }
  }
  if (!write_grad) return;
  __syncthreads();  // the target logit was read above; now overwrite the row
  const float gs = ignored ? 0.f : (gscale_ptr ? *gscale_ptr * gscale : gscale);
  const float nl = -lse * 1.4426950408889634f;
#pragma unroll
  for (int k = 0; k < RV; ++k) {
    const int c = (k * BLOCK + threadIdx.x) * 8;
    if (c < Vv) {
      float f[8];
      unpack8<T>(raw[k], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float gr = ex2_fast(fmaf(f[i], 1.4426950408889634f, nl));
        if (c + i == tgt) gr -= 1.f;
        f[i] = gr * gs;
      }
      *reinterpret_cast<uint4*>(lrow + c) = pack8<T>(f);
    }
  }
  if (has_tail) {
    float gr = ex2_fast(fmaf(tail, 1.4426950408889634f, nl));
    if (Vv + static_cast<int>(threadIdx.x) == tgt) gr -= 1.f;
    lrow[Vv + threadIdx.x] = from_f32<T>(gr * gs);
  }
}

// dlogits *= scale[0] ... used when the mean-reduction denominator (count of
// non-ignored tokens) is only known after all chunks: out = 1 / max(count, 1)
__global__ void ce_finalize_kernel(const float* __restrict__ sum_count, float* __restrict__ loss_out,
                                   float* __restrict__ inv_count_out) {
  const float cnt = fmaxf(sum_count[1], 1.f);
  if (loss_out) *loss_out = sum_count[0] / cnt;
  if (inv_count_out) *inv_count_out = 1.f / cnt;
}

}  // namespace hyp

using namespace hyp;

HYP_API int hyp_ce_fwd_bwd(void* logits, long long ld, const long long* targets, int rows, int V,
                           long long ignore_index, float* loss_sum_count, float* row_lse, const float* gscale_ptr,
                           float gscale, int write_grad, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && V > 0, "hyp_ce_fwd_bwd: empty input");
  const int threads = V >= 8192 ? 512 : 256;
  // rows that fit in registers (16-bit, 16-byte aligned): single read of the logits
  const bool aligned = dtype != DT_F32 && (ld * 2) % 16 == 0 && reinterpret_cast<uintptr_t>(logits) % 16 == 0;
  if (aligned && V >= 8192 && V <= 512 * 8 * 16) {
    const int rv = (V / 8 + 511) / 512;
    static const bool wide = [] { const char* e = getenv("HYPERION_CE_BLOCK"); return !(e && atoi(e) == 512); }();
    if (wide && (V / 8 + 1023) / 1024 <= 7) {
      if (dtype == DT_BF16)
        ce_fwd_bwd_reg_kernel<__nv_bfloat16, 7, 1024><<<rows, 1024, 0, st>>>((__nv_bfloat16*)logits, ld, targets, rows, V, ignore_index, loss_sum_count, row_lse, gscale_ptr, gscale, write_grad);
      else
        ce_fwd_bwd_reg_kernel<__half, 7, 1024><<<rows, 1024, 0, st>>>((__half*)logits, ld, targets, rows, V, ignore_index, loss_sum_count, row_lse, gscale_ptr, gscale, write_grad);
      HYP_CHECK_CUDA(cudaGetLastError());
      return 0;
    }
#define HYP_CE_REG(T, RV) ce_fwd_bwd_reg_kernel<T, RV><<<rows, 512, 0, st>>>((T*)logits, ld, targets, rows, V, ignore_index, loss_sum_count, row_lse, gscale_ptr, gscale, write_grad)
#define HYP_CE_REG_T(T) \
    if (rv <= 4) HYP_CE_REG(T, 4); else if (rv <= 8) HYP_CE_REG(T, 8); else if (rv <= 13) HYP_CE_REG(T, 13); else HYP_CE_REG(T, 16)
    if (dtype == DT_BF16) { HYP_CE_REG_T(__nv_bfloat16); } else { HYP_CE_REG_T(__half); }
#undef HYP_CE_REG_T
#undef HYP_CE_REG
    HYP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  if (dtype == DT_F32)
    ce_fwd_bwd_kernel<float><<<rows, threads, 0, st>>>((float*)logits, ld, targets, rows, V, ignore_index, loss_sum_count, row_lse, gscale_ptr, gscale, write_grad);
  else if (dtype == DT_BF16)
    ce_fwd_bwd_kernel<__nv_bfloat16><<<rows, threads, 0, st>>>((__nv_bfloat16*)logits, ld, targets, rows, V, ignore_index, loss_sum_count, row_lse, gscale_ptr, gscale, write_grad);
  else
    ce_fwd_bwd_kernel<__half><<<rows, threads, 0, st>>>((__half*)logits, ld, targets, rows, V, ignore_index, loss_sum_count, row_lse, gscale_ptr, gscale, write_grad);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_ce_finalize(const float* sum_count, float* loss_out, float* inv_count_out, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  ce_finalize_kernel<<<1, 1, 0, st>>>(sum_count, loss_out, inv_count_out);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
