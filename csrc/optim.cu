// fused_adamw: one launch updates a whole flat parameter buffer.
//
// The framework keeps parameters in flat buffers (fp32 master + optional 16-bit
// compute copy, fp32 exp_avg / exp_avg_sq), so "multi-tensor AdamW" is one
// streaming pass: read g, p, m, v; write p, m, v and the bf16/fp16 copy.  The
// AMP unscale (1/loss_scale), the global-norm clip coefficient and the
// found_inf skip are all read from device scalars, so no host sync is needed.
//
// Reference call sites: torch.optim.AdamW(...).step() (02_development/
// distributed_utils.py:161, 231, 334, 503), GradScaler.unscale_/step (:178-180)
// and torch.nn.utils.clip_grad_norm_ (:351, 522).
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay;
  float bias_c1, bias_c2;  // 1 - beta^step
  float max_norm;          // <= 0: no clipping
  const float* sumsq;      // device: sum of squares of the (unscaled) global gradient, or null
  const float* inv_scale;  // device: 1 / loss_scale, or null
  const float* found_inf;  // device: != 0 -> skip the step, or null
  const float* step_ptr;   // device: step count (>= 1) when the step number must not be baked into the launch (CUDA graphs)
};

template <typename TG, typename TC>
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             TC* __restrict__ p_copy, long long n, AdamArgs a) {
  if (a.found_inf && *a.found_inf != 0.f) return;
  float gscale = a.inv_scale ? *a.inv_scale : 1.f;
  if (a.max_norm > 0.f && a.sumsq) {
    const float norm = sqrtf(*a.sumsq) * (a.inv_scale ? *a.inv_scale : 1.f);
    const float coef = a.max_norm / (norm + 1e-6f);
    if (coef < 1.f) gscale *= coef;
  }
  float c1 = a.bias_c1, c2 = a.bias_c2;
  if (a.step_ptr) {
    const float t = fmaxf(*a.step_ptr, 1.f);
    c1 = 1.f - powf(a.beta1, t);
    c2 = 1.f - powf(a.beta2, t);
  }
  const float step_size = a.lr / c1;
  const float inv_sqrt_c2 = rsqrtf(c2);
  const float decay = 1.f - a.lr * a.weight_decay;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 4;
  long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  for (; i + 3 < n; i += stride) {
    float4 pp = *reinterpret_cast<float4*>(p + i);
    float4 mm = *reinterpret_cast<float4*>(m + i);
    float4 vv = *reinterpret_cast<float4*>(v + i);
    float gg[4];
    if constexpr (sizeof(TG) == 4) {
      const float4 t = *reinterpret_cast<const float4*>(g + i);
      gg[0] = t.x; gg[1] = t.y; gg[2] = t.z; gg[3] = t.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) gg[k] = to_f32<TG>(g[i + k]);
    }
    float* pf = reinterpret_cast<float*>(&pp);
    float* mf = reinterpret_cast<float*>(&mm);
    float* vf = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = gg[k] * gscale;
      mf[k] = a.beta1 * mf[k] + (1.f - a.beta1) * gr;
      vf[k] = a.beta2 * vf[k] + (1.f - a.beta2) * gr * gr;
      const float denom = sqrtf(vf[k]) * inv_sqrt_c2 + a.eps;
      pf[k] = pf[k] * decay - step_size * mf[k] / denom;
    }
    *reinterpret_cast<float4*>(p + i) = pp;
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    if (p_copy) {
#pragma unroll
      for (int k = 0; k < 4; ++k) p_copy[i + k] = from_f32<TC>(pf[k]);
    }
  }
  for (; i < n; ++i) {  // at most 3 tail elements, handled by the one thread that lands here
    const float gr = to_f32<TG>(g[i]) * gscale;
    const float mk = a.beta1 * m[i] + (1.f - a.beta1) * gr;
    const float vk = a.beta2 * v[i] + (1.f - a.beta2) * gr * gr;
    const float pk = p[i] * decay - step_size * mk / (sqrtf(vk) * inv_sqrt_c2 + a.eps);
    m[i] = mk; v[i] = vk; p[i] = pk;
    if (p_copy) p_copy[i] = from_f32<TC>(pk);
  }
}

// out[0] += sum(x^2); out[1] = 1 if any non-finite value was seen.
template <typename T>
__global__ void __launch_bounds__(256) sumsq_kernel(const T* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.f;
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  constexpr int VE = 16 / sizeof(T);
  const bool vec = (reinterpret_cast<uintptr_t>(x) % 16) == 0;
  const long long nv = vec ? n / VE : 0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nv; i += stride) {
    const uint4 u = ld_nc_v4(reinterpret_cast<const uint4*>(x) + i);
    if constexpr (sizeof(T) == 4) {
      const float* f = reinterpret_cast<const float*>(&u);
#pragma unroll
      for (int k = 0; k < 4; ++k) { s += f[k] * f[k]; bad |= !isfinite(f[k]); }
    } else {
      float f[8];
      unpack8<T>(u, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) { s += f[k] * f[k]; bad |= !isfinite(f[k]); }
    }
  }
  for (long long i = nv * VE + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float f = to_f32<T>(x[i]);
    s += f * f;
    bad |= !isfinite(f);
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) atomicAdd(out, s);
  if (bad) out[1] = 1.f;
}

}  // namespace hyp

using namespace hyp;

HYP_API int hyp_adamw(float* p, const void* g, float* m, float* v, void* p_copy, long long n, int grad_dtype,
                      int copy_dtype, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                      float max_norm, const float* sumsq, const float* inv_scale, const float* found_inf,
                      const float* step_ptr, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (n == 0) return 0;
  HYP_REQUIRE(step >= 1 || step_ptr, "hyp_adamw: step counts from 1");
  if (step < 1) step = 1;
  HYP_REQUIRE(((uintptr_t)p | (uintptr_t)m | (uintptr_t)v) % 16 == 0, "hyp_adamw: state pointers must be 16B aligned");
  HYP_REQUIRE(grad_dtype != DT_F32 || (uintptr_t)g % 16 == 0, "hyp_adamw: fp32 grad pointer must be 16B aligned");
  AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bias_c1 = 1.f - powf(beta1, (float)step);
  a.bias_c2 = 1.f - powf(beta2, (float)step);
  a.max_norm = max_norm; a.sumsq = sumsq; a.inv_scale = inv_scale; a.found_inf = found_inf; a.step_ptr = step_ptr;
  long long blocks = (n / 4 + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  const int grid = (int)blocks;
#define HYP_ADAM(TG, TC) adamw_kernel<TG, TC><<<grid, 256, 0, st>>>(p, (const TG*)g, m, v, (TC*)p_copy, n, a)
  if (grad_dtype == DT_F32) {
    if (copy_dtype == DT_F16) HYP_ADAM(float, __half); else HYP_ADAM(float, __nv_bfloat16);
  } else if (grad_dtype == DT_BF16) {
    if (copy_dtype == DT_F16) HYP_ADAM(__nv_bfloat16, __half); else HYP_ADAM(__nv_bfloat16, __nv_bfloat16);
  } else {
    if (copy_dtype == DT_F16) HYP_ADAM(__half, __half); else HYP_ADAM(__half, __nv_bfloat16);
  }
#undef HYP_ADAM
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_sumsq(const void* x, long long n, int dtype, float* out2, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (n == 0) return 0;
  long long blocks = (n + 2047) / 2048;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  const int grid = (int)blocks;
  if (dtype == DT_F32) sumsq_kernel<float><<<grid, 256, 0, st>>>((const float*)x, n, out2);
  else if (dtype == DT_BF16) sumsq_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, n, out2);
  else sumsq_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, n, out2);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ loss scaler
// GradScaler's bookkeeping as two one-thread kernels instead of ~15 tiny elementwise launches per step
// (torch.cuda.amp.GradScaler in the reference's DDP trainers, 02_development/distributed_utils.py:163, 178-180).
namespace hyp {
__global__ void scaler_note_kernel(float* __restrict__ found_inf, const float* __restrict__ sumsq_flag) {
  // sumsq_flag: (sum of squares, non-finite flag) from the norm / reduce kernels
  const float s = sumsq_flag[0], f = sumsq_flag[1];
  if (f != 0.f || !isfinite(s)) found_inf[0] = 1.f;
}
__global__ void scaler_update_kernel(float* __restrict__ scale, float* __restrict__ inv_scale, float* __restrict__ growth_tracker,
                                     float* __restrict__ found_inf, float growth, float backoff, float interval) {
  float sc = scale[0], gt = growth_tracker[0];
  if (found_inf[0] != 0.f) {
    sc *= backoff;
    gt = 0.f;
  } else {
    gt += 1.f;
    if (gt >= interval) { sc *= growth; gt = 0.f; }
  }
  scale[0] = sc;
  inv_scale[0] = 1.f / sc;
  growth_tracker[0] = gt;
  found_inf[0] = 0.f;
}
}  // namespace hyp

HYP_API int hyp_scaler_note(float* found_inf, const float* sumsq_flag, void* stream_) {
  hyp::scaler_note_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(found_inf, sumsq_flag);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_scaler_update(float* scale, float* inv_scale, float* growth_tracker, float* found_inf, float growth_factor,
                              float backoff_factor, float growth_interval, void* stream_) {
  hyp::scaler_update_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(scale, inv_scale, growth_tracker, found_inf,
                                                                                   growth_factor, backoff_factor, growth_interval);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
