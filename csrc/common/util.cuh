// Host/device helpers shared by every translation unit of libhyperion_b200.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#define HYP_API extern "C" __attribute__((visibility("default")))

namespace hyp {

// Per-CTA budgets of sm_100a, checked at compile time wherever a kernel configuration is launched.
constexpr int kMaxSmemPerCta = 227 * 1024;   // opt-in dynamic shared memory per CTA
constexpr int kTmemColumns = 512;            // TMEM columns per SM (128 lanes x 512 x 32 bit = 256 KB)
constexpr bool tmem_alloc_ok(int cols) { return cols >= 32 && cols <= kTmemColumns && (cols & (cols - 1)) == 0; }

// dtype codes of the C ABI (include/hyperion_b200.h)
enum : int { DT_F32 = 0, DT_F16 = 1, DT_BF16 = 2 };

// Last-error slot. Every entry point returns 0 on success or a negative code /
// cudaError_t; the text lives here and is read with hyp_last_error().
void set_error(const char* fmt, ...);
const char* get_error();

#define HYP_CHECK_CUDA(expr)                                                          \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) {                                                          \
      ::hyp::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,           \
                       cudaGetErrorString(_e));                                       \
      return static_cast<int>(_e);                                                    \
    }                                                                                 \
  } while (0)

#define HYP_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      ::hyp::set_error(__VA_ARGS__);      \
      return -1;                          \
    }                                     \
  } while (0)

inline int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = n > 0 ? n : 148;
  }
  return cached[dev];
}

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blocks of up to 1024 threads; `red` is 32 floats of smem.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float t = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
  if (warp == 0) {
    t = warp_sum(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float t = (threadIdx.x < nw) ? red[threadIdx.x] : -INFINITY;
  if (warp == 0) {
    t = warp_max(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// 8 x 16-bit values packed in a uint4 <-> 8 floats.
template <typename T>
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]);
template <>
__device__ __forceinline__ void unpack8<__nv_bfloat16>(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(p[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <>
__device__ __forceinline__ void unpack8<__half>(const uint4& u, float (&f)[8]) {
  const __half2* p = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __half22float2(p[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]);
template <>
__device__ __forceinline__ uint4 pack8<__nv_bfloat16>(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
template <>
__device__ __forceinline__ uint4 pack8<__half>(const float (&f)[8]) {
  uint4 u;
  __half2* p = reinterpret_cast<__half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) p[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  return u;
}

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.f + erff(x * 0.70710678118654752f));
}
__device__ __forceinline__ float dgelu_erf(float x) {
  const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752f));
  const float pdf = 0.39894228040143268f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}
__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float dgelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
__device__ __forceinline__ float silu(float x) { return x / (1.f + __expf(-x)); }
__device__ __forceinline__ float dsilu(float x) {
  const float s = 1.f / (1.f + __expf(-x));
  return s * (1.f + x * (1.f - s));
}


__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---- fast activation variants for GEMM epilogues (outputs are rounded to 16 bits anyway) -------------
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// erf via Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7), one MUFU.EX2 + one MUFU.RCP
__device__ __forceinline__ float erf_fast(float x) {
  const float ax = fabsf(x);
  const float t = __frcp_rn(1.f + 0.3275911f * ax);
  const float poly = t * (0.254829592f + t * (-0.284496736f + t * (1.421413741f + t * (-1.453152027f + t * 1.061405429f))));
  const float y = 1.f - poly * __expf(-ax * ax);
  return copysignf(y, x);
}
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// logistic function on the two MUFU ops (ex2, rcp); __frcp_rn would expand to a correctly rounded software sequence
__device__ __forceinline__ float sigmoid_fast(float x) { return rcp_fast(1.f + ex2_fast(-1.4426950408889634f * x)); }

// ---- packed fp32 pairs (FFMA2 / FMUL2 / FADD2 on sm_100): one issue slot for two lanes of fp32 math.  The fused GEMM
// epilogues are issue-bound (ncu: 42% of all issue slots with only two epilogue warps per scheduler), so halving the
// instruction count of the activation polynomials shows up directly in the GEMM time.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 splat2(float v) { return make_float2(v, v); }

template <int ACT>
__device__ __forceinline__ float act_fast(float x) {
  if constexpr (ACT == 1) return fmaxf(x, 0.f);
  else if constexpr (ACT == 2) return 0.5f * x * (1.f + erf_fast(x * 0.70710678118654752f));
  else if constexpr (ACT == 3) return 0.5f * x * (1.f + tanh_fast(0.7978845608028654f * (x + 0.044715f * x * x * x)));
  else if constexpr (ACT == 4) return x * sigmoid_fast(x);
  else return x;
}
template <int ACT>
__device__ __forceinline__ float dact_fast(float x) {
  if constexpr (ACT == 1) return x > 0.f ? 1.f : 0.f;
  else if constexpr (ACT == 2) {
    const float cdf = 0.5f * (1.f + erf_fast(x * 0.70710678118654752f));
    return cdf + x * 0.39894228040143268f * __expf(-0.5f * x * x);
  } else if constexpr (ACT == 3) {
    const float t = tanh_fast(0.7978845608028654f * (x + 0.044715f * x * x * x));
    return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * 0.7978845608028654f * (1.f + 0.134145f * x * x);
  } else if constexpr (ACT == 4) {
    const float s = sigmoid_fast(x);
    return s * (1.f + x * (1.f - s));
  } else return 1.f;
}

// pair versions: GELU(tanh) runs on packed ops (5 packed + 2 MUFU per pair forward, 10 + 2 for the derivative);
// the other activations apply the scalar form to each half
template <int ACT>
__device__ __forceinline__ float2 act_fast2(float2 x) {
  if constexpr (ACT == 3) {
    const float2 x2 = fmul2(x, x);
    const float2 w = fmul2(x, ffma2(x2, splat2(0.7978845608028654f * 0.044715f), splat2(0.7978845608028654f)));
    const float2 t = make_float2(tanh_fast(w.x), tanh_fast(w.y));
    return fmul2(x, ffma2(t, splat2(0.5f), splat2(0.5f)));
  } else {
    return make_float2(act_fast<ACT>(x.x), act_fast<ACT>(x.y));
  }
}
template <int ACT>
__device__ __forceinline__ float2 dact_fast2(float2 x) {
  if constexpr (ACT == 3) {
    const float2 x2 = fmul2(x, x);
    const float2 w = fmul2(x, ffma2(x2, splat2(0.7978845608028654f * 0.044715f), splat2(0.7978845608028654f)));
    const float2 t = make_float2(tanh_fast(w.x), tanh_fast(w.y));
    const float2 a = ffma2(t, splat2(0.5f), splat2(0.5f));                       // 0.5 (1 + t)
    const float2 b = ffma2(fmul2(t, t), splat2(-1.f), splat2(1.f));              // 1 - t^2
    const float2 c = ffma2(x2, splat2(0.5f * 0.7978845608028654f * 0.134145f), splat2(0.5f * 0.7978845608028654f));
    return ffma2(fmul2(x, b), c, a);
  } else {
    return make_float2(dact_fast<ACT>(x.x), dact_fast<ACT>(x.y));
  }
}

}  // namespace hyp
