// bn_act: training-mode BatchNorm2d fused with ReLU and the residual add, channels-last (NHWC).
//
//   forward   stats kernel : per-channel sum(x), sum(x^2) over the N*H*W rows (fp32 atomics into [2, C])
//             apply kernel : y = relu((x - mean) * rstd * gamma + beta + residual)      one read of x (+res), one write
//   backward  reduce kernel: g = dy * (y > 0);  dbeta = sum g;  dgamma = sum g * xhat
//             dx kernel    : dx = gamma * rstd * (g - dbeta/n - xhat * dgamma/n);  dresidual = g
//
// Replaces the three ATen launches per block tail (batch_norm, add, relu) of the torchvision ResNets the reference
// trains (02_development/distributed_utils.py:229; Phase 1/baseline_performance.ipynb c0 L23) with two memory passes.
// The convolutions in front of it are conv_igemm.cu (forward and weight gradient; the data gradient defaults to cuDNN after
// the per-part A/B in DESIGN.md); in eval mode BatchNorm is folded into the convolution epilogue instead (ops/conv.py).
// gamma / beta may be 16-bit (the *_p16 entry points), and the backward then rounds dgamma / dbeta into the gradient slots.
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {
namespace bn {

// sums[0][c] += sum_r x[r,c]; sums[1][c] += sum_r x[r,c]^2          (block: 32 x 8 threads, 64 channels)
template <typename T>
__global__ void __launch_bounds__(256) stats_kernel(const T* __restrict__ x, float* __restrict__ sums, long long rows, int C,
                                                    long long rows_per_block) {
  __shared__ float red[2][8][64];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 64 + cx * 2;
  const long long r0 = blockIdx.y * rows_per_block, r1 = min(rows, r0 + rows_per_block);
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
  if (c < C) {
    for (long long r = r0 + ry; r < r1; r += 8) {
      const T* p = x + r * C + c;
      const float a = to_f32<T>(p[0]), b = to_f32<T>(p[1]);
      s0 += a; s1 += b; q0 += a * a; q1 += b * b;
    }
  }
  red[0][ry][cx * 2] = s0; red[0][ry][cx * 2 + 1] = s1;
  red[1][ry][cx * 2] = q0; red[1][ry][cx * 2 + 1] = q1;
  __syncthreads();
  if (threadIdx.x < 128) {
    const int which = threadIdx.x >> 6, k = threadIdx.x & 63;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[which][i][k];
    const int cc = blockIdx.x * 64 + k;
    if (cc < C) atomicAdd(sums + which * C + cc, s);
  }
}

// y = relu?((x - mean) * rstd * gamma + beta (+ res)); also writes mean / rstd and updates the running statistics.
template <typename T, typename P = float>
__global__ void __launch_bounds__(256)
apply_kernel(const T* __restrict__ x, const T* __restrict__ res, const float* __restrict__ sums, const P* __restrict__ gamma,
             const P* __restrict__ beta, T* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out,
             float* __restrict__ running_mean, float* __restrict__ running_var, long long rows, int C, float eps,
             float momentum, int relu) {
  const long long total = rows * C / 8;
  const float inv_n = 1.f / static_cast<float>(rows);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c0 = static_cast<int>((i * 8) % C);
    float v[8], r[8], o[8];
    unpack8<T>(*reinterpret_cast<const uint4*>(x + i * 8), v);
    if (res) unpack8<T>(*reinterpret_cast<const uint4*>(res + i * 8), r);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float m = sums[c0 + k] * inv_n;
      const float var = fmaxf(sums[C + c0 + k] * inv_n - m * m, 0.f);
      const float t = (v[k] - m) * rsqrtf(var + eps) * to_f32<P>(gamma[c0 + k]) + to_f32<P>(beta[c0 + k]) + (res ? r[k] : 0.f);
      o[k] = relu ? fmaxf(t, 0.f) : t;
    }
    *reinterpret_cast<uint4*>(y + i * 8) = pack8<T>(o);
  }
  // one thread per channel publishes mean / rstd and the running statistics (unbiased variance, like torch)
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid < C) {
    const float m = sums[gid] * inv_n;
    const float var = fmaxf(sums[C + gid] * inv_n - m * m, 0.f);
    mean_out[gid] = m;
    rstd_out[gid] = rsqrtf(var + eps);
    if (running_mean) {
      const float unbiased = rows > 1 ? var * static_cast<float>(rows) / static_cast<float>(rows - 1) : var;
      running_mean[gid] = (1.f - momentum) * running_mean[gid] + momentum * m;
      running_var[gid] = (1.f - momentum) * running_var[gid] + momentum * unbiased;
    }
  }
}

// dsums[0][c] += sum g, dsums[1][c] += sum g * xhat,  g = dy * (relu ? y > 0 : 1)
template <typename T>
__global__ void __launch_bounds__(256)
bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ y, const float* __restrict__ mean,
                  const float* __restrict__ rstd, float* __restrict__ dsums, long long rows, int C, long long rows_per_block,
                  int relu) {
  __shared__ float red[2][8][64];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 64 + cx * 2;
  const long long r0 = blockIdx.y * rows_per_block, r1 = min(rows, r0 + rows_per_block);
  float b0 = 0.f, b1 = 0.f, g0 = 0.f, g1 = 0.f;
  if (c < C) {
    const float m0 = mean[c], m1 = mean[c + 1], s0 = rstd[c], s1 = rstd[c + 1];
    for (long long r = r0 + ry; r < r1; r += 8) {
      const long long o = r * C + c;
      float d0 = to_f32<T>(dy[o]), d1 = to_f32<T>(dy[o + 1]);
      if (relu) {
        if (!(to_f32<T>(y[o]) > 0.f)) d0 = 0.f;
        if (!(to_f32<T>(y[o + 1]) > 0.f)) d1 = 0.f;
      }
      b0 += d0; b1 += d1;
      g0 += d0 * (to_f32<T>(x[o]) - m0) * s0;
      g1 += d1 * (to_f32<T>(x[o + 1]) - m1) * s1;
    }
  }
  red[0][ry][cx * 2] = b0; red[0][ry][cx * 2 + 1] = b1;
  red[1][ry][cx * 2] = g0; red[1][ry][cx * 2 + 1] = g1;
  __syncthreads();
  if (threadIdx.x < 128) {
    const int which = threadIdx.x >> 6, k = threadIdx.x & 63;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[which][i][k];
    const int cc = blockIdx.x * 64 + k;
    if (cc < C) atomicAdd(dsums + which * C + cc, s);
  }
}

template <typename T, typename P = float>
__global__ void __launch_bounds__(256)
bwd_dx_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ y, const float* __restrict__ mean,
              const float* __restrict__ rstd, const P* __restrict__ gamma, const float* __restrict__ dsums,
              T* __restrict__ dx, T* __restrict__ dres, long long rows, int C, int relu,
              P* __restrict__ dgamma_out = nullptr, P* __restrict__ dbeta_out = nullptr) {
  // the reductions are complete (previous kernel on the stream): block 0 also rounds them into the parameter-dtype
  // gradient tensors, so the caller needs no cast launches (and can hand the kernel the flat-gradient slots directly)
  if (blockIdx.x == 0 && dgamma_out != nullptr) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      dgamma_out[c] = from_f32<P>(dsums[C + c]);
      dbeta_out[c] = from_f32<P>(dsums[c]);
    }
  }
  const long long total = rows * C / 8;
  const float inv_n = 1.f / static_cast<float>(rows);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c0 = static_cast<int>((i * 8) % C);
    float d[8], xv[8], yv[8], o[8];
    unpack8<T>(*reinterpret_cast<const uint4*>(dy + i * 8), d);
    unpack8<T>(*reinterpret_cast<const uint4*>(x + i * 8), xv);
    if (relu) unpack8<T>(*reinterpret_cast<const uint4*>(y + i * 8), yv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float g = (relu && !(yv[k] > 0.f)) ? 0.f : d[k];
      d[k] = g;
      const float xh = (xv[k] - mean[c0 + k]) * rstd[c0 + k];
      o[k] = to_f32<P>(gamma[c0 + k]) * rstd[c0 + k] * (g - dsums[c0 + k] * inv_n - xh * dsums[C + c0 + k] * inv_n);
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8<T>(o);
    if (dres) *reinterpret_cast<uint4*>(dres + i * 8) = pack8<T>(d);
  }
}

static void reduce_grid(long long rows, int C, dim3* grid, long long* rpb) {
  const int gx = (C + 63) / 64;
  long long gy = (static_cast<long long>(num_sms()) * 4 + gx - 1) / gx;
  if (gy > (rows + 63) / 64) gy = (rows + 63) / 64;
  if (gy < 1) gy = 1;
  *rpb = (rows + gy - 1) / gy;
  gy = (rows + *rpb - 1) / *rpb;
  *grid = dim3(gx, static_cast<unsigned>(gy));
}
static int ew_grid(long long total_vec) {
  long long b = (total_vec + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace bn
}  // namespace hyp

using namespace hyp;

// x, res, y: NHWC-contiguous [rows = N*H*W, C], 16-bit; gamma/beta/running stats fp32 [C]; sums: fp32 [2*C] ZEROED by the
// caller (scratch for the batch statistics); mean/rstd: fp32 [C] saved for the backward.
HYP_API int hyp_bn_act_fwd(const void* x, const void* res, const float* gamma, const float* beta, void* y, float* sums,
                           float* mean, float* rstd, float* running_mean, float* running_var, long long rows, int C,
                           float eps, float momentum, int relu, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_BF16 || dtype == DT_F16, "hyp_bn_act_fwd: 16-bit dtypes only");
  HYP_REQUIRE(rows > 0 && C > 0 && C % 8 == 0, "hyp_bn_act_fwd: need C %% 8 == 0 (C=%d)", C);
  dim3 grid; long long rpb;
  bn::reduce_grid(rows, C, &grid, &rpb);
  const int eg = bn::ew_grid(rows * C / 8);
  const int eg2 = eg * 256 < C ? (C + 255) / 256 : eg;
  if (dtype == DT_BF16) {
    bn::stats_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, sums, rows, C, rpb);
    bn::apply_kernel<__nv_bfloat16><<<eg2, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res, sums, gamma, beta, (__nv_bfloat16*)y, mean, rstd, running_mean, running_var, rows, C, eps, momentum, relu);
  } else {
    bn::stats_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, sums, rows, C, rpb);
    bn::apply_kernel<__half><<<eg2, 256, 0, st>>>((const __half*)x, (const __half*)res, sums, gamma, beta, (__half*)y, mean, rstd, running_mean, running_var, rows, C, eps, momentum, relu);
  }
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dsums: fp32 [2*C] ZEROED by the caller; on return dsums[0:C] = dbeta, dsums[C:2C] = dgamma.
HYP_API int hyp_bn_act_bwd(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                           const float* gamma, float* dsums, void* dx, void* dres, long long rows, int C, int relu,
                           int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_BF16 || dtype == DT_F16, "hyp_bn_act_bwd: 16-bit dtypes only");
  HYP_REQUIRE(rows > 0 && C > 0 && C % 8 == 0, "hyp_bn_act_bwd: need C %% 8 == 0 (C=%d)", C);
  dim3 grid; long long rpb;
  bn::reduce_grid(rows, C, &grid, &rpb);
  const int eg = bn::ew_grid(rows * C / 8);
  if (dtype == DT_BF16) {
    bn::bwd_reduce_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, mean, rstd, dsums, rows, C, rpb, relu);
    bn::bwd_dx_kernel<__nv_bfloat16><<<eg, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, mean, rstd, gamma, dsums, (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, rows, C, relu);
  } else {
    bn::bwd_reduce_kernel<__half><<<grid, 256, 0, st>>>((const __half*)dy, (const __half*)x, (const __half*)y, mean, rstd, dsums, rows, C, rpb, relu);
    bn::bwd_dx_kernel<__half><<<eg, 256, 0, st>>>((const __half*)dy, (const __half*)x, (const __half*)y, mean, rstd, gamma, dsums, (__half*)dx, (__half*)dres, rows, C, relu);
  }
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}


// Same two entry points with gamma / beta in the activations' 16-bit dtype (the layout a mixed-precision module holds them
// in), so the caller does not launch two cast kernels per layer and step.  The backward also rounds dgamma / dbeta into
// 16-bit outputs (dgamma16 / dbeta16, both or neither; e.g. the parameters' ranges of a flat gradient buffer).
HYP_API int hyp_bn_act_fwd_p16(const void* x, const void* res, const void* gamma, const void* beta, void* y, float* sums,
                               float* mean, float* rstd, float* running_mean, float* running_var, long long rows, int C,
                               float eps, float momentum, int relu, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_BF16 || dtype == DT_F16, "hyp_bn_act_fwd_p16: 16-bit dtypes only");
  HYP_REQUIRE(rows > 0 && C > 0 && C % 8 == 0, "hyp_bn_act_fwd_p16: need C %% 8 == 0 (C=%d)", C);
  dim3 grid; long long rpb;
  bn::reduce_grid(rows, C, &grid, &rpb);
  const int eg = bn::ew_grid(rows * C / 8);
  const int eg2 = eg * 256 < C ? (C + 255) / 256 : eg;
  if (dtype == DT_BF16) {
    using T = __nv_bfloat16;
    bn::stats_kernel<T><<<grid, 256, 0, st>>>((const T*)x, sums, rows, C, rpb);
    bn::apply_kernel<T, T><<<eg2, 256, 0, st>>>((const T*)x, (const T*)res, sums, (const T*)gamma, (const T*)beta, (T*)y, mean, rstd, running_mean, running_var, rows, C, eps, momentum, relu);
  } else {
    using T = __half;
    bn::stats_kernel<T><<<grid, 256, 0, st>>>((const T*)x, sums, rows, C, rpb);
    bn::apply_kernel<T, T><<<eg2, 256, 0, st>>>((const T*)x, (const T*)res, sums, (const T*)gamma, (const T*)beta, (T*)y, mean, rstd, running_mean, running_var, rows, C, eps, momentum, relu);
  }
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_bn_act_bwd_p16(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                               const void* gamma, float* dsums, void* dx, void* dres, long long rows, int C, int relu,
                               int dtype, void* dgamma16, void* dbeta16, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_BF16 || dtype == DT_F16, "hyp_bn_act_bwd_p16: 16-bit dtypes only");
  HYP_REQUIRE(rows > 0 && C > 0 && C % 8 == 0, "hyp_bn_act_bwd_p16: need C %% 8 == 0 (C=%d)", C);
  dim3 grid; long long rpb;
  bn::reduce_grid(rows, C, &grid, &rpb);
  const int eg = bn::ew_grid(rows * C / 8);
  if (dtype == DT_BF16) {
    using T = __nv_bfloat16;
    bn::bwd_reduce_kernel<T><<<grid, 256, 0, st>>>((const T*)dy, (const T*)x, (const T*)y, mean, rstd, dsums, rows, C, rpb, relu);
    bn::bwd_dx_kernel<T, T><<<eg, 256, 0, st>>>((const T*)dy, (const T*)x, (const T*)y, mean, rstd, (const T*)gamma, dsums, (T*)dx, (T*)dres, rows, C, relu, (T*)dgamma16, (T*)dbeta16);
  } else {
    using T = __half;
    bn::bwd_reduce_kernel<T><<<grid, 256, 0, st>>>((const T*)dy, (const T*)x, (const T*)y, mean, rstd, dsums, rows, C, rpb, relu);
    bn::bwd_dx_kernel<T, T><<<eg, 256, 0, st>>>((const T*)dy, (const T*)x, (const T*)y, mean, rstd, (const T*)gamma, dsums, (T*)dx, (T*)dres, rows, C, relu, (T*)dgamma16, (T*)dbeta16);
  }
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
