// Streaming elementwise kernels: the HBM bandwidth probe (z = x + y, copy),
// column sums for bias gradients, SwiGLU, dtype casts, stand-alone Philox dropout.  All are 128-bit
// vectorised grid-stride loops sized to a multiple of the SM count.
//
// Reference call sites: `z = x + y` in the bandwidth sweep
// (Phase 1/01_hardware_exploration.ipynb c1 L148), autocast casts
// (02_development/distributed_utils.py:175).
#include "common/philox.cuh"
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {

// One block streams a contiguous run of 4 x 512 sixteen-byte vectors: every thread issues its four loads of both operands
// (8 x 16 B in flight) before the first add, and consecutive blocks touch consecutive DRAM pages.  A persistent
// grid-stride version with two loads in flight per operand reached 6.36 TB/s on 500 M fp32 elements where
// torch's elementwise kernel reached 7.0; this shape closes that gap.
template <typename T>
__global__ void __launch_bounds__(512) stream_add_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                         T* __restrict__ z, long long n) {
  constexpr int V = 16 / sizeof(T);
  constexpr int U = 4;
  const long long nv = n / V;
  const uint4* x4 = reinterpret_cast<const uint4*>(x);
  const uint4* y4 = reinterpret_cast<const uint4*>(y);
  uint4* z4 = reinterpret_cast<uint4*>(z);
  const long long base = static_cast<long long>(blockIdx.x) * (U * 512) + threadIdx.x;
  uint4 a[U], b[U];
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const long long i = base + k * 512;
    if (i < nv) { a[k] = ld_nc_v4(x4 + i); b[k] = ld_nc_v4(y4 + i); }
  }
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const long long i = base + k * 512;
    if (i < nv) {
      uint4 r;
      if constexpr (sizeof(T) == 4) {
        const float* fa = reinterpret_cast<const float*>(&a[k]);
        const float* fb = reinterpret_cast<const float*>(&b[k]);
        float* fr = reinterpret_cast<float*>(&r);
#pragma unroll
        for (int e = 0; e < 4; ++e) fr[e] = fa[e] + fb[e];
      } else {
        float fa[8], fb[8];
        unpack8<T>(a[k], fa); unpack8<T>(b[k], fb);
#pragma unroll
        for (int e = 0; e < 8; ++e) fa[e] += fb[e];
        r = pack8<T>(fa);
      }
      z4[i] = r;
    }
  }
  // scalar tail (fewer than V elements): first block
  if (blockIdx.x == 0) {
    const long long t = nv * V + threadIdx.x;
    if (t < n) z[t] = from_f32<T>(to_f32<T>(x[t]) + to_f32<T>(y[t]));
  }
}

__global__ void __launch_bounds__(512) stream_copy_kernel(const uint4* __restrict__ x, uint4* __restrict__ z,
                                                          long long nv) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < nv; i += 4 * stride) {
    const uint4 a0 = ld_nc_v4(x + i), a1 = ld_nc_v4(x + i + stride), a2 = ld_nc_v4(x + i + 2 * stride),
                a3 = ld_nc_v4(x + i + 3 * stride);
    z[i] = a0; z[i + stride] = a1; z[i + 2 * stride] = a2; z[i + 3 * stride] = a3;
  }
  for (; i < nv; i += stride) z[i] = ld_nc_v4(x + i);
}

// out[c] (+)= sum_r x[r, c]  (fp32 atomics; out must be zeroed by the caller unless accumulating)
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, float* __restrict__ out, int M, int N,
                                                     long long ld, int rows_per_block) {
  __shared__ float red[8][64];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 64 + cx * 2;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(M, r0 + rows_per_block);
  float s0 = 0.f, s1 = 0.f;
  if (col < N) {
    const bool pair = (col + 1 < N);
    for (int r = r0 + ry; r < r1; r += 8) {
      const T* p = x + static_cast<long long>(r) * ld + col;
      s0 += to_f32<T>(p[0]);
      if (pair) s1 += to_f32<T>(p[1]);
    }
  }
  red[ry][cx * 2] = s0;
  red[ry][cx * 2 + 1] = s1;
  __syncthreads();
  if (threadIdx.x < 64) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][threadIdx.x];
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c < N) atomicAdd(out + c, s);
  }
}

// Same sum for 16-bit inputs whose rows are 16-byte aligned: each thread owns 8 adjacent columns (one 16-byte load
// per row), a block covers 256 columns x rows_per_block rows, four rows in flight per thread.
template <typename T>
__global__ void __launch_bounds__(256) colsum8_kernel(const T* __restrict__ x, float* __restrict__ out, int M, int N,
                                                      long long ld, int rows_per_block) {
  __shared__ float red[8][256 + 8];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + cx * 8;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(M, r0 + rows_per_block);
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = 0.f;
  if (col < N) {
    const T* base = x + col;
    int r = r0 + ry;
    for (; r + 24 < r1; r += 32) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ld_nc_v4(reinterpret_cast<const uint4*>(base + static_cast<long long>(r + 8 * u) * ld));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8<T>(v[u], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] += f[i];
      }
    }
    for (; r < r1; r += 8) {
      float f[8];
      unpack8<T>(ld_nc_v4(reinterpret_cast<const uint4*>(base + static_cast<long long>(r) * ld)), f);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] += f[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[ry][cx * 8 + i] = s[i];
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < N) atomicAdd(out + c, t);
}

// Column sums rounded to 16 bits in ONE launch (bias gradients): blocks add their partial sums into `acc` (fp32, all
// zero on entry), count themselves in `counter`, and the last block to finish rounds acc into out16 and puts acc and
// counter back to zero -- so the caller keeps one persistent workspace and never has to clear or convert anything.
template <typename T>
__global__ void __launch_bounds__(256) colsum8_cast_kernel(const T* __restrict__ x, T* __restrict__ out16,
                                                           float* __restrict__ acc, unsigned* __restrict__ counter, int M,
                                                           int N, long long ld, int rows_per_block) {
  __shared__ float red[8][256 + 8];
  __shared__ bool is_last;
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + cx * 8;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(M, r0 + rows_per_block);
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = 0.f;
  if (col < N) {
    const T* base = x + col;
    int r = r0 + ry;
    for (; r + 24 < r1; r += 32) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ld_nc_v4(reinterpret_cast<const uint4*>(base + static_cast<long long>(r + 8 * u) * ld));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8<T>(v[u], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] += f[i];
      }
    }
    for (; r < r1; r += 8) {
      float f[8];
      unpack8<T>(ld_nc_v4(reinterpret_cast<const uint4*>(base + static_cast<long long>(r) * ld)), f);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] += f[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[ry][cx * 8 + i] = s[i];
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < N) atomicAdd(acc + c, t);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x * gridDim.y - 1;
  __syncthreads();
  if (is_last) {
    __threadfence();
    for (int i = threadIdx.x; i < N; i += 256) {
      out16[i] = from_f32<T>(__ldcg(acc + i));
      acc[i] = 0.f;
    }
    if (threadIdx.x == 0) *counter = 0u;
  }
}

// SwiGLU: out = silu(gate) * up, and its backward.
template <typename T>
__global__ void swiglu_fwd_kernel(const T* __restrict__ gate, const T* __restrict__ up, T* __restrict__ out,
                                  long long n) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 8;
  for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
    float g[8], u[8];
    unpack8<T>(*reinterpret_cast<const uint4*>(gate + i), g);
    unpack8<T>(*reinterpret_cast<const uint4*>(up + i), u);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = silu(g[k]) * u[k];
    *reinterpret_cast<uint4*>(out + i) = pack8<T>(g);
  }
}
template <typename T>
__global__ void swiglu_bwd_kernel(const T* __restrict__ gate, const T* __restrict__ up, const T* __restrict__ dout,
                                  T* __restrict__ dgate, T* __restrict__ dup, long long n) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 8;
  for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
    float g[8], u[8], d[8], dg[8], du[8];
    unpack8<T>(*reinterpret_cast<const uint4*>(gate + i), g);
    unpack8<T>(*reinterpret_cast<const uint4*>(up + i), u);
    unpack8<T>(*reinterpret_cast<const uint4*>(dout + i), d);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      dg[k] = d[k] * u[k] * dsilu(g[k]);
      du[k] = d[k] * silu(g[k]);
    }
    *reinterpret_cast<uint4*>(dgate + i) = pack8<T>(dg);
    *reinterpret_cast<uint4*>(dup + i) = pack8<T>(du);
  }
}

template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ x, TO* __restrict__ y, long long n, float scale) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    y[i] = from_f32<TO>(to_f32<TI>(x[i]) * scale);
}

// fp32 -> 16-bit, 8 elements per thread (two 16-byte loads, one 16-byte store); scale = host value * optional device value
template <typename TO>
__global__ void __launch_bounds__(256) cast_f32_vec_kernel(const float* __restrict__ x, TO* __restrict__ y, long long n8,
                                                           float scale, const float* __restrict__ scale_ptr) {
  const float sc = scale_ptr ? scale * *scale_ptr : scale;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const float4 a = *reinterpret_cast<const float4*>(x + i * 8);
    const float4 b = *reinterpret_cast<const float4*>(x + i * 8 + 4);
    float f[8] = {a.x * sc, a.y * sc, a.z * sc, a.w * sc, b.x * sc, b.y * sc, b.z * sc, b.w * sc};
    *reinterpret_cast<uint4*>(y + i * 8) = pack8<TO>(f);
  }
}
// 16-bit -> same 16-bit type with a scale, 8 elements per thread
template <typename T>
__global__ void __launch_bounds__(256) scale16_vec_kernel(const T* __restrict__ x, T* __restrict__ y, long long n8, float scale,
                                                          const float* __restrict__ scale_ptr) {
  const float sc = scale_ptr ? scale * *scale_ptr : scale;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float f[8];
    unpack8<T>(*reinterpret_cast<const uint4*>(x + i * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] *= sc;
    *reinterpret_cast<uint4*>(y + i * 8) = pack8<T>(f);
  }
}
template <typename TI, typename TO>
__global__ void cast_dscale_kernel(const TI* __restrict__ x, TO* __restrict__ y, long long n, float scale,
                                   const float* __restrict__ scale_ptr) {
  const float sc = scale_ptr ? scale * *scale_ptr : scale;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    y[i] = from_f32<TO>(to_f32<TI>(x[i]) * sc);
}

static inline int stream_grid(long long work_items, int threads) {
  long long b = (work_items + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace hyp

using namespace hyp;

HYP_API int hyp_stream_add(const void* x, const void* y, void* z, long long n, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(n >= 0, "hyp_stream_add: negative size");
  if (n == 0) return 0;
  HYP_REQUIRE(((uintptr_t)x | (uintptr_t)y | (uintptr_t)z) % 16 == 0, "hyp_stream_add: pointers must be 16B aligned");
  const int threads = 512;
  auto add_grid = [](long long nv) { return static_cast<unsigned>((nv + 2047) / 2048 > 0 ? (nv + 2047) / 2048 : 1); };
  if (dtype == DT_F32) {
    stream_add_kernel<float><<<add_grid(n / 4), threads, 0, st>>>((const float*)x, (const float*)y, (float*)z, n);
  } else if (dtype == DT_BF16) {
    stream_add_kernel<__nv_bfloat16><<<add_grid(n / 8), threads, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (__nv_bfloat16*)z, n);
  } else {
    stream_add_kernel<__half><<<add_grid(n / 8), threads, 0, st>>>((const __half*)x, (const __half*)y, (__half*)z, n);
  }
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_stream_copy(const void* x, void* z, long long nbytes, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(nbytes % 16 == 0 && ((uintptr_t)x | (uintptr_t)z) % 16 == 0, "hyp_stream_copy: needs 16B alignment/size");
  if (nbytes == 0) return 0;
  const long long nv = nbytes / 16;
  stream_copy_kernel<<<stream_grid(nv / 4 + 1, 512), 512, 0, st>>>((const uint4*)x, (uint4*)z, nv);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_colsum(const void* x, float* out, int M, int N, long long ld, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(M > 0 && N > 0, "hyp_colsum: empty input");
  // 16-byte loads need aligned rows; a ragged N is fine as long as the last 8-column group stays inside the row pitch
  // (the padded logits of a 50257-word vocabulary: N = 50257, ld = 50304) -- columns >= N are summed but never stored
  const bool vec = dtype != DT_F32 && ld % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 &&
                   (static_cast<long long>(N) + 7) / 8 * 8 <= ld;
  const int cpb = vec ? 256 : 64;  // columns per block
  const int gx = (N + cpb - 1) / cpb;
  int gy = (num_sms() * 4 + gx - 1) / gx;
  if (gy > (M + 31) / 32) gy = (M + 31) / 32;
  if (gy < 1) gy = 1;
  const int rpb = (M + gy - 1) / gy;
  gy = (M + rpb - 1) / rpb;
  dim3 grid(gx, gy);
  if (dtype == DT_F32) colsum_kernel<float><<<grid, 256, 0, st>>>((const float*)x, out, M, N, ld, rpb);
  else if (dtype == DT_BF16) {
    if (vec) colsum8_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, out, M, N, ld, rpb);
    else colsum_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, out, M, N, ld, rpb);
  } else {
    if (vec) colsum8_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, out, M, N, ld, rpb);
    else colsum_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, out, M, N, ld, rpb);
  }
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// out16[c] = round(sum_r x[r, c]) for 16-bit x; acc_ws: >= N fp32 zeros, counter: one zero word -- both are returned
// zeroed, so the same workspace serves every call on a stream.  Needs N % 8 == 0 and 16-byte aligned rows.
HYP_API int hyp_colsum_cast(const void* x, void* out16, int M, int N, long long ld, int dtype, float* acc_ws,
                            unsigned* counter, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(M > 0 && N > 0, "hyp_colsum_cast: empty input");
  HYP_REQUIRE(dtype == DT_F16 || dtype == DT_BF16, "hyp_colsum_cast: 16-bit inputs only");
  HYP_REQUIRE(N % 8 == 0 && ld % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0,
              "hyp_colsum_cast: needs N %% 8 == 0 and 16-byte aligned rows (N=%d ld=%lld)", N, ld);
  const int gx = (N + 255) / 256;
  int gy = (num_sms() * 4 + gx - 1) / gx;
  if (gy > (M + 31) / 32) gy = (M + 31) / 32;
  if (gy < 1) gy = 1;
  const int rpb = (M + gy - 1) / gy;
  gy = (M + rpb - 1) / rpb;
  dim3 grid(gx, gy);
  if (dtype == DT_BF16)
    colsum8_cast_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out16, acc_ws, counter, M, N, ld, rpb);
  else
    colsum8_cast_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, (__half*)out16, acc_ws, counter, M, N, ld, rpb);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_swiglu_fwd(const void* gate, const void* up, void* out, long long n, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(n % 8 == 0 && dtype != DT_F32, "hyp_swiglu_fwd: n %% 8 == 0 and a 16-bit dtype are required");
  const int g = stream_grid(n / 8, 256);
  if (dtype == DT_BF16) swiglu_fwd_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (__nv_bfloat16*)out, n);
  else swiglu_fwd_kernel<__half><<<g, 256, 0, st>>>((const __half*)gate, (const __half*)up, (__half*)out, n);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_swiglu_bwd(const void* gate, const void* up, const void* dout, void* dgate, void* dup, long long n,
                           int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(n % 8 == 0 && dtype != DT_F32, "hyp_swiglu_bwd: n %% 8 == 0 and a 16-bit dtype are required");
  const int g = stream_grid(n / 8, 256);
  if (dtype == DT_BF16)
    swiglu_bwd_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (const __nv_bfloat16*)dout,
                                                        (__nv_bfloat16*)dgate, (__nv_bfloat16*)dup, n);
  else
    swiglu_bwd_kernel<__half><<<g, 256, 0, st>>>((const __half*)gate, (const __half*)up, (const __half*)dout, (__half*)dgate, (__half*)dup, n);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int cast_impl(const void* x, void* y, long long n, int in_dtype, int out_dtype, float scale, const float* scale_ptr,
                     cudaStream_t st) {
  if (n == 0) return 0;
  const bool vec = in_dtype == DT_F32 && out_dtype != DT_F32 && n % 8 == 0 &&
                   reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0;
  if (vec) {
    const int gv = stream_grid(n / 8, 256);
    if (out_dtype == DT_BF16) cast_f32_vec_kernel<__nv_bfloat16><<<gv, 256, 0, st>>>((const float*)x, (__nv_bfloat16*)y, n / 8, scale, scale_ptr);
    else cast_f32_vec_kernel<__half><<<gv, 256, 0, st>>>((const float*)x, (__half*)y, n / 8, scale, scale_ptr);
    HYP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  if (in_dtype == out_dtype && in_dtype != DT_F32 && n % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 &&
      reinterpret_cast<uintptr_t>(y) % 16 == 0) {
    const int gv = stream_grid(n / 8, 256);
    if (in_dtype == DT_BF16) scale16_vec_kernel<__nv_bfloat16><<<gv, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n / 8, scale, scale_ptr);
    else scale16_vec_kernel<__half><<<gv, 256, 0, st>>>((const __half*)x, (__half*)y, n / 8, scale, scale_ptr);
    HYP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  const int g = stream_grid(n, 256);
#define HYP_CAST_CASE(TI, TO) cast_dscale_kernel<TI, TO><<<g, 256, 0, st>>>((const TI*)x, (TO*)y, n, scale, scale_ptr)
  if (in_dtype == DT_F32 && out_dtype == DT_BF16) HYP_CAST_CASE(float, __nv_bfloat16);
  else if (in_dtype == DT_F32 && out_dtype == DT_F16) HYP_CAST_CASE(float, __half);
  else if (in_dtype == DT_BF16 && out_dtype == DT_F32) HYP_CAST_CASE(__nv_bfloat16, float);
  else if (in_dtype == DT_F16 && out_dtype == DT_F32) HYP_CAST_CASE(__half, float);
  else if (in_dtype == DT_F32 && out_dtype == DT_F32) HYP_CAST_CASE(float, float);
  else if (in_dtype == DT_BF16 && out_dtype == DT_BF16) HYP_CAST_CASE(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == DT_F16 && out_dtype == DT_F16) HYP_CAST_CASE(__half, __half);
  else { set_error("hyp_cast: unsupported dtype pair %d -> %d", in_dtype, out_dtype); return -1; }
#undef HYP_CAST_CASE
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_cast(const void* x, void* y, long long n, int in_dtype, int out_dtype, float scale, void* stream_) {
  return cast_impl(x, y, n, in_dtype, out_dtype, scale, nullptr, reinterpret_cast<cudaStream_t>(stream_));
}

// y = cast(x * scale * *scale_ptr): the extra factor is read on the device (e.g. the incoming gradient of a scalar loss),
// so the caller never has to bring it to the host.
HYP_API int hyp_cast_dscale(const void* x, void* y, long long n, int in_dtype, int out_dtype, float scale,
                            const float* scale_ptr, void* stream_) {
  return cast_impl(x, y, n, in_dtype, out_dtype, scale, scale_ptr, reinterpret_cast<cudaStream_t>(stream_));
}


// ------------------------------------------------------------------------------------------------ dropout
// y[row, col] = x[row, col] * keep(row, col) * scale, mask from common/philox.cuh (2-D plane).  The same call with the
// same (rng_state, salt, p) applied to a gradient is the backward.  Replaces F.dropout on the FFN hidden activation
// (nn.TransformerEncoderLayer, 02_development/distributed_utils.py:79) and LoRA's lora_dropout (:469-473), and undoes
// the dropout folded into hyp_norm_fwd_drop for that op's gradient.  In place (y == x) is fine.
namespace hyp { namespace ew {
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long rows, int cols,
                                                       DropoutArgs drop) {
  const int groups = (cols + 15) >> 4;                    // 16-column groups per row
  const long long total = rows * groups;
  const uint2 key = dropout_key(drop.state);
  for (long long idx = blockIdx.x * 256ll + threadIdx.x; idx < total; idx += 256ll * gridDim.x) {
    const long long row = idx / groups;
    const int g = static_cast<int>(idx - row * groups);
    const uint4 by = dropout_bytes16(key, drop.salt, static_cast<uint32_t>(g), static_cast<uint32_t>(row), kDropoutPlane2D);
    const uint32_t w[4] = {by.x, by.y, by.z, by.w};
    const int c0 = g * 16;
    const T* xr = x + row * cols + c0;
    T* yr = y + row * cols + c0;
    if (c0 + 16 <= cols && sizeof(T) == 2 && (cols & 7) == 0) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float f[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(xr + h * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t byte = (w[h * 2 + (i >> 2)] >> ((i & 3) * 8)) & 0xFFu;
          f[i] = byte >= drop.thresh ? f[i] * drop.scale : 0.f;
        }
        *reinterpret_cast<uint4*>(yr + h * 8) = pack8<T>(f);
      }
    } else {
      for (int i = 0; i < 16 && c0 + i < cols; ++i) {
        const uint32_t byte = (w[i >> 2] >> ((i & 3) * 8)) & 0xFFu;
        yr[i] = byte >= drop.thresh ? from_f32<T>(to_f32<T>(xr[i]) * drop.scale) : from_f32<T>(0.f);
      }
    }
  }
}
}}  // namespace hyp::ew

HYP_API int hyp_dropout(const void* x, void* y, long long rows, int cols, int dtype, const unsigned long long* rng_state,
                        int salt, float dropout_p, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && cols > 0, "hyp_dropout: empty tensor");
  HYP_REQUIRE(rng_state != nullptr && dropout_p > 0.f && dropout_p < 1.f, "hyp_dropout: need an rng state and 0 < p < 1 (p=%f)", dropout_p);
  int t = static_cast<int>(dropout_p * 256.f + 0.5f);
  t = t < 1 ? 1 : (t > 255 ? 255 : t);
  DropoutArgs drop{rng_state, static_cast<unsigned>(salt), static_cast<unsigned>(t), 256.f / static_cast<float>(256 - t)};
  const long long total = rows * ((cols + 15) >> 4);
  long long blocks = (total + 255) / 256;
  const long long cap = 148ll * 16;
  const int grid = static_cast<int>(blocks < cap ? blocks : cap);
  if (dtype == DT_F32) ew::dropout_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (float*)y, rows, cols, drop);
  else if (dtype == DT_BF16) ew::dropout_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, rows, cols, drop);
  else ew::dropout_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, (__half*)y, rows, cols, drop);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
