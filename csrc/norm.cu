// fused_norm: residual-add + LayerNorm / RMSNorm, forward and backward.
//
//   z = x (+ residual);  y = (z - mean) * rstd * gamma + beta        (LayerNorm)
//   z = x (+ residual);  y = z * rsqrt(mean(z^2) + eps) * gamma      (RMSNorm)
//
// 16-bit or fp32 IO, fp32 math, one pass over the row held in registers.
// Backward produces dz in the same pass and per-block partial dgamma/dbeta in a
// caller-provided fp32 scratch, reduced by a second small kernel.
//
// Reference call sites: nn.LayerNorm inside nn.TransformerEncoderLayer
// (02_development/distributed_utils.py:79; post-norm with residual) and the
// HF LlamaRMSNorm reached through distributed_utils.py:465.
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {
namespace norm {

constexpr int MAXC = 4;  // 8-element chunks per thread -> d <= TPR * 32

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&f)[8]) {
  if constexpr (sizeof(T) == 4) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    const float4 b = *reinterpret_cast<const float4*>(p + 4);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  } else {
    unpack8<T>(*reinterpret_cast<const uint4*>(p), f);
  }
}
template <typename T>
__device__ __forceinline__ void store8(T* p, const float (&f)[8]) {
  if constexpr (sizeof(T) == 4) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
  } else {
    *reinterpret_cast<uint4*>(p) = pack8<T>(f);
  }
}

template <int TPR>
__device__ __forceinline__ float group_sum(float v, float* red) {
  if constexpr (TPR == 32) return warp_sum(v);
  else return block_sum(v, red);
}

template <typename T, int TPR>
__global__ void __launch_bounds__(TPR < 128 ? 128 : TPR)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ residual, const T* __restrict__ gamma,
                const T* __restrict__ beta, T* __restrict__ y, T* __restrict__ sum_out, float* __restrict__ mean_out,
                float* __restrict__ rstd_out, int rows, int d, float eps, int rms) {
  __shared__ float red[32];
  constexpr int BLOCK = TPR < 128 ? 128 : TPR;
  constexpr int RPB = BLOCK / TPR;
  const int row = blockIdx.x * RPB + threadIdx.x / TPR;
  const int t = threadIdx.x % TPR;
  if (row >= rows) return;  // only reachable when TPR == 32 (whole warp exits)
  const long long base = static_cast<long long>(row) * d;
  float v[MAXC][8];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < MAXC; ++k) {
    const int c = (t + k * TPR) * 8;
    if (c < d) {
      load8<T>(x + base + c, v[k]);
      if (residual) {
        float r[8];
        load8<T>(residual + base + c, r);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[k][i] += r[i];
      }
      if (sum_out) store8<T>(sum_out + base + c, v[k]);
#pragma unroll
      for (int i = 0; i < 8; ++i) s += rms ? v[k][i] * v[k][i] : v[k][i];
    }
  }
  s = group_sum<TPR>(s, red);
  float mean = 0.f, rstd;
  if (rms) {
    rstd = rsqrtf(s / d + eps);
  } else {
    mean = s / d;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < MAXC; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float u = v[k][i] - mean; q += u * u; }
      }
    }
    q = group_sum<TPR>(q, red);
    rstd = rsqrtf(q / d + eps);
  }
  if (t == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
#pragma unroll
  for (int k = 0; k < MAXC; ++k) {
    const int c = (t + k * TPR) * 8;
    if (c < d) {
      float g[8], b[8], o[8];
      load8<T>(gamma + c, g);
      if (beta) load8<T>(beta + c, b);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = (v[k][i] - mean) * rstd * g[i] + (beta ? b[i] : 0.f);
      store8<T>(y + base + c, o);
    }
  }
}

// Raw (still packed) 8-element chunk: kept in registers between the two passes over a row so that the 16-bit
// types cost 4 registers per chunk instead of 8.
template <typename T> struct Raw8 { uint4 v; };
template <> struct Raw8<float> { float4 lo, hi; };
template <typename T>
__device__ __forceinline__ Raw8<T> load_raw(const T* p) {
  Raw8<T> r;
  if constexpr (sizeof(T) == 4) {
    r.lo = *reinterpret_cast<const float4*>(p);
    r.hi = *reinterpret_cast<const float4*>(p + 4);
  } else {
    r.v = ld_nc_v4(p);
  }
  return r;
}
template <typename T>
__device__ __forceinline__ void unpack_raw(const Raw8<T>& r, float (&f)[8]) {
  if constexpr (sizeof(T) == 4) {
    f[0] = r.lo.x; f[1] = r.lo.y; f[2] = r.lo.z; f[3] = r.lo.w; f[4] = r.hi.x; f[5] = r.hi.y; f[6] = r.hi.z; f[7] = r.hi.w;
  } else {
    unpack8<T>(r.v, f);
  }
}

__host__ __device__ constexpr int bwd_block(int tpr) { return tpr == 32 ? 256 : tpr; }

// dz and per-block partial dgamma / dbeta.  partial layout: [gridDim.x][2][d] fp32.
// TPR == 32 (d <= 1024): eight warps per block, one row per warp per trip; the warps' running sums are folded
// through shared memory at the end so a block writes one partial row.
template <typename T, int TPR>
__global__ void __launch_bounds__(bwd_block(TPR), TPR == 32 ? 2 : 1)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ z, const T* __restrict__ gamma,
                const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const T* __restrict__ dres,
                T* __restrict__ dz, float* __restrict__ partial, int rows, int d, int rms) {
  __shared__ float red[32];
  constexpr int BLOCK = bwd_block(TPR);
  constexpr int RPB = BLOCK / TPR;
  __shared__ __align__(16) float fold[RPB > 1 ? 2 * TPR * 8 * MAXC : 4];
  const int sub = threadIdx.x / TPR;
  const int t = threadIdx.x % TPR;
  float dg[MAXC][8], db[MAXC][8];
#pragma unroll
  for (int k = 0; k < MAXC; ++k) {
#pragma unroll
    for (int i = 0; i < 8; ++i) { dg[k][i] = 0.f; db[k][i] = 0.f; }
  }
  const float inv_d = 1.f / d;
  for (int row = blockIdx.x * RPB + sub; row < rows; row += gridDim.x * RPB) {
    const long long base = static_cast<long long>(row) * d;
    const float mean = rms ? 0.f : mean_in[row];
    const float rstd = rstd_in[row];
    Raw8<T> ra[MAXC], rb[MAXC];
#pragma unroll
    for (int k = 0; k < MAXC; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        ra[k] = load_raw<T>(dy + base + c);
        rb[k] = load_raw<T>(z + base + c);
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < MAXC; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        float a[8], b[8], g[8];
        unpack_raw<T>(ra[k], a);
        unpack_raw<T>(rb[k], b);
        load8<T>(gamma + c, g);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xh = (b[i] - mean) * rstd;
          const float wdy = a[i] * g[i];
          dg[k][i] += a[i] * xh;
          db[k][i] += a[i];
          s1 += wdy;
          s2 += wdy * xh;
        }
      }
    }
    if (rms) s1 = 0.f; else s1 = group_sum<TPR>(s1, red) * inv_d;
    s2 = group_sum<TPR>(s2, red) * inv_d;
#pragma unroll
    for (int k = 0; k < MAXC; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        float a[8], b[8], g[8], o[8];
        unpack_raw<T>(ra[k], a);
        unpack_raw<T>(rb[k], b);
        load8<T>(gamma + c, g);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xh = (b[i] - mean) * rstd;
          o[i] = rstd * (a[i] * g[i] - s1 - xh * s2);
        }
        if (dres) {  // gradient arriving through the skip connection that forked off the norm's input
          float e[8];
          load8<T>(dres + base + c, e);
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] += e[i];
        }
        store8<T>(dz + base + c, o);
      }
    }
  }
  float* pg = partial + static_cast<long long>(blockIdx.x) * 2 * d;
  if constexpr (RPB > 1) {
    for (int w = 0; w < RPB; ++w) {
      if (sub == w) {
#pragma unroll
        for (int k = 0; k < MAXC; ++k) {
          const int c = (t + k * TPR) * 8;
          if (c < d) {
            if (w > 0) {
#pragma unroll
              for (int i = 0; i < 8; ++i) { dg[k][i] += fold[c + i]; db[k][i] += fold[d + c + i]; }
            }
            if (w == RPB - 1) {
              store8<float>(pg + c, dg[k]);
              store8<float>(pg + d + c, db[k]);
            } else {
              store8<float>(fold + c, dg[k]);
              store8<float>(fold + d + c, db[k]);
            }
          }
        }
      }
      __syncthreads();
    }
  } else {
#pragma unroll
    for (int k = 0; k < MAXC; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        store8<float>(pg + c, dg[k]);
        store8<float>(pg + d + c, db[k]);
      }
    }
  }
}

// Column reduce of the partials: out[0][c] = sum_p partial[p][0][c] etc.
// out_dtype: DT_F32 (dgamma / dbeta are float*, optional accumulate) or a 16-bit code (they point at T storage).
__global__ void __launch_bounds__(256) norm_bwd_reduce_kernel(const float* __restrict__ partial, int P, int d,
                                                              void* __restrict__ dgamma, void* __restrict__ dbeta,
                                                              int accumulate, int out_dtype) {
  __shared__ float red[2][8][32];
  const int cx = threadIdx.x & 31, py = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cx;
  float a = 0.f, b = 0.f;
  if (col < d) {
#pragma unroll 4
    for (int p = py; p < P; p += 8) {
      const float* row = partial + static_cast<long long>(p) * 2 * d;
      a += row[col];
      b += row[d + col];
    }
  }
  red[0][py][cx] = a;
  red[1][py][cx] = b;
  __syncthreads();
  if (threadIdx.x < 64) {
    const int which = threadIdx.x >> 5;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[which][k][cx];
    if (col < d) {
      void* out = which ? dbeta : dgamma;
      if (out) {
        if (out_dtype == DT_F32) {
          float* o = reinterpret_cast<float*>(out);
          o[col] = accumulate ? o[col] + s : s;
        } else if (out_dtype == DT_BF16) {
          reinterpret_cast<__nv_bfloat16*>(out)[col] = __float2bfloat16(s);
        } else {
          reinterpret_cast<__half*>(out)[col] = __float2half(s);
        }
      }
    }
  }
}

inline int pick_tpr(int d) { return d <= 1024 ? 32 : (d <= 4096 ? 128 : 256); }
inline int bwd_grid(int rows, int tpr) {
  const int rpb = bwd_block(tpr) / tpr;
  int g = (rows + rpb - 1) / rpb;
  const int cap = num_sms() * 2;
  return g > cap ? cap : g;
}

}  // namespace norm
}  // namespace hyp

using namespace hyp;

HYP_API int hyp_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                         void* sum_out, float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype,
                         void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && d > 0 && d % 8 == 0 && d <= 8192, "hyp_norm_fwd: need d %% 8 == 0 and d <= 8192 (d=%d)", d);
  const int tpr = norm::pick_tpr(d);
#define HYP_NORM_FWD(T, TPR)                                                                                         \
  norm::norm_fwd_kernel<T, TPR><<<(rows + (TPR < 128 ? 128 / TPR : 1) - 1) / (TPR < 128 ? 128 / TPR : 1),            \
                                  (TPR < 128 ? 128 : TPR), 0, st>>>((const T*)x, (const T*)residual, (const T*)gamma, \
                                                                   (const T*)beta, (T*)y, (T*)sum_out, mean, rstd,    \
                                                                   rows, d, eps, rms)
#define HYP_NORM_FWD_T(T)            \
  if (tpr == 32) HYP_NORM_FWD(T, 32); \
  else if (tpr == 128) HYP_NORM_FWD(T, 128); \
  else HYP_NORM_FWD(T, 256)
  if (dtype == DT_F32) { HYP_NORM_FWD_T(float); }
  else if (dtype == DT_BF16) { HYP_NORM_FWD_T(__nv_bfloat16); }
  else { HYP_NORM_FWD_T(__half); }
#undef HYP_NORM_FWD_T
#undef HYP_NORM_FWD
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Bytes of fp32 scratch hyp_norm_bwd needs for (rows, d).
HYP_API long long hyp_norm_bwd_scratch_bytes(int rows, int d) {
  const long long P = norm::bwd_grid(rows, norm::pick_tpr(d));  // one partial row per block
  return P * 2 * d * static_cast<long long>(sizeof(float));
}

namespace hyp { namespace norm {
int launch_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd, void* dz,
               void* dgamma, void* dbeta, float* scratch, int rows, int d, int rms, int dtype, int accumulate,
               cudaStream_t st, const void* dres = nullptr, int param_dtype = DT_F32) {
  const int tpr = pick_tpr(d);
  const int grid = bwd_grid(rows, tpr);
  const int block = bwd_block(tpr);
  const int P = grid;
#define HYP_NORM_BWD(T, TPR)                                                                                  \
  norm_bwd_kernel<T, TPR><<<grid, block, 0, st>>>((const T*)dy, (const T*)z, (const T*)gamma, mean, rstd,     \
                                                  (const T*)dres, (T*)dz, scratch, rows, d, rms)
#define HYP_NORM_BWD_T(T)             \
  if (tpr == 32) HYP_NORM_BWD(T, 32); \
  else if (tpr == 128) HYP_NORM_BWD(T, 128); \
  else HYP_NORM_BWD(T, 256)
  if (dtype == DT_F32) { HYP_NORM_BWD_T(float); }
  else if (dtype == DT_BF16) { HYP_NORM_BWD_T(__nv_bfloat16); }
  else { HYP_NORM_BWD_T(__half); }
#undef HYP_NORM_BWD_T
#undef HYP_NORM_BWD
  HYP_CHECK_CUDA(cudaGetLastError());
  norm_bwd_reduce_kernel<<<(d + 31) / 32, 256, 0, st>>>(scratch, P, d, dgamma, dbeta, accumulate, param_dtype);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
}}  // namespace hyp::norm

HYP_API int hyp_norm_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                         void* dz, float* dgamma, float* dbeta, void* scratch, long long scratch_bytes, int rows, int d,
                         int rms, int dtype, int accumulate, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && d > 0 && d % 8 == 0 && d <= 8192, "hyp_norm_bwd: need d %% 8 == 0 and d <= 8192 (d=%d)", d);
  HYP_REQUIRE(scratch != nullptr && scratch_bytes >= hyp_norm_bwd_scratch_bytes(rows, d),
              "hyp_norm_bwd: scratch too small (%lld < %lld bytes)", scratch_bytes, hyp_norm_bwd_scratch_bytes(rows, d));
  return norm::launch_bwd(dy, z, gamma, mean, rstd, dz, dgamma, dbeta, reinterpret_cast<float*>(scratch), rows, d, rms,
                          dtype, accumulate, st);
}

// hyp_norm_bwd with two fusions for pre-norm residual blocks: `dres` (same shape/dtype as dz, may be null) is added to
// dz -- the gradient that reaches the norm's input through the skip connection, which autograd would otherwise add
// with a separate elementwise pass -- and dgamma / dbeta are written in `param_dtype` (DT_F32, or the 16-bit dtype of
// the parameters so they can land directly in the flat gradient buffer).
HYP_API int hyp_norm_bwd_fused(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                               const void* dres, void* dz, void* dgamma, void* dbeta, int param_dtype, void* scratch,
                               long long scratch_bytes, int rows, int d, int rms, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && d > 0 && d % 8 == 0 && d <= 8192, "hyp_norm_bwd_fused: need d %% 8 == 0 and d <= 8192 (d=%d)", d);
  HYP_REQUIRE(param_dtype == DT_F32 || param_dtype == DT_F16 || param_dtype == DT_BF16, "hyp_norm_bwd_fused: bad param dtype");
  HYP_REQUIRE(scratch != nullptr && scratch_bytes >= hyp_norm_bwd_scratch_bytes(rows, d),
              "hyp_norm_bwd_fused: scratch too small (%lld < %lld bytes)", scratch_bytes, hyp_norm_bwd_scratch_bytes(rows, d));
  return norm::launch_bwd(dy, z, gamma, mean, rstd, dz, dgamma, dbeta, reinterpret_cast<float*>(scratch), rows, d, rms,
                          dtype, 0, st, dres, param_dtype);
}

// ---------------------------------------------------------------------------------------------------------
// hyp_norm_bwd_noalloc: the same backward as hyp_norm_bwd for callers that have no way to obtain scratch.
//
// Who needs this: code that calls the library from a place where device memory cannot be allocated -- a
// plugin running inside another runtime's execution callback (an inference-server backend, a custom-call
// hook of a graph executor), a real-time loop with a fixed memory budget, or a thin C wrapper that owns no
// allocator at all.  Such callers can pass their inputs and outputs, but they cannot call cudaMalloc (the
// host runtime forbids it, or it would synchronise the device at a bad moment) and they have nobody to ask
// for a temporary buffer.
//
// The partial dgamma/dbeta sums therefore live in a fixed array compiled into the library.  Every GPU that
// loads the library holds HYP_NORM_BWD_NOALLOC_SCRATCH_BYTES (16 MiB) of device memory for it until the
// process exits.  Shapes whose partial sums do not fit (hyp_norm_bwd_scratch_bytes(rows, d) > 16 MiB, which
// only happens for d > 4096 with many rows) are refused with an error; use hyp_norm_bwd for those.
//
// Set-up: the first call on each device looks up the device address of the array (cudaGetSymbolAddress,
// with that device current) and caches it; make that first call outside any stream capture.  After that a
// call only launches the two kernels.  No call allocates or frees anything.
//
// Sharing: all calls on a device use the same array, and it keeps whatever the previous call left in it.
// Nothing has to be cleared -- every word the reduce kernel reads is written by the backward kernel of the
// same call first -- but two calls that overlap on the GPU (different streams, no dependency) would write
// the same words and corrupt each other's dgamma/dbeta: serialise them.
namespace hyp { namespace norm {
constexpr long long kNoAllocScratchBytes = 16ll << 20;
static __device__ __align__(256) unsigned char g_bwd_scratch[kNoAllocScratchBytes];
// device address of g_bwd_scratch per device ordinal; the symbol itself is only a host-side stand-in
static void* g_bwd_scratch_addr[64] = {nullptr};

static int resolve_scratch(float** out) {
  int dev = 0;
  HYP_CHECK_CUDA(cudaGetDevice(&dev));
  HYP_REQUIRE(dev >= 0 && dev < 64, "hyp_norm_bwd_noalloc: device ordinal %d out of range", dev);
  if (g_bwd_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    HYP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bwd_scratch));
    g_bwd_scratch_addr[dev] = p;
  }
  *out = reinterpret_cast<float*>(g_bwd_scratch_addr[dev]);
  return 0;
}
}}  // namespace hyp::norm

HYP_API long long hyp_norm_bwd_noalloc_max_scratch_bytes() { return norm::kNoAllocScratchBytes; }

HYP_API int hyp_norm_bwd_noalloc(const void* dy, const void* z, const void* gamma, const float* mean,
                                 const float* rstd, void* dz, float* dgamma, float* dbeta, int rows, int d, int rms,
                                 int dtype, int accumulate, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && d > 0 && d % 8 == 0 && d <= 8192,
              "hyp_norm_bwd_noalloc: need d %% 8 == 0 and d <= 8192 (d=%d)", d);
  const long long need = hyp_norm_bwd_scratch_bytes(rows, d);
  HYP_REQUIRE(need <= norm::kNoAllocScratchBytes,
              "hyp_norm_bwd_noalloc: rows=%d d=%d needs %lld bytes of partial sums but the built-in scratch holds %lld; "
              "call hyp_norm_bwd with your own scratch for this shape", rows, d, need, norm::kNoAllocScratchBytes);
  float* scratch = nullptr;
  const int rc = norm::resolve_scratch(&scratch);
  if (rc) return rc;
  return norm::launch_bwd(dy, z, gamma, mean, rstd, dz, dgamma, dbeta, scratch, rows, d, rms, dtype, accumulate, st);
}
