// fused_norm: residual-add + LayerNorm / RMSNorm, forward and backward.
//
//   z = x (+ residual);  y = (z - mean) * rstd * gamma + beta        (LayerNorm)
//   z = x (+ residual);  y = z * rsqrt(mean(z^2) + eps) * gamma      (RMSNorm)
//
// 16-bit or fp32 IO, fp32 math, one pass over the row held in registers.
// Forward can drop elements of x before the residual add (hyp_norm_fwd_drop: the
// Philox mask of common/philox.cuh, regenerated in backward -- the residual-branch
// dropout of nn.TransformerEncoderLayer costs no extra pass).
// Backward produces dz in the same pass and per-block partial dgamma/dbeta in a
// caller-provided fp32 scratch, reduced by a second small kernel (hyp_norm_bwd), or
// by the last block to finish (hyp_norm_bwd_fused: one launch, also adds the
// skip-connection gradient); hyp_norm_bwd_noalloc is the variant that allocates
// nothing (scratch in a __device__ array of this library).
//
// Reference call sites: nn.LayerNorm inside nn.TransformerEncoderLayer
// (02_development/distributed_utils.py:79; post-norm with residual) and the
// HF LlamaRMSNorm reached through distributed_utils.py:465.
#include <atomic>
#include "common/philox.cuh"
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {
namespace norm {

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

// Work decomposition.  A row is shared by TPR threads (a "group"), each owning CH chunks of 8 adjacent elements
// (chunk k of thread t covers columns (t + k*TPR)*8 ...), and a block holds RPB = BLOCK / TPR groups that walk the
// rows with a grid stride.  The point of the narrow per-thread slice is occupancy: at d = 1024 one chunk per thread
// keeps the backward (which carries dgamma / dbeta accumulators for every owned column) under 64 registers, so 32
// warps per SM keep enough loads in flight to saturate HBM -- the first version gave a warp a whole row (4 chunks,
// 128 registers, 16 warps per SM) and ran at a third of the memory bandwidth.
//
// Sums over a row: warp shuffles, then (TPR > 32) the group's warps meet at a named barrier (id 1 + group index) and
// combine through shared memory; the slots are double-buffered by call parity so one barrier per sum suffices.
constexpr int MAX_RPB = 8;

// one-kernel finish of the backward (see norm_bwd_kernel): fp32 accumulator [2][d] + arrival counter, both zero on
// entry and zero again on exit; acc == nullptr selects the two-kernel mode
struct FinishArgs {
  float* acc;
  unsigned* counter;
  void* dgamma;
  void* dbeta;
  void* dzsum;   // optional third output (one-kernel mode, one chunk per thread): column sums of dz, i.e. the bias
                 // gradient of the linear layer whose output fed this norm's input -- saves that layer a column-sum pass
  int param_dtype;
};
template <int TPR>
__device__ __forceinline__ float2 group_sum2(float2 v, float2* red, int sub, int t, unsigned& parity) {
  v.x = warp_sum(v.x);
  v.y = warp_sum(v.y);
  if constexpr (TPR == 32) {
    return v;
  } else {
    constexpr int W = TPR / 32;
    float2* slot = red + ((parity & 1u) * MAX_RPB + sub) * 8;
    parity ^= 1u;
    if ((t & 31) == 0) slot[t >> 5] = v;
    named_bar_sync(1 + sub, TPR);
    float2 r = make_float2(0.f, 0.f);
#pragma unroll
    for (int k = 0; k < W; ++k) { r.x += slot[k].x; r.y += slot[k].y; }
    return r;
  }
}

// DROP: z = dropout(x) + residual -- the dropout of a residual branch (nn.TransformerEncoderLayer's dropout1 / dropout2)
// folded into the add + norm pass; mask addressed by (row, col), see common/philox.cuh.
template <typename T, int TPR, int CH, int BLOCK, bool DROP = false>
__global__ void __launch_bounds__(BLOCK)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ residual, const T* __restrict__ gamma,
                const T* __restrict__ beta, T* __restrict__ y, T* __restrict__ sum_out, float* __restrict__ mean_out,
                float* __restrict__ rstd_out, int rows, int d, float eps, int rms, DropoutArgs drop = DropoutArgs{}) {
  __shared__ float2 red[2 * MAX_RPB * 8];
  constexpr int RPB = BLOCK / TPR;
  static_assert(RPB <= MAX_RPB && TPR / 32 <= 8, "bad norm config");
  const int sub = threadIdx.x / TPR;
  const int t = threadIdx.x % TPR;
  unsigned parity = 0;
  const float inv_d = 1.f / d;
  // gamma / beta are requested first, so their latency overlaps the row loads and the two reductions
  Raw8<T> rg[CH], rbt[CH];
#pragma unroll
  for (int k = 0; k < CH; ++k) {
    const int c = (t + k * TPR) * 8;
    if (c < d) {
      rg[k] = load_raw<T>(gamma + c);
      if (beta) rbt[k] = load_raw<T>(beta + c);
    }
  }
  for (int row = blockIdx.x * RPB + sub; row < rows; row += gridDim.x * RPB) {
    const long long base = static_cast<long long>(row) * d;
    float v[CH][8];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        load8<T>(x + base + c, v[k]);
        if constexpr (DROP) {
          const uint4 by = dropout_bytes16(dropout_key(drop.state), drop.salt, static_cast<uint32_t>(c) >> 4,
                                           static_cast<uint32_t>(row), kDropoutPlane2D);
          const uint32_t w0 = (c & 8) ? by.z : by.x, w1 = (c & 8) ? by.w : by.y;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t byte = ((i < 4 ? w0 : w1) >> ((i & 3) * 8)) & 0xFFu;
            v[k][i] = byte >= drop.thresh ? v[k][i] * drop.scale : 0.f;
          }
        }
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
    s = group_sum2<TPR>(make_float2(s, 0.f), red, sub, t, parity).x;
    float mean = 0.f, rstd;
    if (rms) {
      rstd = rsqrtf(s * inv_d + eps);
    } else {
      mean = s * inv_d;
      float q = 0.f;
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        const int c = (t + k * TPR) * 8;
        if (c < d) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { const float u = v[k][i] - mean; q += u * u; }
        }
      }
      q = group_sum2<TPR>(make_float2(q, 0.f), red, sub, t, parity).x;
      rstd = rsqrtf(q * inv_d + eps);
    }
    if (t == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        float g[8], b[8], o[8];
        unpack_raw<T>(rg[k], g);
        if (beta) unpack_raw<T>(rbt[k], b);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = (v[k][i] - mean) * rstd * g[i] + (beta ? b[i] : 0.f);
        store8<T>(y + base + c, o);
      }
    }
  }
}

// dz and per-block partial dgamma / dbeta.  partial layout: [gridDim.x][2][d] fp32.  The groups' running sums are
// folded through shared memory at the end so a block writes one partial row.
template <typename T, int TPR, int CH, int BLOCK>
__global__ void __launch_bounds__(BLOCK, CH == 1 ? 2 : 1)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ z, const T* __restrict__ gamma,
                const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const T* __restrict__ dres,
                T* __restrict__ dz, float* __restrict__ partial, int rows, int d, int rms,
                FinishArgs fin) {
  __shared__ float2 red[2 * MAX_RPB * 8];
  constexpr int RPB = BLOCK / TPR;
  static_assert(RPB <= MAX_RPB && TPR / 32 <= 8, "bad norm config");
  constexpr bool DZS = (CH == 1);   // dz column sums are only carried by the narrow configurations (d <= 2048)
  constexpr int NACC = DZS ? 3 : 2;
  __shared__ __align__(16) float fold[RPB > 1 ? NACC * TPR * 8 * CH : 4];
  __shared__ bool is_last;
  const int sub = threadIdx.x / TPR;
  const int t = threadIdx.x % TPR;
  unsigned parity = 0;
  float dg[CH][8], db[CH][8], g[CH][8], dzs[DZS ? 8 : 1];
#pragma unroll
  for (int i = 0; i < (DZS ? 8 : 1); ++i) dzs[i] = 0.f;
#pragma unroll
  for (int k = 0; k < CH; ++k) {
    const int c = (t + k * TPR) * 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) { dg[k][i] = 0.f; db[k][i] = 0.f; g[k][i] = 0.f; }
    if (CH <= 2 && c < d) load8<T>(gamma + c, g[k]);   // few chunks: gamma stays in registers
  }
  const float inv_d = 1.f / d;
  for (int row = blockIdx.x * RPB + sub; row < rows; row += gridDim.x * RPB) {
    const long long base = static_cast<long long>(row) * d;
    const float mean = rms ? 0.f : mean_in[row];
    const float rstd = rstd_in[row];
    Raw8<T> ra[CH], rb[CH], rr[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        ra[k] = load_raw<T>(dy + base + c);
        rb[k] = load_raw<T>(z + base + c);
        if (dres) rr[k] = load_raw<T>(dres + base + c);
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        float a[8], b[8];
        unpack_raw<T>(ra[k], a);
        unpack_raw<T>(rb[k], b);
        if (CH > 2) load8<T>(gamma + c, g[k]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xh = (b[i] - mean) * rstd;
          const float wdy = a[i] * g[k][i];
          dg[k][i] += a[i] * xh;
          db[k][i] += a[i];
          s1 += wdy;
          s2 += wdy * xh;
        }
      }
    }
    const float2 tot = group_sum2<TPR>(make_float2(s1, s2), red, sub, t, parity);
    s1 = rms ? 0.f : tot.x * inv_d;
    s2 = tot.y * inv_d;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int c = (t + k * TPR) * 8;
      if (c < d) {
        float a[8], b[8], o[8];
        unpack_raw<T>(ra[k], a);
        unpack_raw<T>(rb[k], b);
        if (CH > 2) load8<T>(gamma + c, g[k]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xh = (b[i] - mean) * rstd;
          o[i] = rstd * (a[i] * g[k][i] - s1 - xh * s2);
        }
        if (dres) {  // gradient arriving through the skip connection that forked off the norm's input
          float e[8];
          unpack_raw<T>(rr[k], e);
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] += e[i];
        }
        if constexpr (DZS) {
#pragma unroll
          for (int i = 0; i < 8; ++i) dzs[i] += o[i];
        }
        store8<T>(dz + base + c, o);
      }
    }
  }
  // Block totals: folded through shared memory group by group; the last group then either stores the block's
  // partial row (two-kernel mode, `partial`) or adds it to the fp32 accumulator `fin.acc` (one-kernel mode).
  float* pg = fin.acc ? fin.acc : partial + static_cast<long long>(blockIdx.x) * 2 * d;
  const bool want_dzs = DZS && fin.acc != nullptr && fin.dzsum != nullptr;
  for (int w = 0; w < RPB; ++w) {
    if (sub == w) {
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        const int c = (t + k * TPR) * 8;
        if (c < d) {
          if (w > 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { dg[k][i] += fold[c + i]; db[k][i] += fold[d + c + i]; }
            if constexpr (DZS) {
#pragma unroll
              for (int i = 0; i < 8; ++i) dzs[i] += fold[2 * d + c + i];
            }
          }
          if (w == RPB - 1) {
            if (fin.acc) {
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                atomicAdd(reinterpret_cast<float4*>(pg + c) + h, make_float4(dg[k][4 * h], dg[k][4 * h + 1], dg[k][4 * h + 2], dg[k][4 * h + 3]));
                atomicAdd(reinterpret_cast<float4*>(pg + d + c) + h, make_float4(db[k][4 * h], db[k][4 * h + 1], db[k][4 * h + 2], db[k][4 * h + 3]));
                if constexpr (DZS) {
                  if (want_dzs)
                    atomicAdd(reinterpret_cast<float4*>(pg + 2 * d + c) + h, make_float4(dzs[4 * h], dzs[4 * h + 1], dzs[4 * h + 2], dzs[4 * h + 3]));
                }
              }
            } else {
              store8<float>(pg + c, dg[k]);
              store8<float>(pg + d + c, db[k]);
            }
          } else {
            store8<float>(fold + c, dg[k]);
            store8<float>(fold + d + c, db[k]);
            if constexpr (DZS) store8<float>(fold + 2 * d + c, dzs);
          }
        }
      }
    }
    if (RPB > 1) __syncthreads();
  }
  if (fin.acc == nullptr) return;
  // one-kernel mode: the last block to arrive rounds the accumulator into dgamma / dbeta (/ dz sums) and clears it
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(fin.counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int n_out = (want_dzs ? 3 : 2) * d;
  for (int i = threadIdx.x; i < n_out; i += BLOCK) {
    const float v = __ldcg(fin.acc + i);
    fin.acc[i] = 0.f;
    void* out = i < d ? fin.dgamma : (i < 2 * d ? fin.dbeta : fin.dzsum);
    const int col = i < d ? i : (i < 2 * d ? i - d : i - 2 * d);
    if (out == nullptr) continue;
    if (fin.param_dtype == DT_F32) reinterpret_cast<float*>(out)[col] = v;
    else if (fin.param_dtype == DT_BF16) reinterpret_cast<__nv_bfloat16*>(out)[col] = __float2bfloat16(v);
    else reinterpret_cast<__half*>(out)[col] = __float2half(v);
  }
  if (threadIdx.x == 0) *fin.counter = 0u;
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

struct Cfg { int idx, tpr, ch, block; };
inline Cfg pick_cfg(int d) {
  if (d <= 256) return {0, 32, 1, 256};
  if (d <= 512) return {1, 64, 1, 512};
  if (d <= 1024) return {2, 128, 1, 512};
  if (d <= 2048) return {3, 256, 1, 512};
  if (d <= 4096) return {4, 256, 2, 512};
  return {5, 256, 4, 256};
}
// one partial dgamma/dbeta row per block: the grid is what fits on the GPU at once
inline int bwd_grid(int rows, const Cfg& c) {
  const int rpb = c.block / c.tpr;
  const int g = (rows + rpb - 1) / rpb;
  const int cap = num_sms() * (c.ch == 1 ? 2 : 1);
  return g > cap ? cap : g;
}
inline int fwd_grid(int rows, const Cfg& c) {
  const int block = c.tpr < 128 ? 128 : c.tpr;   // see HYP_NORM_FWD_SWITCH
  const int rpb = block / c.tpr;
  return (rows + rpb - 1) / rpb;
}
// forward: small blocks, one row per group and no grid cap -- nothing is carried across rows, so many short blocks
// hide the load -> reduce -> reduce -> store chain better than a few long-lived ones
#define HYP_NORM_FWD_SWITCH(idx, CALL)   \
  switch (idx) {                          \
    case 0: CALL(32, 1, 128); break;      \
    case 1: CALL(64, 1, 128); break;      \
    case 2: CALL(128, 1, 128); break;     \
    case 3: CALL(256, 1, 256); break;     \
    case 4: CALL(256, 2, 256); break;     \
    default: CALL(256, 4, 256); break;    \
  }
#define HYP_NORM_CFG_SWITCH(idx, CALL)   \
  switch (idx) {                          \
    case 0: CALL(32, 1, 256); break;      \
    case 1: CALL(64, 1, 512); break;      \
    case 2: CALL(128, 1, 512); break;     \
    case 3: CALL(256, 1, 512); break;     \
    case 4: CALL(256, 2, 512); break;     \
    default: CALL(256, 4, 256); break;    \
  }

}  // namespace norm
}  // namespace hyp

using namespace hyp;

static int norm_fwd_impl(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                         void* sum_out, float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype,
                         const DropoutArgs& drop, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && d > 0 && d % 8 == 0 && d <= 8192, "hyp_norm_fwd: need d %% 8 == 0 and d <= 8192 (d=%d)", d);
  const norm::Cfg cfg = norm::pick_cfg(d);
  const int grid = norm::fwd_grid(rows, cfg);
#define HYP_NORM_FWD(TPR, CH, BLOCK)                                                                                  \
  norm::norm_fwd_kernel<TT, TPR, CH, BLOCK><<<grid, BLOCK, 0, st>>>((const TT*)x, (const TT*)residual, (const TT*)gamma, \
                                                                     (const TT*)beta, (TT*)y, (TT*)sum_out, mean, rstd, \
                                                                     rows, d, eps, rms)
#define HYP_NORM_FWD_DROP(TPR, CH, BLOCK)                                                                             \
  norm::norm_fwd_kernel<TT, TPR, CH, BLOCK, true><<<grid, BLOCK, 0, st>>>((const TT*)x, (const TT*)residual,          \
                                                                           (const TT*)gamma, (const TT*)beta, (TT*)y, \
                                                                           (TT*)sum_out, mean, rstd, rows, d, eps, rms, drop)
  if (drop.state) {
    if (dtype == DT_F32) { using TT = float; HYP_NORM_FWD_SWITCH(cfg.idx, HYP_NORM_FWD_DROP) }
    else if (dtype == DT_BF16) { using TT = __nv_bfloat16; HYP_NORM_FWD_SWITCH(cfg.idx, HYP_NORM_FWD_DROP) }
    else { using TT = __half; HYP_NORM_FWD_SWITCH(cfg.idx, HYP_NORM_FWD_DROP) }
  } else {
    if (dtype == DT_F32) { using TT = float; HYP_NORM_FWD_SWITCH(cfg.idx, HYP_NORM_FWD) }
    else if (dtype == DT_BF16) { using TT = __nv_bfloat16; HYP_NORM_FWD_SWITCH(cfg.idx, HYP_NORM_FWD) }
    else { using TT = __half; HYP_NORM_FWD_SWITCH(cfg.idx, HYP_NORM_FWD) }
  }
#undef HYP_NORM_FWD
#undef HYP_NORM_FWD_DROP
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                         void* sum_out, float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype,
                         void* stream_) {
  return norm_fwd_impl(x, residual, gamma, beta, y, sum_out, mean, rstd, rows, d, eps, rms, dtype, DropoutArgs{}, stream_);
}

// z = dropout(x, p) + residual, then the norm: the residual-branch dropout fused into the add + norm pass.  The gradient of
// x is hyp_dropout(dz) with the same (rng_state, salt, p); the gradient of `residual` is dz itself.
HYP_API int hyp_norm_fwd_drop(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                              void* sum_out, float* mean, float* rstd, int rows, int d, float eps, int rms, int dtype,
                              const unsigned long long* rng_state, int salt, float dropout_p, void* stream_) {
  DropoutArgs drop{};
  if (rng_state != nullptr && dropout_p > 0.f) {
    HYP_REQUIRE(dropout_p < 1.f, "hyp_norm_fwd_drop: dropout probability must be in [0, 1) (got %f)", dropout_p);
    int t = static_cast<int>(dropout_p * 256.f + 0.5f);
    t = t < 1 ? 1 : (t > 255 ? 255 : t);
    drop.state = rng_state; drop.salt = static_cast<unsigned>(salt); drop.thresh = static_cast<unsigned>(t);
    drop.scale = 256.f / static_cast<float>(256 - t);
  }
  return norm_fwd_impl(x, residual, gamma, beta, y, sum_out, mean, rstd, rows, d, eps, rms, dtype, drop, stream_);
}

// Bytes of fp32 scratch hyp_norm_bwd needs for (rows, d).
HYP_API long long hyp_norm_bwd_scratch_bytes(int rows, int d) {
  const long long P = norm::bwd_grid(rows, norm::pick_cfg(d));  // one partial row per block
  return P * 2 * d * static_cast<long long>(sizeof(float));
}

namespace hyp { namespace norm {
int launch_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd, void* dz,
               void* dgamma, void* dbeta, float* scratch, int rows, int d, int rms, int dtype, int accumulate,
               cudaStream_t st, const void* dres = nullptr, int param_dtype = DT_F32, bool one_kernel = false,
               void* dzsum = nullptr) {
  const Cfg cfg = pick_cfg(d);
  const int grid = bwd_grid(rows, cfg);
  const int P = grid;
  // one-kernel mode: `scratch` is the zeroed accumulator [2][d] followed by the arrival counter
  FinishArgs fin{nullptr, nullptr, dgamma, dbeta, dzsum, param_dtype};
  if (one_kernel) { fin.acc = scratch; fin.counter = reinterpret_cast<unsigned*>(scratch + 3 * d); }
#define HYP_NORM_BWD(TPR, CH, BLOCK)                                                                                \
  norm_bwd_kernel<TT, TPR, CH, BLOCK><<<grid, BLOCK, 0, st>>>((const TT*)dy, (const TT*)z, (const TT*)gamma, mean, rstd, \
                                                              (const TT*)dres, (TT*)dz, scratch, rows, d, rms, fin)
  if (dtype == DT_F32) { using TT = float; HYP_NORM_CFG_SWITCH(cfg.idx, HYP_NORM_BWD) }
  else if (dtype == DT_BF16) { using TT = __nv_bfloat16; HYP_NORM_CFG_SWITCH(cfg.idx, HYP_NORM_BWD) }
  else { using TT = __half; HYP_NORM_CFG_SWITCH(cfg.idx, HYP_NORM_BWD) }
#undef HYP_NORM_BWD
  HYP_CHECK_CUDA(cudaGetLastError());
  if (one_kernel) return 0;
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

// hyp_norm_bwd with three fusions, used by the framework's own ops:
//  * `dres` (same shape/dtype as dz, may be null) is added to dz -- the gradient that reaches the norm's input through
//    the skip connection of a pre-norm block, which autograd would otherwise add with a separate elementwise pass;
//  * dgamma / dbeta are written in `param_dtype` (DT_F32, or the parameters' 16-bit dtype so they can land directly in
//    the flat gradient buffer); `dz_colsum` (may be null; d <= 2048) receives the column sums of dz in the same dtype --
//    the bias gradient of the linear layer that produced this norm's input;
//  * there is no second kernel: blocks add their partial sums into the fp32 accumulator at the start of `ws`, and the
//    last block to finish rounds it into dgamma / dbeta.  `ws` (hyp_norm_bwd_fused_ws_bytes(d) bytes) must be all zero
//    on entry and is all zero again on return, so one persistent workspace per stream serves every call.  The sums
//    are accumulated with floating-point atomics, i.e. not bit-reproducible from run to run.
HYP_API long long hyp_norm_bwd_fused_ws_bytes(int d) { return (3ll * d + 8) * static_cast<long long>(sizeof(float)); }

HYP_API int hyp_norm_bwd_fused(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                               const void* dres, void* dz, void* dgamma, void* dbeta, void* dz_colsum, int param_dtype,
                               void* ws, long long ws_bytes, int rows, int d, int rms, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(rows > 0 && d > 0 && d % 8 == 0 && d <= 8192, "hyp_norm_bwd_fused: need d %% 8 == 0 and d <= 8192 (d=%d)", d);
  HYP_REQUIRE(param_dtype == DT_F32 || param_dtype == DT_F16 || param_dtype == DT_BF16, "hyp_norm_bwd_fused: bad param dtype");
  HYP_REQUIRE(ws != nullptr && ws_bytes >= hyp_norm_bwd_fused_ws_bytes(d) && reinterpret_cast<uintptr_t>(ws) % 16 == 0,
              "hyp_norm_bwd_fused: workspace too small or misaligned (%lld < %lld bytes)", ws_bytes, hyp_norm_bwd_fused_ws_bytes(d));
  HYP_REQUIRE(dz_colsum == nullptr || d <= 2048, "hyp_norm_bwd_fused: dz column sums are only produced for d <= 2048 (d=%d)", d);
  return norm::launch_bwd(dy, z, gamma, mean, rstd, dz, dgamma, dbeta, reinterpret_cast<float*>(ws), rows, d, rms,
                          dtype, 0, st, dres, param_dtype, true, dz_colsum);
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
// loads the library holds HYP_NORM_BWD_NOALLOC_SCRATCH_BYTES (8 MiB) of device memory for it until the
// process exits.  Shapes whose partial sums do not fit (hyp_norm_bwd_scratch_bytes(rows, d) > 8 MiB, which
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
constexpr long long kNoAllocScratchBytes = 8ll << 20;
static __device__ __align__(256) unsigned char g_bwd_scratch[kNoAllocScratchBytes];
// device address of g_bwd_scratch per device ordinal; the symbol itself is only a host-side stand-in.  Two host threads
// making their first call on one device at the same time both look the address up and store the same value, hence atomics
// (relaxed is enough: the value is a constant of the device).
static std::atomic<void*> g_bwd_scratch_addr[64];

static int resolve_scratch(float** out) {
  int dev = 0;
  HYP_CHECK_CUDA(cudaGetDevice(&dev));
  HYP_REQUIRE(dev >= 0 && dev < 64, "hyp_norm_bwd_noalloc: device ordinal %d out of range", dev);
  void* p = g_bwd_scratch_addr[dev].load(std::memory_order_relaxed);
  if (p == nullptr) {
    HYP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bwd_scratch));
    g_bwd_scratch_addr[dev].store(p, std::memory_order_relaxed);
  }
  *out = reinterpret_cast<float*>(p);
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
