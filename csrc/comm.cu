// Communication kernels over NVLink 5 / NVSwitch peer memory.
//
// Every rank maps every peer's symmetric buffers (torch symmetric memory / CUDA
// VMM), so a kernel addresses remote data with plain loads/stores on peer
// pointers, or -- when a multicast (NVLS) mapping exists -- with
// multimem.ld_reduce / multimem.st, which make the switch do the reduction and
// the broadcast.  Cross-GPU ordering uses release/acquire flags at system scope
// in a symmetric flag region; there is no host round-trip and no NCCL call.
//
//   hyp_comm_allreduce      K-C1  DDP bucket all-reduce fused with the AMP
//                                 unscale (x 1/(W*scale)), inf/nan check and cast
//   hyp_comm_reduce_scatter K-C2  FSDP gradient reduce-scatter fused with the
//                                 sharded AdamW update (or, when global-norm
//                                 clipping is on, with the sum-of-squares pass)
//   hyp_comm_allgather      K-C3  FSDP parameter all-gather (the chunk flags it
//                                 publishes let the consumer GEMM start on tiles
//                                 whose weights have landed)
//
// Reference call sites: the implicit collectives of DDP / FSDP1 in
// 02_development/distributed_utils.py:159/178 (all-reduce), :328-332/349
// (reduce-scatter, all-gather), executed there by RCCL.
//
// Synchronisation model: "block b of rank r" only ever depends on "block b of
// every peer" (same grid on all ranks, same index->block mapping), so barriers
// are per block: slot[b*W + src] in the destination's flag region is written
// with a monotonically increasing epoch.  All blocks of a launch must be
// co-resident (grid <= 64), and launches that wait on peers are serialised on
// one stream per process.
#include "common/ptx.cuh"
#include "common/util.cuh"
#include <string.h>

namespace hyp {
namespace comm {

constexpr int MAX_W = 8;
constexpr int MAX_BLOCKS = 64;
constexpr int THREADS = 512;
constexpr int BARRIER_SLOTS = 256 * MAX_W;     // per-block barrier slots; the all-gather-fused GEMM uses up to 148 blocks
constexpr int INF_WORD = BARRIER_SLOTS;        // float word: 1.0 when any rank saw a non-finite gradient
constexpr long long SPIN_LIMIT = 1ll << 31;   // ~ several seconds, then trap instead of hanging the GPU

struct Peers {
  void* buf[MAX_W];        // base of the symmetric data buffer on each rank
  uint32_t* flags[MAX_W];  // base of the symmetric flag region on each rank
};

// Per-block barrier across ranks. All threads of the block must call it.
__device__ __forceinline__ void block_barrier(const Peers& P, int rank, int W, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < W) {
    const int t = threadIdx.x;
    __threadfence_system();
    st_release_sys(P.flags[t] + blockIdx.x * W + rank, epoch);
    const uint32_t* mine = P.flags[rank] + blockIdx.x * W + t;
    long long spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if (++spins > SPIN_LIMIT) {
        printf("hyperion comm: rank %d block %d timed out waiting for rank %d (epoch %u)\n", rank, blockIdx.x, t, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
}

template <typename T>
struct Vec;  // 16-byte vector of T
template <>
struct Vec<float> {
  static constexpr int N = 4;
  __device__ static void load(const void* p, float (&f)[8]) {
    const uint4 u = ld_volatile_v4(p);
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
  }
  __device__ static void store(void* p, const float (&f)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  }
  __device__ static void mc_reduce(const void* mc, float (&f)[8]) {
    const float4 v = multimem_ld_reduce_f32x4(mc);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
};
template <>
struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const void* p, float (&f)[8]) { unpack8<__nv_bfloat16>(ld_volatile_v4(p), f); }
  __device__ static void store(void* p, const float (&f)[8]) { *reinterpret_cast<uint4*>(p) = pack8<__nv_bfloat16>(f); }
  __device__ static void mc_reduce(const void* mc, float (&f)[8]) { unpack8<__nv_bfloat16>(multimem_ld_reduce_bf16x8(mc), f); }
};
template <>
struct Vec<__half> {
  static constexpr int N = 8;
  __device__ static void load(const void* p, float (&f)[8]) { unpack8<__half>(ld_volatile_v4(p), f); }
  __device__ static void store(void* p, const float (&f)[8]) { *reinterpret_cast<uint4*>(p) = pack8<__half>(f); }
  __device__ static void mc_reduce(const void* mc, float (&f)[8]) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    unpack8<__half>(v, f);
  }
};

// Sum of the 16-byte vector at byte offset `off` over all ranks (fp32 accumulation).
template <typename T>
__device__ __forceinline__ void reduce_vec(const Peers& P, const char* mc, int rank, int W, long long off, float (&acc)[8]) {
  constexpr int N = Vec<T>::N;
  if (mc) {
    Vec<T>::mc_reduce(mc + off, acc);
    return;
  }
  float part[MAX_W][8];
#pragma unroll
  for (int p = 0; p < MAX_W; ++p)
    if (p < W) Vec<T>::load(static_cast<const char*>(P.buf[(rank + p) % W]) + off, part[p]);
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i] = part[0][i];
#pragma unroll
  for (int p = 1; p < MAX_W; ++p)
    if (p < W) {
#pragma unroll
      for (int i = 0; i < N; ++i) acc[i] += part[p][i];
    }
}

__device__ __forceinline__ void raise_inf(const Peers& P, int W) {
  for (int p = 0; p < W; ++p) reinterpret_cast<float*>(P.flags[p])[INF_WORD] = 1.0f;
}

// ------------------------------------------------------------------ K-C1
// In-place average of buf[off_bytes : off_bytes + n*sizeof(T)] across ranks.
template <typename T>
__global__ void __launch_bounds__(THREADS, 1)
allreduce_kernel(Peers P, char* mc, int rank, int W, long long off_bytes, long long n, uint32_t epoch_arg,
                 const uint32_t* __restrict__ epoch_base, const float* __restrict__ inv_scale, float extra_scale) {
  const uint32_t epoch = epoch_arg + (epoch_base ? *epoch_base : 0u);
  constexpr int N = Vec<T>::N;
  const long long nvec = (n + N - 1) / N;                 // region is padded to 16 bytes by the caller
  const long long per = (nvec + W - 1) / W;               // vectors per rank slice
  const long long v0 = static_cast<long long>(rank) * per;
  const long long v1 = v0 + per < nvec ? v0 + per : nvec;
  const float scale = (inv_scale ? *inv_scale : 1.f) * extra_scale;
  char* mine = static_cast<char*>(P.buf[rank]);

  block_barrier(P, rank, W, epoch);  // every rank's bucket is complete
  bool bad = false;
  // UNR independent 16-byte vectors per thread are in flight at a time (NVLink round trips are ~2 us)
  constexpr int UNR = 4;
  const long long tstride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long vb = v0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; vb < v1; vb += UNR * tstride) {
    float acc[UNR][8];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long v = vb + u * tstride;
      if (v < v1) reduce_vec<T>(P, mc, rank, W, off_bytes + v * 16, acc[u]);
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long v = vb + u * tstride;
      if (v >= v1) continue;
      const long long off = off_bytes + v * 16;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        acc[u][i] *= scale;
        bad |= !isfinite(acc[u][i]);
      }
      if (mc) {
        uint4 q;
        float tmp[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) tmp[i] = i < N ? acc[u][i] : 0.f;
        if constexpr (sizeof(T) == 4) q = make_uint4(__float_as_uint(tmp[0]), __float_as_uint(tmp[1]), __float_as_uint(tmp[2]), __float_as_uint(tmp[3]));
        else q = pack8<T>(tmp);
        multimem_st_u32x4(mc + off, q);  // the switch broadcasts the reduced vector to every rank
      } else {
        Vec<T>::store(mine + off, acc[u]);
      }
    }
  }
  if (bad) raise_inf(P, W);
  block_barrier(P, rank, W, epoch + 1);  // all slices reduced (and, with NVLS, already broadcast)
  if (!mc) {
    // all-gather phase: pull every other rank's reduced slice
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < per;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
      uint4 tmp[MAX_W];
#pragma unroll
      for (int p = 1; p < MAX_W; ++p)
        if (p < W) {
          const int q = (rank + p) % W;
          const long long v = static_cast<long long>(q) * per + i;
          if (v < nvec) tmp[p] = ld_volatile_v4(static_cast<const char*>(P.buf[q]) + off_bytes + v * 16);
        }
#pragma unroll
      for (int p = 1; p < MAX_W; ++p)
        if (p < W) {
          const int q = (rank + p) % W;
          const long long v = static_cast<long long>(q) * per + i;
          if (v < nvec) *reinterpret_cast<uint4*>(mine + off_bytes + v * 16) = tmp[p];
        }
    }
    block_barrier(P, rank, W, epoch + 2);  // peers are done reading my slice; the buffer may be reused
  }
}

// ------------------------------------------------------------------ K-C2
struct AdamShard {
  float* master;  // fp32 master shard (n_shard)
  float* m;
  float* v;
  void* p16;      // 16-bit parameter shard written for the next all-gather (may be null)
  float lr, beta1, beta2, eps, weight_decay, bias_c1, bias_c2;
  const float* step_ptr;  // device step count (CUDA-graph safe) or null
};

// Rank r reduces slice r of the gradient region across ranks (average).
//   fused_adam = 1: apply AdamW to the local fp32 master shard in the same pass
//   fused_adam = 0: write the fp32 reduced shard to `gshard` and add its sum of squares to sumsq[0]
template <typename T, typename TP>
__global__ void __launch_bounds__(THREADS, 1)
reduce_scatter_kernel(Peers P, char* mc, int rank, int W, long long off_bytes, long long per_vec, uint32_t epoch_arg,
                      const uint32_t* __restrict__ epoch_base, float scale, float* __restrict__ gshard,
                      float* __restrict__ sumsq, int fused_adam, AdamShard A) {
  const uint32_t epoch = epoch_arg + (epoch_base ? *epoch_base : 0u);
  constexpr int N = Vec<T>::N;
  __shared__ float red[32];
  block_barrier(P, rank, W, epoch);
  float ss = 0.f;
  bool bad = false;
  float c1 = A.bias_c1, c2 = A.bias_c2;
  if (A.step_ptr) {
    const float t = fmaxf(*A.step_ptr, 1.f);
    c1 = 1.f - powf(A.beta1, t);
    c2 = 1.f - powf(A.beta2, t);
  }
  const float step_size = A.lr / c1;
  const float inv_sqrt_c2 = rsqrtf(c2);
  const float decay = 1.f - A.lr * A.weight_decay;
  constexpr int UNR = 2;
  const long long tstride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long ib = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; ib < per_vec; ib += UNR * tstride) {
    float gg[UNR][8];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long i = ib + u * tstride;
      if (i < per_vec) reduce_vec<T>(P, mc, rank, W, off_bytes + (static_cast<long long>(rank) * per_vec + i) * 16, gg[u]);
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long i = ib + u * tstride;
      if (i >= per_vec) continue;
      float (&g)[8] = gg[u];
      const long long e0 = i * N;  // element index inside the shard
#pragma unroll
      for (int k = 0; k < N; ++k) {
        g[k] *= scale;
        ss += g[k] * g[k];
        bad |= !isfinite(g[k]);
      }
      if (!fused_adam) {
#pragma unroll
        for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(gshard + e0 + k) = make_float4(g[k], g[k + 1], g[k + 2], g[k + 3]);
      } else {
        float out[8];
#pragma unroll
        for (int k = 0; k < N; k += 4) {
          float4 mm = *reinterpret_cast<float4*>(A.m + e0 + k), vv = *reinterpret_cast<float4*>(A.v + e0 + k),
                 pp = *reinterpret_cast<float4*>(A.master + e0 + k);
          float* mf = reinterpret_cast<float*>(&mm); float* vf = reinterpret_cast<float*>(&vv); float* pf = reinterpret_cast<float*>(&pp);
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            mf[t] = A.beta1 * mf[t] + (1.f - A.beta1) * g[k + t];
            vf[t] = A.beta2 * vf[t] + (1.f - A.beta2) * g[k + t] * g[k + t];
            pf[t] = pf[t] * decay - step_size * mf[t] / (sqrtf(vf[t]) * inv_sqrt_c2 + A.eps);
            out[k + t] = pf[t];
          }
          *reinterpret_cast<float4*>(A.m + e0 + k) = mm;
          *reinterpret_cast<float4*>(A.v + e0 + k) = vv;
          *reinterpret_cast<float4*>(A.master + e0 + k) = pp;
        }
        if (A.p16) {
#pragma unroll
          for (int k = 0; k < N; ++k) reinterpret_cast<TP*>(A.p16)[e0 + k] = from_f32<TP>(out[k]);
        }
      }
    }
  }
  if (sumsq) {
    ss = block_sum(ss, red);
    if (threadIdx.x == 0) atomicAdd(sumsq, ss);
  }
  if (bad) raise_inf(P, W);
  block_barrier(P, rank, W, epoch + 1);  // peers are done reading my gradients
}

// ------------------------------------------------------------------ K-C3 (gather part)
// dst[q*per_vec + i] = peer q's shard[i] for every rank q (own shard included).
// After a block has copied its part of rank q's shard it bumps chunk_flags[q] (local
// memory) so a consumer kernel can start on weights that have landed.
__global__ void __launch_bounds__(THREADS, 1)
allgather_kernel(Peers P, int rank, int W, long long shard_off_bytes, long long per_vec, char* __restrict__ dst,
                 uint32_t epoch_arg, const uint32_t* __restrict__ epoch_base, uint32_t* __restrict__ chunk_flags) {
  const uint32_t epoch = epoch_arg + (epoch_base ? *epoch_base : 0u);
  block_barrier(P, rank, W, epoch);  // every rank's shard holds this step's parameters
  for (int p = 0; p < W; ++p) {
    const int q = (rank + p) % W;
    const char* src = static_cast<const char*>(P.buf[q]) + shard_off_bytes;
    char* d = dst + static_cast<long long>(q) * per_vec * 16;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < per_vec; i += 4 * stride) {
      const uint4 a0 = ld_volatile_v4(src + i * 16), a1 = ld_volatile_v4(src + (i + stride) * 16),
                  a2 = ld_volatile_v4(src + (i + 2 * stride) * 16), a3 = ld_volatile_v4(src + (i + 3 * stride) * 16);
      *reinterpret_cast<uint4*>(d + i * 16) = a0;
      *reinterpret_cast<uint4*>(d + (i + stride) * 16) = a1;
      *reinterpret_cast<uint4*>(d + (i + 2 * stride) * 16) = a2;
      *reinterpret_cast<uint4*>(d + (i + 3 * stride) * 16) = a3;
    }
    for (; i < per_vec; i += stride) *reinterpret_cast<uint4*>(d + i * 16) = ld_volatile_v4(src + i * 16);
    if (chunk_flags) {
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(chunk_flags + q, 1u);
      }
    }
  }
  block_barrier(P, rank, W, epoch + 1);  // my shard may be overwritten by the next optimizer step
}

static int fill_peers(Peers& P, const long long* bufs, const long long* flags, int W) {
  HYP_REQUIRE(W >= 1 && W <= MAX_W, "comm: world size %d not in [1, %d]", W, MAX_W);
  for (int i = 0; i < MAX_W; ++i) {
    P.buf[i] = i < W ? reinterpret_cast<void*>(bufs[i]) : nullptr;
    P.flags[i] = i < W ? reinterpret_cast<uint32_t*>(flags[i]) : nullptr;
  }
  return 0;
}

static int pick_grid(long long work_vecs, int max_blocks) {
  long long b = (work_vecs + THREADS - 1) / THREADS;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return static_cast<int>(b);
}

}  // namespace comm
}  // namespace hyp

using namespace hyp;
using namespace hyp::comm;

// Words (uint32) the symmetric flag region must hold.
HYP_API int hyp_comm_flag_words() { return INF_WORD + 64; }
HYP_API int hyp_comm_inf_word() { return INF_WORD; }
HYP_API int hyp_comm_max_blocks() { return MAX_BLOCKS; }

// peer_bufs / peer_flags: HOST arrays of W device addresses (this rank's view of every rank's
// symmetric buffer / flag region).  mc_ptr: multicast address of the buffer or 0.  The region
// [off_bytes, off_bytes + n*elem) must be 16-byte aligned and padded to 16 bytes.  `epoch` must be the
// same on every rank and grow by 4 per launch.  Uses 3 barrier epochs.
HYP_API int hyp_comm_allreduce(const long long* peer_bufs, const long long* peer_flags, long long mc_ptr, int rank, int W,
                               long long off_bytes, long long n, int dtype, unsigned epoch, const unsigned* epoch_base,
                               const float* inv_scale, float extra_scale, int max_blocks, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  Peers P;
  if (fill_peers(P, peer_bufs, peer_flags, W)) return -1;
  HYP_REQUIRE(off_bytes % 16 == 0, "hyp_comm_allreduce: region must be 16-byte aligned");
  if (n <= 0) return 0;
  if (max_blocks <= 0 || max_blocks > MAX_BLOCKS) max_blocks = MAX_BLOCKS;
  const int es = dtype == DT_F32 ? 4 : 2;
  const long long nvec = (n * es + 15) / 16;
  const int grid = pick_grid((nvec + W - 1) / W, max_blocks);
  char* mc = reinterpret_cast<char*>(mc_ptr);
  if (dtype == DT_F32) allreduce_kernel<float><<<grid, THREADS, 0, st>>>(P, mc, rank, W, off_bytes, n, epoch, epoch_base, inv_scale, extra_scale);
  else if (dtype == DT_BF16) allreduce_kernel<__nv_bfloat16><<<grid, THREADS, 0, st>>>(P, mc, rank, W, off_bytes, n, epoch, epoch_base, inv_scale, extra_scale);
  else allreduce_kernel<__half><<<grid, THREADS, 0, st>>>(P, mc, rank, W, off_bytes, n, epoch, epoch_base, inv_scale, extra_scale);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Gradient region = W slices of per_elems elements starting at off_bytes; rank r reduces slice r.
// fused_adam: master/m/v/p16 describe the local shard (per_elems elements). Uses 2 barrier epochs.
HYP_API int hyp_comm_reduce_scatter(const long long* peer_bufs, const long long* peer_flags, long long mc_ptr, int rank,
                                    int W, long long off_bytes, long long per_elems, int dtype, unsigned epoch,
                                    const unsigned* epoch_base, float scale, float* gshard, float* sumsq, int fused_adam,
                                    float* master, float* m, float* v, void* p16, int p16_dtype, float lr, float beta1,
                                    float beta2, float eps, float weight_decay, int step, const float* step_ptr,
                                    int max_blocks, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  Peers P;
  if (fill_peers(P, peer_bufs, peer_flags, W)) return -1;
  const int es = dtype == DT_F32 ? 4 : 2;
  HYP_REQUIRE(off_bytes % 16 == 0 && (per_elems * es) % 16 == 0, "hyp_comm_reduce_scatter: slices must be 16-byte multiples");
  HYP_REQUIRE(fused_adam ? (master && m && v && (step >= 1 || step_ptr)) : gshard != nullptr, "hyp_comm_reduce_scatter: missing outputs");
  if (per_elems <= 0) return 0;
  if (max_blocks <= 0 || max_blocks > MAX_BLOCKS) max_blocks = MAX_BLOCKS;
  const long long per_vec = per_elems * es / 16;
  const int grid = pick_grid(per_vec, max_blocks);
  AdamShard A;
  A.master = master; A.m = m; A.v = v; A.p16 = p16;
  A.lr = lr; A.beta1 = beta1; A.beta2 = beta2; A.eps = eps; A.weight_decay = weight_decay;
  A.bias_c1 = 1.f - powf(beta1, (float)(step > 0 ? step : 1));
  A.bias_c2 = 1.f - powf(beta2, (float)(step > 0 ? step : 1));
  A.step_ptr = step_ptr;
  char* mc = reinterpret_cast<char*>(mc_ptr);
#define HYP_RS(T, TP) reduce_scatter_kernel<T, TP><<<grid, THREADS, 0, st>>>(P, mc, rank, W, off_bytes, per_vec, epoch, epoch_base, scale, gshard, sumsq, fused_adam, A)
  if (dtype == DT_F32) { if (p16_dtype == DT_F16) HYP_RS(float, __half); else HYP_RS(float, __nv_bfloat16); }
  else if (dtype == DT_BF16) { if (p16_dtype == DT_F16) HYP_RS(__nv_bfloat16, __half); else HYP_RS(__nv_bfloat16, __nv_bfloat16); }
  else { if (p16_dtype == DT_F16) HYP_RS(__half, __half); else HYP_RS(__half, __nv_bfloat16); }
#undef HYP_RS
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dst (local, W*per_bytes) <- every rank's shard bytes [shard_off_bytes, +per_bytes). Uses 2 barrier epochs.
HYP_API int hyp_comm_allgather(const long long* peer_bufs, const long long* peer_flags, int rank, int W,
                               long long shard_off_bytes, long long per_bytes, void* dst, unsigned epoch,
                               const unsigned* epoch_base, unsigned* chunk_flags, int max_blocks, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  Peers P;
  if (fill_peers(P, peer_bufs, peer_flags, W)) return -1;
  HYP_REQUIRE(shard_off_bytes % 16 == 0 && per_bytes % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 16 == 0,
              "hyp_comm_allgather: 16-byte alignment required");
  if (per_bytes <= 0) return 0;
  if (max_blocks <= 0 || max_blocks > MAX_BLOCKS) max_blocks = MAX_BLOCKS;
  const long long per_vec = per_bytes / 16;
  const int grid = pick_grid(per_vec / 4 + 1, max_blocks);
  allgather_kernel<<<grid, THREADS, 0, st>>>(P, rank, W, shard_off_bytes, per_vec, static_cast<char*>(dst), epoch, epoch_base, chunk_flags);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- plain CUDA IPC helpers: the peer-memory provider used when torch symmetric memory is
// unavailable on a box.  hyp_ipc_alloc does cudaMalloc (set-up time only; never on a hot path).
HYP_API int hyp_ipc_alloc(long long bytes, long long* out_ptr, unsigned char* out_handle64) {
  void* p = nullptr;
  HYP_CHECK_CUDA(cudaMalloc(&p, static_cast<size_t>(bytes)));
  HYP_CHECK_CUDA(cudaMemset(p, 0, static_cast<size_t>(bytes)));
  cudaIpcMemHandle_t h;
  HYP_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(out_handle64, &h, 64);
  *out_ptr = reinterpret_cast<long long>(p);
  return 0;
}
HYP_API int hyp_ipc_open(const unsigned char* handle64, long long* out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  HYP_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = reinterpret_cast<long long>(p);
  return 0;
}
HYP_API int hyp_ipc_close(long long ptr) {
  HYP_CHECK_CUDA(cudaIpcCloseMemHandle(reinterpret_cast<void*>(ptr)));
  return 0;
}
HYP_API int hyp_ipc_free(long long ptr) {
  HYP_CHECK_CUDA(cudaFree(reinterpret_cast<void*>(ptr)));
  return 0;
}
