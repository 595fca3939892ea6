// gemm_tcgen05: persistent, warp-specialised GEMM for sm_100a.
//
//   D[M,N] = epilogue( alpha * A[M,K] . B[N,K]^T )       fp32 accumulation in TMEM
//
// * operands are staged by TMA (cp.async.bulk.tensor, SWIZZLE_128B) into a
//   4..6 stage shared-memory ring; one elected thread issues tcgen05.mma
//   (128 x BN x 16, cta_group::1); two TMEM accumulator stages let the
//   epilogue of tile i overlap the main loop of tile i+1.
// * either operand may be "MN-major" (its non-reduction dimension contiguous in
//   global memory), which is what dgrad (dX = dY.W) and wgrad (dW = dY^T.X)
//   need -- no transposed copies are ever made.
// * dtypes: bf16 / fp16 (kind::f16) and fp32 run as tf32 (kind::tf32).
// * epilogue functors: +bias, ReLU / GELU(erf|tanh), +residual, activation
//   derivative (D = acc * act'(aux), optionally with the column sums of D = the
//   bias gradient of the layer below), beta-accumulate, split-K fp32 reduction.
// * gemm2_kernel: the same main loop on CTA PAIRS (tcgen05.mma cta_group::2,
//   256 x BN tiles, each CTA stages half of B, two TMA producer warps) -- the
//   default for shapes that fill the 74 pairs; gemm_kernel<..., AG> adds gather
//   warps that pull the peers' slices of an FSDP unit into this GPU's buffer
//   while the GEMM consumes the slices that have landed (hyp_gemm_ag).
// * tile order and split-K ranges: common/tiling.cuh (host-testable).
//
// Replaces the reference's library GEMM call sites: torch.matmul in the probe
// (Phase 1/01_hardware_exploration.ipynb c1 L92) and every nn.Linear inside
// the models (02_development/distributed_utils.py:75-88).
#include "common/ptx.cuh"
#include "common/tiling.cuh"
#include "common/tmap.cuh"
#include "common/util.cuh"
#include <stdlib.h>
#include <string.h>

namespace hyp {
namespace gemm {

constexpr int BM = 128;
constexpr int ROW_BYTES = 128;  // one swizzle row: 64 x 16-bit or 32 x fp32
constexpr int A_STAGE = BM * ROW_BYTES;
// per-warp shared-memory stage of the coalesced epilogue (see epilogue32s)
constexpr int STG_ROW = 80;
constexpr int STG_WARP_BYTES = 32 * STG_ROW;
constexpr int STG_BYTES = 8 * STG_WARP_BYTES;   // eight epilogue warps
constexpr int NUM_THREADS = 320;  // warp0 TMA, warp1 MMA, warps2-9 epilogue (two warps per TMEM lane quarter)
constexpr int GROUP_M = 8;

// All-gather fusion (FSDP K-C3): the B operand (a weight) lives in a flat "unit" buffer whose W equal slices are
// owned by W ranks.  Four extra warps per CTA pull every rank's slice from peer memory into the local unit buffer
// while the GEMM runs; the TMA producer only waits for the slices that cover the B tile it is about to load, so
// tiles whose weights are local (or have already landed) start immediately.
struct AgParams {
  int enabled, W, rank;
  unsigned epoch;                 // cross-rank barrier epoch (uses epoch and epoch + 1)
  const unsigned* epoch_base;     // optional device word added to `epoch` (CUDA-graph replays)
  const char* src[8];             // peer r's shard of this unit (already offset to the unit's shard)
  uint32_t* peer_flags[8];        // symmetric flag regions (per-block barrier slots)
  char* dst;                      // local unit buffer (W * per_bytes)
  long long per_bytes;            // bytes of one rank's slice
  long long b_off_bytes;          // offset of the B matrix inside the unit buffer
  long long bias_off_bytes;       // offset of the bias inside the unit buffer, or -1
  long long row_bytes;            // bytes of one B row (K * elem)
  uint32_t* arrived;              // local uint32[W]: chunks of rank q copied so far (monotonic across launches)
  unsigned target;                // value arrived[q] reaches when rank q's slice is complete for THIS launch
  const unsigned* target_base;    // optional device word: target = *target_base + target (graph replays)
};

struct Params {
  int M, N, K;
  int esize;       // 2 (f16/bf16) or 4 (tf32)
  int a_mn, b_mn;  // operand majors
  int a_mn_mma, b_mn_mma;  // what the MMA descriptors assume (differs from a_mn/b_mn only under HYPERION_GEMM_DBG)
  int a_mn3, b_mn3;        // MN-major operand addressed through a rank-3 map: one TMA instruction per stage
  // dual-K (pair kernel only): after the K loop over (A, B), k2_blocks more 64-wide k-blocks come from a second operand
  // pair (A2 [M, 64*k2], B2 [N, 64*k2], both K-major) and accumulate into the same tile -- a LoRA update
  // y = x W^T + t B^T runs as one GEMM instead of a second pass over y
  int k2_blocks;
  uint32_t idesc2;
  int one_producer;   // HYPERION_GEMM_PRODUCERS=1: warp 0 issues both loads of a stage (A/B timing)
  alignas(64) CUtensorMap tmA2;
  alignas(64) CUtensorMap tmB2;
  uint32_t idesc;
  int splits;      // split-K factor (fp32 atomic epilogue when > 1)
  // epilogue
  void* D;
  long long ldd;
  int out_dtype;
  const void* bias;
  int bias_dtype;
  const void* aux;  // residual (mode 0) or pre-activation (mode 1)
  long long ld_aux;
  int aux_dtype;
  void* preact_out;  // optional: alpha*acc + bias before the activation (same ld / dtype as D)
  float* colsum;     // optional, EPI 2 only: fp32 [N], column sums of the ROUNDED output tile are added to it (the bias
                     // gradient of the layer whose pre-activation gradient this GEMM produces), so no kernel re-reads D
  int act;   // 0 none, 1 relu, 2 gelu(erf), 3 gelu(tanh), 4 silu
  int mode;  // 0: D = act(alpha*acc + bias) + aux ; 1: D = alpha*acc * act'(aux)
  float alpha, beta;
  int vec_ok;  // D (and aux) rows are 16-byte aligned -> vector stores
  int simple;  // plain epilogue (alpha, optional 16-bit bias) -> branch-free fast path
  AgParams ag;
};

template <int BN>
struct Cfg {
  static constexpr int STAGES = (BN == 256) ? 4 : 6;
  static constexpr int B_STAGE = BN * ROW_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int SMEM_BYTES = STAGES * (A_STAGE + B_STAGE) + 256 + STG_BYTES + 1024;
};

__device__ __forceinline__ float load_as_f32(const void* p, long long idx, int dt) {
  if (dt == DT_F32) return reinterpret_cast<const float*>(p)[idx];
  if (dt == DT_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[idx]);
  return __half2float(reinterpret_cast<const __half*>(p)[idx]);
}
__device__ __forceinline__ void store_from_f32(void* p, long long idx, int dt, float v) {
  if (dt == DT_F32)
    reinterpret_cast<float*>(p)[idx] = v;
  else if (dt == DT_BF16)
    reinterpret_cast<__nv_bfloat16*>(p)[idx] = __float2bfloat16_rn(v);
  else
    reinterpret_cast<__half*>(p)[idx] = __float2half_rn(v);
}
__device__ __forceinline__ float apply_act(float x, int act) {
  switch (act) {
    case 1: return fmaxf(x, 0.f);
    case 2: return gelu_erf(x);
    case 3: return gelu_tanh(x);
    case 4: return silu(x);
    default: return x;
  }
}
__device__ __forceinline__ float apply_dact(float x, int act) {
  switch (act) {
    case 1: return x > 0.f ? 1.f : 0.f;
    case 2: return dgelu_erf(x);
    case 3: return dgelu_tanh(x);
    case 4: return dsilu(x);
    default: return 1.f;
  }
}

// Epilogue for 8 consecutive columns of one output row.
// General (rare-variant) epilogue; deliberately NOT inlined: inlining 32 copies of this body per tile blew the
// kernel up to ~170 KB of SASS and left the epilogue warps stalled on instruction fetch (ncu: stall_no_inst).
__device__ __noinline__ void epilogue8(const Params* pp, float v0, float v1, float v2, float v3, float v4, float v5,
                                       float v6, float v7, int row, int col) {
  const Params& p = *pp;
  float v[8] = {v0, v1, v2, v3, v4, v5, v6, v7};
  const int nvalid = min(8, p.N - col);
  if (nvalid <= 0) return;
  const long long doff = static_cast<long long>(row) * p.ldd + col;
  const long long aoff = static_cast<long long>(row) * p.ld_aux + col;
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] *= p.alpha;

  if (p.splits > 1) {  // fp32 atomic reduction across K splits
    float* d = reinterpret_cast<float*>(p.D) + doff;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i < nvalid) atomicAdd(d + i, v[i]);
    return;
  }
  const bool vec = p.vec_ok && nvalid == 8;
  if (p.mode == 0) {
    if (p.bias) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (i < nvalid) v[i] += load_as_f32(p.bias, col + i, p.bias_dtype);
    }
    if (p.preact_out) {
      if (vec && p.out_dtype == DT_BF16) {
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.preact_out) + doff) = pack8<__nv_bfloat16>(v);
      } else if (vec && p.out_dtype == DT_F16) {
        *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(p.preact_out) + doff) = pack8<__half>(v);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (i < nvalid) store_from_f32(p.preact_out, doff + i, p.out_dtype, v[i]);
      }
    }
    if (p.act) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = apply_act(v[i], p.act);
    }
  }
  if (p.aux) {
    float a[8];
    if (vec && p.aux_dtype != DT_F32) {
      const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.aux) + aoff);
      if (p.aux_dtype == DT_BF16) unpack8<__nv_bfloat16>(u, a); else unpack8<__half>(u, a);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = (i < nvalid) ? load_as_f32(p.aux, aoff + i, p.aux_dtype) : 0.f;
    }
    if (p.mode == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] += a[i];
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] *= apply_dact(a[i], p.act);
    }
  }
  if (p.beta != 0.f) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i < nvalid) v[i] += p.beta * load_as_f32(p.D, doff + i, p.out_dtype);
  }
  if (vec) {
    if (p.out_dtype == DT_BF16) {
      *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.D) + doff) = pack8<__nv_bfloat16>(v);
    } else if (p.out_dtype == DT_F16) {
      *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(p.D) + doff) = pack8<__half>(v);
    } else {
      float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.D) + doff);
      d[0] = make_float4(v[0], v[1], v[2], v[3]);
      d[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i < nvalid) store_from_f32(p.D, doff + i, p.out_dtype, v[i]);
  }
}

using TileCoord = hyp::tiling::TileCoord;

// tile order (L2 grouping, split-K ranges): common/tiling.cuh, host-testable
__device__ __forceinline__ TileCoord decode_tile(int tile, int tiles_m, int tiles_n, int nkb, int splits) {
  return hyp::tiling::decode_gemm_tile<GROUP_M>(tile, tiles_m, tiles_n, nkb, splits);
}

// 32 consecutive columns of one output row, compile-time epilogue kind:
//   EPI 0: D = alpha*acc (+bias)                              (plain)
//   EPI 1: D = ACT(alpha*acc + bias) (+ residual aux); optional pre-activation side output
//   EPI 2: D = alpha*acc * ACT'(aux)                          (activation backward)
//   EPI 4: D += alpha*acc with red.global.add.v4.f32           (split-K partial tiles, fp32 output)
//   EPI 5: D = alpha*acc + beta*D                              (gradient accumulation)
// Requires 16-byte aligned rows (vec_ok), a full 32-column chunk, 16-bit aux / preact.
// ---- staged (coalesced) epilogue for 16-bit tensors -------------------------------------------------------------
// A thread owns one accumulator row (TMEM lane), so storing its 32 columns directly makes every warp-level store
// touch 32 different rows: 32 L1 transactions for 512 bytes, and an epilogue with two outputs and one input
// (bias + GELU + pre-activation, or act' * aux) runs longer than the 256x256x1024 main loop it should hide behind.
// Instead each warp bounces its 32 x 32 tile through a private 2.5 KB shared-memory stage: the row owner writes /
// reads 64 B rows there, and global memory is accessed with 4 lanes per row (64 B contiguous, 8 rows per
// instruction).  Stage rows are padded to 80 B so 16-byte accesses by 8 consecutive rows hit 32 distinct banks.

// the 16-bit tile this epilogue kind reads back (residual / activation input / beta*D), or null
template <int EPI>
__device__ __forceinline__ const uint16_t* epi_input(const Params& p, long long* ld) {
  if ((EPI == 1 || EPI == 2) && p.aux) { *ld = p.ld_aux; return reinterpret_cast<const uint16_t*>(p.aux); }
  if (EPI == 5 && p.out_dtype != DT_F32) { *ld = p.ldd; return reinterpret_cast<const uint16_t*>(p.D); }
  *ld = 0;
  return nullptr;
}
// issue the coalesced loads of that tile (lane -> row 8i + lane/4, 16-byte segment lane%4)
template <int EPI>
__device__ __forceinline__ void epi_prefetch(const Params& p, uint4 (&pre)[4], int row0, int lane, int col0) {
  long long ld;
  const uint16_t* base = epi_input<EPI>(p, &ld);
  if (base == nullptr) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gr = row0 + i * 8 + (lane >> 2);
    pre[i] = make_uint4(0u, 0u, 0u, 0u);
    if (gr < p.M) pre[i] = *reinterpret_cast<const uint4*>(base + static_cast<long long>(gr) * ld + col0 + (lane & 3) * 8);
  }
}
__device__ __forceinline__ void stage_store_tile(uint8_t* stg, const float (&v)[32], int out_dtype, void* dst, long long ld,
                                                 int row0, int lane, int col0, int M, float* colsum = nullptr) {
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    float t[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = v[g * 8 + i];
    *reinterpret_cast<uint4*>(stg + lane * STG_ROW + g * 16) = (out_dtype == DT_BF16) ? pack8<__nv_bfloat16>(t) : pack8<__half>(t);
  }
  __syncwarp();
  if (colsum != nullptr) {
    // lane j adds up column j of the staged 32 x 32 tile (the 16-bit values that are stored, so the result equals a
    // column sum over D); rows past M hold zeros here (zero-filled operand rows times act' of a zero-filled aux row)
    float cs = 0.f;
#pragma unroll
    for (int rr = 0; rr < 32; ++rr) {
      const uint16_t h = *reinterpret_cast<const uint16_t*>(stg + rr * STG_ROW + lane * 2);
      cs += (out_dtype == DT_BF16) ? __uint_as_float(static_cast<uint32_t>(h) << 16) : __half2float(__ushort_as_half(h));
    }
    atomicAdd(colsum + col0 + lane, cs);
  }
  uint16_t* d = reinterpret_cast<uint16_t*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = i * 8 + (lane >> 2);
    const uint4 u = *reinterpret_cast<const uint4*>(stg + rr * STG_ROW + (lane & 3) * 16);
    if (row0 + rr < M) *reinterpret_cast<uint4*>(d + static_cast<long long>(row0 + rr) * ld + col0 + (lane & 3) * 8) = u;
  }
  __syncwarp();
}

template <int EPI, int ACT>
__device__ __forceinline__ void epilogue32s(const Params& p, const uint32_t (&r)[32], const uint4 (&pre)[4], int row0,
                                            int lane, int col0, uint8_t* stg) {
  // fp32 math on packed pairs (util.cuh: ffma2 / fmul2 / fadd2)
  float2 vv[16];
  float (&v)[32] = reinterpret_cast<float (&)[32]>(vv);
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
  if (p.alpha != 1.f) {
    const float2 al = splat2(p.alpha);
#pragma unroll
    for (int i = 0; i < 16; ++i) vv[i] = fmul2(vv[i], al);
  }
  if (EPI != 2 && EPI != 5 && p.bias) {
    if (p.bias_dtype == DT_F32) {
      const float4* b4 = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.bias) + col0);
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        const float4 t = b4[g];
        vv[2 * g] = fadd2(vv[2 * g], make_float2(t.x, t.y));
        vv[2 * g + 1] = fadd2(vv[2 * g + 1], make_float2(t.z, t.w));
      }
    } else {
      const uint4* b4 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.bias) + col0);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float t[8];
        const uint4 u = b4[g];
        if (p.bias_dtype == DT_BF16) unpack8<__nv_bfloat16>(u, t); else unpack8<__half>(u, t);
#pragma unroll
        for (int i = 0; i < 4; ++i) vv[g * 4 + i] = fadd2(vv[g * 4 + i], make_float2(t[2 * i], t[2 * i + 1]));
      }
    }
  }
  if (EPI == 1 && p.preact_out) stage_store_tile(stg, v, p.out_dtype, p.preact_out, p.ldd, row0, lane, col0, p.M);
  if (EPI == 1) {
#pragma unroll
    for (int i = 0; i < 16; ++i) vv[i] = act_fast2<ACT>(vv[i]);
  }
  long long in_ld;
  const uint16_t* in_base = epi_input<EPI>(p, &in_ld);
  if (in_base != nullptr) {
    // prefetched tile -> stage -> my row
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<uint4*>(stg + (i * 8 + (lane >> 2)) * STG_ROW + (lane & 3) * 16) = pre[i];
    __syncwarp();
    const int in_dt = (EPI == 5) ? p.out_dtype : p.aux_dtype;
    const float2 be = splat2(p.beta);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float t[8];
      const uint4 u = *reinterpret_cast<const uint4*>(stg + lane * STG_ROW + g * 16);
      if (in_dt == DT_BF16) unpack8<__nv_bfloat16>(u, t); else unpack8<__half>(u, t);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 tt = make_float2(t[2 * i], t[2 * i + 1]);
        if (EPI == 1) vv[g * 4 + i] = fadd2(vv[g * 4 + i], tt);
        else if (EPI == 2) vv[g * 4 + i] = fmul2(vv[g * 4 + i], dact_fast2<ACT>(tt));
        else vv[g * 4 + i] = ffma2(be, tt, vv[g * 4 + i]);
      }
    }
    __syncwarp();
  }
  stage_store_tile(stg, v, p.out_dtype, p.D, p.ldd, row0, lane, col0, p.M, EPI == 2 ? p.colsum : nullptr);
}

template <int EPI, int ACT>
__device__ __forceinline__ void epilogue32(const Params& p, const uint32_t (&r)[32], int row, int col0) {
  const long long doff = static_cast<long long>(row) * p.ldd + col0;
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
  if (EPI == 4) {  // split-K: vector fp32 reductions into the (zero-initialised) output
    float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.D) + doff);
#pragma unroll
    for (int g = 0; g < 8; ++g) atomicAdd(d + g, make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]));
    return;
  }
  if (EPI != 2 && EPI != 5 && p.bias) {
    if (p.bias_dtype == DT_F32) {
      const float4* b4 = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.bias) + col0);
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        const float4 t = b4[g];
        v[4 * g] += t.x; v[4 * g + 1] += t.y; v[4 * g + 2] += t.z; v[4 * g + 3] += t.w;
      }
    } else {
      const uint4* b4 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.bias) + col0);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float t[8];
        const uint4 u = b4[g];
        if (p.bias_dtype == DT_BF16) unpack8<__nv_bfloat16>(u, t); else unpack8<__half>(u, t);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[g * 8 + i] += t[i];
      }
    }
  }
  if (EPI == 1 && p.preact_out) {
    uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.preact_out) + doff);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float t[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) t[i] = v[g * 8 + i];
      d[g] = (p.out_dtype == DT_BF16) ? pack8<__nv_bfloat16>(t) : pack8<__half>(t);
    }
  }
  if (EPI == 1) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = act_fast<ACT>(v[i]);
  }
  if ((EPI == 1 || EPI == 2) && p.aux) {
    const uint4* a4 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.aux) +
                                                     static_cast<long long>(row) * p.ld_aux + col0);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float t[8];
      const uint4 u = a4[g];
      if (p.aux_dtype == DT_BF16) unpack8<__nv_bfloat16>(u, t); else unpack8<__half>(u, t);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (EPI == 1) v[g * 8 + i] += t[i];
        else v[g * 8 + i] *= dact_fast<ACT>(t[i]);
      }
    }
  }
  if (EPI == 5) {  // D = alpha*acc + beta*D
    if (p.out_dtype == DT_F32) {
      const float4* d = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.D) + doff);
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        const float4 t = d[g];
        v[4 * g] += p.beta * t.x; v[4 * g + 1] += p.beta * t.y; v[4 * g + 2] += p.beta * t.z; v[4 * g + 3] += p.beta * t.w;
      }
    } else {
      const uint4* d = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.D) + doff);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float t[8];
        const uint4 u = d[g];
        if (p.out_dtype == DT_BF16) unpack8<__nv_bfloat16>(u, t); else unpack8<__half>(u, t);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[g * 8 + i] += p.beta * t[i];
      }
    }
  }
  if (p.out_dtype == DT_F32) {
    float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.D) + doff);
#pragma unroll
    for (int g = 0; g < 8; ++g) d[g] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
  } else {
    uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.D) + doff);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float t[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) t[i] = v[g * 8 + i];
      d[g] = (p.out_dtype == DT_BF16) ? pack8<__nv_bfloat16>(t) : pack8<__half>(t);
    }
  }
}

constexpr int AG_THREADS = 128;     // 4 gather warps (warps 10-13) in the all-gather-fused variant
constexpr int AG_CHUNK = 32768;     // bytes copied per work item
constexpr long long AG_SPIN_LIMIT = 1ll << 31;

__device__ __forceinline__ void ag_wait(const uint32_t* ctr, unsigned target) {
  long long spins = 0;
  unsigned v;
  do {
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    if (++spins > AG_SPIN_LIMIT) { printf("hyperion gemm_ag: timed out waiting for a gathered slice\n"); __trap(); }
  } while (static_cast<int>(v - target) < 0);
}

// Barrier between the gather warps of block b on every rank (same protocol as csrc/comm.cu).
__device__ __forceinline__ void ag_rank_barrier(const AgParams& g, int tid, unsigned epoch) {
  named_bar_sync(2, AG_THREADS);
  if (tid < g.W) {
    __threadfence_system();
    // slots [64*W, ...) : disjoint from the ones csrc/comm.cu kernels (<= 64 blocks) use on the communication stream
    st_release_sys(g.peer_flags[tid] + (64 + blockIdx.x) * g.W + g.rank, epoch);
    const uint32_t* mine = g.peer_flags[g.rank] + (64 + blockIdx.x) * g.W + tid;
    long long spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if (++spins > AG_SPIN_LIMIT) { printf("hyperion gemm_ag: rank %d timed out waiting for rank %d\n", g.rank, tid); __trap(); }
    }
  }
  named_bar_sync(2, AG_THREADS);
}

template <int BN, int EPI, int ACT, bool AG = false>
__global__ void __launch_bounds__(NUM_THREADS + (AG ? AG_THREADS : 0), 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ Params p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + C::STAGES * A_STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + C::STAGES * C::B_STAGE);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty + 2);
  uint8_t* stg_base = reinterpret_cast<uint8_t*>(full) + 256;  // 8 x STG_WARP_BYTES, one stage per epilogue warp

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int epr = ROW_BYTES / p.esize;          // elements per 128-byte row == BK
  const int nkb = (p.K + epr - 1) / epr;        // k blocks
  const int tiles_m = (p.M + BM - 1) / BM;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int total = tiles_m * tiles_n * p.splits;
  int ag_rot = 0;
  if (AG && p.ag.enabled) {
    const long long first = static_cast<long long>(p.ag.rank) * p.ag.per_bytes - p.ag.b_off_bytes;
    if (first > 0) ag_rot = static_cast<int>(min(static_cast<long long>(tiles_n - 1), first / (p.ag.row_bytes * BN)));
  }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < C::STAGES; ++i) {
        mbar_init(&full[i], 1);
        mbar_init(&empty[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull[i], 1);
        mbar_init(&tempty[i], 8);
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<C::TMEM_COLS>(tmem_holder);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        TileCoord tc = decode_tile(tile, tiles_m, tiles_n, nkb, p.splits);
        if (AG && p.ag.enabled) {
          // rotate the column-block order so this rank's own rows come first, then wait for the slices
          // (one counter per owner rank) that cover rows [n0, n0 + BN) of the weight
          tc.n_blk = (tc.n_blk + ag_rot) % tiles_n;
          const long long lo = p.ag.b_off_bytes + static_cast<long long>(tc.n_blk) * BN * p.ag.row_bytes;
          long long hi = p.ag.b_off_bytes + static_cast<long long>(min(p.N, (tc.n_blk + 1) * BN)) * p.ag.row_bytes;
          const int q_lo = static_cast<int>(lo / p.ag.per_bytes);
          const int q_hi = static_cast<int>((hi - 1) / p.ag.per_bytes);
          for (int q = q_lo; q <= q_hi && q < p.ag.W; ++q) ag_wait(p.ag.arrived + q, p.ag.target + (p.ag.target_base ? *p.ag.target_base : 0u));
          asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], A_STAGE + C::B_STAGE);
          uint8_t* a_dst = sA + stage * A_STAGE;
          uint8_t* b_dst = sB + stage * C::B_STAGE;
          if (!p.a_mn) {
            tma_load_2d(a_dst, &tmA, &full[stage], kb * epr, tc.m_blk * BM);
          } else if (p.a_mn3) {
            tma_load_3d(a_dst, &tmA, &full[stage], 0, kb * epr, tc.m_blk * (BM / 64));
          } else {
            for (int j = 0; j < BM / epr; ++j)
              tma_load_2d(a_dst + j * epr * ROW_BYTES, &tmA, &full[stage], tc.m_blk * BM + j * epr, kb * epr);
          }
          if (!p.b_mn) {
            tma_load_2d(b_dst, &tmB, &full[stage], kb * epr, tc.n_blk * BN);
          } else if (p.b_mn3) {
            tma_load_3d(b_dst, &tmB, &full[stage], 0, kb * epr, tc.n_blk * (BN / 64));
          } else {
            for (int j = 0; j < BN / epr; ++j)
              tma_load_2d(b_dst + j * epr * ROW_BYTES, &tmB, &full[stage], tc.n_blk * BN + j * epr, kb * epr);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // --------------------------------------------------------- MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      // byte advance of the descriptor start address per UMMA_K step
      const uint32_t a_kstep = p.a_mn ? (p.esize == 2 ? 16u : 8u) * ROW_BYTES : 32u;
      const uint32_t b_kstep = p.b_mn ? (p.esize == 2 ? 16u : 8u) * ROW_BYTES : 32u;
      const uint32_t a_lbo = p.a_mn ? epr * ROW_BYTES : 16u;
      const uint32_t b_lbo = p.b_mn ? epr * ROW_BYTES : 16u;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const TileCoord tc = decode_tile(tile, tiles_m, tiles_n, nkb, p.splits);  // n_blk unused here
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * A_STAGE);
          const uint32_t b_base = smem_u32(sB + stage * C::B_STAGE);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t ad = make_smem_desc_sw128(a_base + k * a_kstep, a_lbo, 1024);
            const uint64_t bd = make_smem_desc_sw128(b_base + k * b_kstep, b_lbo, 1024);
            const uint32_t accum = (kb > tc.kb0 || k > 0) ? 1u : 0u;
            if (p.esize == 2) umma_f16(d_tmem, ad, bd, p.idesc, accum);
            else umma_tf32(d_tmem, ad, bd, p.idesc, accum);
          }
          umma_commit(&empty[stage]);
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (AG && warp >= NUM_THREADS / 32) {
    // ----------------------------------------------------------- all-gather warps (K-C3)
    if (p.ag.enabled) {
      const AgParams& g = p.ag;
      const int tid = threadIdx.x - NUM_THREADS;
      const unsigned ep = g.epoch + (g.epoch_base ? *g.epoch_base : 0u);
      ag_rank_barrier(g, tid, ep);  // every rank's shard holds this step's parameters
      const long long nch = (g.per_bytes + AG_CHUNK - 1) / AG_CHUNK;
      for (int pi = 0; pi < g.W; ++pi) {
        const int q = (g.rank + pi) % g.W;
        const char* src = g.src[q];
        char* dst = g.dst + static_cast<long long>(q) * g.per_bytes;
        for (long long c = blockIdx.x; c < nch; c += gridDim.x) {
          const long long b0 = c * AG_CHUNK;
          const long long nvec = (min(static_cast<long long>(AG_CHUNK), g.per_bytes - b0)) / 16;
          long long i = tid;
          for (; i + 3 * AG_THREADS < nvec; i += 4 * AG_THREADS) {
            const uint4 a0 = ld_volatile_v4(src + b0 + i * 16), a1 = ld_volatile_v4(src + b0 + (i + AG_THREADS) * 16),
                        a2 = ld_volatile_v4(src + b0 + (i + 2 * AG_THREADS) * 16), a3 = ld_volatile_v4(src + b0 + (i + 3 * AG_THREADS) * 16);
            *reinterpret_cast<uint4*>(dst + b0 + i * 16) = a0;
            *reinterpret_cast<uint4*>(dst + b0 + (i + AG_THREADS) * 16) = a1;
            *reinterpret_cast<uint4*>(dst + b0 + (i + 2 * AG_THREADS) * 16) = a2;
            *reinterpret_cast<uint4*>(dst + b0 + (i + 3 * AG_THREADS) * 16) = a3;
          }
          for (; i < nvec; i += AG_THREADS) *reinterpret_cast<uint4*>(dst + b0 + i * 16) = ld_volatile_v4(src + b0 + i * 16);
          named_bar_sync(2, AG_THREADS);
          if (tid == 0) {
            __threadfence();
            atomicAdd(g.arrived + q, 1u);
          }
        }
      }
      ag_rank_barrier(g, tid, ep + 1);  // peers are done reading my shard
    }
  } else {
    // ----------------------------------------------------------- epilogue
    const int quarter = warp & 3;  // TMEM lanes [32q, 32q+32) are visible to this warp
    const int chalf = (warp - 2) >> 2;  // which half of the tile's column chunks this warp drains
    uint8_t* stg = stg_base + (warp - 2) * STG_WARP_BYTES;
    int acc = 0;
    uint32_t acc_phase = 0;
    if (AG && p.ag.enabled && p.bias && p.ag.bias_off_bytes >= 0) {
      // the bias read in this epilogue lives in the same unit: wait for its owner slice(s)
      const long long hi = p.ag.bias_off_bytes + static_cast<long long>(p.N) * (p.bias_dtype == DT_F32 ? 4 : 2) - 1;
      for (int q = static_cast<int>(p.ag.bias_off_bytes / p.ag.per_bytes); q <= static_cast<int>(hi / p.ag.per_bytes) && q < p.ag.W; ++q)
        ag_wait(p.ag.arrived + q, p.ag.target + (p.ag.target_base ? *p.ag.target_base : 0u));
    }
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
      TileCoord tc = decode_tile(tile, tiles_m, tiles_n, nkb, p.splits);
      if (AG && p.ag.enabled) tc.n_blk = (tc.n_blk + ag_rot) % tiles_n;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const int row = tc.m_blk * BM + quarter * 32 + lane;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      const bool has_k = tc.kb1 > tc.kb0;
#pragma unroll 1
      for (int c = chalf * (BN / 64); c < (chalf + 1) * (BN / 64); ++c) {
        const int col0 = tc.n_blk * BN + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        constexpr bool kStageable = (EPI == 0 || EPI == 1 || EPI == 2 || EPI == 5);
        const bool staged = kStageable && p.out_dtype != DT_F32 && has_k && col0 + 32 <= p.N;  // warp-uniform
        if (kStageable && staged) {
          uint4 pre[4];
          epi_prefetch<EPI>(p, pre, row - lane, lane, col0);  // global loads in flight while TMEM is read
          tmem_ld_32x32(t_row + c * 32, r);
          tmem_ld_wait();
          epilogue32s<EPI, ACT>(p, r, pre, row - lane, lane, col0, stg);
          continue;
        }
        tmem_ld_32x32(t_row + c * 32, r);
        tmem_ld_wait();
        if (row < p.M && has_k) {
          if (EPI != 3 && col0 + 32 <= p.N) {
            epilogue32<EPI, ACT>(p, r, row, col0);
          } else {
#pragma unroll
            for (int g = 0; g < 4; ++g)
              epilogue8(&p, __uint_as_float(r[g * 8]), __uint_as_float(r[g * 8 + 1]), __uint_as_float(r[g * 8 + 2]),
                        __uint_as_float(r[g * 8 + 3]), __uint_as_float(r[g * 8 + 4]), __uint_as_float(r[g * 8 + 5]),
                        __uint_as_float(r[g * 8 + 6]), __uint_as_float(r[g * 8 + 7]), row, col0 + g * 8);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------------------------
// 2-CTA variant (cta_group::2): a CTA pair on one TPC computes a 256 x 256 tile.  Each CTA stages its own
// 128 rows of A and HALF of the B tile (128 of the 256 columns); the leader's single thread issues
// tcgen05.mma.cta_group::2 (M = 256), which reads both halves of B from the two CTAs' shared memory, so per-SM
// operand traffic from L2 drops from 48 KB to 32 KB per k-block and the ring deepens to 6 stages.  Each CTA keeps
// its 128 accumulator rows in its own TMEM and runs its own epilogue.
constexpr int G2_BN = 256;
// warp 0: TMA producer for A, warp 1: MMA issuer, warps 2-9: epilogue, warp 10: TMA producer for B.  One thread
// issuing both loads of a stage was the limiter of the main loop (each cp.async.bulk.tensor costs the issuing thread
// a few hundred cycles: 2 per stage -> ~790 cycles per k-block against 540 of MMA time); two producers issue in parallel.
constexpr int G2_THREADS = 352;
// BN = 256: 16 KB A + 16 KB B-half per stage, 6 stages.  BN = 128 (small-M problems: twice the tiles to fill the pairs):
// 16 KB + 8 KB, 8 stages.
template <int BN>
struct G2Cfg {
  static constexpr int STAGES = (BN == 256) ? 6 : 8;
  static constexpr int STAGE_BYTES = A_STAGE + (BN / 2) * ROW_BYTES;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 256 + STG_BYTES + 1024;
};

template <int EPI, int ACT, int BN = G2_BN>
__global__ void __launch_bounds__(G2_THREADS, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __grid_constant__ Params p) {
  constexpr int G2_STAGES = G2Cfg<BN>::STAGES;
  constexpr int G2_STAGE_BYTES = G2Cfg<BN>::STAGE_BYTES;
  constexpr int BH = BN / 2;          // B columns staged by one CTA
  constexpr int BM2 = 2 * BM;         // tile rows of the pair
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + G2_STAGES * A_STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + G2_STAGES * BH * ROW_BYTES);
  uint64_t* empty = full + G2_STAGES;
  uint64_t* tfull = empty + G2_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty + 2);
  uint8_t* stg_base = reinterpret_cast<uint8_t*>(full) + 256;  // 8 x STG_WARP_BYTES, one stage per epilogue warp

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // cluster = one CTA pair
  const uint32_t rank = cluster_ctarank();
  const uint32_t cta = rank & 1u;
  constexpr int csize = 2;
  const bool leader = cta == 0;
  const int epr = ROW_BYTES / p.esize;
  const int nkb = (p.K + epr - 1) / epr;
  const int tiles_m = (p.M + BM2 - 1) / BM2;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int total = tiles_m * tiles_n * p.splits;
  const int cluster_id = blockIdx.x / csize, num_clusters = gridDim.x / csize;
  const uint32_t lead_rank = 0;                                            // the pair's leader
  const uint16_t all_mask = 3, pair_mask = 3;                              // both CTAs of the pair

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < G2_STAGES; ++i) {
        mbar_init(&full[i], 1);    // the leader's arrive.expect_tx covers the bytes of every TMA load into the pair
        mbar_init(&empty[i], 1);   // commit from the leader's MMA thread
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull[i], 1);
        mbar_init(&tempty[i], 16); // 8 epilogue warps of each CTA (only the leader's copy is waited on)
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc_2sm<512>(tmem_holder);
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0 || warp == 10) {
    // ------------------------------------------------------- TMA producers: warp 0 loads A, warp 10 loads B
    if (lane == 0 && !(p.one_producer && warp == 10)) {
      const bool load_a = (warp == 0);
      const bool load_b = (warp == 10) || p.one_producer;
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < total; tile += num_clusters) {
        const TileCoord tc = decode_tile(tile, tiles_m, tiles_n, nkb, p.splits);
        const int m0 = tc.m_blk * BM2 + cta * BM;
        const int n0 = tc.n_blk * BN + cta * BH;
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          const uint32_t lead_bar = mapa_u32(smem_u32(&full[stage]), lead_rank);
          if (load_a) {
            // the peer only contributes transaction bytes (no arrival: a remote release-arrive per k-block costs a
            // cluster-scope fence in the single producer thread and halves the throughput); the A producer of the
            // leader announces the bytes of all loads that land in the pair during this stage
            if (leader) mbar_expect_tx(&full[stage], 2 * G2_STAGE_BYTES);
            uint8_t* a_dst = sA + stage * A_STAGE;
            if (!p.a_mn) {
              tma_load_2d_2sm(a_dst, &tmA, lead_bar, kb * epr, m0);
            } else if (p.a_mn3) {
              tma_load_3d_2sm(a_dst, &tmA, lead_bar, 0, kb * epr, m0 >> 6);
            } else {
              for (int j = 0; j < BM / epr; ++j) tma_load_2d_2sm(a_dst + j * epr * ROW_BYTES, &tmA, lead_bar, m0 + j * epr, kb * epr);
            }
          }
          if (load_b) {
            uint8_t* b_dst = sB + stage * BH * ROW_BYTES;
            if (!p.b_mn) {
              tma_load_2d_2sm(b_dst, &tmB, lead_bar, kb * epr, n0);
            } else if (p.b_mn3) {
              tma_load_3d_2sm(b_dst, &tmB, lead_bar, 0, kb * epr, n0 >> 6);
            } else {
              for (int j = 0; j < BH / epr; ++j) tma_load_2d_2sm(b_dst + j * epr * ROW_BYTES, &tmB, lead_bar, n0 + j * epr, kb * epr);
            }
          }
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
        for (int kb2 = 0; kb2 < p.k2_blocks; ++kb2) {   // dual-K tail: second operand pair, K-major
          mbar_wait(&empty[stage], phase ^ 1);
          const uint32_t lead_bar = mapa_u32(smem_u32(&full[stage]), lead_rank);
          if (load_a) {
            if (leader) mbar_expect_tx(&full[stage], 2 * G2_STAGE_BYTES);
            tma_load_2d_2sm(sA + stage * A_STAGE, &p.tmA2, lead_bar, kb2 * epr, m0);
          }
          if (load_b) tma_load_2d_2sm(sB + stage * BH * ROW_BYTES, &p.tmB2, lead_bar, kb2 * epr, n0);
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // --------------------------------------------------------- MMA issuer (leader CTA only)
    if (lane == 0 && leader) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t a_kstep = p.a_mn_mma ? 16u * ROW_BYTES : 32u;
      const uint32_t b_kstep = p.b_mn_mma ? 16u * ROW_BYTES : 32u;
      const uint32_t a_lbo = p.a_mn_mma ? epr * ROW_BYTES : 16u;
      const uint32_t b_lbo = p.b_mn_mma ? epr * ROW_BYTES : 16u;
      for (int tile = cluster_id; tile < total; tile += num_clusters) {
        const TileCoord tc = decode_tile(tile, tiles_m, tiles_n, nkb, p.splits);
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * A_STAGE);
          const uint32_t b_base = smem_u32(sB + stage * BH * ROW_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t ad = make_smem_desc_sw128(a_base + k * a_kstep, a_lbo, 1024);
            const uint64_t bd = make_smem_desc_sw128(b_base + k * b_kstep, b_lbo, 1024);
            umma_f16_2sm(d_tmem, ad, bd, p.idesc, (kb > tc.kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit_2sm(&empty[stage], all_mask);
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
        for (int kb2 = 0; kb2 < p.k2_blocks; ++kb2) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * A_STAGE);
          const uint32_t b_base = smem_u32(sB + stage * BH * ROW_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_2sm(d_tmem, make_smem_desc_sw128(a_base + k * 32, 16, 1024), make_smem_desc_sw128(b_base + k * 32, 16, 1024),
                         p.idesc2, (tc.kb1 > tc.kb0 || kb2 > 0 || k > 0) ? 1u : 0u);
          umma_commit_2sm(&empty[stage], all_mask);
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&tfull[acc], pair_mask);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ----------------------------------------------------------- epilogue (both CTAs, own 128 rows)
    const int quarter = warp & 3;
    const int chalf = (warp - 2) >> 2;
    uint8_t* stg = stg_base + (warp - 2) * STG_WARP_BYTES;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = cluster_id; tile < total; tile += num_clusters) {
      const TileCoord tc = decode_tile(tile, tiles_m, tiles_n, nkb, p.splits);
      const int row = tc.m_blk * BM2 + cta * BM + quarter * 32 + lane;
      const bool has_k = tc.kb1 > tc.kb0;
      constexpr bool kStageable = (EPI == 0 || EPI == 1 || EPI == 2 || EPI == 5);
      constexpr int NCH = BN / 64;   // 32-column chunks per warp
      const int c_begin = chalf * NCH;
      // whole-tile fast path: every chunk of this warp is inside N and goes through the staged epilogue
      const bool staged_tile = kStageable && p.out_dtype != DT_F32 && has_k && tc.n_blk * BN + (c_begin + NCH) * 32 <= p.N;
      uint4 pre[NCH][4];
      if (kStageable && staged_tile) {
        // the tile this epilogue reads back (residual / act' input / beta*D) is requested for ALL chunks before
        // waiting for the accumulator, so the loads fly while the tile's MMAs finish (ncu: with a per-chunk prefetch the
        // epilogue sat in long-scoreboard stalls on these loads and throttled the main loop)
#pragma unroll
        for (int k = 0; k < NCH; ++k) epi_prefetch<EPI>(p, pre[k], row - lane, lane, tc.n_blk * BN + (c_begin + k) * 32);
      }
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      if (kStageable && staged_tile) {
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
          uint32_t r[32];
          tmem_ld_32x32(t_row + (c_begin + k) * 32, r);
          tmem_ld_wait();
          if (k == NCH - 1) {
            // accumulator fully read: hand the TMEM buffer back before this chunk's math and stores
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster_relaxed(mapa_u32(smem_u32(&tempty[acc]), lead_rank));
          }
          epilogue32s<EPI, ACT>(p, r, pre[k], row - lane, lane, tc.n_blk * BN + (c_begin + k) * 32, stg);
        }
      } else {
#pragma unroll 1
        for (int c = c_begin; c < c_begin + NCH; ++c) {
          const int col0 = tc.n_blk * BN + c * 32;
          if (col0 >= p.N) break;
          uint32_t r[32];
          tmem_ld_32x32(t_row + c * 32, r);
          tmem_ld_wait();
          if (row < p.M && has_k) {
            if (EPI != 3 && col0 + 32 <= p.N) {
              epilogue32<EPI, ACT>(p, r, row, col0);
            } else {
#pragma unroll
              for (int g = 0; g < 4; ++g)
                epilogue8(&p, __uint_as_float(r[g * 8]), __uint_as_float(r[g * 8 + 1]), __uint_as_float(r[g * 8 + 2]),
                          __uint_as_float(r[g * 8 + 3]), __uint_as_float(r[g * 8 + 4]), __uint_as_float(r[g * 8 + 5]),
                          __uint_as_float(r[g * 8 + 6]), __uint_as_float(r[g * 8 + 7]), row, col0 + g * 8);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster_relaxed(mapa_u32(smem_u32(&tempty[acc]), lead_rank));  // the leader's MMA thread waits on its copy
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_2sm<512>(tmem_base);
}

// ------------------------------------------------------------------ host side
static EncodeTiledFn get_encode() { return tmap_encode_fn(); }

// 2D map over a row-major [outer, inner] matrix with leading dimension ld (elements).
int make_tmap_2d(CUtensorMap* tm, const void* ptr, int dtype, long long inner, long long outer, long long ld,
                 int box_inner, int box_outer) {
  EncodeTiledFn enc = get_encode();
  HYP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
  // The encoder is a driver-API call: make sure the primary context is bound to THIS thread first (a fresh autograd
  // worker thread has only had cudaSetDevice, which binds lazily) -- otherwise it fails with CUDA_ERROR_INVALID_CONTEXT.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { cudaFree(nullptr); ctx_bound = true; }
  const int es = (dtype == DT_F32) ? 4 : 2;
  const CUtensorMapDataType dt = (dtype == DT_F32)   ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : (dtype == DT_BF16) ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                      : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  HYP_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand pointer must be 16-byte aligned");
  HYP_REQUIRE((ld * es) % 16 == 0, "TMA operand leading dimension must be a multiple of 16 bytes (ld=%lld)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * es};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HYP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return 0;
}

// Rank-3 map over an MN-major 16-bit operand stored [K, MN] (leading dimension ld): (64 MN elements, K, MN / 64) with
// strides (1, ld, 64), so a box (64, 64, chunks) lands in shared memory as `chunks` consecutive [64 k][64 mn] swizzled
// blocks -- the layout the MN-major UMMA descriptor walks -- with ONE TMA instruction.  Issuing one 64 x 64 box per
// chunk instead costs 15-40% of the GEMM's throughput (the producer thread becomes the bottleneck).  With a ragged MN the
// last chunk runs past the row end, so the caller must guarantee ld >= roundup(MN, 64).
int make_tmap_mn3(CUtensorMap* tm, const void* ptr, int dtype, long long mn, long long k, long long ld, int chunks) {
  const unsigned long long dims[3] = {64ull, static_cast<unsigned long long>(k), static_cast<unsigned long long>((mn + 63) / 64)};
  const unsigned long long strides[2] = {static_cast<unsigned long long>(ld) * 2ull, 128ull};
  const unsigned box[3] = {64u, 64u, static_cast<unsigned>(chunks)};
  return make_tmap_nd(tm, ptr, dtype, 3, dims, strides, box);
}

template <int BN, int EPI, int ACT, bool AG = false>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, cudaStream_t stream) {
  using C = Cfg<BN>;
  static_assert(C::SMEM_BYTES <= kMaxSmemPerCta, "GEMM stage ring + staging exceeds the 227 KB shared memory of an sm_100a CTA");
  static_assert(tmem_alloc_ok(C::TMEM_COLS), "TMEM allocations are powers of two between 32 and 512 columns");
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(gemm_kernel<BN, EPI, ACT, AG>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int tiles = ((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN) * p.splits;
  int grid = tiles < num_sms() ? tiles : num_sms();
  if (AG) grid = num_sms();  // the same grid on every rank: gather warps pair up with their peer blocks
  gemm_kernel<BN, EPI, ACT, AG><<<grid, NUM_THREADS + (AG ? AG_THREADS : 0), C::SMEM_BYTES, stream>>>(ta, tb, p);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int BN>
static int dispatch_ag(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int epi, cudaStream_t st) {
  if (epi == 0) return launch<BN, 0, 0, true>(ta, tb, p, st);
  if (epi == 1) {
    switch (p.act) {
      case 0: return launch<BN, 1, 0, true>(ta, tb, p, st);
      case 1: return launch<BN, 1, 1, true>(ta, tb, p, st);
      case 2: return launch<BN, 1, 2, true>(ta, tb, p, st);
      case 3: return launch<BN, 1, 3, true>(ta, tb, p, st);
      default: return launch<BN, 1, 4, true>(ta, tb, p, st);
    }
  }
  set_error("hyp_gemm_ag: the all-gather fusion supports the forward epilogues only (plain / bias / activation / residual)");
  return -1;
}

template <int EPI, int ACT, int BN = G2_BN>
static int launch2(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, cudaStream_t stream) {
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  auto kernel = gemm2_kernel<EPI, ACT, BN>;
  static_assert(G2Cfg<BN>::SMEM <= kMaxSmemPerCta, "pair-GEMM stage ring + staging exceeds the 227 KB shared memory of an sm_100a CTA");
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G2Cfg<BN>::SMEM));
    attr_set[dev & 63] = true;
  }
  const int csize = 2;
  const int tiles_m = (p.M + 2 * BM - 1) / (2 * BM);
  const int tiles = tiles_m * ((p.N + BN - 1) / BN) * p.splits;
  int clusters = num_sms() / 2;
  if (tiles < clusters) clusters = tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(csize * clusters);
  cfg.blockDim = dim3(G2_THREADS);
  cfg.dynamicSmemBytes = G2Cfg<BN>::SMEM;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  HYP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kernel, ta, tb, p));
  return 0;
}

// the 256 x 128 pair tile is instantiated for the epilogues small-M problems use (plain, bias/residual, beta)
static bool pair128_supported(int epi, int act) { return epi == 0 || (epi == 1 && act == 0) || epi == 5; }
static int dispatch2_n128(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int epi, cudaStream_t st) {
  if (epi == 0) return launch2<0, 0, 128>(ta, tb, p, st);
  if (epi == 1) return launch2<1, 0, 128>(ta, tb, p, st);
  return launch2<5, 0, 128>(ta, tb, p, st);
}

static int dispatch2(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int epi, cudaStream_t st) {
  if (epi == 0) return launch2<0, 0>(ta, tb, p, st);
  if (epi == 1) {
    switch (p.act) {
      case 0: return launch2<1, 0>(ta, tb, p, st);
      case 1: return launch2<1, 1>(ta, tb, p, st);
      case 2: return launch2<1, 2>(ta, tb, p, st);
      case 3: return launch2<1, 3>(ta, tb, p, st);
      default: return launch2<1, 4>(ta, tb, p, st);
    }
  }
  if (epi == 2) {
    switch (p.act) {
      case 1: return launch2<2, 1>(ta, tb, p, st);
      case 2: return launch2<2, 2>(ta, tb, p, st);
      case 3: return launch2<2, 3>(ta, tb, p, st);
      case 4: return launch2<2, 4>(ta, tb, p, st);
      default: break;
    }
  }
  if (epi == 4) return launch2<4, 0>(ta, tb, p, st);
  if (epi == 5) return launch2<5, 0>(ta, tb, p, st);
  return launch2<3, 0>(ta, tb, p, st);
}

// epilogue kind -> kernel instantiation
template <int BN>
static int dispatch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int epi, cudaStream_t st) {
  if (epi == 0) return launch<BN, 0, 0>(ta, tb, p, st);
  if (epi == 1) {
    switch (p.act) {
      case 0: return launch<BN, 1, 0>(ta, tb, p, st);
      case 1: return launch<BN, 1, 1>(ta, tb, p, st);
      case 2: return launch<BN, 1, 2>(ta, tb, p, st);
      case 3: return launch<BN, 1, 3>(ta, tb, p, st);
      default: return launch<BN, 1, 4>(ta, tb, p, st);
    }
  }
  if (epi == 2) {
    switch (p.act) {
      case 1: return launch<BN, 2, 1>(ta, tb, p, st);
      case 2: return launch<BN, 2, 2>(ta, tb, p, st);
      case 3: return launch<BN, 2, 3>(ta, tb, p, st);
      case 4: return launch<BN, 2, 4>(ta, tb, p, st);
      default: break;
    }
  }
  if (epi == 4) return launch<BN, 4, 0>(ta, tb, p, st);
  if (epi == 5) return launch<BN, 5, 0>(ta, tb, p, st);
  return launch<BN, 3, 0>(ta, tb, p, st);
}

}  // namespace gemm
}  // namespace hyp

using namespace hyp;

// See include/hyperion_b200.h for the contract.
struct DualK { const void* A2; const void* B2; int K2; long long lda2, ldb2; };

static int gemm_impl(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                     long long ldd, int in_dtype, int out_dtype, int a_mn_major, int b_mn_major, const void* bias,
                     int bias_dtype, const void* aux, long long ld_aux, int aux_dtype, void* preact_out, int act, int mode,
                     float alpha, float beta, int splits, const gemm::AgParams* ag, void* stream_,
                     const DualK* dual = nullptr) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(M > 0 && N > 0 && K > 0, "hyp_gemm: empty problem %dx%dx%d", M, N, K);
  HYP_REQUIRE(in_dtype >= 0 && in_dtype <= 2 && out_dtype >= 0 && out_dtype <= 2, "hyp_gemm: bad dtype");
  if (splits < 1) splits = 1;
  HYP_REQUIRE(splits == 1 || (out_dtype == DT_F32 && !bias && !aux && !preact_out && act == 0 && beta == 0.f),
              "hyp_gemm: split-K needs a zero-initialised fp32 output and a plain epilogue");
  // tf32 operands must be K-major (the MN-major path needs the 32-bit-base swizzle, not built yet);
  // callers transpose fp32 operands explicitly (ops.native.gemm does).
  HYP_REQUIRE(in_dtype != DT_F32 || (!a_mn_major && !b_mn_major), "hyp_gemm: fp32/tf32 operands must be K-major");
  const int es = (in_dtype == DT_F32) ? 4 : 2;
  const int epr = gemm::ROW_BYTES / es;
  // 2-CTA (cta_group::2) path: 256 x 256 tiles on CTA pairs; 16-bit operands, large-enough problems
  static int use_2cta = -1;
  if (use_2cta < 0) {
    const char* e = getenv("HYPERION_GEMM_2CTA");
    use_2cta = e ? atoi(e) : 1;
  }
  // Tile shape.  Candidates: 256x256 and 256x128 on a CTA pair, 128x256 and 128x128 on one CTA.  Cost model = waves x relative time
  // of one tile (a 128x128 tile measures ~1.15x a 256-wide one: half the MMA work, but it is shared-memory-bandwidth bound;
  // the 1-CTA 128x256 tile runs ~1.35x slower than a pair's 256x256 per SM).  Small-M problems (Llama-7B at 512 tokens:
  // 32 pair tiles for 74 pairs) fill the machine only with the small tile.
  static int tile_force = -1;   // HYPERION_GEMM_TILE: 0 auto, 1 pair 256x256, 2 one-CTA 128x256, 3 one-CTA 128x128, 4 pair 256x128
  if (tile_force < 0) {
    const char* e = getenv("HYPERION_GEMM_TILE");
    tile_force = e ? atoi(e) : 0;
  }
  const bool pair_ok = use_2cta && !ag && es == 2 && N > 128 && M >= 256;
  int BN = (N > 128) ? 256 : 128;   // final choice below, once the epilogue kind is known

  gemm::Params p;
  p.M = M; p.N = N; p.K = K; p.esize = es; p.a_mn = a_mn_major; p.b_mn = b_mn_major;
  const uint32_t fmt = (in_dtype == DT_F16) ? 0u : (in_dtype == DT_BF16 ? 1u : 2u);
  p.idesc = make_idesc(fmt, gemm::BM, BN, a_mn_major != 0, b_mn_major != 0);
  const int nkb = (K + epr - 1) / epr;
  if (splits > nkb) splits = nkb;
  const int per_split = (nkb + splits - 1) / splits;
  p.splits = (nkb + per_split - 1) / per_split;  // no split is ever empty
  p.D = D; p.ldd = ldd; p.out_dtype = out_dtype;
  p.bias = bias; p.bias_dtype = bias_dtype;
  p.aux = aux; p.ld_aux = ld_aux; p.aux_dtype = aux_dtype;
  // mode 1 (D = alpha*acc * act'(aux)) has no pre-activation output; there the same argument carries the optional fp32
  // column-sum accumulator (see Params::colsum and the header)
  float* colsum = nullptr;
  if (mode == 1 && preact_out) { colsum = reinterpret_cast<float*>(preact_out); preact_out = nullptr; }
  p.preact_out = preact_out;
  p.colsum = colsum;
  HYP_REQUIRE(!preact_out || reinterpret_cast<uintptr_t>(preact_out) % 16 == 0, "hyp_gemm: preact_out must be 16B aligned");
  p.act = act; p.mode = mode; p.alpha = alpha; p.beta = beta;
  const int oes = (out_dtype == DT_F32) ? 4 : 2;
  bool vec = (reinterpret_cast<uintptr_t>(D) % 16 == 0) && ((ldd * oes) % 16 == 0);
  if (aux) {
    const int aes = (aux_dtype == DT_F32) ? 4 : 2;
    vec = vec && (reinterpret_cast<uintptr_t>(aux) % 16 == 0) && ((ld_aux * aes) % 16 == 0);
  }
  p.vec_ok = vec ? 1 : 0;
  memset(&p.ag, 0, sizeof(p.ag));
  if (ag) p.ag = *ag;
  // pick the compile-time epilogue: 0 plain, 1 bias/act/residual/preact, 2 act', 3 general (split-K, beta, unaligned)
  const bool fast = vec && p.splits == 1 && beta == 0.f && (!bias || reinterpret_cast<uintptr_t>(bias) % 16 == 0) &&
                    (!aux || aux_dtype != DT_F32) && (!preact_out || out_dtype != DT_F32);
  int epi = 3;
  if (fast && mode == 0 && act == 0 && !aux && !preact_out) epi = 0;
  else if (fast && mode == 0) epi = 1;
  else if (fast && mode == 1 && aux && act != 0) epi = 2;
  else if (vec && p.splits > 1) epi = 4;
  else if (vec && p.splits == 1 && beta != 0.f && mode == 0 && act == 0 && !bias && !aux && !preact_out) epi = 5;
  p.simple = epi;
  HYP_REQUIRE(!colsum || (epi == 2 && out_dtype != DT_F32 && N % 32 == 0),
              "hyp_gemm: the fused column sum needs the act' epilogue on aligned 16-bit tensors with N %% 32 == 0 (N=%d)", N);

  bool want_pair = pair_ok;
  bool pair128 = false;
  if (N > 128) {
    const int sms = num_sms();
    const long long sp = p.splits;
    auto waves = [](long long tiles, int slots) { return static_cast<double>((tiles + slots - 1) / slots); };
    const long long tm2 = (M + 255) / 256, tm1 = (M + 127) / 128;
    const double c_pair = pair_ok ? waves(tm2 * ((N + 255) / 256) * sp, sms / 2) : 1e30;
    const double c_p128 = (pair_ok && gemm::pair128_supported(epi, act)) ? 0.80 * waves(tm2 * ((N + 127) / 128) * sp, sms / 2) : 1e30;
    const double c_256 = 1.35 * waves(tm1 * ((N + 255) / 256) * sp, sms);
    const double c_128 = 1.15 * waves(tm1 * ((N + 127) / 128) * sp, sms);
    int pick = 1;
    double best = c_pair;
    if (c_p128 < best - 1e-9) { best = c_p128; pick = 4; }
    if (c_256 < best - 1e-9) { best = c_256; pick = 2; }
    if (c_128 < best - 1e-9) { best = c_128; pick = 3; }
    if (ag && (pick == 1 || pick == 4)) pick = 2;
    if (tile_force >= 1 && tile_force <= 4) {
      const bool ok = (tile_force == 1 && pair_ok) || (tile_force == 4 && pair_ok && gemm::pair128_supported(epi, act)) ||
                      tile_force == 2 || tile_force == 3;
      if (ok) pick = tile_force;
    }
    want_pair = (pick == 1 || pick == 4);
    pair128 = (pick == 4);
    BN = (pick == 3 || pick == 4) ? 128 : 256;
  }
  p.k2_blocks = 0;
  p.idesc2 = 0;
  static const int one_prod = [] { const char* e = getenv("HYPERION_GEMM_PRODUCERS"); return (e && atoi(e) == 1) ? 1 : 0; }();
  p.one_producer = one_prod;
  if (dual) {
    HYP_REQUIRE(pair_ok && p.splits == 1 && dual->K2 > 0 && dual->K2 % 64 == 0,
                "hyp_gemm_dualk: needs 16-bit operands, M >= 256, N > 128, no split-K and K2 %% 64 == 0 (M=%d N=%d K2=%d)", M, N, dual->K2);
    want_pair = true; pair128 = false; BN = 256;
  }
  p.idesc = make_idesc(fmt, gemm::BM, BN, a_mn_major != 0, b_mn_major != 0);

  CUtensorMap ta, tb;
  int rc;
  p.a_mn_mma = p.a_mn; p.b_mn_mma = p.b_mn;
  static const bool mn3_off = [] { const char* e = getenv("HYPERION_GEMM_MN3"); return e && e[0] == '0'; }();
  // ragged MN is fine when the row padding covers the last chunk's over-read (what it reads there only reaches output
  // rows / columns that are never stored)
  p.a_mn3 = (a_mn_major && es == 2 && (M % 64 == 0 || lda >= (M + 63) / 64 * 64) && !mn3_off) ? 1 : 0;
  p.b_mn3 = (b_mn_major && es == 2 && (N % 64 == 0 || ldb >= (N + 63) / 64 * 64) && !mn3_off) ? 1 : 0;
  if (want_pair) {
    static int dbg = -1;  // timing experiments only (results are wrong): 1 = TMA loads K-major boxes, 2 = MMA assumes K-major
    if (dbg < 0) { const char* e = getenv("HYPERION_GEMM_DBG"); dbg = e ? atoi(e) : 0; }
    if (dbg & 1) { a_mn_major = 0; b_mn_major = 0; p.a_mn = 0; p.b_mn = 0; }
    if (dbg & 2) { p.a_mn_mma = 0; p.b_mn_mma = 0; }
    p.idesc = make_idesc(fmt, 2 * gemm::BM, BN, p.a_mn_mma != 0, p.b_mn_mma != 0);
    if (!a_mn_major)   rc = gemm::make_tmap_2d(&ta, A, in_dtype, K, M, lda, epr, gemm::BM);
    else if (p.a_mn3)  rc = gemm::make_tmap_mn3(&ta, A, in_dtype, M, K, lda, gemm::BM / 64);
    else               rc = gemm::make_tmap_2d(&ta, A, in_dtype, M, K, lda, epr, epr);
    if (rc) return rc;
    if (!b_mn_major)   rc = gemm::make_tmap_2d(&tb, B, in_dtype, K, N, ldb, epr, BN / 2);
    else if (p.b_mn3)  rc = gemm::make_tmap_mn3(&tb, B, in_dtype, N, K, ldb, BN / 128);
    else               rc = gemm::make_tmap_2d(&tb, B, in_dtype, N, K, ldb, epr, epr);
    if (rc) return rc;
    if (dual) {
      p.k2_blocks = dual->K2 / 64;
      p.idesc2 = make_idesc(fmt, 2 * gemm::BM, BN, false, false);
      rc = gemm::make_tmap_2d(&p.tmA2, dual->A2, in_dtype, dual->K2, M, dual->lda2, epr, gemm::BM);
      if (rc) return rc;
      rc = gemm::make_tmap_2d(&p.tmB2, dual->B2, in_dtype, dual->K2, N, dual->ldb2, epr, BN / 2);
      if (rc) return rc;
    }
    if (pair128) return gemm::dispatch2_n128(ta, tb, p, epi, stream);
    return gemm::dispatch2(ta, tb, p, epi, stream);
  }
  if (!a_mn_major)   rc = gemm::make_tmap_2d(&ta, A, in_dtype, K, M, lda, epr, gemm::BM);   // A: [M, K]
  else if (p.a_mn3)  rc = gemm::make_tmap_mn3(&ta, A, in_dtype, M, K, lda, gemm::BM / 64);
  else               rc = gemm::make_tmap_2d(&ta, A, in_dtype, M, K, lda, epr, epr);        // A^T: [K, M]
  if (rc) return rc;
  if (!b_mn_major)   rc = gemm::make_tmap_2d(&tb, B, in_dtype, K, N, ldb, epr, BN);         // B: [N, K]
  else if (p.b_mn3)  rc = gemm::make_tmap_mn3(&tb, B, in_dtype, N, K, ldb, BN / 64);
  else               rc = gemm::make_tmap_2d(&tb, B, in_dtype, N, K, ldb, epr, epr);        // B^T: [K, N]
  if (rc) return rc;
  if (ag) {
    HYP_REQUIRE(!b_mn_major && splits <= 1, "hyp_gemm_ag: B must be the [N, K] weight layout and split-K is not supported");
    return (BN == 256) ? gemm::dispatch_ag<256>(ta, tb, p, epi, stream) : gemm::dispatch_ag<128>(ta, tb, p, epi, stream);
  }
  return (BN == 256) ? gemm::dispatch<256>(ta, tb, p, epi, stream) : gemm::dispatch<128>(ta, tb, p, epi, stream);
}

HYP_API int hyp_gemm(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                     long long ldd, int in_dtype, int out_dtype, int a_mn_major, int b_mn_major, const void* bias,
                     int bias_dtype, const void* aux, long long ld_aux, int aux_dtype, void* preact_out, int act, int mode,
                     float alpha, float beta, int splits, void* stream_) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, in_dtype, out_dtype, a_mn_major, b_mn_major, bias, bias_dtype, aux,
                   ld_aux, aux_dtype, preact_out, act, mode, alpha, beta, splits, nullptr, stream_);
}

// K-C3: D = epilogue(A . B^T) where B (and the optional bias) live in a flat FSDP unit buffer `unit_dst` of W equal
// slices; slice q is pulled from peer_src[q] (peer q's shard, mapped through NVLink) by gather warps inside the same
// kernel, and the GEMM consumes tiles as their weights land.  peer_src / peer_flags: HOST arrays of W device addresses.
// `arrived`: device uint32[W], zero at creation, never reset; `target` = launches so far (incl. this one) * number of
// 32 KiB chunks per slice.  Every rank must launch this with the same shapes and epoch (uses epoch, epoch + 1).
HYP_API int hyp_gemm_ag(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                        long long ldd, int in_dtype, int out_dtype, const void* bias, int bias_dtype, const void* aux,
                        long long ld_aux, int aux_dtype, void* preact_out, int act, float alpha,
                        const long long* peer_src, const long long* peer_flags, int rank, int W, unsigned epoch,
                        const unsigned* epoch_base, void* unit_dst, long long per_bytes, unsigned* arrived,
                        unsigned target, const unsigned* target_base, void* stream_) {
  HYP_REQUIRE(W >= 1 && W <= 8, "hyp_gemm_ag: world size %d not in [1, 8]", W);
  HYP_REQUIRE(per_bytes > 0 && per_bytes % 16 == 0, "hyp_gemm_ag: slice size must be a positive multiple of 16 bytes");
  gemm::AgParams g;
  memset(&g, 0, sizeof(g));
  g.enabled = 1; g.W = W; g.rank = rank; g.epoch = epoch; g.epoch_base = epoch_base; g.target_base = target_base;
  for (int i = 0; i < W; ++i) {
    g.src[i] = reinterpret_cast<const char*>(peer_src[i]);
    g.peer_flags[i] = reinterpret_cast<uint32_t*>(peer_flags[i]);
  }
  g.dst = static_cast<char*>(unit_dst);
  g.per_bytes = per_bytes;
  const long long unit_bytes = per_bytes * W;
  const long long b_off = static_cast<const char*>(B) - static_cast<const char*>(unit_dst);
  HYP_REQUIRE(b_off >= 0 && b_off < unit_bytes, "hyp_gemm_ag: B does not live inside the unit buffer");
  g.b_off_bytes = b_off;
  g.row_bytes = ldb * (in_dtype == DT_F32 ? 4 : 2);
  g.bias_off_bytes = -1;
  if (bias) {
    const long long o = static_cast<const char*>(bias) - static_cast<const char*>(unit_dst);
    if (o >= 0 && o < unit_bytes) g.bias_off_bytes = o;
  }
  g.arrived = arrived; g.target = target;
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, in_dtype, out_dtype, 0, 0, bias, bias_dtype, aux, ld_aux, aux_dtype,
                   preact_out, act, 0, alpha, 0.f, 1, &g, stream_);
}

// D = epilogue(alpha * (A B^T + A2 B2^T)): one GEMM whose reduction runs over two operand pairs (the second K-major,
// K2 a multiple of 64 -- pad with zero columns).  The LoRA projection y = x W^T + t B_l^T (t = s * x A_l^T) is this
// with (A, B) = (x, W) and (A2, B2) = (t, B_l): the rank-r update lands in the accumulator tile in TMEM instead of a
// second read-modify-write pass over y.  Pair kernel only (16-bit, M >= 256, N > 128); bias / residual epilogue.
HYP_API int hyp_gemm_dualk(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                           long long ldd, int in_dtype, int out_dtype, int a_mn_major, int b_mn_major, const void* A2,
                           const void* B2, int K2, long long lda2, long long ldb2, const void* bias, int bias_dtype,
                           const void* aux, long long ld_aux, int aux_dtype, float alpha, void* stream_) {
  DualK d{A2, B2, K2, lda2, ldb2};
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, in_dtype, out_dtype, a_mn_major, b_mn_major, bias, bias_dtype, aux,
                   ld_aux, aux_dtype, nullptr, 0, 0, alpha, 0.f, 1, nullptr, stream_, &d);
}
