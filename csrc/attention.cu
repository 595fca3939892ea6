// flash_attn_sm100: scaled-dot-product attention on the 5th-gen tensor cores.
//
//   O = softmax(Q K^T * scale + mask) V        per (batch, head), never materialising the S x S scores
//
// One CTA owns a 128-row query block of one (batch, head) and streams 128-key blocks:
//   warp 0     TMA producer: Q once, then a 2-stage ring of K/V tiles (4-D tensor maps over [D, S, H, B] with
//              arbitrary strides, so the packed [B, S, 3, H, D] projection output is consumed in place)
//   warp 1     one thread issues tcgen05.mma:  S = Q K^T  (128x128xD, accumulator in TMEM columns [0,128))
//                                              O_blk = P V (128xDx128, accumulator in TMEM columns [128,128+D))
//   warps 2-5  softmax: thread r owns query row r; two passes over the S row in TMEM (row max, then exp2 / row sum),
//              P is rounded to 16 bits and written to shared memory in the 128B-swizzled K-major layout the PV
//              MMA reads; the running output is rescaled and accumulated in registers (online softmax)
// S_{j+1} = Q K_{j+1}^T is issued right behind P_j V_j, so it overlaps the softmax threads' accumulation, and two
// CTAs fit per SM at D = 64 so one CTA's softmax overlaps the other's MMAs.
// Masks: none / causal / causal + key padding (bool [B, Sk]); keys beyond Sk are masked, rows with no visible key
// produce zeros.  The per-row log-sum-exp is saved for the backward.
// Dropout (template DROP): the keep mask of P is drawn from Philox4x32-10 keyed by a device-side (seed, step) state and
// addressed by (batch * H + head, query row, key column, per-call salt) -- common/philox.cuh; nothing is stored, the
// backward kernels regenerate the same bits.
//
// Reference call sites: nn.MultiheadAttention inside nn.TransformerEncoderLayer (02_development/
// distributed_utils.py:79; the reference ran the *math* SDPA path that materialises the scores)
// and HF Llama SDPA with causal + padding mask (distributed_utils.py:465).
#include "common/philox.cuh"
#include "common/ptx.cuh"
#include "common/tmap.cuh"
#include "common/util.cuh"

namespace hyp {
namespace attn {

constexpr int BQ = 128, BKV = 128;
constexpr int THREADS = 192;
constexpr float LOG2E = 1.4426950408889634f;

struct FwdParams {
  int B, H, Sq, Sk, D;
  int causal;
  float scale_log2;          // softmax scale * log2(e)
  const uint8_t* kpm;        // [B, Sk] 1 = attend, or null
  void* o;                   // output, element strides below
  long long o_sb, o_sh, o_ss;
  float* lse;                // [B, H, Sq] natural-log LSE, or null
  int dtype;
  DropoutArgs drop;          // dropout on the attention probabilities (state == nullptr: none)
};

template <int D>
struct Smem {
  static constexpr int TILE = 128 * D * 2;        // one [128 x D] 16-bit tile
  static constexpr int P_BYTES = 128 * 128 * 2;
  static constexpr int Q_OFF = 0;
  static constexpr int K_OFF = TILE;               // 2 stages
  static constexpr int V_OFF = 3 * TILE;           // 2 stages
  static constexpr int P_OFF = 5 * TILE;
  static constexpr int BAR_OFF = P_OFF + P_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256;  // no alignment slack: the dynamic window is declared 1024-aligned
};

// Softmax passes over one 32-column chunk of a score row.  Masking is a compare + select against the row's key limit
// (and the key-padding bytes) -- no per-element branches; the unmasked instantiation has no mask arithmetic at all.
template <bool MASK, bool KPM>
__device__ __forceinline__ float fwd_chunk_max(const uint32_t (&v)[32], int kcol0, int k_limit, const uint8_t* kpm_chunk) {
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    float x = __uint_as_float(v[i]);
    if constexpr (MASK) {
      bool ok = (kcol0 + i) < k_limit;
      if constexpr (KPM) ok = ok & (kpm_chunk[i] != 0);
      x = ok ? x : -INFINITY;
    }
    mx = fmaxf(mx, x);
  }
  return mx;
}
template <bool MASK, bool KPM>
__device__ __forceinline__ float fwd_chunk_exp(const uint32_t (&v)[32], float (&pr)[32], float scale_log2, float neg_m,
                                               int kcol0, int k_limit, const uint8_t* kpm_chunk) {
  // scale and shift on packed fp32 pairs (one FFMA2 per two scores), exponentials on the MUFU, pairwise partial sums
  const float2 sc = splat2(scale_log2), nm = splat2(neg_m);
  float2 l2 = make_float2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    float2 x = ffma2(make_float2(__uint_as_float(v[i]), __uint_as_float(v[i + 1])), sc, nm);
    if constexpr (MASK) {
      bool ok0 = (kcol0 + i) < k_limit, ok1 = (kcol0 + i + 1) < k_limit;
      if constexpr (KPM) { ok0 = ok0 & (kpm_chunk[i] != 0); ok1 = ok1 & (kpm_chunk[i + 1] != 0); }
      x.x = ok0 ? x.x : -INFINITY;
      x.y = ok1 ? x.y : -INFINITY;
    }
    const float2 e = make_float2(ex2_fast(x.x), ex2_fast(x.y));
    pr[i] = e.x;
    pr[i + 1] = e.y;
    l2 = fadd2(l2, e);
  }
  return l2.x + l2.y;
}

template <typename T, int D, bool DROP>
__global__ void __launch_bounds__(THREADS, (D == 64) ? 2 : 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ FwdParams p) {
  using L = Smem<D>;
  extern __shared__ __align__(1024) uint8_t smem_fwd[];
  uint8_t* smem = smem_fwd;
  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u)) __trap();  // SWIZZLE_128B tiles need a 1024-byte aligned base
  uint8_t* sQ = smem + L::Q_OFF;
  uint8_t* sK = smem + L::K_OFF;
  uint8_t* sV = smem + L::V_OFF;
  uint8_t* sP = smem + L::P_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* q_full = bars;          // 1
  uint64_t* kv_full = bars + 1;     // 2
  uint64_t* kv_empty = bars + 3;    // 2
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 8);
  uint8_t* s_kpm = reinterpret_cast<uint8_t*>(bars + 10);  // 128 bytes of key-padding mask for the current block

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq_blk = (p.Sq + BQ - 1) / BQ;
  // 1-D grid, heaviest first: all (batch, head) pairs of the last query block (the most key blocks under a causal
  // mask), then the next one, ... so the final wave is made of the shortest CTAs
  const int bh_count = p.H * p.B;
  const int q_blk = nq_blk - 1 - static_cast<int>(blockIdx.x / bh_count);
  const int bh_idx = static_cast<int>(blockIdx.x % bh_count);
  const int h = bh_idx % p.H, b = bh_idx / p.H;
  const int q0 = q_blk * BQ;
  const int shift = p.Sk - p.Sq;  // causal alignment: query i sees keys <= i + shift
  int kv_end = p.Sk;
  if (p.causal) kv_end = min(p.Sk, q0 + BQ + shift);
  const int nblk = max(0, (kv_end + BKV - 1) / BKV);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < 2; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
      mbar_init(s_full, 1);
      mbar_init(p_full, 128);
      mbar_init(o_full, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<256>(tmem_holder);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tS = tmem_base, tO = tmem_base + 128;

  if (warp == 0) {
    if (lane == 0 && nblk > 0) {
      mbar_expect_tx(q_full, L::TILE);
      for (int c = 0; c < D / 64; ++c) tma_load_4d(sQ + c * 16384, &tmQ, q_full, c * 64, q0, h, b);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&kv_full[st], 2 * L::TILE);
        for (int c = 0; c < D / 64; ++c) {
          tma_load_4d(sK + st * L::TILE + c * 16384, &tmK, &kv_full[st], c * 64, j * BKV, h, b);
          tma_load_4d(sV + st * L::TILE + c * 16384, &tmV, &kv_full[st], c * 64, j * BKV, h, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && nblk > 0) {
      const uint32_t fmt = (p.dtype == DT_F16) ? 0u : 1u;
      const uint32_t idesc_s = make_idesc(fmt, 128, 128, false, false);
      const uint32_t idesc_o = make_idesc(fmt, 128, D, false, true);  // B = V is N(=D)-major
      const uint32_t q_base = smem_u32(sQ), p_base = smem_u32(sP);
      auto issue_s = [&](int j) {
        const uint32_t k_base = smem_u32(sK + (j & 1) * L::TILE);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (k >> 2) * 16384 + (k & 3) * 32;
          umma_f16(tS, make_smem_desc_sw128(q_base + off, 16, 1024), make_smem_desc_sw128(k_base + off, 16, 1024),
                   idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(s_full);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        if (j + 1 < nblk) mbar_wait(&kv_full[(j + 1) & 1], ((j + 1) >> 1) & 1);
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        const uint32_t v_base = smem_u32(sV + st * L::TILE);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          const uint64_t ad = make_smem_desc_sw128(p_base + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024);
          const uint64_t bd = make_smem_desc_sw128(v_base + k * 2048, 16384, 1024);
          umma_f16(tO, ad, bd, idesc_o, k > 0 ? 1u : 0u);
        }
        umma_commit(o_full);
        umma_commit(&kv_empty[st]);
        if (j + 1 < nblk) issue_s(j + 1);
      }
    }
  } else {
    // ------------------------------------------------------------ softmax + output
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;         // row inside the query block == TMEM lane
    const int tid128 = (warp - 2) * 32 + lane; // 0..127 among the softmax threads
    const int qi = q0 + r;
    const int k_limit = p.causal ? min(p.Sk, qi + shift + 1) : p.Sk;  // first key this row may NOT see
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    float o_acc[D];
#pragma unroll
    for (int i = 0; i < D; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    uint2 dkey = make_uint2(0u, 0u);
    if constexpr (DROP) dkey = dropout_key(p.drop.state);
    for (int j = 0; j < nblk; ++j) {
      const int k0 = j * BKV;
      const bool need_mask = (k0 + BKV > p.Sk) || (p.causal && (k0 + BKV - 1 > q0 + shift)) || (p.kpm != nullptr);
      if (p.kpm) {
        named_bar_sync(1, 128);  // previous block's readers are done with s_kpm
        const int kk = k0 + tid128;
        s_kpm[tid128] = (kk < p.Sk) ? p.kpm[static_cast<long long>(b) * p.Sk + kk] : 0;
        named_bar_sync(1, 128);
      }
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      // pass 1: row maximum of the raw scores (scale > 0, so it commutes with max)
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tS + lane_off + c * 32, v);
        tmem_ld_wait();
        if (!need_mask) mx = fmaxf(mx, fwd_chunk_max<false, false>(v, 0, 0, nullptr));
        else if (p.kpm) mx = fmaxf(mx, fwd_chunk_max<true, true>(v, k0 + c * 32, k_limit, s_kpm + c * 32));
        else mx = fmaxf(mx, fwd_chunk_max<true, false>(v, k0 + c * 32, k_limit, nullptr));
      }
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = ex2_fast(m_run - m_use);  // m_run = -inf -> 0
      // pass 2: probabilities -> shared memory (K-major, 128B swizzle), row sum
      float l_blk = 0.f;
      const float neg_m = -m_use;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tS + lane_off + c * 32, v);
        tmem_ld_wait();
        float pr[32];
        if (!need_mask) l_blk += fwd_chunk_exp<false, false>(v, pr, p.scale_log2, neg_m, 0, 0, nullptr);
        else if (p.kpm) l_blk += fwd_chunk_exp<true, true>(v, pr, p.scale_log2, neg_m, k0 + c * 32, k_limit, s_kpm + c * 32);
        else l_blk += fwd_chunk_exp<true, false>(v, pr, p.scale_log2, neg_m, k0 + c * 32, k_limit, nullptr);
        if constexpr (DROP) {
          // the row sum above is of the un-dropped probabilities (softmax normalisation); dropped entries contribute
          // nothing to O, survivors are rescaled once at the end through inv_l
          const uint32_t km = dropout_keep32(dkey, p.drop.salt, static_cast<uint32_t>(k0 + c * 32), static_cast<uint32_t>(qi),
                                             static_cast<uint32_t>(b * p.H + h), p.drop.thresh);
#pragma unroll
          for (int i = 0; i < 32; ++i) pr[i] = ((km >> i) & 1u) ? pr[i] : 0.f;
        }
        uint8_t* prow = sP + (c >> 1) * 16384 + r * 128;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = pr[t * 8 + i];
          const int cc = (c & 1) * 4 + t;
          *reinterpret_cast<uint4*>(prow + ((cc ^ (r & 7)) << 4)) = pack8<T>(f);
        }
      }
      l_run = l_run * alpha + l_blk;
      m_run = m_new;
      tc_fence_before();          // our tcgen05.ld of S are complete before the MMA warp may overwrite S
      fence_proxy_async_smem();   // P (generic-proxy stores) visible to the tensor core (async proxy)
      mbar_arrive(p_full);
      // accumulate O_blk
      mbar_wait(o_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tO + lane_off + c * 32, v);
        tmem_ld_wait();
        const float2 al2 = splat2(alpha);
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float2 o2 = ffma2(make_float2(o_acc[c * 32 + i], o_acc[c * 32 + i + 1]), al2,
                                  make_float2(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));
          o_acc[c * 32 + i] = o2.x;
          o_acc[c * 32 + i + 1] = o2.y;
        }
      }
      tc_fence_before();
    }
    if (qi < p.Sq) {
      const float inv_l = (l_run > 0.f ? 1.f / l_run : 0.f) * (DROP ? p.drop.scale : 1.f);
      T* orow = reinterpret_cast<T*>(p.o) + b * p.o_sb + h * p.o_sh + static_cast<long long>(qi) * p.o_ss;
#pragma unroll
      for (int g = 0; g < D / 8; ++g) {
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = o_acc[g * 8 + i] * inv_l;
        *reinterpret_cast<uint4*>(orow + g * 8) = pack8<T>(f);
      }
      if (p.lse) {
        const float lse = (l_run > 0.f) ? (m_run + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
        p.lse[(static_cast<long long>(b) * p.H + h) * p.Sq + qi] = lse;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
}

// 4-D map over a [B, H, S, D]-indexed tensor with element strides (sb, sh, ss), unit stride in D.
static int make_qkv_map(CUtensorMap* tm, const void* ptr, int dtype, int B, int H, int S, int D, long long sb, long long sh,
                        long long ss) {
  const unsigned long long dims[4] = {(unsigned long long)D, (unsigned long long)S, (unsigned long long)H, (unsigned long long)B};
  const unsigned long long strides[3] = {(unsigned long long)ss * 2, (unsigned long long)sh * 2, (unsigned long long)sb * 2};
  const unsigned box[4] = {64, 128, 1, 1};
  return make_tmap_nd(tm, ptr, dtype, 4, dims, strides, box);
}

template <typename T, int D, bool DROP>
static int launch_fwd_t(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const FwdParams& p, cudaStream_t st) {
  using L = Smem<D>;
  // D = 64 runs two CTAs per SM (each allocates 256 of the 512 TMEM columns): both shared-memory windows, plus the 1 KB the
  // runtime reserves per CTA, must fit the 228 KB of an SM
  static_assert(L::TOTAL <= kMaxSmemPerCta && (D != 64 || 2 * (L::TOTAL + 1024) <= 228 * 1024), "flash forward shared-memory layout");
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(flash_fwd_kernel<T, D, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set[dev & 63] = true;
  }
  dim3 grid(static_cast<unsigned>((p.Sq + BQ - 1) / BQ) * p.H * p.B);
  flash_fwd_kernel<T, D, DROP><<<grid, THREADS, L::TOTAL, st>>>(tq, tk, tv, p);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
// the dropout instantiation is a separate kernel, so the plain one carries none of its registers
template <typename T, int D>
static int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const FwdParams& p, cudaStream_t st) {
  return p.drop.state ? launch_fwd_t<T, D, true>(tq, tk, tv, p, st) : launch_fwd_t<T, D, false>(tq, tk, tv, p, st);
}

}  // namespace attn
}  // namespace hyp

using namespace hyp;

// q/k/v/o are indexed [b, h, s, d] through element strides (unit stride in d); see include/hyperion_b200.h.
namespace hyp { namespace attn {
// p in (0, 1) -> 8-bit threshold and survivor scale (see common/philox.cuh); p == 0 or a null state -> no dropout
static int make_dropout(DropoutArgs* d, const unsigned long long* state, unsigned salt, float p, const char* who) {
  d->state = nullptr; d->salt = salt; d->thresh = 0; d->scale = 1.f;
  if (state == nullptr || p <= 0.f) return 0;
  HYP_REQUIRE(p < 1.f, "%s: dropout probability must be in [0, 1) (got %f)", who, p);
  int t = static_cast<int>(p * 256.f + 0.5f);
  t = t < 1 ? 1 : (t > 255 ? 255 : t);
  d->state = state; d->thresh = static_cast<unsigned>(t); d->scale = 256.f / static_cast<float>(256 - t);
  return 0;
}
}}  // namespace hyp::attn

static int flash_attn_fwd_impl(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Sk,
                               int D, long long q_sb, long long q_sh, long long q_ss, long long k_sb, long long k_sh,
                               long long k_ss, long long v_sb, long long v_sh, long long v_ss, long long o_sb,
                               long long o_sh, long long o_ss, const unsigned char* key_padding_mask, int causal,
                               float scale, int dtype, const unsigned long long* rng_state, unsigned salt, float dropout_p,
                               void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_F16 || dtype == DT_BF16, "hyp_flash_attn_fwd: 16-bit dtypes only");
  HYP_REQUIRE(D == 64 || D == 128, "hyp_flash_attn_fwd: head dim must be 64 or 128 (got %d)", D);
  HYP_REQUIRE(B > 0 && H > 0 && Sq > 0 && Sk > 0, "hyp_flash_attn_fwd: empty problem");
  HYP_REQUIRE(reinterpret_cast<uintptr_t>(o) % 16 == 0 && (o_sb % 8 | o_sh % 8 | o_ss % 8) == 0,
              "hyp_flash_attn_fwd: output rows must be 16-byte aligned");
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = attn::make_qkv_map(&tq, q, dtype, B, H, Sq, D, q_sb, q_sh, q_ss))) return rc;
  if ((rc = attn::make_qkv_map(&tk, k, dtype, B, H, Sk, D, k_sb, k_sh, k_ss))) return rc;
  if ((rc = attn::make_qkv_map(&tv, v, dtype, B, H, Sk, D, v_sb, v_sh, v_ss))) return rc;
  attn::FwdParams p;
  p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.D = D; p.causal = causal;
  p.scale_log2 = scale * attn::LOG2E;
  p.kpm = key_padding_mask;
  p.o = o; p.o_sb = o_sb; p.o_sh = o_sh; p.o_ss = o_ss; p.lse = lse; p.dtype = dtype;
  if ((rc = attn::make_dropout(&p.drop, rng_state, salt, dropout_p, "hyp_flash_attn_fwd_drop"))) return rc;
  if (dtype == DT_BF16) return D == 64 ? attn::launch_fwd<__nv_bfloat16, 64>(tq, tk, tv, p, st) : attn::launch_fwd<__nv_bfloat16, 128>(tq, tk, tv, p, st);
  return D == 64 ? attn::launch_fwd<__half, 64>(tq, tk, tv, p, st) : attn::launch_fwd<__half, 128>(tq, tk, tv, p, st);
}

HYP_API int hyp_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Sk,
                               int D, long long q_sb, long long q_sh, long long q_ss, long long k_sb, long long k_sh,
                               long long k_ss, long long v_sb, long long v_sh, long long v_ss, long long o_sb,
                               long long o_sh, long long o_ss, const unsigned char* key_padding_mask, int causal,
                               float scale, int dtype, void* stream_) {
  return flash_attn_fwd_impl(q, k, v, o, lse, B, H, Sq, Sk, D, q_sb, q_sh, q_ss, k_sb, k_sh, k_ss, v_sb, v_sh, v_ss, o_sb, o_sh,
                             o_ss, key_padding_mask, causal, scale, dtype, nullptr, 0u, 0.f, stream_);
}

// Same, with dropout on the attention probabilities: rng_state -> device [seed, step] (see common/philox.cuh), salt
// identifies this call within the step; the backward must be given the same three values.
HYP_API int hyp_flash_attn_fwd_drop(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq,
                                    int Sk, int D, long long q_sb, long long q_sh, long long q_ss, long long k_sb,
                                    long long k_sh, long long k_ss, long long v_sb, long long v_sh, long long v_ss,
                                    long long o_sb, long long o_sh, long long o_ss, const unsigned char* key_padding_mask,
                                    int causal, float scale, int dtype, const unsigned long long* rng_state, int salt,
                                    float dropout_p, void* stream_) {
  return flash_attn_fwd_impl(q, k, v, o, lse, B, H, Sq, Sk, D, q_sb, q_sh, q_ss, k_sb, k_sh, k_ss, v_sb, v_sh, v_ss, o_sb, o_sh,
                             o_ss, key_padding_mask, causal, scale, dtype, rng_state, static_cast<unsigned>(salt), dropout_p,
                             stream_);
}

// =====================================================================================================
// Backward.  One CTA owns a 128-key block of one (batch, head), keeps K_j / V_j in shared memory and dK_j / dV_j in
// tensor memory, and streams the query blocks that can see it:
//     S  = Q_i K_j^T            dP = dO_i V_j^T                     (two 128x128xD MMAs into TMEM)
//     P  = exp(S*scale - lse)   dS = P o (dP - delta) * scale        (softmax threads; P, dS -> smem, 16-bit)
//     dV += P^T dO_i            dK += dS^T Q_i                       (A operand = the same smem tile read MN-major)
//     dQ_i = dS K_j             -> fp32 red.global.add into the dQ accumulator
// delta = rowsum(dO o O) comes from a small pre-kernel.  TMEM columns: [0,128) S then dQ_i, [128,256) dP,
// [256,256+D) dV, [256+D,256+2D) dK.
namespace hyp {
namespace attn {

struct BwdParams {
  int B, H, Sq, Sk, D;
  int causal;
  float scale, scale_log2;
  const uint8_t* kpm;
  const float* lse;    // [B,H,Sq]
  const float* delta;  // [B,H,Sq]
  float* dq_acc;       // [B,H,Sq,D] fp32, zero-initialised
  void* dk; void* dv;  // element strides below (unit stride in D)
  long long dk_sb, dk_sh, dk_ss, dv_sb, dv_sh, dv_ss;
  int dtype;
  DropoutArgs drop;    // the forward's dropout (state, salt, thresh): the mask is regenerated, never stored
};

template <int D>
struct BwdSmem {
  static constexpr int TILE = 128 * D * 2;
  static constexpr int NST = (D == 64) ? 2 : 1;
  static constexpr int K_OFF = 0, V_OFF = TILE;
  static constexpr int Q_OFF = 2 * TILE;                 // NST stages
  static constexpr int DO_OFF = Q_OFF + NST * TILE;      // NST stages
  static constexpr int P_OFF = DO_OFF + NST * TILE;
  static constexpr int DS_OFF = P_OFF + 32768;
  static constexpr int BAR_OFF = DS_OFF + 32768;
  static constexpr int TOTAL = BAR_OFF + 384 + 1024;
};

template <typename T, int D, bool DROP>
__global__ void __launch_bounds__(THREADS, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                 const __grid_constant__ BwdParams p) {
  using L = BwdSmem<D>;
  constexpr int NST = L::NST;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem + L::K_OFF;
  uint8_t* sV = smem + L::V_OFF;
  uint8_t* sQ = smem + L::Q_OFF;
  uint8_t* sDO = smem + L::DO_OFF;
  uint8_t* sP = smem + L::P_OFF;
  uint8_t* sDS = smem + L::DS_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* kv_full = bars;            // 1
  uint64_t* qdo_full = bars + 1;       // 2
  uint64_t* qdo_empty = bars + 3;      // 2
  uint64_t* sdp_full = bars + 5;
  uint64_t* pds_full = bars + 6;       // 128 arrivals
  uint64_t* dq_full = bars + 7;
  uint64_t* dq_read = bars + 8;        // 128 arrivals
  uint64_t* done = bars + 9;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 10);
  uint8_t* s_kpm = reinterpret_cast<uint8_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid: key block 0 (the most query blocks under a causal mask) of every (batch, head) first
  const int bh_count = p.H * p.B;
  const int j = static_cast<int>(blockIdx.x / bh_count);
  const int h = static_cast<int>(blockIdx.x % bh_count) % p.H, b = static_cast<int>(blockIdx.x % bh_count) / p.H;
  const int k0 = j * BKV;
  const int shift = p.Sk - p.Sq;
  const int nq_blk = (p.Sq + BQ - 1) / BQ;
  int i_start = 0;
  if (p.causal) i_start = max(0, (k0 - shift) / BQ);  // first query block with a row that sees key k0
  const int niter = max(0, nq_blk - i_start);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(kv_full, 1);
      for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
      mbar_init(sdp_full, 1);
      mbar_init(pds_full, 128);
      mbar_init(dq_full, 1);
      mbar_init(dq_read, 128);
      mbar_init(done, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<512>(tmem_holder);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDV = tmem_base + 256, tDK = tmem_base + 256 + D;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * L::TILE);
      for (int c = 0; c < D / 64; ++c) {
        tma_load_4d(sK + c * 16384, &tmK, kv_full, c * 64, k0, h, b);
        tma_load_4d(sV + c * 16384, &tmV, kv_full, c * 64, k0, h, b);
      }
      for (int it = 0; it < niter; ++it) {
        const int st = it % NST;
        mbar_wait(&qdo_empty[st], ((it / NST) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[st], 2 * L::TILE);
        const int q0 = (i_start + it) * BQ;
        for (int c = 0; c < D / 64; ++c) {
          tma_load_4d(sQ + st * L::TILE + c * 16384, &tmQ, &qdo_full[st], c * 64, q0, h, b);
          tma_load_4d(sDO + st * L::TILE + c * 16384, &tmDO, &qdo_full[st], c * 64, q0, h, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t fmt = (p.dtype == DT_F16) ? 0u : 1u;
      const uint32_t idesc_s = make_idesc(fmt, 128, 128, false, false);
      const uint32_t idesc_t = make_idesc(fmt, 128, D, true, true);    // dV, dK: A = P^T / dS^T (MN-major), B MN-major
      const uint32_t idesc_q = make_idesc(fmt, 128, D, false, true);   // dQ: A = dS (K-major), B = K (MN-major)
      const uint32_t k_base = smem_u32(sK), v_base = smem_u32(sV), p_base = smem_u32(sP), ds_base = smem_u32(sDS);
      mbar_wait(kv_full, 0);
      for (int it = 0; it < niter; ++it) {
        const int st = it % NST;
        const uint32_t q_base = smem_u32(sQ + st * L::TILE), do_base = smem_u32(sDO + st * L::TILE);
        mbar_wait(&qdo_full[st], (it / NST) & 1);
        if (it > 0) mbar_wait(dq_read, (it - 1) & 1);  // dQ_{i-1} has been drained from the columns S reuses
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (k >> 2) * 16384 + (k & 3) * 32;
          umma_f16(tS, make_smem_desc_sw128(q_base + off, 16, 1024), make_smem_desc_sw128(k_base + off, 16, 1024),
                   idesc_s, k > 0 ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (k >> 2) * 16384 + (k & 3) * 32;
          umma_f16(tDP, make_smem_desc_sw128(do_base + off, 16, 1024), make_smem_desc_sw128(v_base + off, 16, 1024),
                   idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(sdp_full);
        mbar_wait(pds_full, it & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BQ / 16; ++k) {  // reduction over the 128 query rows
          umma_f16(tDV, make_smem_desc_sw128(p_base + k * 2048, 16384, 1024),
                   make_smem_desc_sw128(do_base + k * 2048, 16384, 1024), idesc_t, (it > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < BQ / 16; ++k) {
          umma_f16(tDK, make_smem_desc_sw128(ds_base + k * 2048, 16384, 1024),
                   make_smem_desc_sw128(q_base + k * 2048, 16384, 1024), idesc_t, (it > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {  // reduction over the 128 keys
          umma_f16(tS, make_smem_desc_sw128(ds_base + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                   make_smem_desc_sw128(k_base + k * 2048, 16384, 1024), idesc_q, k > 0 ? 1u : 0u);
        }
        umma_commit(dq_full);
        umma_commit(&qdo_empty[st]);
      }
      if (niter > 0) umma_commit(done);
    }
  } else {
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const int tid128 = (warp - 2) * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    if (p.kpm) {
      const int kk = k0 + tid128;
      s_kpm[tid128] = (kk < p.Sk) ? p.kpm[static_cast<long long>(b) * p.Sk + kk] : 0;
      named_bar_sync(1, 128);
    }
    const long long bh = static_cast<long long>(b) * p.H + h;
    uint2 dkey = make_uint2(0u, 0u);
    if constexpr (DROP) dkey = dropout_key(p.drop.state);
    for (int it = 0; it < niter; ++it) {
      const int q0 = (i_start + it) * BQ;
      const int qi = q0 + r;
      const bool row_ok = qi < p.Sq;
      float lse2 = 0.f, delta = 0.f;
      if (row_ok) {
        lse2 = p.lse[bh * p.Sq + qi] * LOG2E;
        delta = p.delta[bh * p.Sq + qi];
      }
      const bool dead = !row_ok || lse2 == -INFINITY;
      const bool need_mask = (k0 + BKV > p.Sk) || (p.causal && (k0 + BKV - 1 > q0 + shift)) || (p.kpm != nullptr);
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t vs[32], vd[32];
        tmem_ld_32x32(tS + lane_off + c * 32, vs);
        tmem_ld_32x32(tDP + lane_off + c * 32, vd);
        tmem_ld_wait();
        float pr[32], ds[32];
        uint32_t km = 0xFFFFFFFFu;
        if constexpr (DROP)
          km = dropout_keep32(dkey, p.drop.salt, static_cast<uint32_t>(k0 + c * 32), static_cast<uint32_t>(qi),
                              static_cast<uint32_t>(bh), p.drop.thresh);
        const float ksc = DROP ? p.drop.scale : 1.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          bool ok = !dead;
          if (need_mask) {
            const int kj = k0 + c * 32 + i;
            ok = ok & (kj < p.Sk) & (!p.causal | (kj <= qi + shift));
            if (p.kpm) ok = ok & (s_kpm[c * 32 + i] != 0);
          }
          const float pv = ok ? ex2_fast(fmaf(__uint_as_float(vs[i]), p.scale_log2, -lse2)) : 0.f;
          // with dropout: dV takes the dropped/rescaled P, and dP = mask * scale * (dO V^T)
          if constexpr (DROP) {
            const float mk = ((km >> i) & 1u) ? ksc : 0.f;
            pr[i] = pv * mk;
            ds[i] = pv * (__uint_as_float(vd[i]) * mk - delta) * p.scale;
          } else {
            pr[i] = pv;
            ds[i] = pv * (__uint_as_float(vd[i]) - delta) * p.scale;
          }
        }
        uint8_t* prow = sP + (c >> 1) * 16384 + r * 128;
        uint8_t* drow = sDS + (c >> 1) * 16384 + r * 128;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          float f[8], g[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) { f[i] = pr[t * 8 + i]; g[i] = ds[t * 8 + i]; }
          const int cc = ((c & 1) * 4 + t) ^ (r & 7);
          *reinterpret_cast<uint4*>(prow + (cc << 4)) = pack8<T>(f);
          *reinterpret_cast<uint4*>(drow + (cc << 4)) = pack8<T>(g);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(pds_full);
      // dQ_i tile -> fp32 accumulator in global memory
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      float* dq_row = p.dq_acc + (bh * p.Sq + qi) * D;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tS + lane_off + c * 32, v);
        tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int g = 0; g < 8; ++g)
            atomicAdd(reinterpret_cast<float4*>(dq_row + c * 32) + g,
                      make_float4(__uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]), __uint_as_float(v[4 * g + 2]),
                                  __uint_as_float(v[4 * g + 3])));
        }
      }
      tc_fence_before();
      mbar_arrive(dq_read);
    }
    // dV_j, dK_j: thread r owns key row k0 + r
    if (niter > 0) mbar_wait(done, 0);
    tc_fence_after();
    const int kj = k0 + r;
    T* dv_row = reinterpret_cast<T*>(p.dv) + b * p.dv_sb + h * p.dv_sh + static_cast<long long>(kj) * p.dv_ss;
    T* dk_row = reinterpret_cast<T*>(p.dk) + b * p.dk_sb + h * p.dk_sh + static_cast<long long>(kj) * p.dk_ss;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(tDV + lane_off + c * 32, v);
      tmem_ld_wait();
      if (kj < p.Sk) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = niter > 0 ? __uint_as_float(v[g * 8 + i]) : 0.f;
          *reinterpret_cast<uint4*>(dv_row + c * 32 + g * 8) = pack8<T>(f);
        }
      }
      tmem_ld_32x32(tDK + lane_off + c * 32, v);
      tmem_ld_wait();
      if (kj < p.Sk) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = niter > 0 ? __uint_as_float(v[g * 8 + i]) : 0.f;
          *reinterpret_cast<uint4*>(dk_row + c * 32 + g * 8) = pack8<T>(f);
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ---------------------------------------------------------------------------------------------------------
// D = 64 backward, pipelined.  Same maths and tile shapes as flash_bwd_kernel, but at D = 64 tensor memory has room
// for a dQ tile of its own ([384,448)) and shared memory for two P/dS buffers, so nothing in an iteration has to wait
// for the previous one to drain:
//   MMA thread, iteration i:   wait P_i/dS_i -> issue S_{i+1}, dP_{i+1} -> issue dV += P_i^T dO_i, dK += dS_i^T Q_i
//                              -> (dQ columns drained?) issue dQ_i = dS_i K
//   softmax warps, iteration i: S_i, dP_i -> P_i, dS_i (buffer i & 1) -> signal -> drain dQ_{i-1} (finished long ago)
// so the tensor pipe works on iteration i's three gradient MMAs while the softmax warps are already on S_{i+1}.
// Eight softmax warps: two per TMEM lane quarter, each taking 64 of the 128 key columns of a query row; the row's
// lse / delta for the next iteration are fetched one iteration ahead; the mask arithmetic exists only in the
// instantiation used for blocks that need it (diagonal, ragged tail, key padding).
constexpr int BWD64_THREADS = 320;

struct Bwd64Smem {
  static constexpr int TILE = 128 * 64 * 2;
  static constexpr int K_OFF = 0, V_OFF = TILE;
  static constexpr int Q_OFF = 2 * TILE;              // 2 stages
  static constexpr int DO_OFF = 4 * TILE;             // 2 stages
  static constexpr int P_OFF = 6 * TILE;              // 2 buffers of 32 KiB
  static constexpr int DS_OFF = P_OFF + 2 * 32768;    // 2 buffers of 32 KiB
  static constexpr int BAR_OFF = DS_OFF + 2 * 32768;
  static constexpr int TOTAL = BAR_OFF + 384;         // the dynamic window is declared 1024-aligned
};

// One 32-column chunk of a query row: P and dS from S and dP.  lse2 is +inf for rows that must produce zeros.
template <typename T, bool MASK, bool DROP = false>
__device__ __forceinline__ void bwd_softmax_chunk(const uint32_t (&vs)[32], const uint32_t (&vd)[32], uint8_t* prow,
                                                  uint8_t* drow, int chunk_in_row, int r, float scale_log2, float scale,
                                                  float lse2, float delta_s, int kcol0, int k_limit,
                                                  const uint8_t* s_kpm_chunk, uint32_t keep = 0xFFFFFFFFu, float ksc = 1.f) {
  const float2 sc2 = splat2(scale_log2), nl2 = splat2(-lse2), s2 = splat2(scale), nd2 = splat2(-delta_s);
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    float f[8], g[8];
#pragma unroll
    for (int i = 0; i < 8; i += 2) {   // packed fp32 pairs: two scores per FFMA2 / FMUL2
      const int e = t * 8 + i;
      float2 x = ffma2(make_float2(__uint_as_float(vs[e]), __uint_as_float(vs[e + 1])), sc2, nl2);
      if constexpr (MASK) {
        bool ok0 = (kcol0 + e) < k_limit, ok1 = (kcol0 + e + 1) < k_limit;
        if (s_kpm_chunk) { ok0 = ok0 & (s_kpm_chunk[e] != 0); ok1 = ok1 & (s_kpm_chunk[e + 1] != 0); }
        x.x = ok0 ? x.x : -INFINITY;
        x.y = ok1 ? x.y : -INFINITY;
      }
      const float2 pv = make_float2(ex2_fast(x.x), ex2_fast(x.y));
      if constexpr (DROP) {
        // dV takes the dropped / rescaled P; dP = mask * keep_scale * (dO V^T)
        const float2 mk = make_float2(((keep >> e) & 1u) ? ksc : 0.f, ((keep >> (e + 1)) & 1u) ? ksc : 0.f);
        const float2 ds = fmul2(pv, ffma2(fmul2(make_float2(__uint_as_float(vd[e]), __uint_as_float(vd[e + 1])), mk), s2, nd2));
        const float2 pd = fmul2(pv, mk);
        f[i] = pd.x; f[i + 1] = pd.y;
        g[i] = ds.x; g[i + 1] = ds.y;
      } else {
        const float2 ds = fmul2(pv, ffma2(make_float2(__uint_as_float(vd[e]), __uint_as_float(vd[e + 1])), s2, nd2));
        f[i] = pv.x; f[i + 1] = pv.y;
        g[i] = ds.x; g[i + 1] = ds.y;
      }
    }
    const int cc = (chunk_in_row * 4 + t) ^ (r & 7);
    *reinterpret_cast<uint4*>(prow + (cc << 4)) = pack8<T>(f);
    *reinterpret_cast<uint4*>(drow + (cc << 4)) = pack8<T>(g);
  }
}

template <typename T, bool DROP>
__global__ void __launch_bounds__(BWD64_THREADS, 1)
flash_bwd64_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                   const __grid_constant__ BwdParams p) {
  using L = Bwd64Smem;
  constexpr int D = 64;
  extern __shared__ __align__(1024) uint8_t smem_bwd64[];
  uint8_t* smem = smem_bwd64;
  uint8_t* sK = smem + L::K_OFF;
  uint8_t* sV = smem + L::V_OFF;
  uint8_t* sQ = smem + L::Q_OFF;
  uint8_t* sDO = smem + L::DO_OFF;
  uint8_t* sP = smem + L::P_OFF;
  uint8_t* sDS = smem + L::DS_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* kv_full = bars;            // 1
  uint64_t* qdo_full = bars + 1;       // 2
  uint64_t* qdo_empty = bars + 3;      // 2
  uint64_t* sdp_full = bars + 5;
  uint64_t* pds_full = bars + 6;       // 256 arrivals
  uint64_t* dq_full = bars + 7;
  uint64_t* dq_read = bars + 8;        // 256 arrivals
  uint64_t* done = bars + 9;
  uint64_t* sdp_read = bars + 10;      // 256 arrivals: S_i / dP_i are in the softmax threads' registers
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 11);
  uint8_t* s_kpm = reinterpret_cast<uint8_t*>(bars + 13);   // 128 bytes

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid: key block 0 (the most query blocks under a causal mask) of every (batch, head) first
  const int bh_count = p.H * p.B;
  const int j = static_cast<int>(blockIdx.x / bh_count);
  const int h = static_cast<int>(blockIdx.x % bh_count) % p.H, b = static_cast<int>(blockIdx.x % bh_count) / p.H;
  const int k0 = j * BKV;
  const int shift = p.Sk - p.Sq;
  const int nq_blk = (p.Sq + BQ - 1) / BQ;
  int i_start = 0;
  if (p.causal) i_start = max(0, (k0 - shift) / BQ);
  const int niter = max(0, nq_blk - i_start);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(kv_full, 1);
      for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
      mbar_init(sdp_full, 1);
      mbar_init(pds_full, 256);
      mbar_init(dq_full, 1);
      mbar_init(dq_read, 256);
      mbar_init(done, 1);
      mbar_init(sdp_read, 256);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<512>(tmem_holder);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDV = tmem_base + 256, tDK = tmem_base + 320,
                 tDQ = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * L::TILE);
      tma_load_4d(sK, &tmK, kv_full, 0, k0, h, b);
      tma_load_4d(sV, &tmV, kv_full, 0, k0, h, b);
      for (int it = 0; it < niter; ++it) {
        const int st = it & 1;
        mbar_wait(&qdo_empty[st], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[st], 2 * L::TILE);
        const int q0 = (i_start + it) * BQ;
        tma_load_4d(sQ + st * L::TILE, &tmQ, &qdo_full[st], 0, q0, h, b);
        tma_load_4d(sDO + st * L::TILE, &tmDO, &qdo_full[st], 0, q0, h, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && niter > 0) {
      const uint32_t fmt = (p.dtype == DT_F16) ? 0u : 1u;
      const uint32_t idesc_s = make_idesc(fmt, 128, 128, false, false);
      const uint32_t idesc_t = make_idesc(fmt, 128, D, true, true);    // dV, dK: A = P^T / dS^T (MN-major), B MN-major
      const uint32_t idesc_q = make_idesc(fmt, 128, D, false, true);   // dQ: A = dS (K-major), B = K (MN-major)
      const uint32_t k_base = smem_u32(sK), v_base = smem_u32(sV);
      auto issue_sdp = [&](int st) {
        const uint32_t q_base = smem_u32(sQ + st * L::TILE), do_base = smem_u32(sDO + st * L::TILE);
#pragma unroll
        for (int k = 0; k < D / 16; ++k)
          umma_f16(tS, make_smem_desc_sw128(q_base + k * 32, 16, 1024), make_smem_desc_sw128(k_base + k * 32, 16, 1024),
                   idesc_s, k > 0 ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < D / 16; ++k)
          umma_f16(tDP, make_smem_desc_sw128(do_base + k * 32, 16, 1024), make_smem_desc_sw128(v_base + k * 32, 16, 1024),
                   idesc_s, k > 0 ? 1u : 0u);
        umma_commit(sdp_full);
      };
      mbar_wait(kv_full, 0);
      mbar_wait(&qdo_full[0], 0);
      tc_fence_after();
      issue_sdp(0);
      for (int it = 0; it < niter; ++it) {
        const int st = it & 1;
        const uint32_t q_base = smem_u32(sQ + st * L::TILE), do_base = smem_u32(sDO + st * L::TILE);
        const uint32_t p_base = smem_u32(sP + st * 32768), ds_base = smem_u32(sDS + st * 32768);
        // S_it / dP_it have been pulled into registers: the next block's S and dP can be computed while the softmax
        // threads are still working on this one (ncu on the previous version: 17 % of all samples sat in the softmax
        // warps' wait for S / dP, whose MMAs were only issued after P / dS had been written)
        mbar_wait(sdp_read, it & 1);
        tc_fence_after();
        if (it + 1 < niter) {
          mbar_wait(&qdo_full[st ^ 1], ((it + 1) >> 1) & 1);
          tc_fence_after();
          issue_sdp(st ^ 1);
        }
        mbar_wait(pds_full, it & 1);   // P_it / dS_it are in shared memory
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BQ / 16; ++k)  // reduction over the 128 query rows
          umma_f16(tDV, make_smem_desc_sw128(p_base + k * 2048, 16384, 1024),
                   make_smem_desc_sw128(do_base + k * 2048, 16384, 1024), idesc_t, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < BQ / 16; ++k)
          umma_f16(tDK, make_smem_desc_sw128(ds_base + k * 2048, 16384, 1024),
                   make_smem_desc_sw128(q_base + k * 2048, 16384, 1024), idesc_t, (it > 0 || k > 0) ? 1u : 0u);
        if (it > 0) {
          mbar_wait(dq_read, (it - 1) & 1);  // dQ_{it-1} has left its columns
          tc_fence_after();
        }
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k)  // reduction over the 128 keys
          umma_f16(tDQ, make_smem_desc_sw128(ds_base + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                   make_smem_desc_sw128(k_base + k * 2048, 16384, 1024), idesc_q, k > 0 ? 1u : 0u);
        umma_commit(dq_full);
        umma_commit(&qdo_empty[st]);
      }
      umma_commit(done);
    }
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;          // which 64 key columns of the row (and which 32 dQ/dK/dV columns)
    const int r = quarter * 32 + lane;
    const int tid256 = (warp - 2) * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    if (p.kpm) {
      if (tid256 < 128) {
        const int kk = k0 + tid256;
        s_kpm[tid256] = (kk < p.Sk) ? p.kpm[static_cast<long long>(b) * p.Sk + kk] : 0;
      }
      named_bar_sync(1, 256);
    }
    const long long bh = static_cast<long long>(b) * p.H + h;
    uint2 dkey = make_uint2(0u, 0u);
    if constexpr (DROP) dkey = dropout_key(p.drop.state);
    auto drain_dq = [&](int it_prev) {
      const int qi = (i_start + it_prev) * BQ + r;
      mbar_wait(dq_full, it_prev & 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32(tDQ + lane_off + half * 32, v);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(dq_read);
      if (qi < p.Sq) {
        float* dq_row = p.dq_acc + (bh * p.Sq + qi) * D + half * 32;
#pragma unroll
        for (int g = 0; g < 8; ++g)
          atomicAdd(reinterpret_cast<float4*>(dq_row) + g,
                    make_float4(__uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]), __uint_as_float(v[4 * g + 2]),
                                __uint_as_float(v[4 * g + 3])));
      }
    };
    float lse_n = INFINITY, delta_n = 0.f;
    if (niter > 0) {
      const int qi = i_start * BQ + r;
      if (qi < p.Sq) { lse_n = p.lse[bh * p.Sq + qi]; delta_n = p.delta[bh * p.Sq + qi]; }
    }
    for (int it = 0; it < niter; ++it) {
      const int q0 = (i_start + it) * BQ;
      const int qi = q0 + r;
      // a row with no visible key has lse = -inf: make the exponent -inf instead of nan
      const float lse2 = (lse_n == -INFINITY || qi >= p.Sq) ? INFINITY : lse_n * LOG2E;
      const float delta_s = delta_n * p.scale;
      if (it + 1 < niter) {
        const int qn = qi + BQ;
        lse_n = INFINITY; delta_n = 0.f;
        if (qn < p.Sq) { lse_n = p.lse[bh * p.Sq + qn]; delta_n = p.delta[bh * p.Sq + qn]; }
      }
      const bool need_mask = (k0 + BKV > p.Sk) || (p.causal && (k0 + BKV - 1 > q0 + shift)) || (p.kpm != nullptr);
      uint8_t* prow = sP + (it & 1) * 32768 + half * 16384 + r * 128;
      uint8_t* drow = sDS + (it & 1) * 32768 + half * 16384 + r * 128;
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
      // both 32-column chunks of S and dP go to registers at once, then TMEM is handed back
      uint32_t vs0[32], vs1[32], vd0[32], vd1[32];
      tmem_ld_32x32(tS + lane_off + half * 64, vs0);
      tmem_ld_32x32(tDP + lane_off + half * 64, vd0);
      tmem_ld_32x32(tS + lane_off + half * 64 + 32, vs1);
      tmem_ld_32x32(tDP + lane_off + half * 64 + 32, vd1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(sdp_read);
      if constexpr (DROP) {
        int k_limit = p.Sk;
        if (p.causal) k_limit = min(k_limit, qi + shift + 1);
        const int col = half * 64;
        const uint32_t km0 = dropout_keep32(dkey, p.drop.salt, static_cast<uint32_t>(k0 + col), static_cast<uint32_t>(qi),
                                            static_cast<uint32_t>(bh), p.drop.thresh);
        const uint32_t km1 = dropout_keep32(dkey, p.drop.salt, static_cast<uint32_t>(k0 + col + 32), static_cast<uint32_t>(qi),
                                            static_cast<uint32_t>(bh), p.drop.thresh);
        bwd_softmax_chunk<T, true, true>(vs0, vd0, prow, drow, 0, r, p.scale_log2, p.scale, lse2, delta_s, k0 + col, k_limit,
                                         p.kpm ? s_kpm + col : nullptr, km0, p.drop.scale);
        bwd_softmax_chunk<T, true, true>(vs1, vd1, prow, drow, 1, r, p.scale_log2, p.scale, lse2, delta_s, k0 + col + 32,
                                         k_limit, p.kpm ? s_kpm + col + 32 : nullptr, km1, p.drop.scale);
      } else if (need_mask) {
        int k_limit = p.Sk;
        if (p.causal) k_limit = min(k_limit, qi + shift + 1);
        const int col = half * 64;
        bwd_softmax_chunk<T, true>(vs0, vd0, prow, drow, 0, r, p.scale_log2, p.scale, lse2, delta_s, k0 + col, k_limit,
                                   p.kpm ? s_kpm + col : nullptr);
        bwd_softmax_chunk<T, true>(vs1, vd1, prow, drow, 1, r, p.scale_log2, p.scale, lse2, delta_s, k0 + col + 32, k_limit,
                                   p.kpm ? s_kpm + col + 32 : nullptr);
      } else {
        bwd_softmax_chunk<T, false>(vs0, vd0, prow, drow, 0, r, p.scale_log2, p.scale, lse2, delta_s, 0, 0, nullptr);
        bwd_softmax_chunk<T, false>(vs1, vd1, prow, drow, 1, r, p.scale_log2, p.scale, lse2, delta_s, 0, 0, nullptr);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(pds_full);
      if (it > 0) drain_dq(it - 1);
    }
    if (niter > 0) {
      drain_dq(niter - 1);
      mbar_wait(done, 0);
    }
    tc_fence_after();
    // dV_j, dK_j: thread r owns key row k0 + r, columns [half*32, half*32+32)
    const int kj = k0 + r;
    T* dv_row = reinterpret_cast<T*>(p.dv) + b * p.dv_sb + h * p.dv_sh + static_cast<long long>(kj) * p.dv_ss + half * 32;
    T* dk_row = reinterpret_cast<T*>(p.dk) + b * p.dk_sb + h * p.dk_sh + static_cast<long long>(kj) * p.dk_ss + half * 32;
    uint32_t v[32], w[32];
    if (niter > 0) {
      tmem_ld_32x32(tDV + lane_off + half * 32, v);
      tmem_ld_32x32(tDK + lane_off + half * 32, w);
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) { v[i] = 0u; w[i] = 0u; }
    }
    if (kj < p.Sk) {
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float f[8], e[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { f[i] = __uint_as_float(v[g * 8 + i]); e[i] = __uint_as_float(w[g * 8 + i]); }
        *reinterpret_cast<uint4*>(dv_row + g * 8) = pack8<T>(f);
        *reinterpret_cast<uint4*>(dk_row + g * 8) = pack8<T>(e);
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

template <typename T, bool DROP>
static int launch_bwd64_t(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
                          const BwdParams& p, cudaStream_t st) {
  using L = Bwd64Smem;
  static_assert(L::TOTAL <= kMaxSmemPerCta, "flash backward (D = 64) shared-memory layout exceeds 227 KB");
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(flash_bwd64_kernel<T, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set[dev & 63] = true;
  }
  dim3 grid(static_cast<unsigned>((p.Sk + BKV - 1) / BKV) * p.H * p.B);
  flash_bwd64_kernel<T, DROP><<<grid, BWD64_THREADS, L::TOTAL, st>>>(tq, tk, tv, tdo, p);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
template <typename T>
static int launch_bwd64(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
                        const BwdParams& p, cudaStream_t st) {
  return p.drop.state ? launch_bwd64_t<T, true>(tq, tk, tv, tdo, p, st) : launch_bwd64_t<T, false>(tq, tk, tv, tdo, p, st);
}

// delta[b,h,q] = sum_d dO[b,h,q,d] * O[b,h,q,d]; D/8 adjacent lanes share a row (one 16-byte load of each operand per
// lane, so a row is read as one contiguous 128/256-byte segment) and combine with shuffles.
template <typename T>
__global__ void __launch_bounds__(256)
attn_delta_kernel(const T* __restrict__ o, const T* __restrict__ d_o, float* __restrict__ delta, float* __restrict__ dq_acc,
                  int B, int H, int S, int D, long long o_sb, long long o_sh, long long o_ss, long long g_sb, long long g_sh,
                  long long g_ss) {
  const int lpr = D >> 3;  // lanes per row: 8 or 16
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long idx = gid / lpr;
  const int c = static_cast<int>(gid % lpr) * 8;
  const long long total = static_cast<long long>(B) * H * S;
  const bool valid = idx < total;
  float acc = 0.f;
  if (valid) {
    // the same threads clear the fp32 dQ accumulator the backward kernel reduces into (saves the caller a fill pass)
    float4* z = reinterpret_cast<float4*>(dq_acc + idx * D + c);
    z[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    z[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int s = static_cast<int>(idx % S);
    const int h = static_cast<int>((idx / S) % H);
    const int b = static_cast<int>(idx / (static_cast<long long>(S) * H));
    const T* po = o + b * o_sb + h * o_sh + static_cast<long long>(s) * o_ss + c;
    const T* pg = d_o + b * g_sb + h * g_sh + static_cast<long long>(s) * g_ss + c;
    float a[8], g[8];
    unpack8<T>(ld_nc_v4(po), a);
    unpack8<T>(ld_nc_v4(pg), g);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += a[i] * g[i];
  }
  for (int off = lpr >> 1; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (valid && c == 0) delta[idx] = acc;
}

// dq[b,h,s,:] (16-bit, arbitrary b/h/s strides) = round(dq_acc[b,h,s,:]) -- the fp32 accumulator the backward reduces
// into, converted into its slot of the packed gradient buffer; 8 elements per thread.
template <typename T>
__global__ void __launch_bounds__(256)
dq_convert_kernel(const float* __restrict__ acc, T* __restrict__ dq, int B, int H, int S, int D, long long sb, long long sh,
                  long long ss) {
  const int lpr = D >> 3;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long row = gid / lpr;
  const int c = static_cast<int>(gid % lpr) * 8;
  if (row >= static_cast<long long>(B) * H * S) return;
  const int s = static_cast<int>(row % S);
  const int h = static_cast<int>((row / S) % H);
  const int b = static_cast<int>(row / (static_cast<long long>(S) * H));
  const float4 a0 = *reinterpret_cast<const float4*>(acc + row * D + c);
  const float4 a1 = *reinterpret_cast<const float4*>(acc + row * D + c + 4);
  const float f[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
  *reinterpret_cast<uint4*>(dq + b * sb + h * sh + static_cast<long long>(s) * ss + c) = pack8<T>(f);
}

template <typename T, int D, bool DROP>
static int launch_bwd_t(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
                      const BwdParams& p, cudaStream_t st) {
  using L = BwdSmem<D>;
  static_assert(L::TOTAL <= kMaxSmemPerCta, "flash backward shared-memory layout exceeds 227 KB");
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(flash_bwd_kernel<T, D, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set[dev & 63] = true;
  }
  dim3 grid(static_cast<unsigned>((p.Sk + BKV - 1) / BKV) * p.H * p.B);
  flash_bwd_kernel<T, D, DROP><<<grid, THREADS, L::TOTAL, st>>>(tq, tk, tv, tdo, p);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <typename T, int D>
static int launch_bwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
                      const BwdParams& p, cudaStream_t st) {
  return p.drop.state ? launch_bwd_t<T, D, true>(tq, tk, tv, tdo, p, st) : launch_bwd_t<T, D, false>(tq, tk, tv, tdo, p, st);
}

}  // namespace attn
}  // namespace hyp

// dq_acc: fp32 [B,H,Sq,D], cleared here and then accumulated with reductions; delta: fp32 [B,H,Sq] scratch
// written here; dk / dv are written in full.  All tensors are indexed [b,h,s,d] through element strides.
static int flash_attn_bwd_impl(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                               float* dq_acc, void* dk, void* dv, float* delta, const unsigned char* key_padding_mask,
                               int B, int H, int Sq, int Sk, int D, long long q_sb, long long q_sh, long long q_ss,
                               long long k_sb, long long k_sh, long long k_ss, long long v_sb, long long v_sh, long long v_ss,
                               long long o_sb, long long o_sh, long long o_ss, long long g_sb, long long g_sh, long long g_ss,
                               long long dkv_sb, long long dkv_sh, long long dkv_ss, int causal, float scale, int dtype,
                               const unsigned long long* rng_state, unsigned salt, float dropout_p, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_F16 || dtype == DT_BF16, "hyp_flash_attn_bwd: 16-bit dtypes only");
  HYP_REQUIRE(D == 64 || D == 128, "hyp_flash_attn_bwd: head dim must be 64 or 128 (got %d)", D);
  const long long rows = static_cast<long long>(B) * H * Sq;
  const int dgrid = static_cast<int>((rows * (D / 8) + 255) / 256);
  if (dtype == DT_BF16)
    attn::attn_delta_kernel<__nv_bfloat16><<<dgrid, 256, 0, st>>>((const __nv_bfloat16*)o, (const __nv_bfloat16*)d_o, delta, dq_acc, B, H, Sq, D, o_sb, o_sh, o_ss, g_sb, g_sh, g_ss);
  else
    attn::attn_delta_kernel<__half><<<dgrid, 256, 0, st>>>((const __half*)o, (const __half*)d_o, delta, dq_acc, B, H, Sq, D, o_sb, o_sh, o_ss, g_sb, g_sh, g_ss);
  HYP_CHECK_CUDA(cudaGetLastError());
  CUtensorMap tq, tk, tv, tdo;
  int rc;
  if ((rc = attn::make_qkv_map(&tq, q, dtype, B, H, Sq, D, q_sb, q_sh, q_ss))) return rc;
  if ((rc = attn::make_qkv_map(&tk, k, dtype, B, H, Sk, D, k_sb, k_sh, k_ss))) return rc;
  if ((rc = attn::make_qkv_map(&tv, v, dtype, B, H, Sk, D, v_sb, v_sh, v_ss))) return rc;
  if ((rc = attn::make_qkv_map(&tdo, d_o, dtype, B, H, Sq, D, g_sb, g_sh, g_ss))) return rc;
  attn::BwdParams p;
  p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.D = D; p.causal = causal;
  p.scale = scale; p.scale_log2 = scale * attn::LOG2E;
  p.kpm = key_padding_mask; p.lse = lse; p.delta = delta; p.dq_acc = dq_acc; p.dk = dk; p.dv = dv;
  p.dk_sb = dkv_sb; p.dk_sh = dkv_sh; p.dk_ss = dkv_ss; p.dv_sb = dkv_sb; p.dv_sh = dkv_sh; p.dv_ss = dkv_ss;
  p.dtype = dtype;
  if ((rc = attn::make_dropout(&p.drop, rng_state, salt, dropout_p, "hyp_flash_attn_bwd_drop"))) return rc;
  // HYPERION_FLASH_BWD=v1 keeps D = 64 on the non-pipelined kernel (A/B timing only)
  static const bool v1 = [] { const char* e = getenv("HYPERION_FLASH_BWD"); return e && e[0] == 'v' && e[1] == '1'; }();
  if (D == 64 && !v1)
    return dtype == DT_BF16 ? attn::launch_bwd64<__nv_bfloat16>(tq, tk, tv, tdo, p, st)
                            : attn::launch_bwd64<__half>(tq, tk, tv, tdo, p, st);
  if (dtype == DT_BF16) return D == 64 ? attn::launch_bwd<__nv_bfloat16, 64>(tq, tk, tv, tdo, p, st) : attn::launch_bwd<__nv_bfloat16, 128>(tq, tk, tv, tdo, p, st);
  return D == 64 ? attn::launch_bwd<__half, 64>(tq, tk, tv, tdo, p, st) : attn::launch_bwd<__half, 128>(tq, tk, tv, tdo, p, st);
}

HYP_API int hyp_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                               float* dq_acc, void* dk, void* dv, float* delta, const unsigned char* key_padding_mask,
                               int B, int H, int Sq, int Sk, int D, long long q_sb, long long q_sh, long long q_ss,
                               long long k_sb, long long k_sh, long long k_ss, long long v_sb, long long v_sh, long long v_ss,
                               long long o_sb, long long o_sh, long long o_ss, long long g_sb, long long g_sh, long long g_ss,
                               long long dkv_sb, long long dkv_sh, long long dkv_ss, int causal, float scale, int dtype,
                               void* stream_) {
  return flash_attn_bwd_impl(q, k, v, o, d_o, lse, dq_acc, dk, dv, delta, key_padding_mask, B, H, Sq, Sk, D, q_sb, q_sh, q_ss,
                             k_sb, k_sh, k_ss, v_sb, v_sh, v_ss, o_sb, o_sh, o_ss, g_sb, g_sh, g_ss, dkv_sb, dkv_sh, dkv_ss,
                             causal, scale, dtype, nullptr, 0u, 0.f, stream_);
}

// Backward of hyp_flash_attn_fwd_drop: same arguments plus the forward's (rng_state, salt, dropout_p); must run before the
// engine advances the step word of rng_state.
HYP_API int hyp_flash_attn_bwd_drop(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                    const float* lse, float* dq_acc, void* dk, void* dv, float* delta,
                                    const unsigned char* key_padding_mask, int B, int H, int Sq, int Sk, int D, long long q_sb,
                                    long long q_sh, long long q_ss, long long k_sb, long long k_sh, long long k_ss,
                                    long long v_sb, long long v_sh, long long v_ss, long long o_sb, long long o_sh,
                                    long long o_ss, long long g_sb, long long g_sh, long long g_ss, long long dkv_sb,
                                    long long dkv_sh, long long dkv_ss, int causal, float scale, int dtype,
                                    const unsigned long long* rng_state, int salt, float dropout_p, void* stream_) {
  return flash_attn_bwd_impl(q, k, v, o, d_o, lse, dq_acc, dk, dv, delta, key_padding_mask, B, H, Sq, Sk, D, q_sb, q_sh, q_ss,
                             k_sb, k_sh, k_ss, v_sb, v_sh, v_ss, o_sb, o_sh, o_ss, g_sb, g_sh, g_ss, dkv_sb, dkv_sh, dkv_ss,
                             causal, scale, dtype, rng_state, static_cast<unsigned>(salt), dropout_p, stream_);
}

// dq (16-bit, [b,h,s,d] through element strides, unit stride in d) = dq_acc (fp32 contiguous [B,H,S,D]).
HYP_API int hyp_flash_dq_convert(const float* dq_acc, void* dq, int B, int H, int S, int D, long long sb, long long sh,
                                 long long ss, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_F16 || dtype == DT_BF16, "hyp_flash_dq_convert: 16-bit outputs only");
  HYP_REQUIRE(D % 8 == 0 && (sb % 8 == 0) && (sh % 8 == 0) && (ss % 8 == 0) && reinterpret_cast<uintptr_t>(dq) % 16 == 0,
              "hyp_flash_dq_convert: rows must be 16-byte aligned");
  const long long threads = static_cast<long long>(B) * H * S * (D / 8);
  const int grid = static_cast<int>((threads + 255) / 256);
  if (dtype == DT_BF16)
    attn::dq_convert_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(dq_acc, (__nv_bfloat16*)dq, B, H, S, D, sb, sh, ss);
  else
    attn::dq_convert_kernel<__half><<<grid, 256, 0, st>>>(dq_acc, (__half*)dq, B, H, S, D, sb, sh, ss);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
