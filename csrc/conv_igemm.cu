// conv_igemm_tcgen05: NHWC convolution forward as an implicit GEMM on the 5th-gen tensor cores.
//
//   y[n, ho, wo, k] = sum_{r, s, c} x[n, ho*stride + r - pad, wo*stride + s - pad, c] * w[k, r, s, c]
//
// GEMM view: M = output pixels, N = output channels, K = R*S*C.  An M tile is a patch of NB images x TH x TW output
// pixels (NB*TH*TW = 128).  For one filter tap (r, s) and one block of 64 input channels, the A operand of that patch
// is a 4-D box of the input tensor -- (64 channels, TW, TH, NB) starting at (c0, wo0*stride + s - pad,
// ho0*stride + r - pad, n0), traversed with the convolution stride -- so ONE TMA instruction fetches it, zero-filling
// everything that falls into the padding or past the image.  No im2col buffer exists anywhere.  It lands in shared
// memory as 128 rows x 128 B, the same K-major 128B-swizzled tile the plain GEMM uses; the B operand is a 2-D box of
// the weights stored [K][R][S][C] (what PyTorch calls channels_last for a conv weight).
//
//   warp 0     TMA producer, 6-stage ring over (r, s, c-block)
//   warp 1     one thread issues tcgen05.mma (128 x BN x 16), two accumulator stages in TMEM
//   warps 2-5  epilogue: thread = output pixel; optional per-channel scale/shift (an eval-mode BatchNorm folded into
//              the convolution), residual add and ReLU; 16-bit NHWC stores
//
// Needs C % 64 == 0, K % 32 == 0, 16-bit data; the 3-channel stem convolution stays on cuDNN.
// Reference call sites: torchvision resnet18/50 Conv2d (02_development/distributed_utils.py:229; Phase 1
// baseline_performance.ipynb c0 L23; Phase 5 compilation_optimization.py:89), SURVEY.md K-O7.
#include "common/ptx.cuh"
#include "common/tiling.cuh"
#include "common/tmap.cuh"
#include "common/util.cuh"
#include <stdlib.h>

namespace hyp {
namespace conv {

constexpr int BM = 128;
constexpr int ROW_BYTES = 128;
constexpr int A_STAGE = BM * ROW_BYTES;
constexpr int STAGES = 6;
constexpr int THREADS = 224;   // warp 0: TMA producer (activations), warp 1: MMA, warps 2-5: epilogue, warp 6: TMA producer (second operand)

struct Params {
  int N, H, W, C;        // input NHWC
  int K, R, S;           // output channels, filter
  int stride, pad;
  int Ho, Wo;
  int TW, TH, NB;        // output patch of one M tile
  int tiles_w, tiles_h, tiles_n, tiles_k;
  void* y;               // [N, Ho, Wo, K]
  int dtype;
  const float* scale;    // optional per-output-channel affine (folded BatchNorm): y = acc * scale + shift
  const float* shift;
  const void* residual;  // optional, same shape/dtype as y
  int relu;
  uint32_t idesc;
};

template <int BN>
struct Cfg {
  static constexpr int B_STAGE = BN * ROW_BYTES;
  static constexpr int SMEM_BYTES = STAGES * (A_STAGE + B_STAGE) + 256 + 1024;
  static constexpr int TMEM_COLS = 2 * BN;
};

using Tile = hyp::tiling::ConvTile;   // tile order: common/tiling.cuh, host-testable
__device__ __forceinline__ Tile decode(int t, const Params& p) {
  return hyp::tiling::decode_conv_tile(t, p.tiles_k, p.tiles_w, p.tiles_h);
}

template <int BN>
__global__ void __launch_bounds__(THREADS, 1)
conv_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                const __grid_constant__ Params p) {
  using C = Cfg<BN>;
  extern __shared__ __align__(1024) uint8_t smem_conv[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_conv) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * C::B_STAGE);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cblocks = p.C / 64;
  const int nkb = p.R * p.S * cblocks;
  const int total = p.tiles_k * p.tiles_w * p.tiles_h * p.tiles_n;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
      for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<C::TMEM_COLS>(tmem_holder);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0 || warp == 6) {
    // ------------------------------------------------------------------ TMA producers: warp 0 the input boxes (and the
    // expect_tx of the stage), warp 6 the weight tiles -- a single issuing thread is the bottleneck of such short k-blocks
    if (lane == 0) {
      const bool load_a = (warp == 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const Tile tc = decode(t, p);
        const int w_in0 = tc.pw * p.TW * p.stride - p.pad;
        const int h_in0 = tc.ph * p.TH * p.stride - p.pad;
        for (int r = 0; r < p.R; ++r) {
          for (int s = 0; s < p.S; ++s) {
            for (int cb = 0; cb < cblocks; ++cb) {
              mbar_wait(&empty[stage], phase ^ 1);
              if (load_a) {
                mbar_expect_tx(&full[stage], A_STAGE + C::B_STAGE);
                // the input box of this tap: out-of-image coordinates (padding, ragged patches) are zero-filled
                tma_load_4d(sA + stage * A_STAGE, &tmX, &full[stage], cb * 64, w_in0 + s, h_in0 + r, tc.pn * p.NB);
              } else {
                tma_load_2d(sB + stage * C::B_STAGE, &tmW, &full[stage], (r * p.S + s) * p.C + cb * 64, tc.kb * BN);
              }
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * A_STAGE);
          const uint32_t b_base = smem_u32(sB + stage * C::B_STAGE);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(d_tmem, make_smem_desc_sw128(a_base + k * 32, 16, 1024), make_smem_desc_sw128(b_base + k * 32, 16, 1024),
                     p.idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue: thread = output pixel
    const int quarter = warp & 3;
    const int i = quarter * 32 + lane;          // row of the tile == TMEM lane
    const int tw = i % p.TW, th = (i / p.TW) % p.TH, nb = i / (p.TW * p.TH);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const Tile tc = decode(t, p);
      const int n = tc.pn * p.NB + nb, ho = tc.ph * p.TH + th, wo = tc.pw * p.TW + tw;
      const bool valid = n < p.N && ho < p.Ho && wo < p.Wo;
      const long long pix = (static_cast<long long>(n) * p.Ho + ho) * p.Wo + wo;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int k0 = tc.kb * BN + c * 32;
        if (k0 >= p.K) break;
        uint32_t r[32];
        tmem_ld_32x32(t_row + c * 32, r);
        tmem_ld_wait();
        if (!valid) continue;
        float v[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(r[e]);
        if (p.scale) {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const float4 sc = *reinterpret_cast<const float4*>(p.scale + k0 + 4 * g);
            const float4 sh = *reinterpret_cast<const float4*>(p.shift + k0 + 4 * g);
            v[4 * g] = fmaf(v[4 * g], sc.x, sh.x); v[4 * g + 1] = fmaf(v[4 * g + 1], sc.y, sh.y);
            v[4 * g + 2] = fmaf(v[4 * g + 2], sc.z, sh.z); v[4 * g + 3] = fmaf(v[4 * g + 3], sc.w, sh.w);
          }
        }
        if (p.residual) {
          const uint4* r4 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.residual) + pix * p.K + k0);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float f[8];
            const uint4 u = r4[g];
            if (p.dtype == DT_BF16) unpack8<__nv_bfloat16>(u, f); else unpack8<__half>(u, f);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[g * 8 + e] += f[e];
          }
        }
        if (p.relu) {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = fmaxf(v[e], 0.f);
        }
        uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.y) + pix * p.K + k0);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = v[g * 8 + e];
          d[g] = (p.dtype == DT_BF16) ? pack8<__nv_bfloat16>(f) : pack8<__half>(f);
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

// Output patch of a tile: powers of two with TW * TH * NB = pixels (128 forward, 64 wgrad), chosen to waste the fewest tile rows on the ragged
// right / bottom edges (ties go to the wider patch: longer contiguous runs for the TMA box).
static void pick_patch(int Ho, int Wo, int N, int* TW, int* TH, int* NB, int pixels = 128) {
  double best = 1e30;
  for (int tw = 16; tw >= 1; tw >>= 1) {
    for (int th = pixels / tw; th >= 1; th >>= 1) {
      if (th > 16) continue;
      const int nb = pixels / (tw * th);
      const double cover = static_cast<double>((Wo + tw - 1) / tw * tw) * ((Ho + th - 1) / th * th) * ((N + nb - 1) / nb * nb);
      const double waste = cover / (static_cast<double>(Wo) * Ho * N);
      if (waste < best - 1e-9) { best = waste; *TW = tw; *TH = th; *NB = nb; }
    }
  }
}

template <int BN>
static int launch(const CUtensorMap& tx, const CUtensorMap& tw, const Params& p, cudaStream_t st) {
  using C = Cfg<BN>;
  static_assert(C::SMEM_BYTES <= kMaxSmemPerCta && tmem_alloc_ok(C::TMEM_COLS), "conv forward tile exceeds the sm_100a CTA budget");
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(conv_fwd_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int total = p.tiles_k * p.tiles_w * p.tiles_h * p.tiles_n;
  const int grid = total < num_sms() ? total : num_sms();
  conv_fwd_kernel<BN><<<grid, THREADS, C::SMEM_BYTES, st>>>(tx, tw, p);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace conv
}  // namespace hyp

using namespace hyp;

// x: [N, H, W, C] (NHWC, contiguous), w: [K, R, S, C] (contiguous), y: [N, Ho, Wo, K]; all `dtype` (1 fp16, 2 bf16).
// scale / shift: optional fp32 [K] (y = conv * scale + shift, e.g. a folded eval-mode BatchNorm or a bias with
// scale = 1); residual: optional tensor shaped like y, added before the optional ReLU.
HYP_API int hyp_conv2d_fwd(const void* x, const void* w, void* y, int N, int H, int W, int C, int K, int R, int S,
                           int stride, int pad, const float* scale, const float* shift, const void* residual, int relu,
                           int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_F16 || dtype == DT_BF16, "hyp_conv2d_fwd: 16-bit dtypes only");
  HYP_REQUIRE(C % 64 == 0 && K % 32 == 0, "hyp_conv2d_fwd: needs C %% 64 == 0 and K %% 32 == 0 (C=%d K=%d)", C, K);
  HYP_REQUIRE(stride >= 1 && stride <= 2 && pad >= 0 && R >= 1 && S >= 1, "hyp_conv2d_fwd: stride 1 or 2 only");
  HYP_REQUIRE((scale == nullptr) == (shift == nullptr), "hyp_conv2d_fwd: scale and shift come together");
  conv::Params p;
  p.N = N; p.H = H; p.W = W; p.C = C; p.K = K; p.R = R; p.S = S; p.stride = stride; p.pad = pad;
  p.Ho = (H + 2 * pad - R) / stride + 1;
  p.Wo = (W + 2 * pad - S) / stride + 1;
  HYP_REQUIRE(p.Ho > 0 && p.Wo > 0, "hyp_conv2d_fwd: empty output");
  conv::pick_patch(p.Ho, p.Wo, N, &p.TW, &p.TH, &p.NB);
  const int BN = (K % 128 == 0) ? 128 : 64;
  p.tiles_w = (p.Wo + p.TW - 1) / p.TW;
  p.tiles_h = (p.Ho + p.TH - 1) / p.TH;
  p.tiles_n = (N + p.NB - 1) / p.NB;
  p.tiles_k = (K + BN - 1) / BN;
  p.y = y; p.dtype = dtype; p.scale = scale; p.shift = shift; p.residual = residual; p.relu = relu;
  p.idesc = make_idesc(dtype == DT_F16 ? 0u : 1u, conv::BM, BN, false, false);
  // input: (C, W, H, N), traversed with the convolution stride in W and H
  CUtensorMap tx, tw;
  const unsigned long long xd[4] = {(unsigned long long)C, (unsigned long long)W, (unsigned long long)H, (unsigned long long)N};
  const unsigned long long xs[3] = {(unsigned long long)C * 2, (unsigned long long)W * C * 2, (unsigned long long)H * W * C * 2};
  // with a traversal stride e the box extent counts source elements: ceil(box / e) of them are fetched
  const unsigned xb[4] = {64u, (unsigned)(p.TW * stride), (unsigned)(p.TH * stride), (unsigned)p.NB};
  const unsigned xe[4] = {1u, (unsigned)stride, (unsigned)stride, 1u};
  int rc = make_tmap_nd(&tx, x, dtype, 4, xd, xs, xb, xe);
  if (rc) return rc;
  const unsigned long long wd[2] = {(unsigned long long)R * S * C, (unsigned long long)K};
  const unsigned long long ws[1] = {(unsigned long long)R * S * C * 2};
  const unsigned wb[2] = {64u, (unsigned)BN};
  rc = make_tmap_nd(&tw, w, dtype, 2, wd, ws, wb);
  if (rc) return rc;
  return BN == 128 ? conv::launch<128>(tx, tw, p, st) : conv::launch<64>(tx, tw, p, st);
}

// =====================================================================================================
// Weight gradient.   dw[k, r, s, c] = sum over output pixels  dy[pix, k] * x[pix shifted by (r, s), c]
//
// GEMM view: M = output channels (128 per tile), N = input channels of ONE filter tap (BN per tile), reduction over
// the output pixels in blocks of 64 (a patch of NB x TH x TW pixels).  Both operands are MN-major -- the reduction
// index (the pixel) is the slow one in memory for dy [pix, K] and for x [pix, C] alike -- so a stage is
//   A: 2 chunks of [64 pixels][64 out-channels]   one 5-D TMA box of dy   (64, TW, TH, NB, 2 channel chunks)
//   B: BN/64 chunks of [64 pixels][64 in-channels] one 5-D TMA box of x, shifted by the tap and traversed with the
//      convolution stride (padding / ragged patches are zero-filled and add nothing)
// A work item is (out-channel block, tap, in-channel block, pixel range): few output tiles but a very long
// reduction, so the pixel blocks are split across CTAs and the partial tiles reduced into an fp32 dw (zeroed by the
// caller) with vector atomics.
namespace hyp {
namespace conv {

struct WParams {
  int N, H, W, C, K, R, S, stride, pad, Ho, Wo;
  int TW, TH, NB;                 // 64-pixel patch
  int tiles_w, tiles_h, tiles_n;  // pixel blocks
  int kblocks, cblocks;           // output tiles: ceil(K / 128) x (R*S) x ceil(C / BN)
  int splits, pb_per_split;
  float* dw;                      // [K, R, S, C] fp32
  uint32_t idesc;
};

template <int BN>
struct WCfg {
  static constexpr int A_BYTES = 2 * 8192;
  static constexpr int B_BYTES = (BN / 64) * 8192;
  static constexpr int WSTAGES = (BN == 128) ? 6 : 8;
  static constexpr int SMEM_BYTES = WSTAGES * (A_BYTES + B_BYTES) + 256 + 1024;
  static constexpr int TMEM_COLS = 2 * BN;
};

using WItem = hyp::tiling::WgradItem;
__device__ __forceinline__ WItem wdecode(int t, const WParams& p) {
  return hyp::tiling::decode_wgrad_item(t, p.splits, p.cblocks, p.R * p.S, p.tiles_w * p.tiles_h * p.tiles_n, p.pb_per_split);
}

template <int BN>
__global__ void __launch_bounds__(THREADS, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                  const __grid_constant__ WParams p) {
  using C = WCfg<BN>;
  constexpr int WSTAGES = C::WSTAGES;
  extern __shared__ __align__(1024) uint8_t smem_conv[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_conv) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + WSTAGES * C::A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + WSTAGES * C::B_BYTES);
  uint64_t* empty = full + WSTAGES;
  uint64_t* tfull = empty + WSTAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.kblocks * p.R * p.S * p.cblocks * p.splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < WSTAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
      for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<C::TMEM_COLS>(tmem_holder);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0 || warp == 6) {
    // ------------------------------------------------------------------ TMA producers: warp 0 dy, warp 6 x
    if (lane == 0) {
      const bool load_a = (warp == 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const WItem w = wdecode(t, p);
        const int r = w.tap / p.S, s = w.tap % p.S;
        for (int pb = w.pb0; pb < w.pb1; ++pb) {
          const int pw = pb % p.tiles_w, ph = (pb / p.tiles_w) % p.tiles_h, pn = pb / (p.tiles_w * p.tiles_h);
          mbar_wait(&empty[stage], phase ^ 1);
          if (load_a) {
            mbar_expect_tx(&full[stage], C::A_BYTES + C::B_BYTES);
            tma_load_5d(sA + stage * C::A_BYTES, &tmDY, &full[stage], 0, pw * p.TW, ph * p.TH, pn * p.NB, w.kb * 2);
          } else {
            tma_load_5d(sB + stage * C::B_BYTES, &tmX, &full[stage], 0, pw * p.TW * p.stride + s - p.pad,
                        ph * p.TH * p.stride + r - p.pad, pn * p.NB, w.cb * (BN / 64));
          }
          if (++stage == WSTAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (both operands MN-major)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const WItem w = wdecode(t, p);
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int pb = w.pb0; pb < w.pb1; ++pb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * C::A_BYTES);
          const uint32_t b_base = smem_u32(sB + stage * C::B_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k)   // 16 pixels per MMA: 16 rows of 128 B inside each chunk
            umma_f16(d_tmem, make_smem_desc_sw128(a_base + k * 2048, 8192, 1024), make_smem_desc_sw128(b_base + k * 2048, 8192, 1024),
                     p.idesc, (pb > w.pb0 || k > 0) ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == WSTAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue: thread = output channel
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const WItem w = wdecode(t, p);
      const int k = w.kb * 128 + quarter * 32 + lane;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      float* out = p.dw + (static_cast<long long>(k) * p.R * p.S + w.tap) * p.C + w.cb * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        if (w.cb * BN + c * 32 >= p.C) break;
        uint32_t r[32];
        tmem_ld_32x32(t_row + c * 32, r);
        tmem_ld_wait();
        if (k < p.K && w.pb1 > w.pb0) {
#pragma unroll
          for (int g = 0; g < 8; ++g)
            atomicAdd(reinterpret_cast<float4*>(out + c * 32) + g,
                      make_float4(__uint_as_float(r[4 * g]), __uint_as_float(r[4 * g + 1]), __uint_as_float(r[4 * g + 2]),
                                  __uint_as_float(r[4 * g + 3])));
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

template <int BN>
static int launch_wgrad(const CUtensorMap& tdy, const CUtensorMap& tx, const WParams& p, cudaStream_t st) {
  using C = WCfg<BN>;
  static_assert(C::SMEM_BYTES <= kMaxSmemPerCta && tmem_alloc_ok(C::TMEM_COLS), "conv wgrad tile exceeds the sm_100a CTA budget");
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    HYP_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int total = p.kblocks * p.R * p.S * p.cblocks * p.splits;
  const int grid = total < num_sms() ? total : num_sms();
  conv_wgrad_kernel<BN><<<grid, THREADS, C::SMEM_BYTES, st>>>(tdy, tx, p);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace conv
}  // namespace hyp

// dw [K, R, S, C] fp32, ZEROED by the caller, += the weight gradient of the convolution y = conv(x, w) given
// x [N, H, W, C] and dy [N, Ho, Wo, K] (NHWC, 16-bit).  Needs C % 64 == 0 and K % 64 == 0.
HYP_API int hyp_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int K, int R, int S,
                             int stride, int pad, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(dtype == DT_F16 || dtype == DT_BF16, "hyp_conv2d_wgrad: 16-bit dtypes only");
  HYP_REQUIRE(C % 64 == 0 && K % 64 == 0, "hyp_conv2d_wgrad: needs C %% 64 == 0 and K %% 64 == 0 (C=%d K=%d)", C, K);
  HYP_REQUIRE(stride >= 1 && stride <= 2 && pad >= 0 && R >= 1 && S >= 1, "hyp_conv2d_wgrad: stride 1 or 2 only");
  conv::WParams p;
  p.N = N; p.H = H; p.W = W; p.C = C; p.K = K; p.R = R; p.S = S; p.stride = stride; p.pad = pad;
  p.Ho = (H + 2 * pad - R) / stride + 1;
  p.Wo = (W + 2 * pad - S) / stride + 1;
  HYP_REQUIRE(p.Ho > 0 && p.Wo > 0, "hyp_conv2d_wgrad: empty output");
  conv::pick_patch(p.Ho, p.Wo, N, &p.TW, &p.TH, &p.NB, 64);
  p.tiles_w = (p.Wo + p.TW - 1) / p.TW;
  p.tiles_h = (p.Ho + p.TH - 1) / p.TH;
  p.tiles_n = (N + p.NB - 1) / p.NB;
  const int BN = (C % 128 == 0) ? 128 : 64;
  p.kblocks = (K + 127) / 128;
  p.cblocks = C / BN;
  const int items = p.kblocks * R * S * p.cblocks;
  const int total_pb = p.tiles_w * p.tiles_h * p.tiles_n;
  int splits = (3 * num_sms() + items - 1) / items;
  if (splits > total_pb / 4) splits = total_pb / 4;
  if (splits < 1) splits = 1;
  p.pb_per_split = (total_pb + splits - 1) / splits;
  p.splits = (total_pb + p.pb_per_split - 1) / p.pb_per_split;   // no empty split
  p.dw = dw;
  p.idesc = make_idesc(dtype == DT_F16 ? 0u : 1u, 128, BN, true, true);
  CUtensorMap tdy, tx;
  {
    const unsigned long long d[5] = {64ull, (unsigned long long)p.Wo, (unsigned long long)p.Ho, (unsigned long long)N, (unsigned long long)(K / 64)};
    const unsigned long long sb[4] = {(unsigned long long)K * 2, (unsigned long long)p.Wo * K * 2, (unsigned long long)p.Ho * p.Wo * K * 2, 128ull};
    const unsigned b[5] = {64u, (unsigned)p.TW, (unsigned)p.TH, (unsigned)p.NB, 2u};
    const int rc = make_tmap_nd(&tdy, dy, dtype, 5, d, sb, b);
    if (rc) return rc;
  }
  {
    const unsigned long long d[5] = {64ull, (unsigned long long)W, (unsigned long long)H, (unsigned long long)N, (unsigned long long)(C / 64)};
    const unsigned long long sb[4] = {(unsigned long long)C * 2, (unsigned long long)W * C * 2, (unsigned long long)H * W * C * 2, 128ull};
    const unsigned b[5] = {64u, (unsigned)(p.TW * stride), (unsigned)(p.TH * stride), (unsigned)p.NB, (unsigned)(BN / 64)};
    const unsigned e[5] = {1u, (unsigned)stride, (unsigned)stride, 1u, 1u};
    const int rc = make_tmap_nd(&tx, x, dtype, 5, d, sb, b, e);
    if (rc) return rc;
  }
  return BN == 128 ? conv::launch_wgrad<128>(tdy, tx, p, st) : conv::launch_wgrad<64>(tdy, tx, p, st);
}
