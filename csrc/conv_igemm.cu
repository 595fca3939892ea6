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
#include "common/tmap.cuh"
#include "common/util.cuh"
#include <stdlib.h>

namespace hyp {
namespace conv {

constexpr int BM = 128;
constexpr int ROW_BYTES = 128;
constexpr int A_STAGE = BM * ROW_BYTES;
constexpr int STAGES = 6;
constexpr int THREADS = 192;

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

struct Tile { int kb, pw, ph, pn; };
__device__ __forceinline__ Tile decode(int t, const Params& p) {
  Tile c;
  c.kb = t % p.tiles_k; t /= p.tiles_k;   // output-channel blocks of one patch run back to back (input patch stays in L2)
  c.pw = t % p.tiles_w; t /= p.tiles_w;
  c.ph = t % p.tiles_h; t /= p.tiles_h;
  c.pn = t;
  return c;
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

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
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
              mbar_expect_tx(&full[stage], A_STAGE + C::B_STAGE);
              // the input box of this tap: out-of-image coordinates (padding, ragged patches) are zero-filled
              tma_load_4d(sA + stage * A_STAGE, &tmX, &full[stage], cb * 64, w_in0 + s, h_in0 + r, tc.pn * p.NB);
              tma_load_2d(sB + stage * C::B_STAGE, &tmW, &full[stage], (r * p.S + s) * p.C + cb * 64, tc.kb * BN);
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

// Output patch of an M tile: powers of two with TW * TH * NB = 128, chosen to waste the fewest tile rows on the ragged
// right / bottom edges (ties go to the wider patch: longer contiguous runs for the TMA box).
static void pick_patch(int Ho, int Wo, int N, int* TW, int* TH, int* NB) {
  double best = 1e30;
  for (int tw = 16; tw >= 1; tw >>= 1) {
    for (int th = 128 / tw; th >= 1; th >>= 1) {
      if (th > 16) continue;
      const int nb = 128 / (tw * th);
      const double cover = static_cast<double>((Wo + tw - 1) / tw * tw) * ((Ho + th - 1) / th * th) * ((N + nb - 1) / nb * nb);
      const double waste = cover / (static_cast<double>(Wo) * Ho * N);
      if (waste < best - 1e-9) { best = waste; *TW = tw; *TH = th; *NB = nb; }
    }
  }
}

template <int BN>
static int launch(const CUtensorMap& tx, const CUtensorMap& tw, const Params& p, cudaStream_t st) {
  using C = Cfg<BN>;
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
