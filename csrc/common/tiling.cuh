// Tile-index arithmetic of the persistent kernels (GEMM tile order with its L2 grouping and split-K ranges, convolution forward
// patches, convolution wgrad work items), kept free of PTX and CUDA types so that the SAME code compiles for the host:
// tests/test_cpu_kernel_host.py builds it with g++ and checks, without a GPU, that every schedule covers its iteration space exactly
// once and that the orders have the locality the kernels rely on.
#pragma once

#if defined(__CUDACC__)
#define HYP_TILING_FN __host__ __device__ __forceinline__
#else
#define HYP_TILING_FN inline
#endif

namespace hyp {
namespace tiling {

HYP_TILING_FN int imin(int a, int b) { return a < b ? a : b; }

// ---- GEMM: tile -> (m block, n block, k-block range) ------------------------------------------------------------
// Tiles of one split are walked in groups of GROUP_M row blocks: inside a group the m index runs fastest, so GROUP_M
// consecutive CTAs share one B panel and the group re-reads its A panels from L2 (126 MB) rather than HBM.  With
// `splits` > 1 the whole (m, n) walk repeats once per K range.
struct TileCoord {
  int m_blk, n_blk, kb0, kb1;
};

template <int GROUP_M>
HYP_TILING_FN TileCoord decode_gemm_tile(int tile, int tiles_m, int tiles_n, int nkb, int splits) {
  const int per_split = tiles_m * tiles_n;
  const int s = tile / per_split;
  const int t = tile - s * per_split;
  const int group_size = GROUP_M * tiles_n;
  const int g = t / group_size;
  const int first_m = g * GROUP_M;
  const int gm = imin(tiles_m - first_m, GROUP_M);
  const int r = t - g * group_size;
  TileCoord c;
  c.m_blk = first_m + r % gm;
  c.n_blk = r / gm;
  const int per = (nkb + splits - 1) / splits;
  c.kb0 = s * per;
  c.kb1 = imin(nkb, c.kb0 + per);
  return c;
}

// ---- convolution forward: tile -> (output-channel block, patch x, patch y, image block) -------------------------
// The output-channel blocks of one spatial patch run back to back, so the input patch they all read stays in L2.
struct ConvTile {
  int kb, pw, ph, pn;
};

HYP_TILING_FN ConvTile decode_conv_tile(int t, int tiles_k, int tiles_w, int tiles_h) {
  ConvTile c;
  c.kb = t % tiles_k; t /= tiles_k;
  c.pw = t % tiles_w; t /= tiles_w;
  c.ph = t % tiles_h; t /= tiles_h;
  c.pn = t;
  return c;
}

// ---- convolution wgrad: item -> (output-channel block, filter tap, input-channel block, patch range of this split) ----
struct WgradItem {
  int kb, tap, cb, pb0, pb1;
};

HYP_TILING_FN WgradItem decode_wgrad_item(int t, int splits, int cblocks, int taps, int total_pb, int pb_per_split) {
  WgradItem w;
  const int split = t % splits; t /= splits;
  w.cb = t % cblocks; t /= cblocks;
  w.tap = t % taps; t /= taps;
  w.kb = t;
  w.pb0 = split * pb_per_split;
  w.pb1 = imin(total_pb, w.pb0 + pb_per_split);
  return w;
}

}  // namespace tiling
}  // namespace hyp
