// Counter-based dropout masks (Philox4x32-10), shared by every kernel that drops elements: flash attention (forward and
// the regenerated mask in backward), the residual-dropout inside fused_norm, and the stand-alone dropout kernel.
//
// A mask element is addressed by (plane, row, col) and a per-call `salt`; one Philox call yields 16 bytes = the bytes of
// 16 consecutive columns (col >> 4 is part of the counter), and an element is KEPT when its byte >= thresh, where
// thresh = round(p * 256).  So the effective drop probability is thresh / 256 and survivors are scaled by
// 256 / (256 - thresh) -- eight bits per element keep the generator at ~4 integer ops per element.
// The key is (seed, step), both read from device memory (`state[0]`, `state[1]`): `step` is advanced by the training
// engine once per optimizer step, so a CUDA graph that bakes `salt` still draws fresh masks on every replay, and a
// backward (or an activation-checkpoint recompute) launched before that advance regenerates exactly the forward's mask.
#pragma once
#include <cstdint>

namespace hyp {

struct DropoutArgs {
  const unsigned long long* state;  // device: [seed, step]; nullptr = no dropout
  unsigned salt;                    // distinguishes the calls within a step
  unsigned thresh;                  // drop when byte < thresh (1..255)
  float scale;                      // 256 / (256 - thresh)
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

__device__ __forceinline__ uint2 dropout_key(const unsigned long long* state) {
  return make_uint2(static_cast<uint32_t>(state[0]), static_cast<uint32_t>(state[1]));
}

// 16 mask bytes for columns [16 * group, 16 * group + 16) of (plane, row)
__device__ __forceinline__ uint4 dropout_bytes16(uint2 key, uint32_t salt, uint32_t group, uint32_t row, uint32_t plane) {
  return philox4x32_10(make_uint4(group, row, plane, salt), key);
}

// bit i of the result = keep column (16 * group + i)
__device__ __forceinline__ uint32_t dropout_keep16(uint2 key, uint32_t salt, uint32_t group, uint32_t row, uint32_t plane,
                                                   uint32_t thresh) {
  const uint4 b = dropout_bytes16(key, salt, group, row, plane);
  const uint32_t w[4] = {b.x, b.y, b.z, b.w};
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) m |= (((w[i >> 2] >> ((i & 3) * 8)) & 0xFFu) >= thresh ? 1u : 0u) << i;
  return m;
}
// keep bits of 32 consecutive columns starting at col0 (a multiple of 32)
__device__ __forceinline__ uint32_t dropout_keep32(uint2 key, uint32_t salt, uint32_t col0, uint32_t row, uint32_t plane,
                                                   uint32_t thresh) {
  return dropout_keep16(key, salt, col0 >> 4, row, plane, thresh) |
         (dropout_keep16(key, salt, (col0 >> 4) + 1, row, plane, thresh) << 16);
}

constexpr uint32_t kDropoutPlane2D = 0x80000000u;  // plane id of the 2-D (row, col) masks; attention uses b * H + h

}  // namespace hyp
