// tcgen05 operand descriptors: pure bit arithmetic, kept free of PTX so the same code compiles for the host.
// tests/test_cpu_kernel_host.py builds it with g++ and compares every encoding the kernels use with the bit-field unions of the
// vendored CUTLASS headers (cute/arch/mma_sm100_desc.hpp: UMMA::InstrDescriptor, UMMA::SmemDescriptor) when they are present, and
// with the field table below otherwise.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define HYP_DESC_FN __host__ __device__ __forceinline__ constexpr
#else
#define HYP_DESC_FN inline constexpr
#endif

namespace hyp {

// Shared-memory matrix descriptor (SWIZZLE_128B, sm_100 version field = 1).  lbo / sbo are byte offsets (multiples of 16).
//   [0,14) start address >> 4   [16,30) leading-dimension byte offset >> 4   [32,46) stride-dimension byte offset >> 4
//   [46,48) version = 1 (Blackwell)   [61,64) layout type: 2 = SWIZZLE_128B
HYP_DESC_FN uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;  // descriptor version (Blackwell)
  d |= 2ull << 61;  // LayoutType::SWIZZLE_128B
  return d;
}

// Instruction descriptor for kind::f16 / kind::tf32 with fp32 accumulation.
// fmt: 0 = f16, 1 = bf16, 2 = tf32.  *_mn_major: operand's non-reduction dimension is the contiguous one in shared memory.
//   [4,6) accumulator format: 1 = F32   [7,10) A format   [10,13) B format   [15] A major   [16] B major (0 = K-major, 1 = MN-major)
//   [17,23) N >> 3   [24,29) M >> 4
HYP_DESC_FN uint32_t make_idesc(uint32_t fmt, uint32_t m, uint32_t n, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((a_mn_major ? 1u : 0u) << 15) |
         ((b_mn_major ? 1u : 0u) << 16) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

}  // namespace hyp
