// Host-side TMA descriptor construction (cuTensorMapEncodeTiled resolved at run time through the runtime's
// driver-entry-point query, so the library needs no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "util.cuh"

namespace hyp {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn tmap_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  }
  return fn;
}

// rank-N (<= 5) tiled map, SWIZZLE_128B, zero fill out of bounds.  dims[0] is the contiguous dimension;
// strides_bytes[i] is the stride of dims[i+1]; elem_strides (optional) are the per-dimension traversal strides.
inline int make_tmap_nd(CUtensorMap* tm, const void* ptr, int dtype, int rank, const unsigned long long* dims,
                        const unsigned long long* strides_bytes, const unsigned* box,
                        const unsigned* elem_strides = nullptr) {
  EncodeTiledFn enc = tmap_encode_fn();
  HYP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
  // The encoder is a driver-API call: make sure the primary context is bound to THIS thread first (a fresh autograd
  // worker thread has only had cudaSetDevice, which binds lazily) -- otherwise it fails with CUDA_ERROR_INVALID_CONTEXT.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { cudaFree(nullptr); ctx_bound = true; }
  const CUtensorMapDataType dt = (dtype == DT_F32)   ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : (dtype == DT_BF16) ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                      : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  HYP_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand pointer must be 16-byte aligned");
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = elem_strides ? elem_strides[i] : 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    HYP_REQUIRE(strides_bytes[i] % 16 == 0, "TMA strides must be multiples of 16 bytes (stride %d = %llu)", i, strides_bytes[i]);
    s[i] = strides_bytes[i];
  }
  CUresult r = enc(tm, dt, rank, const_cast<void*>(ptr), d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HYP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return 0;
}

}  // namespace hyp
