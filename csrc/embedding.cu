// embedding_fwd / embedding_bwd: row gather and fp32 scatter-add.
//
// Forward copies one 16-byte vector per thread (a warp moves 512 contiguous
// bytes of a row).  Backward accumulates into an fp32 gradient table with
// red.global.add.f32 -- the framework's weight gradients live in flat fp32
// buffers, so no dense [vocab, d] temporary is created and rows equal to
// padding_idx are skipped.
//
// Reference call site: nn.Embedding(vocab_size, emb_dim) at
// 02_development/distributed_utils.py:78 (and the HF Llama embed_tokens).
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {

template <typename T>
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const T* __restrict__ table, const long long* __restrict__ idx, T* __restrict__ out,
                     long long n_tok, int d, long long vocab) {
  const int vec_per_row = d / (16 / (int)sizeof(T));
  const long long total = n_tok * vec_per_row;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long tok = i / vec_per_row;
    const int v = static_cast<int>(i - tok * vec_per_row);
    long long id = idx[tok];
    uint4 val = make_uint4(0, 0, 0, 0);
    if (id >= 0 && id < vocab) val = reinterpret_cast<const uint4*>(table + id * d)[v];
    reinterpret_cast<uint4*>(out + tok * d)[v] = val;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const T* __restrict__ dout, const long long* __restrict__ idx, float* __restrict__ dtable,
                     long long n_tok, int d, long long vocab, long long padding_idx) {
  const long long total = n_tok * d;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long tok = i / d;
    const int c = static_cast<int>(i - tok * d);
    const long long id = idx[tok];
    if (id < 0 || id >= vocab || id == padding_idx) continue;
    atomicAdd(dtable + id * d + c, to_f32<T>(dout[i]));
  }
}

}  // namespace hyp

using namespace hyp;

HYP_API int hyp_embedding_fwd(const void* table, const long long* idx, void* out, long long n_tok, int d,
                              long long vocab, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (n_tok == 0) return 0;
  const int es = dtype == DT_F32 ? 4 : 2;
  HYP_REQUIRE((d * es) % 16 == 0, "hyp_embedding_fwd: row bytes must be a multiple of 16");
  const long long total = n_tok * (d * es / 16);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (dtype == DT_F32) embedding_fwd_kernel<float><<<(int)blocks, 256, 0, st>>>((const float*)table, idx, (float*)out, n_tok, d, vocab);
  else if (dtype == DT_BF16) embedding_fwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>((const __nv_bfloat16*)table, idx, (__nv_bfloat16*)out, n_tok, d, vocab);
  else embedding_fwd_kernel<__half><<<(int)blocks, 256, 0, st>>>((const __half*)table, idx, (__half*)out, n_tok, d, vocab);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

HYP_API int hyp_embedding_bwd(const void* dout, const long long* idx, float* dtable, long long n_tok, int d,
                              long long vocab, long long padding_idx, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (n_tok == 0) return 0;
  const long long total = n_tok * d;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (dtype == DT_F32) embedding_bwd_kernel<float><<<(int)blocks, 256, 0, st>>>((const float*)dout, idx, dtable, n_tok, d, vocab, padding_idx);
  else if (dtype == DT_BF16) embedding_bwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>((const __nv_bfloat16*)dout, idx, dtable, n_tok, d, vocab, padding_idx);
  else embedding_bwd_kernel<__half><<<(int)blocks, 256, 0, st>>>((const __half*)dout, idx, dtable, n_tok, d, vocab, padding_idx);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
