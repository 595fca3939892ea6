// rope: rotary position embedding (rotate-half convention of HF Llama), in
// place on a [B, S, H, D] or [B, H, S, D] tensor described by strides.  The
// angle is recomputed from (position, pair index) -- no cos/sin tables are read.
// Backward is the same kernel with the sign of sin flipped (the rotation is
// orthogonal), selected by `inverse`.
//
// Reference call site: apply_rotary_pos_emb inside the HF LlamaAttention
// reached through AutoModelForCausalLM (02_development/distributed_utils.py:465-468).
#include "common/ptx.cuh"
#include "common/util.cuh"

namespace hyp {

template <typename T>
__global__ void __launch_bounds__(256)
rope_kernel(T* __restrict__ x, int B, int S, int H, int D, long long sb, long long ss, long long sh,
            const long long* __restrict__ positions, float log2_theta, int inverse) {
  const int half = D / 2;
  const long long total = static_cast<long long>(B) * S * H * half;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int j = static_cast<int>(i % half);
    long long r = i / half;
    const int h = static_cast<int>(r % H); r /= H;
    const int s = static_cast<int>(r % S);
    const int b = static_cast<int>(r / S);
    const float pos = positions ? static_cast<float>(positions[static_cast<long long>(b) * S + s]) : static_cast<float>(s);
    const float inv_freq = exp2f(-log2_theta * (2.f * j) / D);
    float sn, cs;
    sincosf(pos * inv_freq, &sn, &cs);
    if (inverse) sn = -sn;
    T* p = x + b * sb + s * ss + h * sh;
    const float x1 = to_f32<T>(p[j]);
    const float x2 = to_f32<T>(p[j + half]);
    p[j] = from_f32<T>(x1 * cs - x2 * sn);
    p[j + half] = from_f32<T>(x2 * cs + x1 * sn);
  }
}

}  // namespace hyp

using namespace hyp;

HYP_API int hyp_rope(void* x, int B, int S, int H, int D, long long stride_b, long long stride_s, long long stride_h,
                     const long long* positions, float theta, int inverse, int dtype, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  HYP_REQUIRE(D % 2 == 0 && B > 0 && S > 0 && H > 0, "hyp_rope: bad shape");
  const long long total = (long long)B * S * H * (D / 2);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  const float l2t = log2f(theta);
  if (dtype == DT_F32) rope_kernel<float><<<(int)blocks, 256, 0, st>>>((float*)x, B, S, H, D, stride_b, stride_s, stride_h, positions, l2t, inverse);
  else if (dtype == DT_BF16) rope_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>((__nv_bfloat16*)x, B, S, H, D, stride_b, stride_s, stride_h, positions, l2t, inverse);
  else rope_kernel<__half><<<(int)blocks, 256, 0, st>>>((__half*)x, B, S, H, D, stride_b, stride_s, stride_h, positions, l2t, inverse);
  HYP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
