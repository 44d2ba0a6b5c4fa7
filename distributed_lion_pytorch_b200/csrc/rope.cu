// rope.cu — N4: rotary position embedding, in place on q and k, forward and backward.
//
// HF "rotate_half" pairing: element i of the first half of a head is rotated with element
// i + head_dim/2.  One thread handles 8 consecutive first-half elements and their 8 partners
// (two 16-byte loads, two 16-byte stores); cos/sin come from an fp32 [seq, head_dim/2] table,
// so the rotation is done in fp32 with a single rounding (HF rounds cos/sin to bf16 first).
// The backward is the same kernel with the rotation transposed (sin -> -sin).
// reference: HF LlamaAttention.apply_rotary_pos_emb reached through sft_llama2.py:147 /
// dpo_llama2.py:133 models — 6 elementwise kernels + 2 cats per layer upstream.
#include "common.cuh"

namespace dlion {

template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(T* __restrict__ q, T* __restrict__ k,
                                                   const float* __restrict__ cosT, const float* __restrict__ sinT,
                                                   int64_t tokens, int seq_len, int q_heads, int k_heads, int head_dim,
                                                   int64_t q_tok_stride, int64_t k_tok_stride, float sin_sign) {
  const int half = head_dim >> 1;
  const int vph = half >> 3;                          // 8-element vectors per half head
  const int heads = q_heads + k_heads;
  const int64_t total = tokens * heads * vph;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int v = (int)(idx % vph);
    const int64_t th = idx / vph;
    const int h = (int)(th % heads);
    const int64_t tok = th / heads;
    const int pos = (int)(tok % seq_len);
    T* base = h < q_heads ? q + tok * q_tok_stride + (int64_t)h * head_dim
                          : k + tok * k_tok_stride + (int64_t)(h - q_heads) * head_dim;
    float x1[8], x2[8], c[8], s[8];
    load8<T>(base + v * 8, x1);
    load8<T>(base + half + v * 8, x2);
    load8<float>(cosT + (int64_t)pos * half + v * 8, c);
    load8<float>(sinT + (int64_t)pos * half + v * 8, s);
    float o1[8], o2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float sn = s[i] * sin_sign;
      o1[i] = x1[i] * c[i] - x2[i] * sn;
      o2[i] = x2[i] * c[i] + x1[i] * sn;
    }
    store8<T>(base + v * 8, o1);
    store8<T>(base + half + v * 8, o2);
  }
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_rope(void* q, void* k, const float* cos, const float* sin, int64_t tokens, int seq_len,
                          int q_heads, int k_heads, int head_dim, int64_t q_token_stride, int64_t k_token_stride,
                          int inverse, int dtype, void* stream) {
  DLION_CHECK_ARG(tokens >= 0 && seq_len > 0 && q_heads >= 0 && k_heads >= 0, "dlion_rope: bad shape");
  DLION_CHECK_ARG(head_dim > 0 && head_dim % 16 == 0, "dlion_rope: head_dim (%d) must be a multiple of 16", head_dim);
  DLION_CHECK_ARG(q_token_stride % 8 == 0 && k_token_stride % 8 == 0, "dlion_rope: token strides must be multiples of 8");
  DLION_CHECK_ARG(aligned16(q) && aligned16(k) && aligned16(cos) && aligned16(sin), "dlion_rope: unaligned pointer");
  DLION_CHECK_ARG((q_heads == 0 || q) && (k_heads == 0 || k), "dlion_rope: null tensor");
  const int64_t total = tokens * (q_heads + k_heads) * (head_dim / 16);
  if (total == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  int64_t want = (total + 255) / 256;
  const int grid = (int)(want > kNumSMs * 16 ? kNumSMs * 16 : want);
  const float sgn = inverse ? -1.f : 1.f;
  DISPATCH_DTYPE(dtype, rope_kernel<T><<<grid, 256, 0, s>>>((T*)q, (T*)k, cos, sin, tokens, seq_len, q_heads, k_heads,
                                                             head_dim, q_token_stride, k_token_stride, sgn));
  return DLION_LAUNCH_CHECK("dlion_rope");
}
