// quant4.cu — blockwise 4-bit weight codec for frozen LoRA base weights (NF4 / FP4 code books).
//
// Storage layout (the one bitsandbytes uses, so packed tensors interchange): the flattened weight
// is cut into blocks of `blocksize` elements; each block keeps one fp32 absmax and every element
// the index (0..15) of the code-book entry nearest to x / absmax; element 2i sits in the HIGH
// nibble of byte i, element 2i+1 in the low nibble.  The code book is a kernel parameter, so the
// same two kernels serve NF4 (normal-float quantiles, QLoRA) and FP4 (e2m1).
//
//   quantize   one warp per block: warp-shuffle absmax, 16-way nearest search, lanes emit bytes.
//              Runs once at load time.
//   dequantize the hot one (every forward and backward of a 4-bit layer): one thread turns 16
//              packed bytes (one 16 B load) into 32 outputs (four 16 B stores in bf16) through a
//              256-entry byte -> (hi, lo) table in shared memory, so an element costs 0.5 B read
//              + 2 B written and no arithmetic beyond one multiply.
//
// reference: BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_quant_type="nf4",
// bnb_4bit_compute_dtype=bfloat16) at sft_llama2.py:141-145 and load_in_4bit=True at
// dpo_llama2.py:137,151 (bitsandbytes CUDA kernels upstream; not installed on this image).
#include "common.cuh"

namespace dlion {

struct Code16 {
  float v[16];
};

template <typename T>
__global__ void __launch_bounds__(256) q4_quantize_kernel(const T* __restrict__ w, uint8_t* __restrict__ packed,
                                                          float* __restrict__ absmax, int64_t nblocks,
                                                          int blocksize, const __grid_constant__ Code16 code) {
  const int lane = threadIdx.x & 31;
  const int per = blocksize >> 5;                     // elements per lane, even (blocksize % 64 == 0)
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t blk = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; blk < nblocks; blk += warps) {
    const T* src = w + blk * blocksize + lane * per;
    float amax = 0.f;
    for (int i = 0; i < per; ++i) amax = fmaxf(amax, fabsf(DT<T>::to_f(src[i])));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0) absmax[blk] = amax;
    uint8_t* dst = packed + ((blk * blocksize + lane * per) >> 1);
    for (int i = 0; i < per; i += 2) {
      uint32_t nib[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float x = DT<T>::to_f(src[i + h]);
        const float xn = amax > 0.f ? x / amax : 0.f;          // IEEE division: matches the torch oracle
        float best = fabsf(xn - code.v[0]);
        uint32_t idx = 0;
#pragma unroll
        for (int c = 1; c < 16; ++c) {
          const float d = fabsf(xn - code.v[c]);
          if (d < best) { best = d; idx = c; }                 // first minimum wins, like argmin
        }
        nib[h] = idx;
      }
      dst[i >> 1] = (uint8_t)((nib[0] << 4) | nib[1]);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) q4_dequantize_kernel(const uint8_t* __restrict__ packed,
                                                            const float* __restrict__ absmax, T* __restrict__ out,
                                                            int64_t nvec, int blocksize, const __grid_constant__ Code16 code) {
  __shared__ float2 lut[256];
  lut[threadIdx.x] = make_float2(code.v[threadIdx.x >> 4], code.v[threadIdx.x & 15]);
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(packed) + v);        // 16 bytes = 32 elements
    const float a = __ldg(absmax + (v * 32) / blocksize);
    const uint32_t wds[4] = {q.x, q.y, q.z, q.w};
    T* dst = out + v * 32;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float o[8];
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float2 c = lut[(wds[k] >> (8 * b)) & 0xffu];
        o[2 * b] = c.x * a;
        o[2 * b + 1] = c.y * a;
      }
      store8<T>(dst + 8 * k, o);
    }
  }
}

}  // namespace dlion

using namespace dlion;

static int load_code(const float* code16, Code16* c, const char* who) {
  DLION_CHECK_ARG(code16 != nullptr, "%s: code book is null", who);
  for (int i = 0; i < 16; ++i) c->v[i] = code16[i];
  return 0;
}

extern "C" int dlion_q4_quantize(const void* w, void* packed, float* absmax, int64_t n, int blocksize,
                                 const float* code16, int dtype, void* stream) {
  DLION_CHECK_ARG(n >= 0 && blocksize >= 64 && blocksize % 64 == 0 && blocksize <= 4096,
                  "dlion_q4_quantize: blocksize (%d) must be a multiple of 64, at most 4096", blocksize);
  DLION_CHECK_ARG(n % blocksize == 0, "dlion_q4_quantize: n (%lld) must be a multiple of blocksize (%d)",
                  (long long)n, blocksize);
  DLION_CHECK_ARG((n == 0) || (w && packed && absmax), "dlion_q4_quantize: null tensor");
  Code16 c;
  if (int rc = load_code(code16, &c, "dlion_q4_quantize")) return rc;
  if (n == 0) return 0;
  const int64_t nblocks = n / blocksize;
  const int64_t want = (nblocks + 7) / 8;                     // 8 warps per CTA
  const int grid = (int)(want > kNumSMs * 8 ? kNumSMs * 8 : want);
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, q4_quantize_kernel<T><<<grid, 256, 0, s>>>((const T*)w, (uint8_t*)packed, absmax, nblocks,
                                                                   blocksize, c));
  return DLION_LAUNCH_CHECK("dlion_q4_quantize");
}

extern "C" int dlion_q4_dequantize(const void* packed, const float* absmax, void* out, int64_t n, int blocksize,
                                   const float* code16, int dtype, void* stream) {
  DLION_CHECK_ARG(n >= 0 && blocksize >= 32 && blocksize % 32 == 0,
                  "dlion_q4_dequantize: blocksize (%d) must be a multiple of 32", blocksize);
  DLION_CHECK_ARG(n % blocksize == 0, "dlion_q4_dequantize: n (%lld) must be a multiple of blocksize (%d)",
                  (long long)n, blocksize);
  DLION_CHECK_ARG((n == 0) || (packed && absmax && out), "dlion_q4_dequantize: null tensor");
  DLION_CHECK_ARG(aligned16(packed) && aligned16(out), "dlion_q4_dequantize: unaligned pointer");
  Code16 c;
  if (int rc = load_code(code16, &c, "dlion_q4_dequantize")) return rc;
  if (n == 0) return 0;
  const int64_t nvec = n / 32;
  const int64_t want = (nvec + 255) / 256;
  const int grid = (int)(want > kNumSMs * 16 ? kNumSMs * 16 : want);
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, q4_dequantize_kernel<T><<<grid, 256, 0, s>>>((const uint8_t*)packed, absmax, (T*)out, nvec,
                                                                     blocksize, c));
  return DLION_LAUNCH_CHECK("dlion_q4_dequantize");
}
