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
//   dequantize the hot one (every forward and backward of a 4-bit layer): a lane turns one packed
//              word into 8 outputs (one 16 B store in bf16, 512 contiguous bytes per warp) through a
//              bank-conflict-free per-lane copy of the code book in shared memory, so an element
//              costs 0.5 B read + 2 B written, one LDS.32 and one multiply.
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

__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

// One packed word -> 8 outputs.  `col` is the shared-memory byte address of this lane's column of the
// code-book table; the table is 2 KiB-aligned, so bits 7..10 of `col` are zero and the row offset
// (code index * 128 B) is OR-ed in: shift, mask-or, LDS, multiply per element and nothing else.
template <typename T>
__device__ __forceinline__ void q4_expand_word(uint32_t w, float a, uint32_t col, T* __restrict__ dst) {
  float o[8];
#pragma unroll
  for (int b = 0; b < 4; ++b) {                          // byte b of the word: high nibble first
    const uint32_t hi = 8 * b + 4, lo = 8 * b;
    const uint32_t rh = hi >= 7 ? (w >> (hi - 7)) : (w << (7 - hi));
    const uint32_t rl = lo >= 7 ? (w >> (lo - 7)) : (w << (7 - lo));
    o[2 * b] = lds_f32((rh & 0x780u) | col) * a;
    o[2 * b + 1] = lds_f32((rl & 0x780u) | col) * a;
  }
  store8<T>(dst, o);
}

template <typename T>
__global__ void __launch_bounds__(256) q4_dequantize_kernel(const uint8_t* __restrict__ packed,
                                                            const float* __restrict__ absmax, T* __restrict__ out,
                                                            int64_t nwords, int log2_words_per_block,
                                                            const __grid_constant__ Code16 code) {
  // tab[c][lane]: every lane owns a private column of the 16-entry code book, so a warp's 32 look-ups
  // always hit 32 different banks (a byte-indexed table shared by the warp cost ~6 wavefronts per access).
  // The 2 KiB table must start on a 2 KiB boundary of the shared WINDOW (q4_expand_word ORs offsets in);
  // static shared memory starts after the driver's reserved kilobyte, so the boundary is found at run time
  // inside a 4 KiB array instead of trusting __align__.
  __shared__ float raw[2 * 16 * 32];
  const uint32_t tab = ((uint32_t)__cvta_generic_to_shared(raw) + 2047u) & ~2047u;
  for (int i = threadIdx.x; i < 16 * 32; i += blockDim.x)
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(tab + 4u * i), "f"(code.v[i >> 5]) : "memory");
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const uint32_t col = tab + 4u * lane;
  const uint32_t* words = reinterpret_cast<const uint32_t*>(packed);
  // A lane handles one 4-byte word (8 elements) per step, so each warp-wide store covers 512 contiguous
  // bytes of bf16 output (whole sectors); four independent steps per trip, and the next trip prefetched.
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nfull = nwords >> 7;                      // chunks of 128 words with no ragged edge
  // software pipeline: the next trip's words and scales are in flight while this trip is expanded
  uint32_t w[4], wn[4];
  float a[4], an[4];
  int64_t c = warp;
  if (c < nfull) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t wi = (c << 7) + lane + k * 32;
      w[k] = __ldg(words + wi);
      a[k] = __ldg(absmax + (wi >> log2_words_per_block));
    }
  }
  while (c < nfull) {
    const int64_t cn = c + nwarps;
    if (cn < nfull) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t wi = (cn << 7) + lane + k * 32;
        wn[k] = __ldg(words + wi);
        an[k] = __ldg(absmax + (wi >> log2_words_per_block));
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) q4_expand_word<T>(w[k], a[k], col, out + ((c << 7) + lane + k * 32) * 8);
#pragma unroll
    for (int k = 0; k < 4; ++k) { w[k] = wn[k]; a[k] = an[k]; }
    c = cn;
  }
  if (warp == nwarps - 1) {                               // ragged tail (< 128 words), one warp
    for (int64_t wi = (nfull << 7) + lane; wi < nwords; wi += 32)
      q4_expand_word<T>(__ldg(words + wi), __ldg(absmax + (wi >> log2_words_per_block)), col, out + wi * 8);
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
  DLION_CHECK_ARG(n >= 0 && blocksize >= 8 && (blocksize & (blocksize - 1)) == 0,
                  "dlion_q4_dequantize: blocksize (%d) must be a power of two >= 8", blocksize);
  DLION_CHECK_ARG(n % blocksize == 0, "dlion_q4_dequantize: n (%lld) must be a multiple of blocksize (%d)",
                  (long long)n, blocksize);
  DLION_CHECK_ARG((n == 0) || (packed && absmax && out), "dlion_q4_dequantize: null tensor");
  DLION_CHECK_ARG(aligned16(packed) && aligned16(out), "dlion_q4_dequantize: unaligned pointer");
  Code16 c;
  if (int rc = load_code(code16, &c, "dlion_q4_dequantize")) return rc;
  if (n == 0) return 0;
  const int64_t nwords = n / 8;                               // a lane expands one 4-byte word per step
  const int64_t want = (nwords + 1023) / 1024;                // 8 warps x 128 words per CTA trip
  cudaStream_t s = (cudaStream_t)stream;
  int lg = 0;
  while ((8 << lg) < blocksize) ++lg;                         // words (8 elements) per absmax block, as a shift
  DISPATCH_DTYPE(dtype, {
    static int per_sm = 0;                                    // one resident wave: no partial second wave
    if (per_sm == 0) {
      int b = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, q4_dequantize_kernel<T>, 256, 0) != cudaSuccess || b < 1) b = 4;
      per_sm = b;
    }
    const int64_t cap = (int64_t)kNumSMs * per_sm;
    const int grid = (int)(want > cap ? cap : want);
    q4_dequantize_kernel<T><<<grid, 256, 0, s>>>((const uint8_t*)packed, absmax, (T*)out, nwords, lg, c);
  });
  return DLION_LAUNCH_CHECK("dlion_q4_dequantize");
}
