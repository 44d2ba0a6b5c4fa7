// colsum.cu — bias gradient: column sums of a [rows, cols] activation-gradient matrix.
//
// ATen's generic reduce_kernel reaches about a quarter of the HBM roofline on these tall,
// narrow-ish bf16 matrices (ncu launch list, profiles/launches_gpt2_microstep_run1.csv: 1.9 ms
// per GPT-2 micro-step for 48 bias gradients).  Here a warp reads 512 contiguous bytes of a row
// (32 lanes x 8 columns), 8 row lanes per CTA walk the rows, every thread keeps 8 fp32 sums; the
// CTA folds its 8 row lanes through shared memory and writes one fp32 partial row; the partial
// rows are folded by dlion_norm_bwd_finalize's kernel.
// reference: the `grad_output.sum(0)` of every biased HF GPT-2 Conv1D in backward, reached from
// accelerator.backward (async_trainer.py:33).
#include <stdlib.h>

#include "common.cuh"

namespace dlion {

constexpr int kColsumParts = kNumSMs * 4;

// FUSED: the CTA that finishes last on a 256-column tile (ticket per tile, self-resetting) folds the
// tile's partial rows and writes — or accumulates into — the final [cols] vector, so the bias gradient
// needs one launch instead of three (partials, fold, add into .grad).
// What is being summed.  A source yields the 8 values of (row, 8-column group) and may write a by-product while
// it has them in registers — the producing elementwise kernel and the bias-gradient reduction become one pass.
// Ordered 16-byte accesses: `asm volatile` statements keep their program order among themselves, so a batch of
// these loads stays AHEAD of the first of these stores — ptxas otherwise sinks the later rows' loads below the
// earlier rows' math to save registers and one row at a time is in flight (measured: 4.2 TB/s).
__device__ __forceinline__ uint4 ldg_stream16(const void* ptr) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(ptr) : "memory");
  return r;
}
__device__ __forceinline__ void stg_stream16(void* ptr, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
template <typename T>
__device__ __forceinline__ Raw8<T> load_stream8(const T* __restrict__ ptr) {
  Raw8<T> r;
#pragma unroll
  for (int i = 0; i < (int)(sizeof(T) * 8 / 16); ++i) r.q[i] = ldg_stream16(reinterpret_cast<const uint4*>(ptr) + i);
  return r;
}
template <typename T>
__device__ __forceinline__ void store_stream8(T* __restrict__ ptr, const float (&v)[8]) {
  if constexpr (sizeof(T) == 4) {
    stg_stream16(ptr, make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3])));
    stg_stream16(ptr + 4, make_uint4(__float_as_uint(v[4]), __float_as_uint(v[5]), __float_as_uint(v[6]), __float_as_uint(v[7])));
  } else {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (DT<T>::id == DLION_BF16) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
      } else {
        const __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
      }
    }
    stg_stream16(ptr, make_uint4(w[0], w[1], w[2], w[3]));
  }
}

// A source is split in two so that the kernel can put the raw loads of several rows in flight before the first
// dependent instruction (and before any store: a by-product store between two rows' loads would pin their order):
//   load(r, cg)            -> Raw   the 16-byte global loads of one (row, column group), nothing else
//   finish(raw, r, cg, v)           convert / compute, store the by-product, leave the 8 values to sum in v
template <typename T>
struct PlainSource {                     // x[rows, cols] as stored
  const T* __restrict__ x;
  int cols;
  struct Raw { Raw8<T> a; };
  __device__ __forceinline__ Raw load(int64_t r, int cg) const { return Raw{load_stream8<T>(x + r * cols + cg * 8)}; }
  __device__ __forceinline__ void finish(const Raw& raw, int64_t, int, float (&v)[8]) const { unpack_raw8<T>(raw.a, v); }
};
// d(pre) = dy * gelu_tanh'(pre), written to dpre, and summed over rows: GPT-2's c_fc bias gradient falls out of
// the activation backward (reference: HF NewGELUActivation backward + Conv1D bias grad, two passes over dY there)
template <typename T>
struct GeluBwdSource {
  const T* __restrict__ pre;
  const T* __restrict__ dy;
  T* __restrict__ dpre;
  int cols;
  struct Raw { Raw8<T> x, d; };
  __device__ __forceinline__ Raw load(int64_t r, int cg) const {
    const int64_t off = r * cols + cg * 8;
    return Raw{load_stream8<T>(pre + off), load_stream8<T>(dy + off)};
  }
  __device__ __forceinline__ void finish(const Raw& raw, int64_t r, int cg, float (&d)[8]) const {
    constexpr float kK = 0.7978845608028654f, kC = 0.044715f;
    float f[8];
    unpack_raw8<T>(raw.x, f);
    unpack_raw8<T>(raw.d, d);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float x2 = f[i] * f[i];
      float t;
      const float u = kK * fmaf(kC * x2, f[i], f[i]);
      if constexpr (sizeof(T) == 2) asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
      else t = tanhf(u);
      const float du = kK * fmaf(3.0f * kC, x2, 1.0f);
      d[i] *= 0.5f * (1.0f + t) + 0.5f * f[i] * (1.0f - t * t) * du;
    }
    store_stream8<T>(dpre + r * cols + cg * 8, d);
    round8<T>(d);                        // sum what was stored: the same bias gradient a separate pass would give
  }
};
// d(qkv)[B, S, 3, H, D] packed from attention's dq, dk, dv ([B, H, S, D] with arbitrary b/h/s strides, d contiguous)
// and summed over (b, s): replaces the stack/cat kernel of unbind's backward AND c_attn's bias-gradient pass.
template <typename T>
struct QkvPackSource {
  const T* __restrict__ src[3];
  int64_t sb[3], sh[3], ss[3];
  T* __restrict__ out;                   // [B*S, 3*H*D]
  int S, H, D, cols;
  struct Raw { Raw8<T> a; };
  __device__ __forceinline__ Raw load(int64_t r, int cg) const {
    const int col = cg * 8;
    const int hd = H * D;
    const int t = col / hd, rem = col - t * hd;
    const int h = rem / D, d = rem - h * D;
    const int64_t b = r / S, s = r - b * S;
    // selects, not indexing: a runtime index would put the parameter arrays in local memory
    const T* base = t == 0 ? src[0] : (t == 1 ? src[1] : src[2]);
    const int64_t ob = t == 0 ? sb[0] : (t == 1 ? sb[1] : sb[2]);
    const int64_t oh = t == 0 ? sh[0] : (t == 1 ? sh[1] : sh[2]);
    const int64_t os = t == 0 ? ss[0] : (t == 1 ? ss[1] : ss[2]);
    return Raw{load_stream8<T>(base + b * ob + h * oh + s * os + d)};
  }
  __device__ __forceinline__ void finish(const Raw& raw, int64_t r, int cg, float (&v)[8]) const {
    T* dst = out + r * cols + cg * 8;
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) * 8 / 16); ++i) stg_stream16(reinterpret_cast<uint4*>(dst) + i, raw.a.q[i]);   // bit copy
    unpack_raw8<T>(raw.a, v);
  }
};

// rows whose loads a thread puts in flight before the math (DLION_COLSUM_K = 1 | 2 | 4 overrides, for tuning)
static int colsum_rows_in_flight() {
  static const int k = [] {
    const char* v = getenv("DLION_COLSUM_K");
    const int x = v ? atoi(v) : 4;
    return x == 1 || x == 2 ? x : 4;
  }();
  return k;
}

template <typename T, bool FUSED, typename Source, int K>
__global__ void __launch_bounds__(256, K == 2 ? 6 : 4) colsum_kernel(const Source src, float* __restrict__ partial,
                                                     int64_t rows, int cols, unsigned int* __restrict__ tickets,
                                                     T* __restrict__ out, int accumulate) {
  __shared__ float sm[8][32][9];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int cg = blockIdx.x * 32 + lane;          // 8-column group
  const bool live = cg * 8 < cols;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (live) {
    const int64_t step = (int64_t)gridDim.y * 8;
    int64_t r = (int64_t)blockIdx.y * 8 + rl;
    for (; r + (K - 1) * step < rows; r += K * step) {      // K rows' loads in flight per thread, then the math
      typename Source::Raw raw[K];
#pragma unroll
      for (int q = 0; q < K; ++q) raw[q] = src.load(r + q * step, cg);
#pragma unroll
      for (int q = 0; q < K; ++q) {
        float a[8];
        src.finish(raw[q], r + q * step, cg, a);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += a[i];
      }
    }
    for (; r < rows; r += step) {
      float a[8];
      const typename Source::Raw raw = src.load(r, cg);
      src.finish(raw, r, cg, a);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += a[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) sm[rl][lane][i] = acc[i];
  __syncthreads();
  // thread t folds column (t) of this block's 256 columns over the 8 row lanes
  const int c = threadIdx.x;
  const int col = blockIdx.x * 256 + c;
  if (col < cols) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += sm[q][c >> 3][c & 7];
    partial[(int64_t)blockIdx.y * cols + col] = s;
  }
  if constexpr (FUSED) {
    __threadfence();                               // this CTA's partial row is visible before its ticket
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int t = atomicAdd(&tickets[blockIdx.x], 1u);
      is_last = (t == gridDim.y - 1);
      if (is_last) tickets[blockIdx.x] = 0u;       // ready for the next call on this stream
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();                               // acquire: the other CTAs' partial rows
    float tot[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (live) {
      for (int p = rl; p < (int)gridDim.y; p += 8) {
        const float4* src = reinterpret_cast<const float4*>(partial + (int64_t)p * cols + cg * 8);
        const float4 u = __ldcg(src), v = __ldcg(src + 1);          // L2: written by other SMs
        tot[0] += u.x; tot[1] += u.y; tot[2] += u.z; tot[3] += u.w;
        tot[4] += v.x; tot[5] += v.y; tot[6] += v.z; tot[7] += v.w;
      }
    }
    __syncthreads();                               // sm is reused
#pragma unroll
    for (int i = 0; i < 8; ++i) sm[rl][lane][i] = tot[i];
    __syncthreads();
    if (col < cols) {
      float s = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) s += sm[q][c >> 3][c & 7];
      if (accumulate) s += DT<T>::to_f(out[col]);
      out[col] = DT<T>::from_f(s);
    }
  }
}

}  // namespace dlion

using namespace dlion;

// expects `grid`, `s`, `partial`, `rows`, `cols` in scope
#define COLSUM_LAUNCH(FUSED_, SRC_T, SRC, TICKETS, OUT, ACC)                                                          \
  do {                                                                                                                \
    const int k_ = colsum_rows_in_flight();                                                                           \
    if (k_ == 1) colsum_kernel<T, FUSED_, SRC_T, 1><<<grid, 256, 0, s>>>(SRC, partial, rows, cols, TICKETS, OUT, ACC);      \
    else if (k_ == 2) colsum_kernel<T, FUSED_, SRC_T, 2><<<grid, 256, 0, s>>>(SRC, partial, rows, cols, TICKETS, OUT, ACC); \
    else colsum_kernel<T, FUSED_, SRC_T, 4><<<grid, 256, 0, s>>>(SRC, partial, rows, cols, TICKETS, OUT, ACC);              \
  } while (0)

extern "C" int dlion_colsum_parts(int64_t rows, int cols) {
  const int gx = (cols + 255) / 256;
  const int ctas = colsum_rows_in_flight() == 2 ? kNumSMs * 6 : kColsumParts;     // one resident wave
  int gy = ctas / (gx > 0 ? gx : 1);
  const int64_t max_y = (rows + 7) / 8;
  if (gy > max_y) gy = (int)max_y;
  return gy < 1 ? 1 : gy;
}

extern "C" int dlion_colsum(const void* x, int64_t rows, int cols, float* partial, int dtype, void* stream) {
  DLION_CHECK_ARG(rows >= 0 && cols > 0 && (cols & 7) == 0, "dlion_colsum: cols (%d) must be a positive multiple of 8", cols);
  DLION_CHECK_ARG(aligned16(x) && partial != nullptr, "dlion_colsum: bad pointers");
  if (rows == 0) return 0;
  const dim3 grid((cols + 255) / 256, dlion_colsum_parts(rows, cols));
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, COLSUM_LAUNCH(false, PlainSource<T>, (PlainSource<T>{(const T*)x, cols}), nullptr, (T*)nullptr, 0));
  return DLION_LAUNCH_CHECK("dlion_colsum");
}

extern "C" int dlion_colsum_tiles(int cols) { return (cols + 255) / 256; }

extern "C" int dlion_colsum_fused(const void* x, int64_t rows, int cols, float* partial, unsigned int* tickets,
                                  void* out, int accumulate, int dtype, void* stream) {
  DLION_CHECK_ARG(rows > 0 && cols > 0 && (cols & 7) == 0, "dlion_colsum_fused: cols (%d) must be a positive multiple of 8", cols);
  DLION_CHECK_ARG(aligned16(x) && partial != nullptr && tickets != nullptr && out != nullptr,
                  "dlion_colsum_fused: bad pointers");
  const dim3 grid((cols + 255) / 256, dlion_colsum_parts(rows, cols));
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, COLSUM_LAUNCH(true, PlainSource<T>, (PlainSource<T>{(const T*)x, cols}), tickets, (T*)out, accumulate));
  return DLION_LAUNCH_CHECK("dlion_colsum_fused");
}

extern "C" int dlion_gelu_tanh_bwd_colsum(const void* pre, const void* dy, void* dpre, int64_t rows, int cols,
                                          float* partial, unsigned int* tickets, void* out, int accumulate, int dtype,
                                          void* stream) {
  DLION_CHECK_ARG(rows > 0 && cols > 0 && (cols & 7) == 0, "dlion_gelu_tanh_bwd_colsum: cols (%d) must be a positive multiple of 8", cols);
  DLION_CHECK_ARG(aligned16(pre) && aligned16(dy) && aligned16(dpre) && partial != nullptr && tickets != nullptr && out != nullptr,
                  "dlion_gelu_tanh_bwd_colsum: bad pointers");
  const dim3 grid((cols + 255) / 256, dlion_colsum_parts(rows, cols));
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, COLSUM_LAUNCH(true, GeluBwdSource<T>, (GeluBwdSource<T>{(const T*)pre, (const T*)dy, (T*)dpre, cols}),
                                      tickets, (T*)out, accumulate));
  return DLION_LAUNCH_CHECK("dlion_gelu_tanh_bwd_colsum");
}

extern "C" int dlion_qkv_pack_colsum(const void* dq, const void* dk, const void* dv, const int64_t* strides_bhs,
                                     void* dqkv, int batch, int seq, int heads, int head_dim, float* partial,
                                     unsigned int* tickets, void* out, int accumulate, int dtype, void* stream) {
  const int cols = 3 * heads * head_dim;
  const int64_t rows = (int64_t)batch * seq;
  DLION_CHECK_ARG(rows > 0 && heads > 0 && head_dim > 0 && (head_dim & 7) == 0,
                  "dlion_qkv_pack_colsum: head_dim (%d) must be a positive multiple of 8", head_dim);
  DLION_CHECK_ARG(aligned16(dq) && aligned16(dk) && aligned16(dv) && aligned16(dqkv) && strides_bhs != nullptr &&
                      partial != nullptr && tickets != nullptr && out != nullptr,
                  "dlion_qkv_pack_colsum: bad pointers");
  for (int i = 0; i < 9; ++i)
    DLION_CHECK_ARG((strides_bhs[i] & 7) == 0, "dlion_qkv_pack_colsum: strides must be multiples of 8 elements");
  const dim3 grid((cols + 255) / 256, dlion_colsum_parts(rows, cols));
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, {
    QkvPackSource<T> src;
    const void* ptrs[3] = {dq, dk, dv};
    for (int t = 0; t < 3; ++t) {
      src.src[t] = (const T*)ptrs[t];
      src.sb[t] = strides_bhs[3 * t];
      src.sh[t] = strides_bhs[3 * t + 1];
      src.ss[t] = strides_bhs[3 * t + 2];
    }
    src.out = (T*)dqkv; src.S = seq; src.H = heads; src.D = head_dim; src.cols = cols;
    COLSUM_LAUNCH(true, QkvPackSource<T>, src, tickets, (T*)out, accumulate);
  });
  return DLION_LAUNCH_CHECK("dlion_qkv_pack_colsum");
}
