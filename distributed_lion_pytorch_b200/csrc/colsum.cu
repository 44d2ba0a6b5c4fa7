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
#include "common.cuh"

namespace dlion {

constexpr int kColsumParts = kNumSMs * 4;

// FUSED: the CTA that finishes last on a 256-column tile (ticket per tile, self-resetting) folds the
// tile's partial rows and writes — or accumulates into — the final [cols] vector, so the bias gradient
// needs one launch instead of three (partials, fold, add into .grad).
// What is being summed.  A source yields the 8 values of (row, 8-column group) and may write a by-product while
// it has them in registers — the producing elementwise kernel and the bias-gradient reduction become one pass.
template <typename T>
struct PlainSource {                     // x[rows, cols] as stored
  const T* __restrict__ x;
  int cols;
  __device__ __forceinline__ void fetch(int64_t r, int cg, float (&v)[8]) const { load8<T>(x + r * cols + cg * 8, v); }
};
// d(pre) = dy * gelu_tanh'(pre), written to dpre, and summed over rows: GPT-2's c_fc bias gradient falls out of
// the activation backward (reference: HF NewGELUActivation backward + Conv1D bias grad, two passes over dY there)
template <typename T>
struct GeluBwdSource {
  const T* __restrict__ pre;
  const T* __restrict__ dy;
  T* __restrict__ dpre;
  int cols;
  __device__ __forceinline__ void fetch(int64_t r, int cg, float (&d)[8]) const {
    constexpr float kK = 0.7978845608028654f, kC = 0.044715f;
    float f[8];
    const int64_t off = r * cols + cg * 8;
    load8<T>(pre + off, f);
    load8<T>(dy + off, d);
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
    store8<T>(dpre + off, d);
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
  __device__ __forceinline__ void fetch(int64_t r, int cg, float (&v)[8]) const {
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
    load8<T>(base + b * ob + h * oh + s * os + d, v);
    store8<T>(out + r * cols + col, v);
  }
};

template <typename T, bool FUSED, typename Source>
__global__ void __launch_bounds__(256) colsum_kernel(const Source src, float* __restrict__ partial,
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
    for (; r + step < rows; r += 2 * step) {      // two rows in flight per thread
      float a[8], b[8];
      src.fetch(r, cg, a);
      src.fetch(r + step, cg, b);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += a[i] + b[i];
    }
    if (r < rows) {
      float a[8];
      src.fetch(r, cg, a);
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

extern "C" int dlion_colsum_parts(int64_t rows, int cols) {
  const int gx = (cols + 255) / 256;
  int gy = kColsumParts / (gx > 0 ? gx : 1);
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
  DISPATCH_DTYPE(dtype, (colsum_kernel<T, false, PlainSource<T>><<<grid, 256, 0, s>>>(PlainSource<T>{(const T*)x, cols},
                                                                                     partial, rows, cols, nullptr, nullptr, 0)));
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
  DISPATCH_DTYPE(dtype, (colsum_kernel<T, true, PlainSource<T>><<<grid, 256, 0, s>>>(PlainSource<T>{(const T*)x, cols},
                                                                                    partial, rows, cols, tickets, (T*)out,
                                                                                    accumulate)));
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
  DISPATCH_DTYPE(dtype, (colsum_kernel<T, true, GeluBwdSource<T>><<<grid, 256, 0, s>>>(
                            GeluBwdSource<T>{(const T*)pre, (const T*)dy, (T*)dpre, cols}, partial, rows, cols, tickets,
                            (T*)out, accumulate)));
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
    colsum_kernel<T, true, QkvPackSource<T>><<<grid, 256, 0, s>>>(src, partial, rows, cols, tickets, (T*)out, accumulate);
  });
  return DLION_LAUNCH_CHECK("dlion_qkv_pack_colsum");
}
