// cross_entropy.cu — N5: cross-entropy over a chunk of logits rows, forward + backward in place.
//
// The Python side (ops/cross_entropy.py) tiles the token dimension, so only one
// [chunk, vocab] logits block exists at a time and the full [tokens, vocab] logits (and their
// fp32 upcast) are never materialised.  For each row one CTA makes an online (max, sum-exp)
// pass, then a second pass that overwrites the logits with (softmax - onehot) * scale — the
// second pass re-reads the row from L2 (a row is <= 100 KB, 148 CTAs' rows fit in the 126 MB
// L2 many times over).  Rows are 16-byte aligned because the caller pads the vocab stride.
// reference: the HF ForCausalLM loss (shift, fp32 upcast of the full logits,
// F.cross_entropy) inside compute_loss (async_trainer.py:24); TRL's log-prob gather for DPO.
#include <math_constants.h>
#include <stdlib.h>

#include "common.cuh"

namespace dlion {

constexpr int kCEThreads = 512;

// one MUFU.EX2; inputs here are <= 0 (x - max), flush-to-zero on underflow is exactly what softmax wants
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct MaxSum {
  float m, s;
};
__device__ __forceinline__ MaxSum combine(MaxSum a, MaxSum b) {
  const float m = fmaxf(a.m, b.m);
  if (m == -CUDART_INF_F) return {m, 0.f};
  return {m, a.s * __expf(a.m - m) + b.s * __expf(b.m - m)};
}
template <int THREADS = kCEThreads>
__device__ __forceinline__ MaxSum block_maxsum(MaxSum v, MaxSum* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSum other{__shfl_xor_sync(0xffffffffu, v.m, o), __shfl_xor_sync(0xffffffffu, v.s, o)};
    v = combine(v, other);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  v = lane < (THREADS >> 5) ? smem[lane] : MaxSum{-CUDART_INF_F, 0.f};
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSum other{__shfl_xor_sync(0xffffffffu, v.m, o), __shfl_xor_sync(0xffffffffu, v.s, o)};
    v = combine(v, other);
  }
  return v;   // identical in every thread
}

// MODE 0: loss only.  MODE 1: loss + grad with a uniform scale.  MODE 2: grad with per-row scale.
template <typename T, int MODE>
__global__ void __launch_bounds__(kCEThreads) ce_kernel(T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                        float* __restrict__ loss, const float* __restrict__ row_scale,
                                                        int64_t rows, int vocab, int64_t row_stride,
                                                        int64_t ignore_index, float grad_scale) {
  __shared__ MaxSum red[32];
  const int nvec = vocab >> 3;
  const int tail0 = nvec << 3;
  const int padded = (int)row_stride;   // columns [vocab, row_stride) are padding: zero their grad
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    T* x = logits + row * row_stride;
    const int64_t label = labels[row];
    const bool ignored = label == ignore_index || label < 0 || label >= vocab;
    float scale = grad_scale;
    if (MODE == 2) scale = row_scale[row];
    if (ignored) {
      if (MODE != 2 && threadIdx.x == 0 && loss != nullptr) loss[row] = 0.f;
      if (MODE != 0) {
        const float zeros[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = threadIdx.x; j < (padded >> 3); j += kCEThreads) store8<T>(x + j * 8, zeros);
      }
      continue;   // block-uniform branch
    }
    MaxSum acc{-CUDART_INF_F, 0.f};
    for (int j = threadIdx.x; j < nvec; j += kCEThreads) {
      float f[8];
      load8<T>(x + j * 8, f);
      float mx = f[0];
#pragma unroll
      for (int i = 1; i < 8; ++i) mx = fmaxf(mx, f[i]);
      const float m = fmaxf(acc.m, mx);
      const float ms = m * 1.4426950408889634f;
      float s = acc.s * fast_exp2(fmaf(acc.m, 1.4426950408889634f, -ms));
#pragma unroll
      for (int i = 0; i < 8; ++i) s += fast_exp2(fmaf(f[i], 1.4426950408889634f, -ms));   // one fma + one MUFU
      acc = {m, s};
    }
    if (threadIdx.x < vocab - tail0) {
      const float f = DT<T>::to_f(x[tail0 + threadIdx.x]);
      acc = combine(acc, MaxSum{f, 1.f});
    }
    __syncthreads();   // red[] reuse across rows
    const MaxSum tot = block_maxsum(acc, red);
    const float lse = tot.m + logf(tot.s);
    if (MODE != 2 && threadIdx.x == 0 && loss != nullptr) loss[row] = lse - DT<T>::to_f(x[label]);
    if (MODE == 0) continue;
    __syncthreads();   // everyone (incl. thread 0's x[label] read) is done with the raw logits
    // p * scale = exp2(x*log2e - m*log2e) * (scale / sum): one fma + one MUFU + one mul per element;
    // the label and the vocabulary boundary are handled per vector, not per element
    constexpr float kLog2e = 1.4426950408889634f;
    const float inv = scale / tot.s;
    const float mxs = tot.m * kLog2e;
    const int label_vec = (int)(label >> 3), label_lane = (int)(label & 7);
    for (int j = threadIdx.x; j < (padded >> 3); j += kCEThreads) {
      float f[8];
      load8<T>(x + j * 8, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = fast_exp2(fmaf(f[i], kLog2e, -mxs)) * inv;
      if (j >= nvec) {                      // the vector(s) straddling or beyond the real vocabulary
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (j * 8 + i >= vocab) f[i] = 0.f;
      }
      if (j == label_vec) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (i == label_lane) f[i] -= scale;
      }
      store8<T>(x + j * 8, f);
    }
  }
}

// Register-resident variant for 16-bit logits: every thread loads its NV 16-byte slices of the
// row ONCE (all loads in flight together), keeps them packed in registers, and makes the three
// logical passes (max, sum of exp, gradient) without touching memory again: one HBM read and one
// write per element and one exp per element, against two reads, one write and two exps for the
// streaming kernel above (ncu: that one runs at 75 % issue utilisation, MUFU + conversions).
// The exponentials are kept as packed 16-bit values between the sum and the normalisation, which
// costs one extra rounding of a gradient that is stored in 16 bits anyway.
template <typename T, int NV, int MODE>
__global__ void __launch_bounds__(kCEThreads, 1) ce_reg_kernel(T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                               float* __restrict__ loss,
                                                               const float* __restrict__ row_scale, int64_t rows,
                                                               int vocab, int64_t row_stride, int64_t ignore_index,
                                                               float grad_scale) {
  __shared__ float red[32];
  __shared__ float label_logit;
  const int nvec_pad = (int)(row_stride >> 3);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr float kLog2e = 1.4426950408889634f;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    T* x = logits + row * row_stride;
    const int64_t label = labels[row];
    const bool ignored = label == ignore_index || label < 0 || label >= vocab;
    float scale = grad_scale;
    if (MODE == 2) scale = row_scale[row];
    if (ignored) {
      if (MODE != 2 && threadIdx.x == 0 && loss != nullptr) loss[row] = 0.f;
      if (MODE != 0) {
        const float zeros[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = threadIdx.x; j < nvec_pad; j += kCEThreads) store8<T>(x + j * 8, zeros);
      }
      continue;   // block-uniform
    }
    uint4 raw[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int j = threadIdx.x + k * kCEThreads;
      raw[k] = j < nvec_pad ? *reinterpret_cast<const uint4*>(x + j * 8) : make_uint4(0, 0, 0, 0);
    }
    // ---- max over the real vocabulary
    float mx = -CUDART_INF_F;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int j = threadIdx.x + k * kCEThreads;
      if (j < nvec_pad) {
        float f[8];
        load8<T>(reinterpret_cast<const T*>(&raw[k]), f);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (j * 8 + i < vocab) mx = fmaxf(mx, f[i]);
        if (label >= j * 8 && label < j * 8 + 8) label_logit = f[label - j * 8];
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    __syncthreads();                       // red[] / label_logit reuse across rows
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[lane < (kCEThreads >> 5) ? lane : 0];
    const float ll = label_logit;          // read here: the next row may overwrite it before thread 0 writes the loss
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mx = __shfl_sync(0xffffffffu, mx, 0);
    // ---- exp once; keep it packed; sum in fp32
    const float mxs = mx * kLog2e;
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int j = threadIdx.x + k * kCEThreads;
      if (j < nvec_pad) {
        float f[8];
        load8<T>(reinterpret_cast<const T*>(&raw[k]), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          f[i] = j * 8 + i < vocab ? exp2f(fmaf(f[i], kLog2e, -mxs)) : 0.f;
          sum += f[i];
        }
        if (MODE != 0) store8<T>(reinterpret_cast<T*>(&raw[k]), f);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncthreads();
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = red[lane < (kCEThreads >> 5) ? lane : 0];
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    sum = __shfl_sync(0xffffffffu, sum, 0);
    if (MODE != 2 && threadIdx.x == 0 && loss != nullptr) loss[row] = mx + logf(sum) - ll;
    if (MODE == 0) continue;
    // ---- gradient from the registers
    const float inv = scale / sum;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int j = threadIdx.x + k * kCEThreads;
      if (j < nvec_pad) {
        float f[8];
        load8<T>(reinterpret_cast<const T*>(&raw[k]), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          f[i] *= inv;
          if (j * 8 + i == (int)label) f[i] -= scale;
        }
        store8<T>(x + j * 8, f);
      }
    }
  }
}

template <typename T, int MODE>
static bool try_launch_reg(void* logits, const int64_t* labels, float* loss, const float* row_scale, int64_t rows,
                           int vocab, int64_t row_stride, int64_t ignore_index, float grad_scale, int grid,
                           cudaStream_t s) {
  if (sizeof(T) != 2) return false;
  // Measured on B200 (GPT-2 vocab, 4096 rows): 312 us against 254 us for the streaming kernel — one
  // 512-thread CTA per SM at 114 registers cannot overlap a row's load / reduce / store phases.
  // Kept as an opt-in experiment (DLION_CE_REGISTER=1); the streaming kernel is the default.
  static const bool on = getenv("DLION_CE_REGISTER") != nullptr;
  if (!on) return false;
  const int64_t nvec = row_stride >> 3;
  if (nvec <= 8 * kCEThreads) {
    ce_reg_kernel<T, 8, MODE><<<grid, kCEThreads, 0, s>>>((T*)logits, labels, loss, row_scale, rows, vocab, row_stride,
                                                          ignore_index, grad_scale);
    return true;
  }
  if (nvec <= 13 * kCEThreads) {
    ce_reg_kernel<T, 13, MODE><<<grid, kCEThreads, 0, s>>>((T*)logits, labels, loss, row_scale, rows, vocab, row_stride,
                                                           ignore_index, grad_scale);
    return true;
  }
  return false;
}


// ------------------------------------------------------------------ row-resident variant (16-bit logits)
// One persistent CTA per SM keeps TWO whole rows in shared memory (a GPT-2 row is 100 KB, a Llama row 64 KB): while
// the CTA works on row i out of buffer i & 1, the TMA engine (cp.async.bulk, completion counted on an mbarrier)
// already brings row i + 1 into the other buffer.  Every logit is read from HBM exactly once; both passes (online
// max / sum of exp, then the gradient) read shared memory; the gradient goes out as coalesced 16-byte stores.
// Against the streaming kernel above that is 2 passes over HBM instead of 3, with the loads fully overlapped.
__device__ __forceinline__ uint32_t ce_smem_u32(const void* ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }

constexpr int kCERowThreads = 512;     // measured: 1024 threads per CTA is 5-7 % slower (184 vs 175 us at the GPT-2 shape)

template <typename T, int MODE>
__global__ void __launch_bounds__(kCERowThreads, 1) ce_rowres_kernel(T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                                  float* __restrict__ loss,
                                                                  const float* __restrict__ row_scale, int64_t rows,
                                                                  int vocab, int64_t row_stride, int64_t ignore_index,
                                                                  float grad_scale) {
  extern __shared__ __align__(128) uint8_t ce_smem[];
  __shared__ MaxSum red[32];
  __shared__ __align__(8) uint64_t full[2];
  const uint32_t row_bytes = (uint32_t)row_stride * (uint32_t)sizeof(T);
  const uint32_t buf_bytes = (row_bytes + 127u) & ~127u;
  const int nvec = vocab >> 3;
  const int tail0 = nvec << 3;
  const int padded_vecs = (int)(row_stride >> 3);
  constexpr float kLog2e = 1.4426950408889634f;
  if (threadIdx.x == 0) {
    for (int b = 0; b < 2; ++b)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ce_smem_u32(&full[b])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int64_t row, int b) {      // thread 0 only
    const uint32_t bar = ce_smem_u32(&full[b]);
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(bar), "r"(row_bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     ce_smem_u32(ce_smem + (size_t)b * buf_bytes)),
                 "l"(logits + row * row_stride), "r"(row_bytes), "r"(bar)
                 : "memory");
  };
  const int64_t first = blockIdx.x;
  if (threadIdx.x == 0 && first < rows) issue(first, 0);
  int it = 0;
  for (int64_t row = first; row < rows; row += gridDim.x, ++it) {
    const int b = it & 1;
    const int64_t next = row + gridDim.x;
    // buffer b ^ 1 was last read in iteration it - 1, which ended with a block barrier: safe to refill
    if (threadIdx.x == 0 && next < rows) issue(next, b ^ 1);
    {   // wait for this row's bytes (phase parity = number of earlier uses of this barrier, mod 2)
      const uint32_t bar = ce_smem_u32(&full[b]);
      const uint32_t parity = (uint32_t)((it >> 1) & 1);
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    }
    const T* x = reinterpret_cast<const T*>(ce_smem + (size_t)b * buf_bytes);
    T* gx = logits + row * row_stride;
    const int64_t label = labels[row];
    const bool ignored = label == ignore_index || label < 0 || label >= vocab;
    float scale = grad_scale;
    if (MODE == 2) scale = row_scale[row];
    if (ignored) {
      if (MODE != 2 && threadIdx.x == 0 && loss != nullptr) loss[row] = 0.f;
      if (MODE != 0) {
        const float zeros[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = threadIdx.x; j < padded_vecs; j += kCERowThreads) store8<T>(gx + j * 8, zeros);
      }
      __syncthreads();
      continue;   // block-uniform
    }
    MaxSum acc{-CUDART_INF_F, 0.f};
    for (int j = threadIdx.x; j < nvec; j += kCERowThreads) {
      float f[8];
      load8<T>(x + j * 8, f);
      float mx = f[0];
#pragma unroll
      for (int i = 1; i < 8; ++i) mx = fmaxf(mx, f[i]);
      const float m = fmaxf(acc.m, mx);
      const float ms = m * kLog2e;
      float sacc = acc.s * fast_exp2(fmaf(acc.m, kLog2e, -ms));
#pragma unroll
      for (int i = 0; i < 8; ++i) sacc += fast_exp2(fmaf(f[i], kLog2e, -ms));
      acc = {m, sacc};
    }
    if (threadIdx.x < vocab - tail0) {
      const float f = DT<T>::to_f(x[tail0 + threadIdx.x]);
      acc = combine(acc, MaxSum{f, 1.f});
    }
    const MaxSum tot = block_maxsum<kCERowThreads>(acc, red);
    const float lse = tot.m + logf(tot.s);
    if (MODE != 2 && threadIdx.x == 0 && loss != nullptr) loss[row] = lse - DT<T>::to_f(x[label]);
    if (MODE != 0) {
      const float inv = scale / tot.s;
      const float mxs = tot.m * kLog2e;
      const int label_vec = (int)(label >> 3), label_lane = (int)(label & 7);
      for (int j = threadIdx.x; j < padded_vecs; j += kCERowThreads) {
        float f[8];
        load8<T>(x + j * 8, f);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fast_exp2(fmaf(f[i], kLog2e, -mxs)) * inv;
        if (j >= nvec) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (j * 8 + i >= vocab) f[i] = 0.f;
        }
        if (j == label_vec) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (i == label_lane) f[i] -= scale;
        }
        store8<T>(gx + j * 8, f);
      }
    }
    __syncthreads();   // everyone is done with buffer b (and with red[]) before it is refilled / reused
  }
}

template <typename T, int MODE>
static bool try_launch_rowres(void* logits, const int64_t* labels, float* loss, const float* row_scale, int64_t rows,
                              int vocab, int64_t row_stride, int64_t ignore_index, float grad_scale, cudaStream_t s) {
  if (sizeof(T) != 2) return false;
  static const bool off = getenv("DLION_CE_STREAMING") != nullptr;      // force the streaming kernel (A/B)
  if (off) return false;
  const size_t buf = ((size_t)row_stride * sizeof(T) + 127) & ~(size_t)127;
  const size_t smem = 2 * buf;
  if (smem > 220 * 1024 || row_stride * sizeof(T) >= (1u << 20) || rows < 1) return false;
  static thread_local int configured_dev = -1;
  static thread_local size_t configured_smem = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (dev != configured_dev || smem > configured_smem) {
    if (cudaFuncSetAttribute(ce_rowres_kernel<T, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    configured_dev = dev;
    configured_smem = 220 * 1024;
  }
  const int grid = (int)(rows < kNumSMs ? rows : kNumSMs);
  ce_rowres_kernel<T, MODE><<<grid, kCERowThreads, smem, s>>>((T*)logits, labels, loss, row_scale, rows, vocab, row_stride,
                                                           ignore_index, grad_scale);
  return true;
}

template <int MODE>
static int launch(void* logits, const int64_t* labels, float* loss, const float* row_scale, int64_t rows, int vocab,
                  int64_t row_stride, int64_t ignore_index, float grad_scale, int dtype, cudaStream_t s) {
  const int64_t cap = (int64_t)kNumSMs * 4;
  const int grid = (int)(rows < cap ? rows : cap);
  const int64_t cap1 = (int64_t)kNumSMs;              // register-resident kernel: one persistent CTA per SM
  const int grid1 = (int)(rows < cap1 ? rows : cap1);
  DISPATCH_DTYPE(dtype, {
    if (!try_launch_reg<T, MODE>(logits, labels, loss, row_scale, rows, vocab, row_stride, ignore_index, grad_scale,
                                 grid1, s) &&
        !try_launch_rowres<T, MODE>(logits, labels, loss, row_scale, rows, vocab, row_stride, ignore_index, grad_scale, s))
      ce_kernel<T, MODE><<<grid, kCEThreads, 0, s>>>((T*)logits, labels, loss, row_scale, rows, vocab, row_stride,
                                                      ignore_index, grad_scale);
  });
  return DLION_LAUNCH_CHECK("dlion_ce");
}

static int check_ce(const void* logits, const int64_t* labels, int64_t rows, int vocab, int64_t row_stride) {
  DLION_CHECK_ARG(rows >= 0 && vocab > 0 && row_stride >= vocab, "dlion_ce: bad shape");
  DLION_CHECK_ARG(row_stride % 8 == 0 && aligned16(logits), "dlion_ce: rows must be 16-byte aligned (pad the vocab stride to a multiple of 8)");
  DLION_CHECK_ARG(labels != nullptr, "dlion_ce: labels required");
  return 0;
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_ce_fwd_bwd(void* logits, const int64_t* labels, float* loss, int64_t rows, int vocab,
                                int64_t row_stride, int64_t ignore_index, float grad_scale, int write_grad, int dtype,
                                void* stream) {
  int rc = check_ce(logits, labels, rows, vocab, row_stride);
  if (rc) return rc;
  DLION_CHECK_ARG(loss != nullptr, "dlion_ce_fwd_bwd: loss output required");
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  return write_grad ? launch<1>(logits, labels, loss, nullptr, rows, vocab, row_stride, ignore_index, grad_scale, dtype, s)
                    : launch<0>(logits, labels, loss, nullptr, rows, vocab, row_stride, ignore_index, grad_scale, dtype, s);
}

extern "C" int dlion_ce_bwd_rowscale(void* logits, const int64_t* labels, const float* row_scale, int64_t rows,
                                     int vocab, int64_t row_stride, int64_t ignore_index, int dtype, void* stream) {
  int rc = check_ce(logits, labels, rows, vocab, row_stride);
  if (rc) return rc;
  DLION_CHECK_ARG(row_scale != nullptr, "dlion_ce_bwd_rowscale: row_scale required");
  if (rows == 0) return 0;
  return launch<2>(logits, labels, nullptr, row_scale, rows, vocab, row_stride, ignore_index, 0.f, dtype,
                   (cudaStream_t)stream);
}
