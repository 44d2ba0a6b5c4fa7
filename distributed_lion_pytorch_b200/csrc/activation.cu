// activation.cu — MLP activations as single streaming passes.
//   dlion_gelu_tanh_fwd / _bwd : GPT-2's gelu_new (tanh form), y = 0.5 x (1 + tanh(k (x + 0.044715 x^3)))
//   dlion_swiglu_fwd / _bwd    : Llama's silu(gate) * up, forward in one pass (read 2, write 1),
//                                backward in one pass (read gate, up, dy; write dgate, dup)
// reference: HF GPT2MLP (NewGELUActivation: 8 ATen elementwise kernels upstream of torch's fused
// gelu) and LlamaMLP (silu, mul) reached through Trainer.compute_loss (async_trainer.py:24);
// autograd runs silu_backward + two mul backward kernels there (9 tensor passes against 5 here).
#include "common.cuh"

namespace dlion {

constexpr float kGeluK = 0.7978845608028654f;   // sqrt(2/pi)
constexpr float kGeluC = 0.044715f;

template <typename T>
__device__ __forceinline__ float tanh_fast(float x) {
  if constexpr (sizeof(T) == 2) {               // 16-bit storage: MUFU.TANH (rel. error ~2^-11) is below the rounding
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  } else {
    return tanhf(x);
  }
}
template <typename T>
__device__ __forceinline__ float sigmoid_fast(float x) {
  if constexpr (sizeof(T) == 2) return __fdividef(1.0f, 1.0f + __expf(-x));
  else return 1.0f / (1.0f + expf(-x));
}

template <typename T>
__global__ void __launch_bounds__(256) gelu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float f[8];
    load8<T>(x + v * 8, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float u = kGeluK * fmaf(kGeluC * f[i] * f[i], f[i], f[i]);
      f[i] = 0.5f * f[i] * (1.0f + tanh_fast<T>(u));
    }
    store8<T>(y + v * 8, f);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) gelu_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                       T* __restrict__ dx, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float f[8], d[8];
    load8<T>(x + v * 8, f);
    load8<T>(dy + v * 8, d);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float x2 = f[i] * f[i];
      const float t = tanh_fast<T>(kGeluK * fmaf(kGeluC * x2, f[i], f[i]));
      const float du = kGeluK * fmaf(3.0f * kGeluC, x2, 1.0f);
      d[i] *= 0.5f * (1.0f + t) + 0.5f * f[i] * (1.0f - t * t) * du;
    }
    store8<T>(dx + v * 8, d);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const T* __restrict__ gate, const T* __restrict__ up,
                                                         T* __restrict__ y, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float g[8], u[8];
    load8<T>(gate + v * 8, g);
    load8<T>(up + v * 8, u);
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = g[i] * sigmoid_fast<T>(g[i]);
    round8<T>(g);                                  // HF: silu(gate) is a bf16 tensor before the multiply
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] *= u[i];
    store8<T>(y + v * 8, g);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) swiglu_bwd_kernel(const T* __restrict__ gate, const T* __restrict__ up,
                                                         const T* __restrict__ dy, T* __restrict__ dgate,
                                                         T* __restrict__ dup, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float g[8], u[8], d[8], dg[8], du[8];
    load8<T>(gate + v * 8, g);
    load8<T>(up + v * 8, u);
    load8<T>(dy + v * 8, d);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float s = sigmoid_fast<T>(g[i]);
      const float silu = g[i] * s;
      du[i] = d[i] * silu;
      dg[i] = d[i] * u[i] * (s * fmaf(g[i], 1.0f - s, 1.0f));     // d silu / d g = s (1 + g (1 - s))
    }
    store8<T>(dgate + v * 8, dg);
    store8<T>(dup + v * 8, du);
  }
}

static inline int act_grid(int64_t nvec) {
  const int64_t want = (nvec + 255) / 256;
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace dlion

using namespace dlion;

#define ACT_CHECK(n, ...)                                                                              \
  DLION_CHECK_ARG((n) >= 0 && ((n) & 7) == 0, "activation kernels need a multiple of 8 elements");      \
  if ((n) == 0) return 0;

extern "C" int dlion_gelu_tanh_fwd(const void* x, void* y, int64_t n, int dtype, void* stream) {
  ACT_CHECK(n);
  DLION_CHECK_ARG(aligned16(x) && aligned16(y), "dlion_gelu_tanh_fwd: unaligned pointer");
  DISPATCH_DTYPE(dtype, gelu_fwd_kernel<T><<<act_grid(n >> 3), 256, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, n >> 3));
  return DLION_LAUNCH_CHECK("dlion_gelu_tanh_fwd");
}
extern "C" int dlion_gelu_tanh_bwd(const void* x, const void* dy, void* dx, int64_t n, int dtype, void* stream) {
  ACT_CHECK(n);
  DLION_CHECK_ARG(aligned16(x) && aligned16(dy) && aligned16(dx), "dlion_gelu_tanh_bwd: unaligned pointer");
  DISPATCH_DTYPE(dtype, gelu_bwd_kernel<T><<<act_grid(n >> 3), 256, 0, (cudaStream_t)stream>>>((const T*)x, (const T*)dy,
                                                                                              (T*)dx, n >> 3));
  return DLION_LAUNCH_CHECK("dlion_gelu_tanh_bwd");
}
extern "C" int dlion_swiglu_fwd(const void* gate, const void* up, void* y, int64_t n, int dtype, void* stream) {
  ACT_CHECK(n);
  DLION_CHECK_ARG(aligned16(gate) && aligned16(up) && aligned16(y), "dlion_swiglu_fwd: unaligned pointer");
  DISPATCH_DTYPE(dtype, swiglu_fwd_kernel<T><<<act_grid(n >> 3), 256, 0, (cudaStream_t)stream>>>((const T*)gate, (const T*)up,
                                                                                                (T*)y, n >> 3));
  return DLION_LAUNCH_CHECK("dlion_swiglu_fwd");
}
extern "C" int dlion_swiglu_bwd(const void* gate, const void* up, const void* dy, void* dgate, void* dup, int64_t n,
                                int dtype, void* stream) {
  ACT_CHECK(n);
  DLION_CHECK_ARG(aligned16(gate) && aligned16(up) && aligned16(dy) && aligned16(dgate) && aligned16(dup),
                  "dlion_swiglu_bwd: unaligned pointer");
  DISPATCH_DTYPE(dtype, swiglu_bwd_kernel<T><<<act_grid(n >> 3), 256, 0, (cudaStream_t)stream>>>(
                            (const T*)gate, (const T*)up, (const T*)dy, (T*)dgate, (T*)dup, n >> 3));
  return DLION_LAUNCH_CHECK("dlion_swiglu_bwd");
}
