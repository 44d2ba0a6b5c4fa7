// bucket_ops.cu — N2: flat-bucket plumbing kernels.
//   dlion_multi_gather : copy/cast many gradient tensors into their bucket slots, 96 per launch
//   dlion_sumsq        : sum of squares (local grad-norm), two-level reduction through scratch
//   dlion_clip_scale   : x *= min(1, max_norm / (||x|| + 1e-6)) with the norm read on device
// reference: the per-tensor Python loop in distributed_lion.py:179-198 and flatten_and_pad
// (:14-24); gradient clipping is HF Trainer's (max_grad_norm=1.0 default), inherited upstream.
#include <mutex>

#include "common.cuh"

namespace dlion {

constexpr int kMGTensors = 96;     // 96 * 24 B = 2304 B of kernel parameters
constexpr int kMGBlocksPerTensor = 48;

struct MGTable {
  const void* src[kMGTensors];
  void* dst[kMGTensors];
  int64_t n[kMGTensors];
};

template <typename S, typename D>
__global__ void __launch_bounds__(256) multi_gather_kernel(const __grid_constant__ MGTable tab) {
  const int t = blockIdx.y;
  const S* __restrict__ src = (const S*)tab.src[t];
  D* __restrict__ dst = (D*)tab.dst[t];
  const int64_t n = tab.n[t];
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t done = 0;
  if (vec_ok) {
    const int64_t nvec = n >> 3;
    for (int64_t v = tid; v < nvec; v += stride) {
      float f[8];
      load8<S>(src + v * 8, f);
      store8<D>(dst + v * 8, f);
    }
    done = nvec << 3;
  }
  for (int64_t e = done + tid; e < n; e += stride) dst[e] = DT<D>::from_f(DT<S>::to_f(src[e]));
}

// ------------------------------------------------------------------ sum of squares
constexpr int kSumsqBlocks = kNumSMs * 4;
// scratch layout: [0] uint32 ticket, [1] pad, [2 .. 2+kSumsqBlocks) float partials
constexpr int64_t kSumsqScratchBytes = 8 + (int64_t)kSumsqBlocks * 4;

__device__ __forceinline__ float block_sum(float v, float* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = lane < (blockDim.x >> 5) ? smem[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  }
  return v;   // valid in warp 0
}

template <typename T>
__device__ __forceinline__ void sumsq_body(const T* __restrict__ x, int64_t n, unsigned int* ticket,
                                           float* partials, float* __restrict__ out) {
  __shared__ float red[32];
  __shared__ bool last;
  const int64_t nvec = n >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float f[8];
    load8<T>(x + v * 8, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc = fmaf(f[i], f[i], acc);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int64_t e = nvec << 3; e < n; ++e) {
      const float f = DT<T>::to_f(x[e]);
      acc = fmaf(f, f, acc);
    }
  const float bs = block_sum(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = bs;
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {   // the block that arrives last folds the partials in a fixed order: deterministic
    __threadfence();
    float v = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) v += __ldcg(partials + i);
    __syncthreads();
    const float tot = block_sum(v, red);
    if (threadIdx.x == 0) {
      out[0] = tot;
      *ticket = 0u;   // leave the scratch ready for the next call
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) sumsq_kernel(const T* __restrict__ x, int64_t n, unsigned int* ticket,
                                                    float* partials, float* __restrict__ out) {
  sumsq_body<T>(x, n, ticket, partials, out);
}

template <typename T>
__global__ void __launch_bounds__(256) clip_scale_kernel(T* __restrict__ x, int64_t n,
                                                         const float* __restrict__ sumsq, float max_norm) {
  const float norm = sqrtf(__ldg(sumsq));
  const float coef = max_norm / (norm + 1e-6f);
  if (!(coef < 1.0f)) return;   // nothing to clip (also the NaN-norm case: leave the grads alone)
  const int64_t nvec = n >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float f[8];
    load8<T>(x + v * 8, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] *= coef;
    store8<T>(x + v * 8, f);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int64_t e = nvec << 3; e < n; ++e) x[e] = DT<T>::from_f(DT<T>::to_f(x[e]) * coef);
}

template <typename S>
static int launch_gather_dst(const MGTable& tab, int count, int dst_dtype, cudaStream_t s) {
  dim3 grid(kMGBlocksPerTensor, count);
  switch (dst_dtype) {
    case DLION_F32: multi_gather_kernel<S, float><<<grid, 256, 0, s>>>(tab); break;
    case DLION_BF16: multi_gather_kernel<S, __nv_bfloat16><<<grid, 256, 0, s>>>(tab); break;
    case DLION_F16: multi_gather_kernel<S, __half><<<grid, 256, 0, s>>>(tab); break;
    default: set_error("dlion_multi_gather: unsupported dst dtype %d", dst_dtype); return DLION_E_UNSUPPORTED;
  }
  return DLION_LAUNCH_CHECK("dlion_multi_gather");
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_multi_gather(const void* const* srcs, void* const* dsts, const int64_t* numels, int count,
                                  int src_dtype, int dst_dtype, void* stream) {
  DLION_CHECK_ARG(count >= 0, "dlion_multi_gather: negative count");
  cudaStream_t s = (cudaStream_t)stream;
  for (int base = 0; base < count; base += kMGTensors) {
    MGTable tab;
    const int c = count - base < kMGTensors ? count - base : kMGTensors;
    for (int i = 0; i < c; ++i) {
      tab.src[i] = srcs[base + i];
      tab.dst[i] = dsts[base + i];
      tab.n[i] = numels[base + i];
    }
    int rc;
    switch (src_dtype) {
      case DLION_F32: rc = launch_gather_dst<float>(tab, c, dst_dtype, s); break;
      case DLION_BF16: rc = launch_gather_dst<__nv_bfloat16>(tab, c, dst_dtype, s); break;
      case DLION_F16: rc = launch_gather_dst<__half>(tab, c, dst_dtype, s); break;
      default: set_error("dlion_multi_gather: unsupported src dtype %d", src_dtype); return DLION_E_UNSUPPORTED;
    }
    if (rc) return rc;
  }
  return 0;
}

extern "C" int64_t dlion_sumsq_scratch_bytes(void) { return kSumsqScratchBytes; }

extern "C" int dlion_sumsq(const void* x, int64_t n, int dtype, void* scratch, float* out, void* stream) {
  DLION_CHECK_ARG(n >= 0 && scratch && out, "dlion_sumsq: bad arguments");
  DLION_CHECK_ARG(aligned16(x), "dlion_sumsq: x must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  int64_t want = ((n >> 3) + 255) / 256;
  const int grid = (int)(want < 1 ? 1 : (want > kSumsqBlocks ? kSumsqBlocks : want));
  unsigned int* ticket = (unsigned int*)scratch;
  float* partials = (float*)scratch + 2;
  DISPATCH_DTYPE(dtype, sumsq_kernel<T><<<grid, 256, 0, s>>>((const T*)x, n, ticket, partials, out));
  return DLION_LAUNCH_CHECK("dlion_sumsq");
}

extern "C" int dlion_clip_scale(void* x, int64_t n, int dtype, const float* sumsq, float max_norm, void* stream) {
  DLION_CHECK_ARG(n >= 0 && sumsq, "dlion_clip_scale: bad arguments");
  if (n == 0) return 0;
  DLION_CHECK_ARG(aligned16(x), "dlion_clip_scale: x must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  int64_t want = ((n >> 3) + 255) / 256;
  const int grid = (int)(want < 1 ? 1 : (want > kNumSMs * 16 ? kNumSMs * 16 : want));
  DISPATCH_DTYPE(dtype, clip_scale_kernel<T><<<grid, 256, 0, s>>>((T*)x, n, sumsq, max_norm));
  return DLION_LAUNCH_CHECK("dlion_clip_scale");
}

// ====================================================================================
// dlion_sumsq_static — sum of squares for callers that cannot allocate device memory.
//
// dlion_sumsq() needs a caller-provided scratch buffer (a ticket word plus one fp32 partial
// per CTA).  Some callers have no way to obtain one.  Examples: a plugin living inside a host
// application that has handed its whole device-memory budget out at start-up and forbids
// cudaMalloc afterwards (serving stacks and real-time pipelines often work like that); a
// monitoring hook that wants a gradient norm but must not disturb the framework's caching
// allocator; a plain C program that links this library and has no allocator layer of its
// own.  For them the scratch lives in the file-scope array below, which every GPU that loads
// the library keeps resident until the process exits (kSumsqStaticBytes = 16,640 bytes per
// device).
//
// Size limit: the launch uses grid = ceil(n / kSumsqStaticChunk) CTAs (kSumsqStaticChunk = 65,536), every CTA
// owns exactly one partial slot, and the array has kSumsqStaticSlots = 4,096 slots — so a call may cover at most
// 4,096 * 65,536 = 268,435,456 elements (one fused-optimizer chunk).  (The CTAs walk the input grid-stride, in
// interleaved 2,048-element strips, not one contiguous 65,536-element chunk each; only the slot count is fixed.)
// Larger n is refused with DLION_E_TOOBIG and a message; nothing is launched.
//
// Rules of use
//  * The array name is only a host-side handle: the kernel needs the DEVICE address, which is
//    looked up with cudaGetSymbolAddress once per device, with that device current, the first
//    time the entry point is called on it.  Make that first call outside any stream capture.
//  * No cudaMalloc / cudaFree / allocating library routine is called, here or on later calls.
//  * All calls on a device share the one array and it keeps the previous call's contents: the
//    ticket word is cleared (cudaMemsetAsync on `stream`) before every launch, the partial slots
//    are fully overwritten before they are read.  Two calls that overlap on the GPU (different
//    streams, no dependency between them) race on the array and give undefined results —
//    serialise them.
namespace dlion {

constexpr int kSumsqStaticSlots = 4096;
constexpr int64_t kSumsqStaticChunk = 65536;
constexpr int64_t kSumsqStaticBytes = 256 + (int64_t)kSumsqStaticSlots * 4;   // ticket (padded) + partials
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_sumsq_static_scratch[kSumsqStaticBytes];

static void* g_sumsq_static_ptr[kMaxDevices];     // device address of the array, per device ordinal
static bool g_sumsq_static_ready[kMaxDevices];
static std::mutex g_sumsq_static_mu;              // two host threads may make "the first call" at once

static int sumsq_static_address(void** out) {
  int dev = 0;
  int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  if (rc) return rc;
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("dlion_sumsq_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    return DLION_E_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_sumsq_static_mu);
  if (!g_sumsq_static_ready[dev]) {
    void* ptr = nullptr;
    rc = check_cuda(cudaGetSymbolAddress(&ptr, g_sumsq_static_scratch),
                    "cudaGetSymbolAddress(g_sumsq_static_scratch) — first call on a device must be outside stream capture");
    if (rc) return rc;
    g_sumsq_static_ptr[dev] = ptr;
    g_sumsq_static_ready[dev] = true;
  }
  *out = g_sumsq_static_ptr[dev];
  return 0;
}

}  // namespace dlion

extern "C" int64_t dlion_sumsq_static_max_n(void) { return (int64_t)kSumsqStaticSlots * kSumsqStaticChunk; }

extern "C" int dlion_sumsq_static(const void* x, int64_t n, int dtype, float* out, void* stream) {
  DLION_CHECK_ARG(n >= 0 && out != nullptr, "dlion_sumsq_static: bad arguments");
  DLION_CHECK_ARG(aligned16(x), "dlion_sumsq_static: x must be 16-byte aligned");
  if (n > dlion_sumsq_static_max_n()) {
    set_error("dlion_sumsq_static: n = %lld exceeds the built-in scratch (%d slots x %lld elements = %lld); "
              "split the input or use dlion_sumsq with your own scratch buffer",
              (long long)n, kSumsqStaticSlots, (long long)kSumsqStaticChunk, (long long)dlion_sumsq_static_max_n());
    return DLION_E_TOOBIG;
  }
  void* base = nullptr;
  int rc = sumsq_static_address(&base);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  unsigned int* ticket = (unsigned int*)base;
  float* partials = (float*)((unsigned char*)base + 256);
  rc = check_cuda(cudaMemsetAsync(ticket, 0, 256, s), "cudaMemsetAsync(ticket)");   // must start at zero
  if (rc) return rc;
  int64_t want = (n + kSumsqStaticChunk - 1) / kSumsqStaticChunk;
  const int grid = (int)(want < 1 ? 1 : want);                                      // <= kSumsqStaticSlots by the check above
  DISPATCH_DTYPE(dtype, sumsq_kernel<T><<<grid, 256, 0, s>>>((const T*)x, n, ticket, partials, out));
  return DLION_LAUNCH_CHECK("dlion_sumsq_static");
}
