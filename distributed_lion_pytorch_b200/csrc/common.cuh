// common.cuh — shared device helpers for libdlion (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/dlion.h"

namespace dlion {

void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);

#define DLION_CHECK_ARG(cond, ...)      \
  do {                                  \
    if (!(cond)) {                      \
      ::dlion::set_error(__VA_ARGS__);  \
      return DLION_E_BADARG;            \
    }                                   \
  } while (0)

#define DLION_LAUNCH_CHECK(what) ::dlion::check_cuda(cudaGetLastError(), what)

#define DISPATCH_DTYPE(dtype, ...)                                   \
  switch (dtype) {                                                   \
    case DLION_F32: { using T = float; __VA_ARGS__; } break;         \
    case DLION_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;\
    case DLION_F16: { using T = __half; __VA_ARGS__; } break;        \
    default: ::dlion::set_error("unsupported dtype %d", dtype); return DLION_E_UNSUPPORTED; \
  }


constexpr int kNumSMs = 148;  // B200

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ------------------------------------------------------------------ dtype traits
// Every op rounds through the storage type exactly where ATen's elementwise kernels do, so
// the fused kernels reproduce the reference's bf16 arithmetic bit for bit.
template <typename T>
struct DT;
template <>
struct DT<float> {
  static constexpr int id = DLION_F32;
  __device__ __forceinline__ static float rnd(float x) { return x; }
  __device__ __forceinline__ static float to_f(float x) { return x; }
  __device__ __forceinline__ static float from_f(float x) { return x; }
};
template <>
struct DT<__nv_bfloat16> {
  static constexpr int id = DLION_BF16;
  __device__ __forceinline__ static float rnd(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
  __device__ __forceinline__ static float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }
  __device__ __forceinline__ static __nv_bfloat16 from_f(float x) { return __float2bfloat16_rn(x); }
};
template <>
struct DT<__half> {
  static constexpr int id = DLION_F16;
  __device__ __forceinline__ static float rnd(float x) { return __half2float(__float2half_rn(x)); }
  __device__ __forceinline__ static float to_f(__half x) { return __half2float(x); }
  __device__ __forceinline__ static __half from_f(float x) { return __float2half_rn(x); }
};

// 8 consecutive elements <-> float[8].  Pointers must be 16-byte aligned (fp32 uses two 16-byte
// accesses, the 2-byte types one).
template <typename T>
__device__ __forceinline__ void load8(const T* __restrict__ ptr, float (&v)[8]);
template <typename T>
__device__ __forceinline__ void store8(T* __restrict__ ptr, const float (&v)[8]);

template <>
__device__ __forceinline__ void load8<float>(const float* __restrict__ ptr, float (&v)[8]) {
  const float4 a = reinterpret_cast<const float4*>(ptr)[0];
  const float4 b = reinterpret_cast<const float4*>(ptr)[1];
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void store8<float>(float* __restrict__ ptr, const float (&v)[8]) {
  reinterpret_cast<float4*>(ptr)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(ptr)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* __restrict__ ptr, float (&v)[8]) {
  const uint4 raw = *reinterpret_cast<const uint4*>(ptr);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {  // bf16 -> fp32 is a 16-bit shift
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <>
__device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* __restrict__ ptr, const float (&v)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<const uint32_t*>(&h);
  }
  *reinterpret_cast<uint4*>(ptr) = make_uint4(w[0], w[1], w[2], w[3]);
}
template <>
__device__ __forceinline__ void load8<__half>(const __half* __restrict__ ptr, float (&v)[8]) {
  const uint4 raw = *reinterpret_cast<const uint4*>(ptr);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
template <>
__device__ __forceinline__ void store8<__half>(__half* __restrict__ ptr, const float (&v)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<const uint32_t*>(&h);
  }
  *reinterpret_cast<uint4*>(ptr) = make_uint4(w[0], w[1], w[2], w[3]);
}

// Raw (unconverted) 8-element vectors: lets a kernel issue the loads of a later iteration early and pay
// the conversion only when the values are consumed.
template <typename T>
struct Raw8 {
  uint4 q[sizeof(T) * 8 / 16];
};
template <typename T>
__device__ __forceinline__ Raw8<T> load_raw8(const T* __restrict__ ptr) {
  Raw8<T> r;
#pragma unroll
  for (int i = 0; i < (int)(sizeof(T) * 8 / 16); ++i) r.q[i] = reinterpret_cast<const uint4*>(ptr)[i];
  return r;
}
template <typename T>
__device__ __forceinline__ void unpack_raw8(const Raw8<T>& r, float (&v)[8]);
template <>
__device__ __forceinline__ void unpack_raw8<float>(const Raw8<float>& r, float (&v)[8]) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    v[4 * i] = __uint_as_float(r.q[i].x);
    v[4 * i + 1] = __uint_as_float(r.q[i].y);
    v[4 * i + 2] = __uint_as_float(r.q[i].z);
    v[4 * i + 3] = __uint_as_float(r.q[i].w);
  }
}
template <>
__device__ __forceinline__ void unpack_raw8<__nv_bfloat16>(const Raw8<__nv_bfloat16>& r, float (&v)[8]) {
  const uint32_t w[4] = {r.q[0].x, r.q[0].y, r.q[0].z, r.q[0].w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <>
__device__ __forceinline__ void unpack_raw8<__half>(const Raw8<__half>& r, float (&v)[8]) {
  const uint32_t w[4] = {r.q[0].x, r.q[0].y, r.q[0].z, r.q[0].w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}

// ------------------------------------------------------------------ rounding through the storage type
// round8: v[i] <- float(T(v[i])).  Packed converts (F2FP.*.PACK_AB, two values per ALU-pipe
// instruction) instead of eight scalar F2F on the slow conversion pipe: with six roundings per
// element the scalar form made the optimizer kernels conversion-bound (ncu: stall_mio on F2F).
template <typename T>
__device__ __forceinline__ void round8(float (&v)[8]);
template <>
__device__ __forceinline__ void round8<float>(float (&)[8]) {}
template <>
__device__ __forceinline__ void round8<__nv_bfloat16>(float (&v)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    const uint32_t w = *reinterpret_cast<const uint32_t*>(&h);
    v[2 * i] = __uint_as_float(w << 16);
    v[2 * i + 1] = __uint_as_float(w & 0xffff0000u);
  }
}
template <>
__device__ __forceinline__ void round8<__half>(float (&v)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(__floats2half2_rn(v[2 * i], v[2 * i + 1]));
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
// smallest |x| whose rounding to T is non-zero is just above this: round(x) > 0  <=>  x > tie_to_zero
template <typename T> __device__ __forceinline__ float tie_to_zero();
template <> __device__ __forceinline__ float tie_to_zero<float>() { return 0.f; }
template <> __device__ __forceinline__ float tie_to_zero<__nv_bfloat16>() { return 0x1p-134f; }   // half of 2^-133
template <> __device__ __forceinline__ float tie_to_zero<__half>() { return 0x1p-25f; }           // half of 2^-24

// ------------------------------------------------------------------ Lion math (8 elements)
// The reference rounds through the parameter dtype after every ATen op.  Three of those six
// roundings are made by the store (store8 rounds to nearest even exactly like rnd) or only feed
// a sign test (round(c) > 0 <=> c > tie_to_zero), so only m*b1, m*b2 and p*decay are rounded
// explicitly; the results are bit-identical to the op-by-op sequence.
//
// blend: c = round(m*b1) + (1-b1)*g   (reference: distributed_lion.py:54,68 —
// exp_avg.clone().mul_(beta1).add(grad, alpha=1-beta1); ATen's add kernel is one fp32 fma)
template <typename T>
__device__ __forceinline__ void blend8(const float (&m)[8], const float (&g)[8], float beta1, float omb1,
                                       float (&c)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i] = m[i] * beta1;
  round8<T>(c);
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i] = fmaf(omb1, g[i], c[i]);
}
// momentum: m' = round(m*b2) + (1-b2)*g, rounded by the store   (reference: distributed_lion.py:59,96)
template <typename T>
__device__ __forceinline__ void momentum8(float (&m)[8], const float (&g)[8], float beta2, float omb2) {
#pragma unroll
  for (int i = 0; i < 8; ++i) m[i] *= beta2;
  round8<T>(m);
#pragma unroll
  for (int i = 0; i < 8; ++i) m[i] = fmaf(omb2, g[i], m[i]);
}
// parameter: p' = round(p*decay) - lr*u, rounded by the store   (reference: distributed_lion.py:51,55,64,92)
template <typename T>
__device__ __forceinline__ void apply8(float (&p)[8], const float (&u)[8], float lr, float decay) {
#pragma unroll
  for (int i = 0; i < 8; ++i) p[i] *= decay;
  round8<T>(p);
#pragma unroll
  for (int i = 0; i < 8; ++i) p[i] = fmaf(-lr, u[i], p[i]);
}
// ------------------------------------------------------------------ stochastic rounding of the parameter write
// Opt-in (Lion(stochastic_rounding=True)); bf16 only.  At lr ~ 1e-4 a Lion update is below half a bf16 ulp
// of a typical weight, so round-to-nearest leaves most weights unchanged step after step; rounding up with
// probability (discarded bits / 2^16) makes the expected stored value equal the fp32 result.  The random
// bits are a counter hash of (key, element index): every rank draws the same bits for the same element, so
// replicas stay bit-identical, and the split and fused paths agree bit for bit.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {   // "lowbias32" integer hash: full avalanche in 5 ops
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
// four 32-bit hashes = eight 16-bit uniforms for 8-element vector number `vec` of the bucket
__device__ __forceinline__ void hash_bits8(uint64_t key, int64_t vec, uint32_t (&r)[4]) {
  const uint32_t lo = (uint32_t)key, hi = (uint32_t)(key >> 32);
  const uint32_t base = mix32((uint32_t)vec * 0x9E3779B9u + lo) ^ mix32((uint32_t)(vec >> 32) + hi);
#pragma unroll
  for (int i = 0; i < 4; ++i) r[i] = mix32(base + (uint32_t)i * 0x85EBCA6Bu + hi);
}
// Dropout fused into the residual add: keep-mask bits for the 8 elements of vector `vec` of the
// whole [rows, cols] tensor, regenerated (not stored) in the backward from (seed, vec): keep iff a
// 16-bit uniform >= p * 65536.
// mix32 (the counter hash) lives in common.cuh
__device__ __forceinline__ uint32_t keep_mask8(uint64_t seed, int64_t vec, uint32_t thr16) {
  // four counter-based 32-bit hashes -> eight 16-bit uniforms (Philox4x32-10 here made the norm
  // kernels issue-bound: ~70 instructions per 8 elements against ~25 for this)
  const uint32_t lo = (uint32_t)seed, hi = (uint32_t)(seed >> 32);
  const uint32_t base = mix32((uint32_t)vec * 0x9E3779B9u + lo) ^ mix32((uint32_t)(vec >> 32) + hi);
  uint32_t mask = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t w = mix32(base + (uint32_t)i * 0x85EBCA6Bu + hi);
    mask |= ((w & 0xffffu) >= thr16 ? 1u : 0u) << (2 * i);
    mask |= ((w >> 16) >= thr16 ? 1u : 0u) << (2 * i + 1);
  }
  return mask;
}

template <typename T>
struct SrCapable { static constexpr bool value = false; };
template <>
struct SrCapable<__nv_bfloat16> { static constexpr bool value = true; };
// p' = p*decay - lr*u in fp32, no intermediate rounding (the stochastic store is the only rounding)
__device__ __forceinline__ void apply8_fp32(float (&p)[8], const float (&u)[8], float lr, float decay) {
#pragma unroll
  for (int i = 0; i < 8; ++i) p[i] = fmaf(-lr, u[i], p[i] * decay);
}
__device__ __forceinline__ uint32_t sr_add(uint32_t bits, uint32_t r16) {
  return (bits & 0x7f800000u) == 0x7f800000u ? bits : bits + r16;       // leave inf / nan alone
}
__device__ __forceinline__ void store8_sr(__nv_bfloat16* __restrict__ ptr, const float (&v)[8], uint64_t key,
                                          int64_t vec) {
  uint32_t r[4], w[4];
  hash_bits8(key, vec, r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t a = sr_add(__float_as_uint(v[2 * i]), r[i] & 0xffffu);
    const uint32_t b = sr_add(__float_as_uint(v[2 * i + 1]), r[i] >> 16);
    w[i] = (a >> 16) | (b & 0xffff0000u);
  }
  *reinterpret_cast<uint4*>(ptr) = make_uint4(w[0], w[1], w[2], w[3]);
}
// Parameter update + store in either mode; SR falls back to the reference rounding for types without it.
template <typename T>
__device__ __forceinline__ void apply_store8(T* __restrict__ ptr, float (&p)[8], const float (&u)[8], float lr,
                                             float decay, uint64_t sr_key, int64_t vec) {
  if constexpr (SrCapable<T>::value) {
    if (sr_key != 0) {
      apply8_fp32(p, u, lr, decay);
      store8_sr(ptr, p, sr_key, vec);
      return;
    }
  }
  apply8<T>(p, u, lr, decay);
  store8<T>(ptr, p);
}

// scalar forms for ragged tails
template <typename T>
__device__ __forceinline__ float lion_blend(float m, float g, float beta1, float omb1) {
  return fmaf(omb1, g, DT<T>::rnd(m * beta1));
}
template <typename T>
__device__ __forceinline__ float lion_momentum(float m, float g, float beta2, float omb2) {
  return fmaf(omb2, g, DT<T>::rnd(m * beta2));
}
template <typename T>
__device__ __forceinline__ float lion_apply(float p, float u, float lr, float decay) {
  return fmaf(-lr, u, DT<T>::rnd(p * decay));
}

// ------------------------------------------------------------------ Philox4x32-10
struct Philox {
  __device__ __forceinline__ static uint4 gen(uint64_t seed, uint64_t ctr_lo, uint32_t ctr_hi0, uint32_t ctr_hi1) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = ctr_hi0, c3 = ctr_hi1;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      k0 += 0x9E3779B9u;
      k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
  __device__ __forceinline__ static float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
};

// the byte of 8 sign bits for 8 consecutive elements starting at global element index e0
template <typename T>
__device__ __forceinline__ uint32_t sign_byte(const float (&m)[8], const float (&g)[8], float beta1, float omb1,
                                              float stoch_r, uint64_t seed, uint32_t step, uint32_t rank,
                                              int64_t e0) {
  float c[8];
  blend8<T>(m, g, beta1, omb1, c);
  uint32_t byte = 0;
  if (stoch_r > 0.f) {
    const uint4 r0 = Philox::gen(seed, (uint64_t)e0, step, rank);
    const uint4 r1 = Philox::gen(seed, (uint64_t)e0 + 4, step, rank);
    const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    const float inv = 1.0f / (2.0f * stoch_r);
#pragma unroll
    for (int i = 0; i < 8; ++i) byte |= (Philox::u01(rr[i]) < (c[i] + stoch_r) * inv ? 1u : 0u) << i;
  } else {
    const float thr = tie_to_zero<T>();
#pragma unroll
    for (int i = 0; i < 8; ++i) byte |= (c[i] > thr ? 1u : 0u) << i;
  }
  return byte;
}

// norm_tma.cu: TMA tile-ring forward for cols = 256 / 512 / 768 / 1024 (dispatched from dlion_norm_fwd)
bool norm_tma_ok(int cols);
template <typename T>
int launch_norm_fwd_tma(const void* x, const void* res, const void* w, const void* b, void* y, void* sum_out, float* mean,
                        float* rstd, int64_t rows, int cols, float eps, bool rms, bool drop, float inv_keep, uint32_t thr16,
                        uint64_t seed, cudaStream_t s);

// direction u for one element given the +1 count among L live workers
__device__ __forceinline__ float vote_dir(int cnt, int L, int vote) {
  const int twice = 2 * cnt;
  if (vote == DLION_VOTE_AVERAGE) return __fdiv_rn((float)(twice - L), (float)L);
  if (vote == DLION_VOTE_MAJORITY_ZERO && twice == L) return 0.f;
  return twice > L ? 1.f : -1.f;
}

}  // namespace dlion
