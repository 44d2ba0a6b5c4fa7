// lion_elementwise.cu — the non-exchanging Lion kernels:
//   dlion_lion_local   (N6)  single-worker Lion in one pass
//   dlion_sign_pack          sign(b1*m+(1-b1)*g) -> 32 signs / word          } the "nccl_1bit"
//   dlion_vote_update        gathered words -> vote -> p, m                  } baseline pair
// All three are pure streaming kernels: 16-byte vector accesses, 8 elements per thread per
// trip, grid-stride over the bucket; the roofline is HBM bytes (10 B/param in bf16 for the
// update kernels).  reference: distributed_lion.py:47-96.
#include "common.cuh"

namespace dlion {

constexpr int kThreads = 256;

// Gradient clipping folded into the optimizer kernels: the coefficient clip_grad_norm_ would apply, read on
// the device from the sum of squares (no host sync).  1 (and `false`) = nothing to clip; a NaN norm also
// leaves the gradients alone, like dlion_clip_scale.
struct ClipArg {
  const float* sumsq;   // device pointer or nullptr
  float max_norm;
};
__device__ __forceinline__ bool clip_coef(const ClipArg& c, float& coef) {
  coef = 1.f;
  if (c.sumsq == nullptr) return false;
  const float k = c.max_norm / (sqrtf(__ldg(c.sumsq)) + 1e-6f);
  if (!(k < 1.0f)) return false;
  coef = k;
  return true;
}
// g' = round_T(g * coef): what a separate clip kernel would have stored
template <typename T>
__device__ __forceinline__ void clip8(float (&g)[8], float coef) {
#pragma unroll
  for (int i = 0; i < 8; ++i) g[i] *= coef;
  round8<T>(g);
}
template <typename T>
__device__ __forceinline__ void zero8(T* __restrict__ ptr) {
#pragma unroll
  for (int i = 0; i < (int)(sizeof(T) * 8 / 16); ++i) reinterpret_cast<uint4*>(ptr)[i] = make_uint4(0u, 0u, 0u, 0u);
}

static inline int grid_for(int64_t nvec) {
  const int64_t want = (nvec + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)kNumSMs * 16;   // 8 resident CTAs/SM * 2 waves of grid-stride
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

// ------------------------------------------------------------------ N6
template <typename T>
__global__ void __launch_bounds__(kThreads) lion_local_kernel(T* __restrict__ p, T* __restrict__ g,
                                                              T* __restrict__ m, int64_t n, float lr, float decay,
                                                              float beta1, float omb1, float beta2, float omb2,
                                                              uint64_t sr_key, int64_t vec_base, ClipArg clip_arg,
                                                              int zero_g) {
  const int64_t nvec = n >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float coef;
  const bool clip = clip_coef(clip_arg, coef);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float pf[8], gf[8], mf[8];
    load8<T>(g + v * 8, gf);
    load8<T>(m + v * 8, mf);
    load8<T>(p + v * 8, pf);
    if (clip) clip8<T>(gf, coef);
    if (zero_g) zero8<T>(g + v * 8);
    float c[8], u[8];
    blend8<T>(mf, gf, beta1, omb1, c);
    const float thr = tie_to_zero<T>();
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = c[i] > thr ? 1.f : (c[i] < -thr ? -1.f : 0.f);   // torch.sign: sign(0) = 0
    apply_store8<T>(p + v * 8, pf, u, lr, decay, sr_key, vec_base + v);
    momentum8<T>(mf, gf, beta2, omb2);
    store8<T>(m + v * 8, mf);
  }
  // ragged tail (< 8 elements): one thread (always round-to-nearest)
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (int64_t e = nvec << 3; e < n; ++e) {
      float gf = DT<T>::to_f(g[e]);
      const float mf = DT<T>::to_f(m[e]), pf = DT<T>::to_f(p[e]);
      if (clip) gf = DT<T>::rnd(gf * coef);
      if (zero_g) g[e] = DT<T>::from_f(0.f);
      const float c = DT<T>::rnd(lion_blend<T>(mf, gf, beta1, omb1));
      const float u = c > 0.f ? 1.f : (c < 0.f ? -1.f : 0.f);
      p[e] = DT<T>::from_f(lion_apply<T>(pf, u, lr, decay));
      m[e] = DT<T>::from_f(lion_momentum<T>(mf, gf, beta2, omb2));
    }
  }
}

// ------------------------------------------------------------------ sign + pack
template <typename T>
__global__ void __launch_bounds__(kThreads) sign_pack_kernel(const T* __restrict__ m, const T* __restrict__ g,
                                                             uint32_t* __restrict__ words, int64_t n, float beta1,
                                                             float omb1, float stoch_r, uint64_t seed, uint32_t step,
                                                             uint32_t rank, ClipArg clip_arg) {
  float coef;
  const bool clip = clip_coef(clip_arg, coef);
  // nvec4 = number of 32-element groups, rounded up: every lane quad owns one output word
  const int64_t nwords = (n + 31) >> 5;
  const int64_t nvec_padded = nwords << 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;   // multiple of 32
  for (int64_t vb = (int64_t)blockIdx.x * blockDim.x; vb < nvec_padded; vb += stride) {  // block-uniform bound:
    const int64_t v = vb + threadIdx.x;                                                    // shuffles stay converged
    const int64_t e0 = v << 3;
    uint32_t byte = 0;
    if (e0 + 8 <= n) {
      float gf[8], mf[8];
      load8<T>(g + e0, gf);
      load8<T>(m + e0, mf);
      if (clip) clip8<T>(gf, coef);
      byte = sign_byte<T>(mf, gf, beta1, omb1, stoch_r, seed, step, rank, e0);
    } else if (e0 < n) {
      float gf[8], mf[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const bool ok = e0 + i < n;
        gf[i] = ok ? DT<T>::to_f(g[e0 + i]) : 0.f;
        mf[i] = ok ? DT<T>::to_f(m[e0 + i]) : 0.f;
      }
      if (clip) clip8<T>(gf, coef);
      byte = sign_byte<T>(mf, gf, beta1, omb1, stoch_r, seed, step, rank, e0);
      byte &= (1u << (int)(n - e0)) - 1u;   // pad bits never vote
    }
    // quad of lanes -> one little-endian word (byte j = elements 8j..8j+7 of the word)
    uint32_t w = byte << (8 * (threadIdx.x & 3));
    w |= __shfl_xor_sync(0xffffffffu, w, 1);
    w |= __shfl_xor_sync(0xffffffffu, w, 2);
    if ((threadIdx.x & 3) == 0 && v < nvec_padded) words[v >> 2] = w;
  }
}

// ------------------------------------------------------------------ vote + update
template <typename T>
__global__ void __launch_bounds__(kThreads) vote_update_kernel(T* __restrict__ p, T* __restrict__ g,
                                                               T* __restrict__ m, const uint8_t* __restrict__ gathered,
                                                               int64_t row_bytes, int world, uint32_t live_mask,
                                                               int L, int64_t n, float lr, float decay, float beta2,
                                                               float omb2, int vote, uint64_t sr_key, ClipArg clip_arg,
                                                               int zero_g) {
  float coef;
  const bool clip = clip_coef(clip_arg, coef);
  const int64_t nvec = (n + 7) >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const int64_t e0 = v << 3;
    // bit-sliced ballot count over the 8 lanes of this byte: planes c0..c4 hold the binary
    // digits of the per-element count (world <= 16 -> count <= 16 needs 5 bits)
    uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
    for (int r = 0; r < world; ++r) {
      if (!((live_mask >> r) & 1u)) continue;
      uint32_t x = gathered[(int64_t)r * row_bytes + v];
      uint32_t t = c0 & x; c0 ^= x; x = t;
      t = c1 & x; c1 ^= x; x = t;
      t = c2 & x; c2 ^= x; x = t;
      t = c3 & x; c3 ^= x; x = t;
      c4 ^= x;
    }
    const bool full = e0 + 8 <= n;
    float pf[8], gf[8], mf[8];
    if (full) {
      load8<T>(g + e0, gf);
      load8<T>(m + e0, mf);
      load8<T>(p + e0, pf);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const bool ok = e0 + i < n;
        gf[i] = ok ? DT<T>::to_f(g[e0 + i]) : 0.f;
        mf[i] = ok ? DT<T>::to_f(m[e0 + i]) : 0.f;
        pf[i] = ok ? DT<T>::to_f(p[e0 + i]) : 0.f;
      }
    }
    if (clip) clip8<T>(gf, coef);
    if (zero_g) {
      if (full) zero8<T>(g + e0);
      else
        for (int i = 0; i < 8 && e0 + i < n; ++i) g[e0 + i] = DT<T>::from_f(0.f);
    }
    float uf[8];
    if (vote == DLION_VOTE_AVERAGE) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int cnt = ((c0 >> i) & 1) | (((c1 >> i) & 1) << 1) | (((c2 >> i) & 1) << 2) | (((c3 >> i) & 1) << 3) |
                        (((c4 >> i) & 1) << 4);
        uf[i] = vote_dir(cnt, L, vote);
      }
    } else {
      // bit-sliced comparison of the 8 counts with the (uniform) thresholds: one byte of
      // "majority" bits and one of "tie" bits instead of 8 reassembled integers
      const uint32_t pc[5] = {c0, c1, c2, c3, c4};
      const int thr = L / 2 + 1;
      uint32_t gt = 0u, eq = 0xffu, tz = 0xffu;
#pragma unroll
      for (int k = 4; k >= 0; --k) {
        const uint32_t kb = ((thr >> k) & 1) ? 0xffu : 0u;
        const uint32_t hb = (((L / 2) >> k) & 1) ? 0xffu : 0u;
        gt |= eq & pc[k] & ~kb;
        eq &= ~(pc[k] ^ kb);
        tz &= ~(pc[k] ^ hb);
      }
      const uint32_t maj = (gt | eq) & 0xffu;
      const uint32_t tie = (vote == DLION_VOTE_MAJORITY_ZERO && (L & 1) == 0) ? (tz & 0xffu) : 0u;
#pragma unroll
      for (int i = 0; i < 8; ++i) uf[i] = ((tie >> i) & 1u) ? 0.f : (((maj >> i) & 1u) ? 1.f : -1.f);
    }
    if (full) {
      apply_store8<T>(p + e0, pf, uf, lr, decay, sr_key, v);
      momentum8<T>(mf, gf, beta2, omb2);
      store8<T>(m + e0, mf);
    } else {
      apply8<T>(pf, uf, lr, decay);
      momentum8<T>(mf, gf, beta2, omb2);
      for (int i = 0; i < 8 && e0 + i < n; ++i) {
        p[e0 + i] = DT<T>::from_f(pf[i]);
        m[e0 + i] = DT<T>::from_f(mf[i]);
      }
    }
  }
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_lion_local_clip(void* p, void* g, void* m, int64_t n, int dtype, float lr, float decay,
                                     float beta1, float omb1, float beta2, float omb2, uint64_t sr_key,
                                     int64_t elem_base, const float* clip_sumsq, float clip_max_norm, int zero_g,
                                     void* stream) {
  DLION_CHECK_ARG(n >= 0, "dlion_lion_local: negative n");
  DLION_CHECK_ARG(clip_sumsq == nullptr || clip_max_norm > 0.f, "dlion_lion_local: clip_max_norm must be positive");
  DLION_CHECK_ARG(elem_base >= 0 && (elem_base & 7) == 0, "dlion_lion_local: elem_base must be a multiple of 8");
  if (n == 0) return 0;
  DLION_CHECK_ARG(aligned16(p) && aligned16(g) && aligned16(m), "dlion_lion_local: pointers must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = grid_for(n >> 3);
  const ClipArg ca{clip_sumsq, clip_max_norm};
  DISPATCH_DTYPE(dtype, lion_local_kernel<T><<<grid, kThreads, 0, s>>>((T*)p, (T*)g, (T*)m, n, lr, decay, beta1, omb1,
                                                                         beta2, omb2, sr_key, elem_base >> 3, ca,
                                                                         zero_g));
  return DLION_LAUNCH_CHECK("dlion_lion_local");
}

extern "C" int dlion_lion_local_sr(void* p, const void* g, void* m, int64_t n, int dtype, float lr, float decay,
                                   float beta1, float omb1, float beta2, float omb2, uint64_t sr_key,
                                   int64_t elem_base, void* stream) {
  return dlion_lion_local_clip(p, const_cast<void*>(g), m, n, dtype, lr, decay, beta1, omb1, beta2, omb2, sr_key,
                               elem_base, nullptr, 0.f, 0, stream);
}

extern "C" int dlion_lion_local(void* p, const void* g, void* m, int64_t n, int dtype, float lr, float decay,
                                float beta1, float omb1, float beta2, float omb2, void* stream) {
  return dlion_lion_local_sr(p, g, m, n, dtype, lr, decay, beta1, omb1, beta2, omb2, 0, 0, stream);
}

extern "C" int dlion_sign_pack_clip(const void* m, const void* g, uint32_t* words, int64_t n, int dtype, float beta1,
                                    float omb1, float stoch_r, uint64_t seed, uint32_t step, uint32_t rank,
                                    const float* clip_sumsq, float clip_max_norm, void* stream) {
  DLION_CHECK_ARG(n >= 0, "dlion_sign_pack: negative n");
  DLION_CHECK_ARG(clip_sumsq == nullptr || clip_max_norm > 0.f, "dlion_sign_pack: clip_max_norm must be positive");
  if (n == 0) return 0;
  DLION_CHECK_ARG(aligned16(g) && aligned16(m), "dlion_sign_pack: pointers must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = grid_for(((n + 31) >> 5) << 2);
  const ClipArg ca{clip_sumsq, clip_max_norm};
  DISPATCH_DTYPE(dtype, sign_pack_kernel<T><<<grid, kThreads, 0, s>>>((const T*)m, (const T*)g, words, n, beta1, omb1,
                                                                        stoch_r, seed, step, rank, ca));
  return DLION_LAUNCH_CHECK("dlion_sign_pack");
}

extern "C" int dlion_sign_pack(const void* m, const void* g, uint32_t* words, int64_t n, int dtype, float beta1,
                               float omb1, float stoch_r, uint64_t seed, uint32_t step, uint32_t rank, void* stream) {
  return dlion_sign_pack_clip(m, g, words, n, dtype, beta1, omb1, stoch_r, seed, step, rank, nullptr, 0.f, stream);
}

extern "C" int dlion_vote_update_clip(void* p, void* g, void* m, const uint32_t* gathered, int64_t row_words,
                                      int world, uint32_t live_mask, int64_t n, int dtype, float lr, float decay,
                                      float beta2, float omb2, int vote, uint64_t sr_key, const float* clip_sumsq,
                                      float clip_max_norm, int zero_g, void* stream) {
  DLION_CHECK_ARG(n >= 0, "dlion_vote_update: negative n");
  DLION_CHECK_ARG(clip_sumsq == nullptr || clip_max_norm > 0.f, "dlion_vote_update: clip_max_norm must be positive");
  DLION_CHECK_ARG(world >= 1 && world <= DLION_MAX_RANKS, "dlion_vote_update: world %d outside [1, %d]", world,
                  DLION_MAX_RANKS);
  DLION_CHECK_ARG(vote >= 0 && vote <= 2, "dlion_vote_update: bad vote rule %d", vote);
  DLION_CHECK_ARG(row_words * 32 >= n, "dlion_vote_update: rows of %lld words cannot hold %lld signs",
                  (long long)row_words, (long long)n);
  const uint32_t all = world >= 32 ? 0xffffffffu : ((1u << world) - 1u);
  live_mask &= all;
  const int L = __builtin_popcount(live_mask);
  DLION_CHECK_ARG(L > 0, "dlion_vote_update: live mask excludes every rank");
  if (n == 0) return 0;
  DLION_CHECK_ARG(aligned16(p) && aligned16(g) && aligned16(m), "dlion_vote_update: pointers must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = grid_for((n + 7) >> 3);
  const ClipArg ca{clip_sumsq, clip_max_norm};
  DISPATCH_DTYPE(dtype, vote_update_kernel<T><<<grid, kThreads, 0, s>>>(
                            (T*)p, (T*)g, (T*)m, (const uint8_t*)gathered, row_words * 4, world, live_mask, L, n, lr,
                            decay, beta2, omb2, vote, sr_key, ca, zero_g));
  return DLION_LAUNCH_CHECK("dlion_vote_update");
}

extern "C" int dlion_vote_update_sr(void* p, const void* g, void* m, const uint32_t* gathered, int64_t row_words,
                                    int world, uint32_t live_mask, int64_t n, int dtype, float lr, float decay,
                                    float beta2, float omb2, int vote, uint64_t sr_key, void* stream) {
  return dlion_vote_update_clip(p, const_cast<void*>(g), m, gathered, row_words, world, live_mask, n, dtype, lr, decay,
                                beta2, omb2, vote, sr_key, nullptr, 0.f, 0, stream);
}

extern "C" int dlion_vote_update(void* p, const void* g, void* m, const uint32_t* gathered, int64_t row_words,
                                 int world, uint32_t live_mask, int64_t n, int dtype, float lr, float decay,
                                 float beta2, float omb2, int vote, void* stream) {
  return dlion_vote_update_sr(p, g, m, gathered, row_words, world, live_mask, n, dtype, lr, decay, beta2, omb2, vote, 0,
                              stream);
}
