// lion_fused_sm100.cu — N1 / N7: the fused sign-pack-exchange-vote-update kernel.
//
// One persistent, cooperatively launched kernel per bucket chunk does the whole Distributed
// Lion step of every rank at once; the sign exchange happens INSIDE the kernel as peer stores
// over NVLink, there is no NCCL call and no second kernel:
//
//   warp 0  (loader)  one lane streams tiles of g, m, p from HBM into an 8-stage shared-memory
//                     ring with cp.async.bulk (TMA, 1-D) completing on mbarriers;
//   warps 2-9 (compute, software-pipelined LOOKAHEAD tiles deep):
//      produce(t): sign(b1*m + (1-b1)*g) for the tile from smem -> 32 signs/word ->
//                  one warp per destination rank stores the tile's words (16 B / lane, one
//                  coalesced request) into that rank's receive arena -- a peer-mapped pointer,
//                  so the store travels over NVSwitch -- then fence + st.release.sys of a
//                  per-(source rank, CTA) progress counter on that rank;
//      consume(t): after the voter has published the vote planes of tile t: decoupled weight
//                  decay + voted update on p, momentum EMA on m, 16-byte stores to HBM;
//   warp 1  (voter)   ld.acquire.sys-polls the local progress counters of all awaited ranks,
//                     reads the W rows of packed words of the tile from the local arena,
//                     adds them bit-sliced (5 carry-save planes = per-element ballot count
//                     0..16), and publishes either {majority, tie} planes or the count planes.
//
// Between produce(t) and consume(t) the CTA works on LOOKAHEAD other tiles, which hides the
// NVLink store + flag latency; the loader keeps the remaining stages of the ring in flight,
// which hides HBM latency.  HBM traffic is the minimum: read g, m, p once, write p, m once
// (10 B/param in bf16) + W/8 B/param of gathered words.
//
// Liveness: every CTA is co-resident (cooperative launch, 1 CTA/SM) and each CTA issues
// produce(t) strictly before consume(t), so a rank never waits for data whose producer is
// blocked behind it.  Arenas are double-buffered by launch parity: a rank can be at most one
// launch ahead of any peer (it needs the peer's flags of the current launch to finish it).
// All waits are bounded: a flag that never arrives is reported in `status` and the kernel
// runs to completion instead of hanging the GPU.
//
// reference: distributed_lion.py:61-96 (update_fn_distributed: ~52 ATen kernels, a blocking
// dist.all_gather of an int64 tensor and torch.mode per parameter tensor) and :98-136
// (update_fn_distributed_stoc, dead code upstream) for the stochastic variant.
#include <cooperative_groups.h>

#include "common.cuh"

namespace dlion {

constexpr int kStages = 8;
constexpr int kLookahead = 4;
constexpr int kArrayBytes = 8192;            // bytes of g (or m, or p) per tile
constexpr int kStageBytes = 3 * kArrayBytes;
constexpr int kComputeWarps = 8;
constexpr int kComputeThreads = kComputeWarps * 32;
constexpr int kFusedThreads = 64 + kComputeThreads;   // loader warp + voter warp + compute warps
constexpr int64_t kMbarSpinLimit = 1ll << 24;
constexpr int64_t kDefaultFlagSpin = 1ll << 24;

struct FusedParams {
  void* p;
  const void* g;
  void* m;
  int64_t n;
  int64_t elem_base;
  int world, rank;
  uint32_t live_mask, wait_mask;
  int L;
  float lr, decay, beta1, omb1, beta2, omb2;
  int vote;
  float stoch_r;
  uint64_t seed;
  uint32_t step;
  uint64_t epoch;
  uint32_t* arena[DLION_MAX_RANKS];
  uint64_t* flags[DLION_MAX_RANKS];
  int64_t arena_words;
  int grid;
  int emulate_ranks;
  int64_t emu_stride;
  int32_t* status;
  int64_t spin_limit;
};

// ------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void report(int32_t* status, int code, int a, int b) {
  if (status != nullptr && atomicCAS(status, 0, code) == 0) {
    status[1] = a;
    status[2] = b;
    __threadfence_system();
  }
}
// bounded wait: a protocol bug must never hang the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int32_t* status, int who) {
  for (int64_t it = 0; !mbar_try_wait(bar, parity); ++it) {
    if (it > kMbarSpinLimit) {
      report(status, 2, (int)blockIdx.x, who);
      __trap();
    }
  }
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_load(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void st_release_sys(uint64_t* ptr, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(ptr), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* ptr) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(ptr) : "memory");
  return v;
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

template <typename T>
struct Cfg {
  static constexpr int TILE = kArrayBytes / (int)sizeof(T);   // 4096 (2-byte types) / 2048 (fp32)
  static constexpr int UNITS = TILE / 8;                      // 8-element units per tile
  static constexpr int WORDS = TILE / 32;                     // sign words per tile: 128 / 64
  static constexpr int UPT = UNITS / kComputeThreads;         // units per compute thread: 2 / 1
  static constexpr int WPL = WORDS / 32;                      // words per voter lane: 4 / 2
  static constexpr int PLANE_BYTES = WORDS * 4;
  static constexpr int SMEM_PLANES = kStages * 5 * PLANE_BYTES;
  static constexpr int SMEM_WBUF = kStages * PLANE_BYTES;
  static constexpr int SMEM_TOTAL = kStages * kStageBytes + SMEM_PLANES + SMEM_WBUF + 3 * kStages * 8 + 16;
};

// bit-sliced "count >= K" / "count == K" over 5 planes (count in 0..31), K uniform
__device__ __forceinline__ uint32_t planes_ge(const uint32_t (&c)[5], int K) {
  uint32_t gt = 0u, eq = 0xffffffffu;
#pragma unroll
  for (int k = 4; k >= 0; --k) {
    const uint32_t kb = ((K >> k) & 1) ? 0xffffffffu : 0u;
    gt |= eq & c[k] & ~kb;
    eq &= ~(c[k] ^ kb);
  }
  return gt | eq;
}
__device__ __forceinline__ uint32_t planes_eq(const uint32_t (&c)[5], int K) {
  uint32_t eq = 0xffffffffu;
#pragma unroll
  for (int k = 4; k >= 0; --k) {
    const uint32_t kb = ((K >> k) & 1) ? 0xffffffffu : 0u;
    eq &= ~(c[k] ^ kb);
  }
  return eq;
}

template <typename T>
__global__ void __launch_bounds__(kFusedThreads, 1) lion_fused_kernel(const __grid_constant__ FusedParams P) {
  using C = Cfg<T>;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* stage_base = smem;
  uint32_t* planes = reinterpret_cast<uint32_t*>(smem + kStages * kStageBytes);
  uint32_t* wbuf = reinterpret_cast<uint32_t*>(smem + kStages * kStageBytes + C::SMEM_PLANES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes + C::SMEM_PLANES + C::SMEM_WBUF);
  uint64_t* full_bar = bars;                 // TMA landed           (1 arrival + tx bytes)
  uint64_t* empty_bar = bars + kStages;      // tile fully consumed  (kComputeWarps arrivals)
  uint64_t* vote_bar = bars + 2 * kStages;   // vote planes ready    (1 arrival)

  const int G = P.grid;
  const bool emu = P.emulate_ranks > 0;
  const int vrank = emu ? (int)blockIdx.x / G : 0;
  const int cta = emu ? (int)blockIdx.x % G : (int)blockIdx.x;
  const int my_rank = emu ? vrank : P.rank;
  const int W = P.world;
  T* __restrict__ p = reinterpret_cast<T*>(P.p) + (int64_t)vrank * P.emu_stride;
  const T* __restrict__ g = reinterpret_cast<const T*>(P.g) + (int64_t)vrank * P.emu_stride;
  T* __restrict__ m = reinterpret_cast<T*>(P.m) + (int64_t)vrank * P.emu_stride;

  const int64_t ntiles = P.n / C::TILE;
  const int n_my = cta < ntiles ? (int)((ntiles - cta + G - 1) / G) : 0;
  const uint32_t parity_sel = (uint32_t)(P.epoch & 1ull);
  const uint64_t epoch_hi = P.epoch << 32;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      mbar_init(smem_u32(full_bar + s), 1);
      mbar_init(smem_u32(empty_bar + s), kComputeWarps);
      mbar_init(smem_u32(vote_bar + s), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (n_my == 0) return;

  if (warp == 0) {
    // ============================ loader ============================
    if (lane == 0) {
      for (int i = 0; i < n_my; ++i) {
        const int slot = i % kStages, k = i / kStages;
        if (k > 0) mbar_wait(smem_u32(empty_bar + slot), (uint32_t)((k - 1) & 1), P.status, 0);
        const int64_t off = ((int64_t)cta + (int64_t)i * G) * C::TILE;
        const uint32_t bar = smem_u32(full_bar + slot);
        const uint32_t dst = smem_u32(stage_base + slot * kStageBytes);
        mbar_arrive_expect_tx(bar, kStageBytes);
        bulk_load(dst, g + off, kArrayBytes, bar);
        bulk_load(dst + kArrayBytes, m + off, kArrayBytes, bar);
        bulk_load(dst + 2 * kArrayBytes, p + off, kArrayBytes, bar);
      }
    }
  } else if (warp == 1) {
    // ============================ voter ============================
    const uint64_t* my_flags = P.flags[my_rank];
    const uint32_t* my_arena = P.arena[my_rank];
    const int64_t spin_limit = P.spin_limit > 0 ? P.spin_limit : kDefaultFlagSpin;
    bool gave_up = false;
    const int thr = P.L / 2 + 1;             // +1 iff 2*cnt > L  <=>  cnt >= floor(L/2)+1
    for (int i = 0; i < n_my; ++i) {
      const int slot = i % kStages, k = i / kStages;
      if (k > 0) mbar_wait(smem_u32(empty_bar + slot), (uint32_t)((k - 1) & 1), P.status, 1);
      const uint64_t target = epoch_hi | (uint64_t)(i + 1);
      if (lane < W && ((P.wait_mask >> lane) & 1u) && !gave_up) {
        const uint64_t* f = my_flags + (int64_t)lane * DLION_FUSED_MAX_CTAS + cta;
        int64_t it = 0;
        while (ld_acquire_sys(f) < target) {
          if (++it > spin_limit) {
            report(P.status, 1, cta, lane);
            gave_up = true;
            break;
          }
        }
      }
      gave_up = __any_sync(0xffffffffu, gave_up);   // also orders every lane after the acquires
      const int64_t woff = ((int64_t)cta + (int64_t)i * G) * C::WORDS + lane * C::WPL;
      uint32_t c[C::WPL][5];
#pragma unroll
      for (int w = 0; w < C::WPL; ++w)
#pragma unroll
        for (int k2 = 0; k2 < 5; ++k2) c[w][k2] = 0u;
      for (int r = 0; r < W; ++r) {
        if (!((P.live_mask >> r) & 1u)) continue;
        const uint32_t* src = my_arena + ((int64_t)(parity_sel * W + r)) * P.arena_words + woff;
        uint32_t x[C::WPL];
        if constexpr (C::WPL == 4) {
          const uint4 q = __ldcv(reinterpret_cast<const uint4*>(src));
          x[0] = q.x; x[1] = q.y; x[2] = q.z; x[3] = q.w;
        } else {
          const uint2 q = __ldcv(reinterpret_cast<const uint2*>(src));
          x[0] = q.x; x[1] = q.y;
        }
#pragma unroll
        for (int w = 0; w < C::WPL; ++w) {   // carry-save add of one ballot bit per element
          uint32_t carry = x[w];
#pragma unroll
          for (int k2 = 0; k2 < 5; ++k2) {
            const uint32_t t = c[w][k2] & carry;
            c[w][k2] ^= carry;
            carry = t;
          }
        }
      }
      uint32_t* pl = planes + slot * 5 * C::WORDS + lane * C::WPL;
      if (P.vote == DLION_VOTE_AVERAGE) {
#pragma unroll
        for (int k2 = 0; k2 < 5; ++k2)
#pragma unroll
          for (int w = 0; w < C::WPL; ++w) pl[k2 * C::WORDS + w] = c[w][k2];
      } else {
#pragma unroll
        for (int w = 0; w < C::WPL; ++w) {
          pl[w] = planes_ge(c[w], thr);
          pl[C::WORDS + w] =
              (P.vote == DLION_VOTE_MAJORITY_ZERO && (P.L & 1) == 0) ? planes_eq(c[w], P.L / 2) : 0u;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(vote_bar + slot));
    }
  } else {
    // ============================ compute ============================
    const int ct = threadIdx.x - 64;
    const int cw = warp - 2;

    auto produce = [&](int i) {
      const int slot = i % kStages, k = i / kStages;
      mbar_wait(smem_u32(full_bar + slot), (uint32_t)(k & 1), P.status, 2);
      const T* sg = reinterpret_cast<const T*>(stage_base + slot * kStageBytes);
      const T* sm = reinterpret_cast<const T*>(stage_base + slot * kStageBytes + kArrayBytes);
      const int64_t tile = (int64_t)cta + (int64_t)i * G;
      uint32_t* wb = wbuf + slot * C::WORDS;
#pragma unroll
      for (int u = 0; u < C::UPT; ++u) {
        const int unit = ct + u * kComputeThreads;
        float gf[8], mf[8];
        load8<T>(sg + unit * 8, gf);
        load8<T>(sm + unit * 8, mf);
        const uint32_t byte = sign_byte<T>(mf, gf, P.beta1, P.omb1, P.stoch_r, P.seed, P.step, (uint32_t)my_rank,
                                           P.elem_base + tile * C::TILE + unit * 8);
        uint32_t wv = byte << (8 * (lane & 3));
        wv |= __shfl_xor_sync(0xffffffffu, wv, 1);
        wv |= __shfl_xor_sync(0xffffffffu, wv, 2);
        if ((lane & 3) == 0) wb[unit >> 2] = wv;
      }
      named_bar_sync(1, kComputeThreads);
      // one warp per destination rank: a single coalesced 16-byte-per-lane store, then publish
      for (int r = cw; r < W; r += kComputeWarps) {
        uint32_t* dst = P.arena[r] + ((int64_t)(parity_sel * W + my_rank)) * P.arena_words + tile * C::WORDS;
        if (lane < C::WORDS / 4)
          reinterpret_cast<uint4*>(dst)[lane] = reinterpret_cast<const uint4*>(wb)[lane];
        __syncwarp();
        if (lane == 0) {
          __threadfence_system();
          st_release_sys(P.flags[r] + (int64_t)my_rank * DLION_FUSED_MAX_CTAS + cta, epoch_hi | (uint64_t)(i + 1));
        }
      }
    };

    auto consume = [&](int i) {
      const int slot = i % kStages, k = i / kStages;
      mbar_wait(smem_u32(vote_bar + slot), (uint32_t)(k & 1), P.status, 3);
      const T* sg = reinterpret_cast<const T*>(stage_base + slot * kStageBytes);
      const T* sm = reinterpret_cast<const T*>(stage_base + slot * kStageBytes + kArrayBytes);
      const T* sp = reinterpret_cast<const T*>(stage_base + slot * kStageBytes + 2 * kArrayBytes);
      const uint8_t* pl = reinterpret_cast<const uint8_t*>(planes + slot * 5 * C::WORDS);
      const int64_t off = ((int64_t)cta + (int64_t)i * G) * C::TILE;
#pragma unroll
      for (int u = 0; u < C::UPT; ++u) {
        const int unit = ct + u * kComputeThreads;
        float gf[8], mf[8], pf[8];
        load8<T>(sg + unit * 8, gf);
        load8<T>(sm + unit * 8, mf);
        load8<T>(sp + unit * 8, pf);
        if (P.vote == DLION_VOTE_AVERAGE) {
          uint32_t b[5];
#pragma unroll
          for (int k2 = 0; k2 < 5; ++k2) b[k2] = pl[k2 * C::PLANE_BYTES + unit];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int cnt = ((b[0] >> e) & 1) | (((b[1] >> e) & 1) << 1) | (((b[2] >> e) & 1) << 2) |
                            (((b[3] >> e) & 1) << 3) | (((b[4] >> e) & 1) << 4);
            const float uu = __fdiv_rn((float)(2 * cnt - P.L), (float)P.L);
            pf[e] = lion_apply<T>(pf[e], uu, P.lr, P.decay);
            mf[e] = lion_momentum<T>(mf[e], gf[e], P.beta2, P.omb2);
          }
        } else {
          const uint32_t a = pl[unit], z = pl[C::PLANE_BYTES + unit];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float uu = ((z >> e) & 1u) ? 0.f : (((a >> e) & 1u) ? 1.f : -1.f);
            pf[e] = lion_apply<T>(pf[e], uu, P.lr, P.decay);
            mf[e] = lion_momentum<T>(mf[e], gf[e], P.beta2, P.omb2);
          }
        }
        store8<T>(p + off + unit * 8, pf);
        store8<T>(m + off + unit * 8, mf);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(empty_bar + slot));
    };

    const int pro = n_my < kLookahead ? n_my : kLookahead;
    for (int i = 0; i < pro; ++i) produce(i);
    for (int j = 0; j < n_my; ++j) {
      if (j + kLookahead < n_my) produce(j + kLookahead);
      consume(j);
    }
  }
}

template <typename T>
static int configure(int* max_grid) {
  static thread_local int cached_dev = -1;
  static thread_local int cached_grid = 0;
  int dev = 0;
  int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  if (rc) return rc;
  if (dev != cached_dev) {
    rc = check_cuda(cudaFuncSetAttribute(lion_fused_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg<T>::SMEM_TOTAL),
                    "cudaFuncSetAttribute(max dynamic smem)");
    if (rc) return rc;
    int per_sm = 0, sms = 0;
    rc = check_cuda(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lion_fused_kernel<T>, kFusedThreads,
                                                                  Cfg<T>::SMEM_TOTAL),
                    "cudaOccupancyMaxActiveBlocksPerMultiprocessor");
    if (rc) return rc;
    rc = check_cuda(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), "cudaDeviceGetAttribute");
    if (rc) return rc;
    cached_grid = per_sm * sms;
    cached_dev = dev;
  }
  *max_grid = cached_grid;
  return 0;
}

template <typename T>
static int launch_fused(const dlion_fused_args* a, cudaStream_t s) {
  using C = Cfg<T>;
  int max_grid = 0;
  int rc = configure<T>(&max_grid);
  if (rc) return rc;
  const int vr = a->emulate_ranks > 0 ? a->emulate_ranks : 1;
  DLION_CHECK_ARG(a->n % C::TILE == 0, "dlion_fused_step: n (%lld) must be a multiple of the tile (%d)", (long long)a->n,
                  C::TILE);
  DLION_CHECK_ARG(a->n / 32 <= a->arena_words, "dlion_fused_step: arena of %lld words cannot hold %lld signs",
                  (long long)a->arena_words, (long long)a->n);
  DLION_CHECK_ARG(a->grid >= 1 && a->grid <= DLION_FUSED_MAX_CTAS, "dlion_fused_step: grid %d outside [1, %d]", a->grid,
                  DLION_FUSED_MAX_CTAS);
  if (a->grid * vr > max_grid) {
    set_error("dlion_fused_step: %d CTAs requested but only %d can be co-resident on this device", a->grid * vr,
              max_grid);
    return DLION_E_TOOBIG;
  }
  FusedParams P;
  P.p = a->p; P.g = a->g; P.m = a->m; P.n = a->n; P.elem_base = a->elem_base;
  P.world = a->world; P.rank = a->rank;
  const uint32_t all = (1u << a->world) - 1u;
  P.live_mask = a->live_mask & all;
  P.wait_mask = a->wait_mask & all;
  P.L = __builtin_popcount(P.live_mask);
  DLION_CHECK_ARG(P.L > 0, "dlion_fused_step: live mask excludes every rank");
  P.lr = a->lr; P.decay = a->decay; P.beta1 = a->beta1; P.omb1 = a->omb1; P.beta2 = a->beta2; P.omb2 = a->omb2;
  P.vote = a->vote; P.stoch_r = a->stoch_r; P.seed = a->seed; P.step = a->step; P.epoch = a->epoch;
  for (int r = 0; r < DLION_MAX_RANKS; ++r) {
    P.arena[r] = a->arena[r];
    P.flags[r] = a->flags[r];
  }
  for (int r = 0; r < a->world; ++r)
    DLION_CHECK_ARG(P.arena[r] != nullptr && P.flags[r] != nullptr && aligned16(P.arena[r]),
                    "dlion_fused_step: arena/flags pointer of rank %d is null or unaligned", r);
  P.arena_words = a->arena_words; P.grid = a->grid; P.emulate_ranks = a->emulate_ranks; P.emu_stride = a->emu_stride;
  P.status = a->status; P.spin_limit = a->spin_limit;
  void* kargs[] = {(void*)&P};
  return check_cuda(cudaLaunchCooperativeKernel((void*)lion_fused_kernel<T>, dim3(a->grid * vr), dim3(kFusedThreads),
                                                kargs, (size_t)C::SMEM_TOTAL, s),
                    "cudaLaunchCooperativeKernel(lion_fused_kernel)");
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_fused_tile(int dtype) {
  switch (dtype) {
    case DLION_F32: return Cfg<float>::TILE;
    case DLION_BF16: return Cfg<__nv_bfloat16>::TILE;
    case DLION_F16: return Cfg<__half>::TILE;
  }
  return DLION_E_UNSUPPORTED;
}

extern "C" int dlion_fused_max_grid(int dtype, int* out) {
  DLION_CHECK_ARG(out != nullptr, "dlion_fused_max_grid: null out");
  DISPATCH_DTYPE(dtype, return configure<T>(out));
  return 0;
}

extern "C" int dlion_fused_step(const dlion_fused_args* a, void* stream) {
  DLION_CHECK_ARG(a != nullptr, "dlion_fused_step: null args");
  DLION_CHECK_ARG(a->world >= 1 && a->world <= DLION_MAX_RANKS, "dlion_fused_step: world %d outside [1, %d]", a->world,
                  DLION_MAX_RANKS);
  DLION_CHECK_ARG(a->emulate_ranks == 0 || a->emulate_ranks == a->world,
                  "dlion_fused_step: emulate_ranks must be 0 or equal to world");
  DLION_CHECK_ARG(a->emulate_ranks > 0 || (a->rank >= 0 && a->rank < a->world), "dlion_fused_step: bad rank %d", a->rank);
  DLION_CHECK_ARG(a->vote >= 0 && a->vote <= 2, "dlion_fused_step: bad vote rule %d", a->vote);
  DLION_CHECK_ARG(a->epoch >= 1 && a->epoch < (1ull << 31), "dlion_fused_step: epoch must be in [1, 2^31)");
  DLION_CHECK_ARG(aligned16(a->p) && aligned16(a->g) && aligned16(a->m), "dlion_fused_step: p/g/m must be 16-byte aligned");
  if (a->n == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(a->dtype, return launch_fused<T>(a, s));
  return 0;
}
