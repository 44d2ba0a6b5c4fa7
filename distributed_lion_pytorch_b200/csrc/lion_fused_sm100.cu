// lion_fused_sm100.cu — N1 / N7: the fused sign-pack-exchange-vote-update kernel.
//
// One persistent, cooperatively launched kernel per bucket chunk does the whole Distributed
// Lion step of a rank; the sign exchange happens INSIDE the kernel as peer stores over
// NVLink — there is no NCCL call and no second kernel.  The step is split into two streaming
// pipelines that run concurrently in every CTA and are coupled only through the peers' flags:
//
//   P pipeline (needs g, m):  sign(b1*m + (1-b1)*g) -> 32 signs / word -> words to every rank's
//                             receive arena;   m' = b2*m + (1-b2)*g written back immediately.
//   C pipeline (needs p):     wait for the tile's words from all awaited ranks -> bit-sliced
//                             vote -> p' = p*(1 - lr*wd) - lr*u written back.
//
// Nothing is carried from P to C inside the CTA (a rank's own words go through its own arena
// like everybody else's), so P may run arbitrarily far ahead of C: NVLink store + flag latency
// is hidden by distance, not by shared-memory capacity.  HBM traffic is the minimum: read g, m,
// p once, write m, p once (10 B/param in bf16) + W/8 B/param of gathered words.
//
// Two things a training step needs next to the update are folded into the same pass instead of being
// kernels of their own (reference: HF Trainer's clip_grad_norm_ before optimizer.step() and
// model.zero_grad() after it, run_clm.py:588-622 via transformers.Trainer):
//   clip     `clip_sumsq` points at the device-side sum of g^2; g' = round_T(g * coef) is formed in
//            registers exactly where a separate clip kernel would round it, and feeds sign and momentum;
//   zero     `zero_g`: each consumed g tile is overwritten with zeros by ONE cp.async.bulk
//            shared->global copy from a zero page in shared memory (no thread stores anything).
//
// Warp roles (24 warps, 1 CTA per SM; every compute role has 2 warps per SM sub-partition):
//   0      P-loader   one lane: cp.async.bulk (TMA 1-D) of g, m tiles into a 6-stage smem ring
//   1      C-loader   one lane: cp.async.bulk of p tiles into an 8-stage smem ring
//   2-3    senders    drain a smem ring of packed tiles in batches: one coalesced 16 B/lane
//                     store per (tile, destination rank) into the peer-mapped arena, then ONE
//                     fence.acq_rel.sys + progress-counter store per batch and destination.
//                     The fence latency is paid by these warps only and amortised over the batch.
//   4-7    voters     (tile i -> voter i % 4) ld.acquire.sys on the local progress counters, read
//                     the W word rows of the tile from the local arena (all loads issued before
//                     the first use), carry-save add them into 5 bit planes (ballot count 0..16
//                     per element), publish {majority, tie} or the count planes
//   8-15   P-compute  sign bytes + momentum from smem; words into the sender ring
//   16-23  C-compute  decay + voted update from smem p and the vote planes
//
// Liveness: all CTAs are co-resident (cooperative launch); the P pipeline never waits on a
// remote event, so every flag a peer waits for is eventually produced.  Arenas are
// double-buffered by launch parity: a rank can be at most one launch ahead of a peer (it needs
// the peer's flags of the current launch to finish it).  All waits are bounded: a flag that
// never arrives is reported in `status` and the kernel runs to completion.
//
// reference: distributed_lion.py:61-96 (update_fn_distributed: ~52 ATen kernels, a blocking
// dist.all_gather of an int64 tensor and torch.mode per parameter tensor) and :98-136
// (update_fn_distributed_stoc, dead code upstream) for the stochastic variant.
#include <cooperative_groups.h>

#include "common.cuh"

namespace dlion {

constexpr int kArrayBytes = 8192;                 // bytes of g (or m, or p) per tile
constexpr int kPStages = 6;                       // g+m ring
constexpr int kPStageBytes = 2 * kArrayBytes;
constexpr int kCStages = 8;                       // p ring
constexpr int kCStageBytes = kArrayBytes;
constexpr int kWordRing = 32;                     // packed tiles buffered for the senders
constexpr int kSenders = 2;
constexpr int kVoters = 4;
constexpr int kPWarps = 8, kCWarps = 8;
constexpr int kPThreads = kPWarps * 32, kCThreads = kCWarps * 32;
constexpr int kWarpPLoad = 0, kWarpCLoad = 1, kWarpSend0 = 2, kWarpVoter0 = kWarpSend0 + kSenders;
constexpr int kWarpP0 = kWarpVoter0 + kVoters, kWarpC0 = kWarpP0 + kPWarps;
constexpr int kFusedThreads = (kWarpC0 + kCWarps) * 32;     // 24 warps = 768 threads
static_assert(kCStages % kVoters == 0, "a voter must always meet the same ring slots");
constexpr int64_t kMbarSpinLimit = 1ll << 26;
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
  uint32_t* mc_arena;
  int64_t arena_words;
  int grid;
  int emulate_ranks;
  int64_t emu_stride;
  int32_t* status;
  int64_t spin_limit;
  uint64_t sr_key;
  const float* clip_sumsq;
  float clip_max_norm;
  int zero_g;
  uint64_t wait_timeout_ns;     // 0 = poll-count limit only
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
// TMA 1-D bulk copy shared -> global (bulk async-group completion)
__device__ __forceinline__ void bulk_store(void* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void st_relaxed_sys(uint64_t* ptr, uint64_t v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(ptr), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_sys(const uint64_t* ptr) {
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(ptr) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* ptr) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(ptr) : "memory");
  return v;
}
// one 16-byte store to an NVLS multicast address: the switch writes it into every rank's replica
__device__ __forceinline__ void multimem_st_v4(uint32_t* mc_ptr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_ptr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void st_release_cta(uint32_t* smem_ptr, uint32_t v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_u32(smem_ptr)), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_cta(const uint32_t* smem_ptr) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(smem_ptr)) : "memory");
  return v;
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

template <typename T>
struct Cfg {
  static constexpr int TILE = kArrayBytes / (int)sizeof(T);   // 4096 (2-byte types) / 2048 (fp32)
  static constexpr int UNITS = TILE / 8;                      // 8-element units per tile: 512 / 256
  static constexpr int WORDS = TILE / 32;                     // sign words per tile: 128 / 64
  static constexpr int UPT_P = UNITS / kPThreads;             // units per P-compute thread: 4 / 2
  static constexpr int UPT_C = UNITS / kCThreads;
  static constexpr int WPL = WORDS / 32;                      // words per voter lane: 4 / 2
  static constexpr int PLANE_BYTES = WORDS * 4;
  static constexpr int OFF_PRING = 0;
  static constexpr int OFF_CRING = OFF_PRING + kPStages * kPStageBytes;
  static constexpr int OFF_WRING = OFF_CRING + kCStages * kCStageBytes;
  static constexpr int OFF_PLANES = OFF_WRING + kWordRing * PLANE_BYTES;
  static constexpr int OFF_BARS = OFF_PLANES + kCStages * 5 * PLANE_BYTES;
  static constexpr int NUM_BARS = 2 * kPStages + 3 * kCStages;
  static constexpr int OFF_CTRS = OFF_BARS + NUM_BARS * 8;
  static constexpr int OFF_SKIP = OFF_CTRS + 64;                      // [kCStages] u32: tile's vote is void (a wait expired)
  static constexpr int OFF_ZERO = (OFF_SKIP + kCStages * 4 + 127) / 128 * 128;   // zero page: source of the g zero-fill
  static constexpr int SMEM_TOTAL = OFF_ZERO + kArrayBytes;
};

template <int N> struct VecOf;
template <> struct VecOf<4> { using type = uint4; };
template <> struct VecOf<2> { using type = uint2; };

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
  uint8_t* pring = smem + C::OFF_PRING;                                    // [kPStages][g | m]
  uint8_t* cring = smem + C::OFF_CRING;                                    // [kCStages][p]
  uint32_t* wring = reinterpret_cast<uint32_t*>(smem + C::OFF_WRING);      // [kWordRing][WORDS]
  uint32_t* planes = reinterpret_cast<uint32_t*>(smem + C::OFF_PLANES);    // [kCStages][5][WORDS]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BARS);
  uint64_t* fullP = bars;                          // g, m landed              (1 arrival + tx)
  uint64_t* emptyP = fullP + kPStages;             // P-compute done with slot (kPWarps arrivals)
  uint64_t* fullC = emptyP + kPStages;             // p landed                 (1 arrival + tx)
  uint64_t* emptyC = fullC + kCStages;             // C-compute done with slot (kCWarps arrivals)
  uint64_t* voteC = emptyC + kCStages;             // vote planes ready        (1 arrival)
  uint32_t* ctrs = reinterpret_cast<uint32_t*>(smem + C::OFF_CTRS);
  uint32_t* words_done = ctrs;                     // tiles whose packed words are in wring
  uint32_t* sent = ctrs + 1;                       // [kSenders] tiles each sender has pushed out
  uint32_t* skip = reinterpret_cast<uint32_t*>(smem + C::OFF_SKIP);   // [kCStages]

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
  // averaged vote: direction for every possible ballot count, divided once per CTA instead of once per element
  // (8 IEEE divisions per 8-element unit made the C pipeline of the averaged variant twice as slow as the majority one)
  __shared__ float avg_lut[32];
  if (threadIdx.x < 32) avg_lut[threadIdx.x] = __fdiv_rn((float)(2 * (int)threadIdx.x - P.L), (float)P.L);
  if (threadIdx.x == 0) {
    for (int s = 0; s < kPStages; ++s) {
      mbar_init(smem_u32(fullP + s), 1);
      mbar_init(smem_u32(emptyP + s), kPWarps);
    }
    for (int s = 0; s < kCStages; ++s) {
      mbar_init(smem_u32(fullC + s), 1);
      mbar_init(smem_u32(emptyC + s), kCWarps);
      mbar_init(smem_u32(voteC + s), 1);
    }
    *words_done = 0u;
    for (int s = 0; s < kSenders; ++s) sent[s] = 0u;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (P.zero_g) {
    uint4* z = reinterpret_cast<uint4*>(smem + C::OFF_ZERO);
    for (int i = threadIdx.x; i < kArrayBytes / 16; i += kFusedThreads) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to TMA
  __syncthreads();
  if (n_my == 0) return;

  if (warp == kWarpPLoad) {
    // ============================ P-loader: g, m ============================
    if (lane == 0) {
      for (int i = 0; i < n_my; ++i) {
        const int slot = i % kPStages, k = i / kPStages;
        if (k > 0) mbar_wait(smem_u32(emptyP + slot), (uint32_t)((k - 1) & 1), P.status, 0);
        const int64_t off = ((int64_t)cta + (int64_t)i * G) * C::TILE;
        const uint32_t bar = smem_u32(fullP + slot);
        const uint32_t dst = smem_u32(pring + slot * kPStageBytes);
        mbar_arrive_expect_tx(bar, kPStageBytes);
        bulk_load(dst, g + off, kArrayBytes, bar);
        bulk_load(dst + kArrayBytes, m + off, kArrayBytes, bar);
      }
    }
  } else if (warp == kWarpCLoad) {
    // ============================ C-loader: p ============================
    if (lane == 0) {
      for (int i = 0; i < n_my; ++i) {
        const int slot = i % kCStages, k = i / kCStages;
        if (k > 0) mbar_wait(smem_u32(emptyC + slot), (uint32_t)((k - 1) & 1), P.status, 1);
        const int64_t off = ((int64_t)cta + (int64_t)i * G) * C::TILE;
        const uint32_t bar = smem_u32(fullC + slot);
        mbar_arrive_expect_tx(bar, kCStageBytes);
        bulk_load(smem_u32(cring + slot * kCStageBytes), p + off, kArrayBytes, bar);
      }
    }
  } else if (warp >= kWarpVoter0 && warp < kWarpP0) {
    // ============================ voters ============================
    const int v = warp - kWarpVoter0;
    const uint64_t* my_flags = P.flags[my_rank];
    const uint32_t* my_arena = P.arena[my_rank] + (int64_t)parity_sel * W * P.arena_words;
    const int64_t spin_limit = P.spin_limit > 0 ? P.spin_limit : kDefaultFlagSpin;
    const uint64_t deadline = P.wait_timeout_ns ? global_timer_ns() + P.wait_timeout_ns : 0ull;
    // sticky: once any launch has recorded a stalled exchange, later launches neither wait nor apply votes until
    // the host has seen the status and cleared it — the host learns of a stall one step late, and must not find a
    // queue of further launches each spinning through its own time-out
    bool gave_up = P.status != nullptr && *reinterpret_cast<volatile const int32_t*>(P.status) != 0;
    uint64_t seen = 0;                       // last progress value this lane observed for rank `lane`
    const int thr = P.L / 2 + 1;             // +1 iff 2*cnt > L  <=>  cnt >= floor(L/2)+1
    const bool waits = lane < W && ((P.wait_mask >> lane) & 1u);
    const uint64_t* f = my_flags + (int64_t)(lane < W ? lane : 0) * DLION_FUSED_MAX_CTAS + cta;
    using WordVec = typename VecOf<C::WPL>::type;
    for (int i = v; i < n_my; i += kVoters) {
      const int slot = i % kCStages, k = i / kCStages;
      if (k > 0) mbar_wait(smem_u32(emptyC + slot), (uint32_t)((k - 1) & 1), P.status, 2);
      const uint64_t target = epoch_hi | (uint64_t)(i + 1);
      if (waits && !gave_up && seen < target) {
        // poll with relaxed loads (no L1 invalidate per poll), back off, and pay ONE acquire once the
        // counter is there: everything the sender stored before its release is then visible
        int64_t it = 0;
        while (ld_relaxed_sys(f) < target) {
          if (++it > spin_limit || (deadline && (it & 63) == 0 && global_timer_ns() > deadline)) {
            report(P.status, 1, cta, lane);
            gave_up = true;
            break;
          }
          __nanosleep(it < 4 ? 32 : 128);
        }
        if (!gave_up) seen = ld_acquire_sys(f);
      }
      gave_up = __any_sync(0xffffffffu, gave_up);   // also orders every lane after the acquires
      const int64_t woff = ((int64_t)cta + (int64_t)i * G) * C::WORDS + lane * C::WPL;
      uint32_t c[C::WPL][5];
#pragma unroll
      for (int w = 0; w < C::WPL; ++w)
#pragma unroll
        for (int k2 = 0; k2 < 5; ++k2) c[w][k2] = 0u;
      for (int r0 = 0; r0 < W; r0 += 8) {
        WordVec x[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {               // all loads in flight before the first use
          const int r = r0 + j;
          const bool on = r < W && ((P.live_mask >> r) & 1u);
          x[j] = on ? __ldcv(reinterpret_cast<const WordVec*>(my_arena + (int64_t)r * P.arena_words + woff))
                    : WordVec{};
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t* xw = reinterpret_cast<const uint32_t*>(&x[j]);
#pragma unroll
          for (int w = 0; w < C::WPL; ++w) {        // carry-save add of one ballot bit per element
            uint32_t carry = xw[w];
#pragma unroll
            for (int k2 = 0; k2 < 5; ++k2) {
              const uint32_t t = c[w][k2] & carry;
              c[w][k2] ^= carry;
              carry = t;
            }
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
      // a tile whose ballots never (all) arrived is not applied: better an untouched parameter tile, repaired
      // by resync_replicas(), than an update voted from stale words
      if (lane == 0) skip[slot] = gave_up ? 1u : 0u;
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(voteC + slot));
    }
  } else if (warp >= kWarpSend0 && warp < kWarpVoter0) {
    // ============================ senders ============================
    const int s = warp - kWarpSend0;
    uint32_t done_sent = 0;
    int64_t idle = 0;
    while (done_sent < (uint32_t)n_my) {
      uint32_t done = 0;
      if (lane == 0) done = ld_acquire_cta(words_done);
      done = __shfl_sync(0xffffffffu, done, 0);      // the shuffle also orders the ring reads after the acquire
      if (done == done_sent) {
        if (++idle > kMbarSpinLimit) {
          report(P.status, 2, (int)blockIdx.x, 3);
          __trap();
        }
        __nanosleep(64);
        continue;
      }
      idle = 0;
      for (uint32_t t = done_sent; t < done; ++t) {
        const uint4* src = reinterpret_cast<const uint4*>(wring + (t % kWordRing) * C::WORDS);
        const int64_t woff = ((int64_t)cta + (int64_t)t * G) * C::WORDS;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (lane < C::WORDS / 4) v = src[lane];
        if (P.mc_arena != nullptr) {                 // NVLS: sender 0 stores once, the switch replicates
          if (s == 0 && lane < C::WORDS / 4)
            multimem_st_v4(P.mc_arena + ((int64_t)(parity_sel * W + my_rank)) * P.arena_words + woff + lane * 4, v);
        } else {
          for (int r = s; r < W; r += kSenders) {
            if (!((P.wait_mask >> r) & 1u)) continue;   // a rank nobody waits for (declared dead) is not written to
            uint32_t* dst = P.arena[r] + ((int64_t)(parity_sel * W + my_rank)) * P.arena_words + woff;
            if (lane < C::WORDS / 4) reinterpret_cast<uint4*>(dst)[lane] = v;   // one coalesced request per tile
          }
        }
      }
      __syncwarp();
      if (lane == 0) {
        fence_acq_rel_sys();                         // one fence per batch: everything above is visible first
        // progress counters stay unicast (one 8-byte store per destination); with multicast data
        // sender 0 issued every data store, so it also publishes to every destination
        const int r_begin = P.mc_arena != nullptr ? (s == 0 ? 0 : W) : s;
        const int r_step = P.mc_arena != nullptr ? 1 : kSenders;
        for (int r = r_begin; r < W; r += r_step)
          if ((P.wait_mask >> r) & 1u)
            st_relaxed_sys(P.flags[r] + (int64_t)my_rank * DLION_FUSED_MAX_CTAS + cta, epoch_hi | (uint64_t)done);
        st_release_cta(sent + s, done);              // ring slots below `done` may be reused
      }
      __syncwarp();
      done_sent = done;
    }
  } else if (warp < kWarpC0) {
    // ============================ P-compute ============================
    const int pt = threadIdx.x - kWarpP0 * 32;
    // gradient clipping folded in: the coefficient torch.nn.utils.clip_grad_norm_ would apply
    float coef = 1.f;
    bool clip = false;
    if (P.clip_sumsq != nullptr) {
      const float c = P.clip_max_norm / (sqrtf(__ldg(P.clip_sumsq)) + 1e-6f);
      if (c < 1.0f) {          // also false for a NaN norm: leave the gradients alone
        coef = c;
        clip = true;
      }
    }
    const uint32_t zero_page = smem_u32(smem + C::OFF_ZERO);
    for (int i = 0; i < n_my; ++i) {
      const int slot = i % kPStages, k = i / kPStages;
      if (i >= kWordRing) {                           // ring capacity: tile i-kWordRing must have left
        int64_t it = 0;
        for (;;) {
          uint32_t lo = ld_acquire_cta(sent);
#pragma unroll
          for (int s2 = 1; s2 < kSenders; ++s2) {
            const uint32_t v = ld_acquire_cta(sent + s2);
            lo = v < lo ? v : lo;
          }
          if (lo + kWordRing > (uint32_t)i) break;
          if (++it > kMbarSpinLimit) {
            report(P.status, 2, (int)blockIdx.x, 4);
            __trap();
          }
          __nanosleep(64);
        }
      }
      mbar_wait(smem_u32(fullP + slot), (uint32_t)(k & 1), P.status, 5);
      const T* sg = reinterpret_cast<const T*>(pring + slot * kPStageBytes);
      const T* sm = reinterpret_cast<const T*>(pring + slot * kPStageBytes + kArrayBytes);
      const int64_t tile = (int64_t)cta + (int64_t)i * G;
      uint32_t* wb = wring + (i % kWordRing) * C::WORDS;
      if (P.zero_g && pt == 0) {                     // the tile's g is in shared memory: clear it in HBM
        bulk_store(const_cast<T*>(g) + tile * C::TILE, zero_page, kArrayBytes);
        bulk_commit();
      }
#pragma unroll
      for (int u = 0; u < C::UPT_P; ++u) {
        const int unit = pt + u * kPThreads;
        float gf[8], mf[8];
        load8<T>(sg + unit * 8, gf);
        load8<T>(sm + unit * 8, mf);
        if (clip) {
#pragma unroll
          for (int e = 0; e < 8; ++e) gf[e] *= coef;
          round8<T>(gf);                             // a separate clip kernel would round through T here
        }
        const uint32_t byte = sign_byte<T>(mf, gf, P.beta1, P.omb1, P.stoch_r, P.seed, P.step, (uint32_t)my_rank,
                                           P.elem_base + tile * C::TILE + unit * 8);
        uint32_t wv = byte << (8 * (lane & 3));
        wv |= __shfl_xor_sync(0xffffffffu, wv, 1);
        wv |= __shfl_xor_sync(0xffffffffu, wv, 2);
        if ((lane & 3) == 0) wb[unit >> 2] = wv;
        momentum8<T>(mf, gf, P.beta2, P.omb2);
        store8<T>(m + tile * C::TILE + unit * 8, mf);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(emptyP + slot));
      named_bar_sync(1, kPThreads);                  // every word of tile i is in the ring
      if (pt == 0) st_release_cta(words_done, (uint32_t)(i + 1));
    }
    if (P.zero_g && pt == 0) bulk_wait_all();        // the zero page must outlive every copy that reads it
  } else {
    // ============================ C-compute ============================
    const int ct = threadIdx.x - kWarpC0 * 32;
    for (int i = 0; i < n_my; ++i) {
      const int slot = i % kCStages, k = i / kCStages;
      mbar_wait(smem_u32(fullC + slot), (uint32_t)(k & 1), P.status, 6);
      mbar_wait(smem_u32(voteC + slot), (uint32_t)(k & 1), P.status, 7);
      const T* sp = reinterpret_cast<const T*>(cring + slot * kCStageBytes);
      const uint8_t* pl = reinterpret_cast<const uint8_t*>(planes + slot * 5 * C::WORDS);
      const int64_t off = ((int64_t)cta + (int64_t)i * G) * C::TILE;
      const bool void_vote = skip[slot] != 0u;       // ordered by the voteC mbarrier (uniform over the CTA)
#pragma unroll
      for (int u = 0; u < C::UPT_C && !void_vote; ++u) {
        const int unit = ct + u * kCThreads;
        float pf[8];
        load8<T>(sp + unit * 8, pf);
        float uf[8];
        if (P.vote == DLION_VOTE_AVERAGE) {
          uint32_t b[5];
#pragma unroll
          for (int k2 = 0; k2 < 5; ++k2) b[k2] = pl[k2 * C::PLANE_BYTES + unit];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int cnt = ((b[0] >> e) & 1) | (((b[1] >> e) & 1) << 1) | (((b[2] >> e) & 1) << 2) |
                            (((b[3] >> e) & 1) << 3) | (((b[4] >> e) & 1) << 4);
            uf[e] = avg_lut[cnt];
          }
        } else {
          const uint32_t a = pl[unit], z = pl[C::PLANE_BYTES + unit];
#pragma unroll
          for (int e = 0; e < 8; ++e) uf[e] = ((z >> e) & 1u) ? 0.f : (((a >> e) & 1u) ? 1.f : -1.f);
        }
        apply_store8<T>(p + off + unit * 8, pf, uf, P.lr, P.decay, P.sr_key, (P.elem_base + off + unit * 8) >> 3);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(emptyC + slot));
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
  // multicast reaches every member of the group: only usable while every rank is awaited (none declared dead)
  P.mc_arena = (a->emulate_ranks > 0 || P.wait_mask != all) ? nullptr : a->mc_arena;
  P.arena_words = a->arena_words; P.grid = a->grid; P.emulate_ranks = a->emulate_ranks; P.emu_stride = a->emu_stride;
  P.status = a->status; P.spin_limit = a->spin_limit; P.sr_key = a->sr_key;
  P.clip_sumsq = a->clip_sumsq; P.clip_max_norm = a->clip_max_norm; P.zero_g = a->zero_g;
  P.wait_timeout_ns = a->wait_timeout_us * 1000ull;
  DLION_CHECK_ARG(a->clip_sumsq == nullptr || a->clip_max_norm > 0.f, "dlion_fused_step: clip_max_norm must be positive");
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
