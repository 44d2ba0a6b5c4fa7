// gemm_dgelu_sm100.cu — data-gradient GEMM of GPT-2's MLP down-projection with the GELU backward and the
// up-projection's bias gradient in its epilogue, on the 5th-generation tensor cores:
//
//   dlion_dgrad_dgelu   dpre[M, N] = (dy[M, K] · w[N, K]^T) ⊙ gelu_tanh'(pre[M, N])      (bf16 / fp16 out)
//                       part[m_tile, n] = sum over the 128 rows of the tile of dpre[., n]  (fp32; folded by
//   dlion_dgrad_dgelu_bias_finalize into the bias gradient, optionally accumulating into the live gradient)
//
// cuBLASLt's DGELU_BGRAD epilogue takes 314 us at the GPT-2 shape (20480 x 3072 x 768) and GEMM + this package's
// one-pass activation-backward/column-sum kernel 79 + 67 us; here the product never leaves the SM before the
// activation backward has been applied: fp32 accumulators in TMEM -> registers -> ⊙ gelu'(pre) -> bf16 -> HBM once.
//
// One persistent CTA per SM, nineteen warps:
//   warp 0 (one lane)   TMA producer of the operand ring: [128 x 64] of dy and [256 x 64] of w per k-step, 4 stages
//                       (192 KB in flight per SM: the loads are L2 hits at ~1.6 us round trip under this load)
//   warp 1 (one lane)   `tcgen05.mma.cta_group::1.kind::f16` (M 128, N 256, K 16) into one of two TMEM accumulators
//   warps 2-17          epilogue = 4 row quarters x 4 column groups, thread = one row.  The [128 x 256] tile is walked in
//                       eight boxes of 32 columns; a box of `pre` arrives by TMA in a 4-deep staging ring (64-byte
//                       swizzle, 8 KB), the thread multiplies its 32 accumulator values (`tcgen05.ld.32x32b.x32`) by
//                       gelu'(pre), writes bf16 back IN PLACE and its warp arrives on the box's mbarrier; column sums
//                       are reduced inside the warp by recursive halving (31 shuffles per 32 x 32 patch) and stored per
//                       32-row quarter — deterministic, no atomics, no block barrier anywhere in the loop
//   warp 18 (one lane)  store warp: sends each finished box to HBM with a TMA store and re-arms the buffer freed by the
//                       store before with the `pre` box three boxes ahead
// All tensor maps cover buffers of THIS GPU only; no load uses cluster multicast.  Waits are bounded (tc05.cuh).
// reference: HF GPT2MLP backward under autograd (Conv1D -> NewGELUActivation -> Conv1D), reached from
// accelerator.backward (async_trainer.py:33): a GEMM, ~8 elementwise kernels and a column reduction upstream.
#include "tc05.cuh"

#ifndef DLION_DGELU_VARIANT_DEFAULT
#define DLION_DGELU_VARIANT_DEFAULT 0
#endif

namespace dlion {
namespace gd {
using namespace tc;

constexpr int BM = 128, BN = 256, BK = 64, UK = 16;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kBoxCols = 32, kBoxes = BN / kBoxCols, kBoxBytes = BM * kBoxCols * 2;   // [128 x 32] = 8 KB, 64-byte swizzle
constexpr int kStaging = 4;
constexpr int kEpiWarps = 16, kGroups = kEpiWarps / 4;   // 4 row quarters x kGroups column groups
constexpr int kStoreWarp = 2 + kEpiWarps, kThreads = (kStoreWarp + 1) * 32;
constexpr int kTmemCols = 512;
constexpr int smem_bytes(int stages, int nbuf) { return stages * kStageBytes + nbuf * kBoxBytes + 512 + 1024; }

struct Params {
  float* part;        // [m_tiles * 4, n] column sums per 32-row quarter of a row tile
  int* status;
  long long* prof;     // optional [grid][16] cycle counters per wait site (bench/diagnostics), or nullptr
  int m, n, k;
  int m_tiles, n_tiles;
  uint32_t idesc;
};

// gelu_tanh'(x) = s (1 + x (1 - s) (2k + 6kc x^2)),  s = (1 + tanh(u)) / 2 = sigmoid(2u),  u = k (x + c x^3)
// MODE 0: s through tanh.approx (MUFU.TANH);  MODE 1: s = rcp(1 + ex2(-2u log2 e)) (MUFU.EX2 + MUFU.RCP);
// MODE 2: no activation at all (timing floor of the pipeline, wrong numbers — bench only)
template <int MODE>
__device__ __forceinline__ float dgelu(float x) {
  constexpr float kK = 0.7978845608028654f, kC = 0.044715f;
  if constexpr (MODE == 2) return x;
  const float x2 = x * x;
  float s;
  if constexpr (MODE == 0) {
    const float u = x * fmaf(kK * kC, x2, kK);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
    s = fmaf(0.5f, t, 0.5f);
  } else {
    constexpr float kA = -2.0f * kK * kLog2e;
    const float e = ex2(x * fmaf(kA * kC, x2, kA));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(1.0f + e));
  }
  const float w = x * fmaf(6.0f * kK * kC, x2, 2.0f * kK);
  return fmaf(s, w * (1.0f - s), s);
}

template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t w);
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
}
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t w) {
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}
// two floats -> packed 16-bit pair, and the rounded values back (the column sums add what was stored)
template <typename T> __device__ __forceinline__ uint32_t pack2(float& a, float& b);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float& a, float& b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  const uint32_t w = *reinterpret_cast<const uint32_t*>(&h);
  a = __uint_as_float(w << 16);
  b = __uint_as_float(w & 0xFFFF0000u);
  return w;
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float& a, float& b) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  a = f.x;
  b = f.y;
  return *reinterpret_cast<const uint32_t*>(&h);
}

// lane = row: v[c] is (row lane, column c) of a 32 x 32 patch.  Returns the sum over the 32 rows of column `lane`.
__device__ __forceinline__ float colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool upper = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = upper ? v[i] : v[i + s];
      const float keep = upper ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// cycles spent in one wait, added to slot `slot` of this CTA's counters (only when profiling is on)
#define GD_TIMED(slot, stmt)                                  \
  do {                                                        \
    if (p.prof != nullptr) {                                  \
      const long long t0_ = clock64();                        \
      stmt;                                                   \
      prof_acc[slot] += clock64() - t0_;                      \
    } else {                                                  \
      stmt;                                                   \
    }                                                         \
  } while (0)

// STAGES operand stages of 48 KB and NBUF staging boxes of 8 KB: <4, 4> or, with a whole tile of `pre` boxes in flight, <3, 8>
template <typename T, int MODE, int STAGES, int NBUF>
__global__ void __launch_bounds__(kThreads, 1)
dgrad_dgelu_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_aux, const __grid_constant__ CUtensorMap map_out, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + STAGES * kStageBytes;                               // NBUF x [128 x 32] boxes, swizzled
  uint64_t* full = reinterpret_cast<uint64_t*>(staging + NBUF * kBoxBytes);
  uint64_t* empty = full + STAGES;
  uint64_t* acc_full = empty + STAGES;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* aux_full = acc_empty + 2;
  uint64_t* box_done = aux_full + NBUF;                                      // epilogue warps -> store warp
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(box_done + NBUF);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_blocks = (p.k + BK - 1) / BK;
  const int total = p.m_tiles * p.n_tiles;
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long prof_t0 = clock64();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_aux);
    tma_prefetch_desc(&map_out);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(acc_full + a, 1);
      mbar_init(acc_empty + a, kEpiWarps);
    }
    for (int b = 0; b < NBUF; ++b) {
      mbar_init(aux_full + b, 1);
      mbar_init(box_done + b, 4);                 // the four row-quarter warps of one column group
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ operand producer ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles;
        for (int kb = 0; kb < k_blocks; ++kb) {
          GD_TIMED(0, mbar_wait(empty + stage, phase ^ 1, p.status, 1));
          uint8_t* a = smem + stage * kStageBytes;
          mbar_expect_tx(full + stage, kStageBytes);
          tma_load_2d(a, &map_a, kb * BK, mt * BM, full + stage);
          tma_load_2d(a + kABytes, &map_b, kb * BK, nt * BN, full + stage);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        GD_TIMED(1, mbar_wait(acc_empty + acc, acc_phase ^ 1, p.status, 2));
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          GD_TIMED(2, mbar_wait(full + stage, phase, p.status, 3));
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
          const uint64_t da = smem_desc_sw128(a_addr), db = smem_desc_sw128(a_addr + kABytes);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k)
            umma_f16(d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), p.idesc, (uint32_t)((kb | k) != 0));
          umma_commit(empty + stage);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(acc_full + acc);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp == kStoreWarp) {
    // ================================ store warp: finished boxes out, `pre` boxes in ================================
    if (lane == 0) {
      // box jj of this CTA's sequence (8 per tile): `pre` box -> staging buffer jj % 4
      auto issue_aux = [&](int jj) {
        const int t = blockIdx.x + (jj / kBoxes) * gridDim.x;
        if (t >= total) return;
        const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles, b = jj % kBoxes, buf = jj % NBUF;
        mbar_expect_tx(aux_full + buf, kBoxBytes);
        tma_load_2d(staging + buf * kBoxBytes, &map_aux, nt * BN + b * kBoxCols, mt * BM, aux_full + buf);
      };
      for (int jj = 0; jj < NBUF - 1; ++jj) issue_aux(jj);
      int j = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles;
        for (int b = 0; b < kBoxes; ++b, ++j) {
          const int buf = j % NBUF;
          GD_TIMED(3, mbar_wait(box_done + buf, (uint32_t)((j / NBUF) & 1), p.status, 6));   // its four warps have written it
          tma_store_2d(&map_out, staging + buf * kBoxBytes, nt * BN + b * kBoxCols, mt * BM);
          tma_store_commit();
          GD_TIMED(4, tma_store_wait_read<1>());   // the store before this one has drained its buffer ...
          issue_aux(j + NBUF - 1);                 // ... which is the one box j + NBUF - 1 lands in
        }
      }
      tma_store_wait_all();
    }
  } else {
    // ================================ epilogue: 16 warps = 4 row quarters x 4 column groups ================================
    const int quarter = warp & 3;                  // TMEM lanes 32 * (warp % 4) .. + 31 are this warp's
    const int g = (warp - 2) >> 2;                 // column group: boxes g, g + kGroups, ... of every tile
    const int row_in_tile = quarter * 32 + lane;
    const int swz = (row_in_tile >> 1) & 3;        // 64-byte swizzle: 16-byte chunk index ^= address bits 7-8
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int acc = 0, tile_i = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x, ++tile_i) {
      const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles;
      float* prow = p.part + (int64_t)(mt * 4 + quarter) * p.n;
      GD_TIMED(5, mbar_wait(acc_full + acc, acc_phase, p.status, 4));
      tc_fence_after();
#pragma unroll 1
      for (int b = g; b < kBoxes; b += kGroups) {
        const int j = tile_i * kBoxes + b, buf = j % NBUF;
        GD_TIMED(6, mbar_wait(aux_full + buf, (uint32_t)((j / NBUF) & 1), p.status, 5));
        float v[32];
        tmem_ld32(lane_base + (uint32_t)(acc * BN + b * kBoxCols), v);
        if (b + kGroups >= kBoxes) {                     // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_empty + acc);
        }
        uint8_t* rowp = staging + buf * kBoxBytes + row_in_tile * 64;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4* ptr = reinterpret_cast<uint4*>(rowp + ((c ^ swz) << 4));
          const uint4 x = *ptr;
          const uint32_t xin[4] = {x.x, x.y, x.z, x.w};
          uint32_t xo[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = unpack2<T>(xin[e]);
            float d0 = v[c * 8 + 2 * e] * dgelu<MODE>(f.x), d1 = v[c * 8 + 2 * e + 1] * dgelu<MODE>(f.y);
            xo[e] = pack2<T>(d0, d1);
            v[c * 8 + 2 * e] = d0;
            v[c * 8 + 2 * e + 1] = d1;
          }
          *ptr = make_uint4(xo[0], xo[1], xo[2], xo[3]);
        }
        fence_async_smem();                        // this thread's writes -> visible to the TMA store
        __syncwarp();
        if (lane == 0) mbar_arrive(box_done + buf);
        const float cs = colsum32(v, lane);        // this quarter's 32 rows, column b * 32 + lane
        const int col = nt * BN + b * kBoxCols + lane;
        if (col < p.n) prow[col] = cs;
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  if (p.prof != nullptr && lane == 0 && (warp <= 2 || warp == kStoreWarp)) {
    // slots: 0 producer/empty, 1 mma/acc_empty, 2 mma/full, 3 store/box_done, 4 store/wait_read, 5 epi/acc_full, 6 epi/aux_full
    long long* out = p.prof + (int64_t)blockIdx.x * 16;
    for (int i = 0; i < 7; ++i)
      if (prof_acc[i] != 0) out[i] = prof_acc[i];
    if (warp == 2) out[8] = clock64() - prof_t0;       // this epilogue warp's whole loop
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): two CTAs of a cluster share one [256 x 256] tile.  Each loads its own 128 rows of dy
// and HALF of the w tile (128 of the 256 rows), one M 256 x N 256 x K 16 `tcgen05.mma.cta_group::2` issued by the leader
// reads both halves: 8 KB of operands from each CTA's shared memory per MMA instead of 12 KB, and 32 KB instead of 48 KB
// of TMA traffic per k-step and CTA — the shared-memory bandwidth the epilogue's staging traffic needs.  Each CTA's TMEM
// receives its 128 rows of the accumulator; the epilogue and the store warp are those of the single-CTA kernel.
constexpr int kPairStages = 6;
constexpr int kPairStageBytes = kABytes + kABytes;                                  // [128 x 64] of dy + [128 x 64] of w
constexpr int kPairSmemBytes = kPairStages * kPairStageBytes + kStaging * kBoxBytes + 512 + 1024;

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
dgrad_dgelu_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const __grid_constant__ CUtensorMap map_aux, const __grid_constant__ CUtensorMap map_out, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + kPairStages * kPairStageBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(staging + kStaging * kBoxBytes);   // used in the leader only
  uint64_t* empty = full + kPairStages;                                            // one set per CTA (multicast commit)
  uint64_t* acc_full = empty + kPairStages;                                        // one set per CTA (multicast commit)
  uint64_t* acc_empty = acc_full + 2;                                              // leader only: 2 x 16 warps arrive
  uint64_t* aux_full = acc_empty + 2;
  uint64_t* box_done = aux_full + kStaging;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(box_done + kStaging);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();           // 0 = leader
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int k_blocks = (p.k + BK - 1) / BK;
  const int pair_m_tiles = (p.m_tiles + 1) >> 1;
  const int total = pair_m_tiles * p.n_tiles;        // [256 x 256] tiles

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_aux);
    tma_prefetch_desc(&map_out);
    for (int s = 0; s < kPairStages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(acc_full + a, 1);
      mbar_init(acc_empty + a, 2 * kEpiWarps);
    }
    for (int b = 0; b < kStaging; ++b) {
      mbar_init(aux_full + b, 1);
      mbar_init(box_done + b, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {                                   // the same warp in both CTAs
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();                                // both CTAs' barriers exist before anything arrives on them remotely
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ operand producer (both CTAs) ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster_id; t < total; t += n_clusters) {
        const int pm = t / p.n_tiles, nt = t - pm * p.n_tiles;
        const int row0 = pm * 2 * BM + (int)rank * BM, wrow0 = nt * BN + (int)rank * (BN / 2);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(empty + stage, phase ^ 1, p.status, 1);
          uint8_t* a = smem + stage * kPairStageBytes;
          if (rank == 0) mbar_expect_tx(full + stage, 2 * kPairStageBytes);        // both CTAs' bytes land on this barrier
          tma_load_2d_pair(a, &map_a, kb * BK, row0, full + stage);
          tma_load_2d_pair(a + kABytes, &map_b, kb * BK, wrow0, full + stage);
          if (++stage == kPairStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (leader CTA only) ================================
    if (lane == 0 && rank == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = cluster_id; t < total; t += n_clusters) {
        mbar_wait_cluster(acc_empty + acc, acc_phase ^ 1, p.status, 2);            // both CTAs' epilogues have drained it
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait_cluster(full + stage, phase, p.status, 3);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * kPairStageBytes);
          const uint64_t da = smem_desc_sw128(a_addr), db = smem_desc_sw128(a_addr + kABytes);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k)
            umma_f16_pair(d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), p.idesc, (uint32_t)((kb | k) != 0));
          umma_commit_pair(empty + stage);           // frees the stage in both CTAs
          if (++stage == kPairStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(acc_full + acc);            // accumulator complete, both CTAs' epilogues
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp == kStoreWarp) {
    // ================================ store warp: finished boxes out, `pre` boxes in ================================
    if (lane == 0) {
      // box jj of this CTA's sequence (8 per tile): `pre` box -> staging buffer jj % 4
      auto issue_aux = [&](int jj) {
        const int t = cluster_id + (jj / kBoxes) * n_clusters;
        if (t >= total) return;
        const int pm = t / p.n_tiles, nt = t - pm * p.n_tiles, mt = pm * 2 + (int)rank, b = jj % kBoxes, buf = jj % kStaging;
        mbar_expect_tx(aux_full + buf, kBoxBytes);
        tma_load_2d(staging + buf * kBoxBytes, &map_aux, nt * BN + b * kBoxCols, mt * BM, aux_full + buf);
      };
      issue_aux(0);
      issue_aux(1);
      issue_aux(2);
      int j = 0;
      for (int t = cluster_id; t < total; t += n_clusters) {
        const int pm = t / p.n_tiles, nt = t - pm * p.n_tiles, mt = pm * 2 + (int)rank;
        for (int b = 0; b < kBoxes; ++b, ++j) {
          const int buf = j % kStaging;
          mbar_wait(box_done + buf, (uint32_t)((j / kStaging) & 1), p.status, 6);   // its four warps have written it
          tma_store_2d(&map_out, staging + buf * kBoxBytes, nt * BN + b * kBoxCols, mt * BM);
          tma_store_commit();
          tma_store_wait_read<1>();   // the store before this one has drained its buffer ...
          issue_aux(j + 3);                        // ... which is the one box j + 3 lands in
        }
      }
      tma_store_wait_all();
    }
  } else {
    // ================================ epilogue: 16 warps = 4 row quarters x 4 column groups ================================
    const int quarter = warp & 3;                  // TMEM lanes 32 * (warp % 4) .. + 31 are this warp's
    const int g = (warp - 2) >> 2;                 // column group: boxes g, g + kGroups, ... of every tile
    const int row_in_tile = quarter * 32 + lane;
    const int swz = (row_in_tile >> 1) & 3;        // 64-byte swizzle: 16-byte chunk index ^= address bits 7-8
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int acc = 0, tile_i = 0;
    uint32_t acc_phase = 0;
    for (int t = cluster_id; t < total; t += n_clusters, ++tile_i) {
      const int pm = t / p.n_tiles, nt = t - pm * p.n_tiles, mt = pm * 2 + (int)rank;
      float* prow = p.part + (int64_t)(mt * 4 + quarter) * p.n;
      mbar_wait(acc_full + acc, acc_phase, p.status, 4);
      tc_fence_after();
#pragma unroll 1
      for (int b = g; b < kBoxes; b += kGroups) {
        const int j = tile_i * kBoxes + b, buf = j % kStaging;
        mbar_wait(aux_full + buf, (uint32_t)((j / kStaging) & 1), p.status, 5);
        float v[32];
        tmem_ld32(lane_base + (uint32_t)(acc * BN + b * kBoxCols), v);
        if (b + kGroups >= kBoxes) {                     // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(acc_empty + acc, 0);   // the leader's barrier counts both CTAs
        }
        uint8_t* rowp = staging + buf * kBoxBytes + row_in_tile * 64;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4* ptr = reinterpret_cast<uint4*>(rowp + ((c ^ swz) << 4));
          const uint4 x = *ptr;
          const uint32_t xin[4] = {x.x, x.y, x.z, x.w};
          uint32_t xo[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = unpack2<T>(xin[e]);
            float d0 = v[c * 8 + 2 * e] * dgelu<0>(f.x), d1 = v[c * 8 + 2 * e + 1] * dgelu<0>(f.y);
            xo[e] = pack2<T>(d0, d1);
            v[c * 8 + 2 * e] = d0;
            v[c * 8 + 2 * e + 1] = d1;
          }
          *ptr = make_uint4(xo[0], xo[1], xo[2], xo[3]);
        }
        fence_async_smem();                        // this thread's writes -> visible to the TMA store
        __syncwarp();
        if (lane == 0) mbar_arrive(box_done + buf);
        const float cs = colsum32(v, lane);        // this quarter's 32 rows, column b * 32 + lane
        const int col = nt * BN + b * kBoxCols + lane;
        if (col < p.n) prow[col] = cs;
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  __syncwarp();                                      // lanes that had no role wait here for their warp's working lane
  tc_fence_before();
  cluster_sync_all();                                // nobody leaves while its pair may still read its shared memory / TMEM
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// bias gradient = sum over the partial rows (fixed order: deterministic).  32 columns x 32 row groups per CTA: every
// load of a thread is independent of the others, so the ~20 rows it owns are in flight together.
template <typename T>
__global__ void __launch_bounds__(1024) dgelu_bias_finalize_kernel(const float* __restrict__ part, T* __restrict__ out, int rows,
                                                                   int n, int accumulate) {
  __shared__ float red[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + tx;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (col < n) {
    int r = ty;
    for (; r + 96 < rows; r += 128) {
      s0 += __ldg(part + (int64_t)r * n + col);
      s1 += __ldg(part + (int64_t)(r + 32) * n + col);
      s2 += __ldg(part + (int64_t)(r + 64) * n + col);
      s3 += __ldg(part + (int64_t)(r + 96) * n + col);
    }
    for (; r < rows; r += 32) s0 += __ldg(part + (int64_t)r * n + col);
  }
  red[ty][tx] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (ty == 0 && col < n) {
    float r = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) r += red[i][tx];
    if (accumulate) r += DT<T>::to_f(out[col]);
    out[col] = DT<T>::from_f(r);
  }
}

}  // namespace gd
}  // namespace dlion

using namespace dlion;

static long long* g_dgelu_prof = nullptr;
/* diagnostics: device buffer of 148 x 16 int64 that the next launches fill with per-wait-site cycle counts (NULL = off) */
extern "C" int dlion_dgrad_dgelu_set_profile(long long* buf) {
  g_dgelu_prof = buf;
  return 0;
}

// which build of the kernel runs: 0 = single CTA, 4 operand stages + 4 staging boxes; 1 = CTA pair (cta_group::2);
// 2 = single CTA, 3 operand stages + 8 staging boxes (a whole tile of `pre` in flight).
// DLION_DGELU_VARIANT=staged|pair|deep, read once per process.
static int variant() {
  static const int v = [] {
    const char* e = getenv("DLION_DGELU_VARIANT");
    if (e == nullptr) return DLION_DGELU_VARIANT_DEFAULT;
    return e[0] == 'p' ? 1 : (e[0] == 'd' ? 2 : 0);
  }();
  return v;
}
static bool use_pair() { return variant() == 1; }

extern "C" int dlion_dgrad_dgelu_parts(int64_t m) {
  if (use_pair()) return (int)((m + 2 * gd::BM - 1) / (2 * gd::BM)) * 8;      // both halves of every [256 x 256] tile write
  return (int)((m + gd::BM - 1) / gd::BM) * 4;
}

template <typename T, int MODE, int STAGES, int NBUF>
static int launch_dgrad_dgelu_mode(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mx, const CUtensorMap& mo,
                                   const gd::Params& p, cudaStream_t s) {
  constexpr int kSmem = gd::smem_bytes(STAGES, NBUF);
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(gd::dgrad_dgelu_kernel<T, MODE, STAGES, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  });
  if (attr_err != cudaSuccess) return check_cuda(attr_err, "dlion_dgrad_dgelu: shared-memory attribute");
  const int total = p.m_tiles * p.n_tiles;
  const int grid = total < kNumSMs ? total : kNumSMs;
  gd::dgrad_dgelu_kernel<T, MODE, STAGES, NBUF><<<grid, gd::kThreads, kSmem, s>>>(ma, mb, mx, mo, p);
  return DLION_LAUNCH_CHECK("dlion_dgrad_dgelu");
}

template <typename T>
static int launch_dgrad_dgelu_pair(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mx, const CUtensorMap& mo,
                                   const gd::Params& p, cudaStream_t s) {
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(gd::dgrad_dgelu_pair_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, gd::kPairSmemBytes);
  });
  if (attr_err != cudaSuccess) return check_cuda(attr_err, "dlion_dgrad_dgelu: shared-memory attribute (pair)");
  const int total = ((p.m_tiles + 1) / 2) * p.n_tiles;
  const int grid = 2 * total < kNumSMs ? 2 * total : (kNumSMs & ~1);
  gd::dgrad_dgelu_pair_kernel<T><<<grid, gd::kThreads, gd::kPairSmemBytes, s>>>(ma, mb, mx, mo, p);
  return DLION_LAUNCH_CHECK("dlion_dgrad_dgelu (pair)");
}

template <typename T>
static int launch_dgrad_dgelu(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mx, const CUtensorMap& mo,
                              const gd::Params& p, cudaStream_t s, bool deep) {
  static const int mode = [] {
    const char* e = getenv("DLION_DGELU_MODE");      // A/B switch of the microbenchmark
    return e != nullptr ? atoi(e) : 0;
  }();
  if (deep) return launch_dgrad_dgelu_mode<T, 0, 3, 8>(ma, mb, mx, mo, p, s);
  if (mode == 1) return launch_dgrad_dgelu_mode<T, 1, 4, 4>(ma, mb, mx, mo, p, s);
  if (mode == 2) return launch_dgrad_dgelu_mode<T, 2, 4, 4>(ma, mb, mx, mo, p, s);
  return launch_dgrad_dgelu_mode<T, 0, 4, 4>(ma, mb, mx, mo, p, s);
}

extern "C" int dlion_dgrad_dgelu(const void* dy, int64_t ld_dy, const void* w, int64_t ld_w, const void* pre, int64_t ld_pre,
                                 void* dpre, int64_t ld_dpre, float* part, int* status, int m, int n, int k, int dtype,
                                 void* stream) {
  DLION_CHECK_ARG(dtype == DLION_BF16 || dtype == DLION_F16, "dlion_dgrad_dgelu: bf16 / fp16 only (got dtype %d)", dtype);
  DLION_CHECK_ARG(m > 0 && n > 0 && k > 0 && (k & 7) == 0 && (n & 7) == 0, "dlion_dgrad_dgelu: n (%d) and k (%d) must be multiples of 8", n, k);
  DLION_CHECK_ARG((ld_dy & 7) == 0 && (ld_w & 7) == 0 && (ld_pre & 7) == 0 && (ld_dpre & 7) == 0 && ld_dy >= k && ld_w >= k &&
                      ld_pre >= n && ld_dpre >= n,
                  "dlion_dgrad_dgelu: row strides must be multiples of 8 elements and cover the rows");
  DLION_CHECK_ARG(aligned16(dy) && aligned16(w) && aligned16(pre) && aligned16(dpre) && part != nullptr,
                  "dlion_dgrad_dgelu: bad pointers");
  gd::Params p{};
  p.part = part;
  p.status = status;
  p.prof = g_dgelu_prof;
  p.m = m;
  p.n = n;
  p.k = k;
  p.m_tiles = (m + gd::BM - 1) / gd::BM;
  p.n_tiles = (n + gd::BN - 1) / gd::BN;
  const bool pair = use_pair();
  p.idesc = tc::instr_desc(dtype == DLION_BF16 ? 1u : 0u, pair ? 2 * gd::BM : gd::BM, gd::BN);
  CUtensorMap ma, mb, mx, mo;
  if (!tc::make_map(&ma, dy, k, m, ld_dy, gd::BM, dtype) || !tc::make_map(&mb, w, k, n, ld_w, pair ? gd::BN / 2 : gd::BN, dtype) ||
      !tc::make_map(&mx, pre, n, m, ld_pre, gd::BM, dtype, gd::kBoxCols) ||
      !tc::make_map(&mo, dpre, n, m, ld_dpre, gd::BM, dtype, gd::kBoxCols)) {
    set_error("dlion_dgrad_dgelu: cuTensorMapEncodeTiled unavailable or refused the layout");
    return DLION_E_UNSUPPORTED;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (pair) {
    if (dtype == DLION_BF16) return launch_dgrad_dgelu_pair<__nv_bfloat16>(ma, mb, mx, mo, p, s);
    return launch_dgrad_dgelu_pair<__half>(ma, mb, mx, mo, p, s);
  }
  if (dtype == DLION_BF16) return launch_dgrad_dgelu<__nv_bfloat16>(ma, mb, mx, mo, p, s, variant() == 2);
  return launch_dgrad_dgelu<__half>(ma, mb, mx, mo, p, s, variant() == 2);
}

extern "C" int dlion_dgrad_dgelu_bias_finalize(const float* part, void* out, int m_tiles, int n, int accumulate, int dtype,
                                               void* stream) {
  DLION_CHECK_ARG(part != nullptr && out != nullptr && m_tiles > 0 && n > 0, "dlion_dgrad_dgelu_bias_finalize: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, gd::dgelu_bias_finalize_kernel<T><<<(n + 31) / 32, 1024, 0, s>>>(part, (T*)out, m_tiles, n, accumulate));
  return DLION_LAUNCH_CHECK("dlion_dgrad_dgelu_bias_finalize");
}
