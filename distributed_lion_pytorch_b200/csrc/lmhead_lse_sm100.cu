// lmhead_lse_sm100.cu — vocabulary projection fused with log-sum-exp on the 5th-generation tensor cores:
// per-token log-probabilities of a [rows, hidden] activation against a [vocab, hidden] head WITHOUT the
// [rows, vocab] logits ever reaching HBM (SURVEY N5 "vocab-tiled, logits-never-materialised", no-grad side).
//
//   dlion_lmhead_logprobs   lse[r] = log sum_v exp(h[r]·w[v]),  logp[r] = h[r]·w[label[r]] - lse[r]
//
// One persistent CTA per SM, six warps, three roles (the canonical Blackwell GEMM anatomy):
//   warp 0 (one lane)  TMA producer: 2-D tensor-map loads (`cp.async.bulk.tensor.2d`, 128-byte swizzle) of a
//                      [128 x 64] tile of h and a [256 x 64] tile of w per k-step into a 4-stage ring;
//                      both tensor maps cover buffers of THIS GPU only and no load uses cluster multicast
//   warp 1 (one lane)  MMA issuer: 4 x `tcgen05.mma.cta_group::1.kind::f16` (M 128, N 256, K 16) per stage,
//                      accumulating a [128 x 256] fp32 logits tile in TMEM; `tcgen05.commit` frees the smem
//                      stage and, after the last k-step, publishes the accumulator.  TMEM holds two
//                      accumulators (2 x 256 = all 512 columns), so the MMAs of tile i+1 run under the
//                      epilogue of tile i
//   warps 2-5          epilogue: thread = one row of the tile; `tcgen05.ld.32x32b.x32` brings 32 logits at a
//                      time into registers, an online (max, sum exp) is carried across the vocabulary tiles
//                      of the work unit, the label's logit is picked up on the way.  Nothing is stored
//                      but 3 floats per row per unit.
// A work unit is (row tile, vocabulary split); CTAs of one wave share the split, so they walk the same
// slices of w at the same time and the head is read from HBM about once per wave even when it does not
// fit in L2 (Llama: 262 MB).  A second small kernel merges the splits.
//
// Every mbarrier wait is bounded (`%globaltimer`, 10 s): a protocol fault traps instead of hanging the GPU.
// reference: HF `ForCausalLM` loss inside Trainer.compute_loss (run_clm.py:588-639, evaluation at :626-639)
// and TRL DPOTrainer's per-token log-probs of the frozen reference model (dpo_llama2.py:219-231) — upstream
// both materialise fp32 [batch, seq, vocab] logits.
#include "tc05.cuh"

namespace dlion {
namespace lm {
using namespace tc;

constexpr int BM = 128;               // rows per tile (= TMEM lanes, = UMMA M)
constexpr int BN = 256;               // vocabulary entries per tile (= UMMA N, = TMEM columns per accumulator)
constexpr int BK = 64;                // k-step: 64 x 2 B = one 128-byte swizzle atom
constexpr int UK = 16;                // K of one tcgen05.mma for 16-bit inputs
constexpr int kStages = 4;
constexpr int kABytes = BM * BK * 2;  // 16 KB
constexpr int kBBytes = BN * BK * 2;  // 32 KB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 192;
constexpr int kTmemCols = 512;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /* alignment slack */ + 256 /* barriers */;

// online (max, sum exp) over 32 logits of one row; with MASK only the first `valid` of them exist
template <bool MASK>
__device__ __forceinline__ void fold_chunk(const float (&v)[32], int valid, float& run_m, float& run_s) {
  constexpr float kNeg = -3.0e38f;
  float cm = v[0];
#pragma unroll
  for (int j = 1; j < 32; ++j) cm = fmaxf(cm, (MASK && j >= valid) ? kNeg : v[j]);
  const float new_m = fmaxf(run_m, cm);
  const float shift = -new_m * kLog2e;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float e0 = ex2(fmaf(v[j], kLog2e, shift)), e1 = ex2(fmaf(v[j + 1], kLog2e, shift));
    const float e2 = ex2(fmaf(v[j + 2], kLog2e, shift)), e3 = ex2(fmaf(v[j + 3], kLog2e, shift));
    s0 += (MASK && j >= valid) ? 0.f : e0;
    s1 += (MASK && j + 1 >= valid) ? 0.f : e1;
    s2 += (MASK && j + 2 >= valid) ? 0.f : e2;
    s3 += (MASK && j + 3 >= valid) ? 0.f : e3;
  }
  run_s = fmaf(run_s, ex2((run_m - new_m) * kLog2e), (s0 + s1) + (s2 + s3));
  run_m = new_m;
}
// v[idx] for a run-time idx without putting v in local memory: a chain of selects
__device__ __forceinline__ float pick32(const float (&v)[32], int idx) {
  float r = v[0];
#pragma unroll
  for (int j = 1; j < 32; ++j) r = (idx == j) ? v[j] : r;
  return r;
}

struct Params {
  const int64_t* labels;   // [rows] or nullptr
  float* part_m;           // [splits, rows]  running maximum of the unit (natural units)
  float* part_s;           // [splits, rows]  sum of exp(x - maximum)
  float* lab_logit;        // [rows]          written by the unit whose vocabulary range holds the label
  int* status;
  int rows, vocab, hidden;
  int m_tiles, n_tiles, splits;
  uint32_t idesc;
};


// unit u -> (row tile, split); the vocabulary tiles of split s are [n_begin(s), n_begin(s + 1))
__device__ __forceinline__ int n_begin(const Params& p, int split) { return (int)((int64_t)p.n_tiles * split / p.splits); }

__global__ void __launch_bounds__(kThreads, 1)
lmhead_lse_kernel(const __grid_constant__ CUtensorMap map_h, const __grid_constant__ CUtensorMap map_w, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  // 128-byte-swizzled tiles want 1024-byte alignment (the swizzle is a function of address bits 7-9)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);   // TMA -> MMA, one per stage
  uint64_t* empty = full + kStages;                                              // MMA -> TMA
  uint64_t* acc_full = empty + kStages;                                          // MMA -> epilogue, one per accumulator
  uint64_t* acc_empty = acc_full + 2;                                            // epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_blocks = (p.hidden + BK - 1) / BK;
  const int units = p.m_tiles * p.splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_h);
    tma_prefetch_desc(&map_w);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(acc_full + a, 1);
      mbar_init(acc_empty + a, 4);          // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {                          // the whole warp allocates all 512 TMEM columns (one CTA per SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int split = u / p.m_tiles, mt = u - split * p.m_tiles;
        const int n0 = n_begin(p, split), n1 = n_begin(p, split + 1);
        for (int nt = n0; nt < n1; ++nt) {
          for (int kb = 0; kb < k_blocks; ++kb) {
            mbar_wait(empty + stage, phase ^ 1, p.status, 1);
            uint8_t* a = smem + stage * kStageBytes;
            mbar_expect_tx(full + stage, kStageBytes);
            tma_load_2d(a, &map_h, kb * BK, mt * BM, full + stage);
            tma_load_2d(a + kABytes, &map_w, kb * BK, nt * BN, full + stage);
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int split = u / p.m_tiles;
        const int n0 = n_begin(p, split), n1 = n_begin(p, split + 1);
        for (int nt = n0; nt < n1; ++nt) {
          mbar_wait(acc_empty + acc, acc_phase ^ 1, p.status, 2);      // epilogue has drained this accumulator
          tc_fence_after();
          const uint32_t d = tmem_base + (uint32_t)(acc * BN);
          for (int kb = 0; kb < k_blocks; ++kb) {
            mbar_wait(full + stage, phase, p.status, 3);               // this stage's tiles have landed
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
            const uint64_t da = smem_desc_sw128(a_addr), db = smem_desc_sw128(a_addr + kABytes);
#pragma unroll
            for (int k = 0; k < BK / UK; ++k)      // +32 bytes along K inside the swizzle atom = +2 in the address field
              umma_f16(d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), p.idesc, (uint32_t)((kb | k) != 0));
            umma_commit(empty + stage);            // stage is free once these MMAs have read it
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
          umma_commit(acc_full + acc);             // accumulator complete
          if (++acc == 2) {
            acc = 0;
            acc_phase ^= 1;
          }
        }
      }
    }
  } else {
    // ================================ epilogue: 4 warps, thread = row ================================
    const int quarter = warp & 3;                  // a warp may touch TMEM lanes 32 * (warp % 4) .. + 31 only
    const int row_in_tile = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const int split = u / p.m_tiles, mt = u - split * p.m_tiles;
      const int n0 = n_begin(p, split), n1 = n_begin(p, split + 1);
      const int row = mt * BM + row_in_tile;
      const bool live = row < p.rows;
      const int64_t label = (live && p.labels != nullptr) ? p.labels[row] : -1;
      float run_m = -3.0e38f, run_s = 0.f;
      for (int nt = n0; nt < n1; ++nt) {
        mbar_wait(acc_full + acc, acc_phase, p.status, 4);
        tc_fence_after();
        const int col0 = nt * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          const int cbase = col0 + c * 32;
          if (cbase >= p.vocab) break;             // columns past the vocabulary (TMA zero-filled them)
          float v[32];
          tmem_ld32(lane_base + (uint32_t)(acc * BN + c * 32), v);
          const int64_t rel = label - cbase;
          if (rel >= 0 && rel < 32) p.lab_logit[row] = pick32(v, (int)rel);
          if (cbase + 32 > p.vocab)
            fold_chunk<true>(v, p.vocab - cbase, run_m, run_s);
          else
            fold_chunk<false>(v, 32, run_m, run_s);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty + acc);     // this warp's quarter of the accumulator is drained
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      if (live) {
        p.part_m[(int64_t)split * p.rows + row] = run_m;
        p.part_s[(int64_t)split * p.rows + row] = run_s;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// merge the vocabulary splits of each row: lse = M + ln sum_s part_s * exp(part_m - M); logp = label logit - lse
__global__ void __launch_bounds__(256) lmhead_merge_kernel(const float* __restrict__ part_m, const float* __restrict__ part_s,
                                                           const float* __restrict__ lab_logit,
                                                           const int64_t* __restrict__ labels, float* __restrict__ logp,
                                                           float* __restrict__ lse, int rows, int vocab, int splits) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= rows) return;
  float mx = -3.0e38f;
  for (int s = 0; s < splits; ++s) mx = fmaxf(mx, part_m[(int64_t)s * rows + row]);
  float sum = 0.f;
  for (int s = 0; s < splits; ++s)
    sum += part_s[(int64_t)s * rows + row] * ex2((part_m[(int64_t)s * rows + row] - mx) * kLog2e);
  const float l = mx + __logf(sum);
  if (lse != nullptr) lse[row] = l;
  if (logp != nullptr) {
    const int64_t lab = labels != nullptr ? labels[row] : -1;
    logp[row] = (lab >= 0 && lab < vocab) ? lab_logit[row] - l : 0.f;
  }
}

// number of vocabulary splits: minimise (waves of units over the SMs) x (tiles per unit)
static int plan_splits(int m_tiles, int n_tiles, int ctas) {
  int best = 1;
  int64_t best_cost = INT64_MAX;
  const int max_s = n_tiles < 64 ? n_tiles : 64;
  for (int s = 1; s <= max_s; ++s) {
    const int64_t units = (int64_t)m_tiles * s;
    const int64_t waves = (units + ctas - 1) / ctas;
    const int64_t cost = waves * ((n_tiles + s - 1) / s) * 64 + s;       // "+ s": fewer splits on a tie
    if (cost < best_cost) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
}

}  // namespace lm
}  // namespace dlion

using namespace dlion;

extern "C" int dlion_lmhead_plan(int rows, int vocab, int* splits, int64_t* workspace_floats) {
  DLION_CHECK_ARG(rows > 0 && vocab > 0 && splits != nullptr && workspace_floats != nullptr, "dlion_lmhead_plan: bad arguments");
  const int m_tiles = (rows + lm::BM - 1) / lm::BM, n_tiles = (vocab + lm::BN - 1) / lm::BN;
  *splits = lm::plan_splits(m_tiles, n_tiles, kNumSMs);
  *workspace_floats = (int64_t)(2 * *splits + 1) * rows;
  return 0;
}

extern "C" int dlion_lmhead_logprobs(const void* h, int64_t ld_h, const void* w, int64_t ld_w, const int64_t* labels,
                                     float* logp, float* lse, float* workspace, int64_t workspace_floats, int* status,
                                     int rows, int vocab, int hidden, int dtype, void* stream) {
  DLION_CHECK_ARG(dtype == DLION_BF16 || dtype == DLION_F16, "dlion_lmhead_logprobs: bf16 / fp16 inputs only (got dtype %d)", dtype);
  DLION_CHECK_ARG(rows > 0 && vocab > 0 && hidden > 0 && (hidden & 7) == 0 && (ld_h & 7) == 0 && (ld_w & 7) == 0 &&
                      ld_h >= hidden && ld_w >= hidden,
                  "dlion_lmhead_logprobs: hidden (%d) and the row strides must be multiples of 8 elements", hidden);
  DLION_CHECK_ARG(aligned16(h) && aligned16(w) && workspace != nullptr && (logp != nullptr || lse != nullptr),
                  "dlion_lmhead_logprobs: bad pointers");
  DLION_CHECK_ARG(logp == nullptr || labels != nullptr, "dlion_lmhead_logprobs: logp needs labels");
  lm::Params p{};
  p.m_tiles = (rows + lm::BM - 1) / lm::BM;
  p.n_tiles = (vocab + lm::BN - 1) / lm::BN;
  p.splits = lm::plan_splits(p.m_tiles, p.n_tiles, kNumSMs);
  DLION_CHECK_ARG(workspace_floats >= (int64_t)(2 * p.splits + 1) * rows, "dlion_lmhead_logprobs: workspace too small (see dlion_lmhead_plan)");
  CUtensorMap map_h, map_w;
  if (!tc::make_map(&map_h, h, hidden, rows, ld_h, lm::BM, dtype) || !tc::make_map(&map_w, w, hidden, vocab, ld_w, lm::BN, dtype)) {
    set_error("dlion_lmhead_logprobs: cuTensorMapEncodeTiled unavailable or refused the layout");
    return DLION_E_UNSUPPORTED;
  }
  p.labels = labels;
  p.part_m = workspace;
  p.part_s = workspace + (int64_t)p.splits * rows;
  p.lab_logit = workspace + (int64_t)2 * p.splits * rows;
  p.status = status;
  p.rows = rows;
  p.vocab = vocab;
  p.hidden = hidden;
  p.idesc = tc::instr_desc(dtype == DLION_BF16 ? 1u : 0u, lm::BM, lm::BN);
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    attr_err = cudaFuncSetAttribute(lm::lmhead_lse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, lm::kSmemBytes);
  });
  if (attr_err != cudaSuccess) return check_cuda(attr_err, "dlion_lmhead_logprobs: shared-memory attribute");
  const int units = p.m_tiles * p.splits;
  const int grid = units < kNumSMs ? units : kNumSMs;
  cudaStream_t s = (cudaStream_t)stream;
  lm::lmhead_lse_kernel<<<grid, lm::kThreads, lm::kSmemBytes, s>>>(map_h, map_w, p);
  int rc = DLION_LAUNCH_CHECK("dlion_lmhead_logprobs");
  if (rc != 0) return rc;
  lm::lmhead_merge_kernel<<<(rows + 255) / 256, 256, 0, s>>>(p.part_m, p.part_s, p.lab_logit, labels, logp, lse, rows, vocab, p.splits);
  return DLION_LAUNCH_CHECK("dlion_lmhead_logprobs (merge)");
}
