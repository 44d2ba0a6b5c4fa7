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
#include <cuda.h>

#include <mutex>

#include "common.cuh"

namespace dlion {
namespace lm {

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
constexpr float kLog2e = 1.4426950408889634f;
constexpr uint64_t kWaitNs = 10ull * 1000 * 1000 * 1000;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
               : "=r"(ok)
               : "r"(smem_u32(bar)), "r"(parity)
               : "memory");
  return ok != 0;
}
// bounded wait: a fault in the pipeline protocol ends in a trap (the launch fails), never in a hung GPU
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int* status, int code) {
  if (mbar_try(bar, parity)) return;
  const uint64_t t0 = timer_ns();
  uint32_t n = 0;
  while (!mbar_try(bar, parity)) {
    if ((++n & 1023u) == 0 && timer_ns() - t0 > kWaitNs) {
      if (status != nullptr) atomicExch(status, code);
      __threadfence_system();
      asm volatile("trap;");
    }
  }
}
// 2-D tiled TMA load (tensor map over this GPU's memory; no multicast): box lands at `dst`, bytes on `bar`
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c_inner, int c_outer, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                   smem_u32(dst)),
               "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer)
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] · B[smem]^T, both operands K-major, issued by ONE thread for the CTA
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on `bar` once every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// shared-memory matrix descriptor, K-major operand in the canonical 128-byte-swizzle layout TMA writes:
// rows 128 B apart, groups of 8 rows 1024 B apart (stride byte offset), version 1 (sm_100), layout type 2
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: fp32 accumulate, A and B of format `fmt` (0 = f16, 1 = bf16), both K-major, M 128, N 256
__host__ __device__ constexpr uint32_t instr_desc(uint32_t fmt) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 32 consecutive TMEM columns of this thread's lane -> 32 registers
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  // the registers are named as in/out operands so that no use of them can be scheduled above the wait
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

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

// ---- host side -------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

// [outer, inner] 16-bit matrix, `ld` elements between rows, box [box_outer, 64], 128-byte swizzle, zero fill
static bool make_map(CUtensorMap* map, const void* base, int64_t inner, int64_t outer, int64_t ld, int box_outer, int dtype) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_outer};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = dtype == DLION_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  return fn(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
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
  if (!lm::make_map(&map_h, h, hidden, rows, ld_h, lm::BM, dtype) || !lm::make_map(&map_w, w, hidden, vocab, ld_w, lm::BN, dtype)) {
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
  p.idesc = lm::instr_desc(dtype == DLION_BF16 ? 1u : 0u);
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
