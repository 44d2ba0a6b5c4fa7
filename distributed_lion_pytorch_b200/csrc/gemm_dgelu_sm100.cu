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
// One persistent CTA per SM, ten warps:
//   warp 0 (one lane)   TMA producer of the operand ring: [128 x 64] of dy and [256 x 64] of w per k-step, 3 stages
//   warp 1 (one lane)   `tcgen05.mma.cta_group::1.kind::f16` (M 128, N 256, K 16) into one of two TMEM accumulators
//   warps 2-9           epilogue, thread = (row, 32-column group).  The [128 x 256] tile is walked in four boxes of
//                       64 columns; a box of `pre` arrives by TMA in a 4-deep staging ring (128-byte swizzle, 16 KB),
//                       the thread multiplies its 32 accumulator values (`tcgen05.ld`) by gelu'(pre), writes bf16
//                       back IN PLACE, and one elected thread sends the box to HBM with a TMA store and re-arms the
//                       buffer freed by the store before (prefetch distance: three boxes).  Column sums are reduced
//                       inside the warp by recursive halving (31 shuffles per 32 x 32 patch), the four row quarters
//                       are folded through 8 KB of shared memory at the end of the tile: deterministic, no atomics.
// All tensor maps cover buffers of THIS GPU only; no load uses cluster multicast.  Waits are bounded (tc05.cuh).
// reference: HF GPT2MLP backward under autograd (Conv1D -> NewGELUActivation -> Conv1D), reached from
// accelerator.backward (async_trainer.py:33): a GEMM, ~8 elementwise kernels and a column reduction upstream.
#include "tc05.cuh"

namespace dlion {
namespace gd {
using namespace tc;

constexpr int BM = 128, BN = 256, BK = 64, UK = 16;
constexpr int kStages = 3;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kBoxCols = 64, kBoxes = BN / kBoxCols, kBoxBytes = BM * kBoxCols * 2;   // 16 KB
constexpr int kStaging = 4;
constexpr int kEpiWarps = 8, kEpiThreads = kEpiWarps * 32, kThreads = 64 + kEpiThreads;
constexpr int kColPartFloats = 2 * 4 * BN;                                          // [tile parity][row quarter][column]
constexpr int kTmemCols = 512;
constexpr int kSmemBytes = kStages * kStageBytes + kStaging * kBoxBytes + kColPartFloats * 4 + 256 + 1024;

struct Params {
  float* part;        // [m_tiles, n] column sums per row tile
  int* status;
  int m, n, k;
  int m_tiles, n_tiles;
  uint32_t idesc;
};

// gelu_tanh'(x) = s (1 + x (1 - s) (2k + 6kc x^2)),  s = (1 + tanh(k (x + c x^3))) / 2
__device__ __forceinline__ float dgelu(float x) {
  constexpr float kK = 0.7978845608028654f, kC = 0.044715f;
  const float x2 = x * x;
  const float u = x * fmaf(kK * kC, x2, kK);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  const float s = fmaf(0.5f, t, 0.5f);
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

template <typename T>
__global__ void __launch_bounds__(kThreads, 1)
dgrad_dgelu_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_aux, const __grid_constant__ CUtensorMap map_out, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + kStages * kStageBytes;                               // 4 x [128 x 64] boxes, swizzled
  float* colpart = reinterpret_cast<float*>(staging + kStaging * kBoxBytes);
  uint64_t* full = reinterpret_cast<uint64_t*>(colpart + kColPartFloats);
  uint64_t* empty = full + kStages;
  uint64_t* acc_full = empty + kStages;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* aux_full = acc_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux_full + kStaging);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_blocks = (p.k + BK - 1) / BK;
  const int total = p.m_tiles * p.n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_aux);
    tma_prefetch_desc(&map_out);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(acc_full + a, 1);
      mbar_init(acc_empty + a, kEpiWarps);
    }
    for (int b = 0; b < kStaging; ++b) mbar_init(aux_full + b, 1);
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
          mbar_wait(empty + stage, phase ^ 1, p.status, 1);
          uint8_t* a = smem + stage * kStageBytes;
          mbar_expect_tx(full + stage, kStageBytes);
          tma_load_2d(a, &map_a, kb * BK, mt * BM, full + stage);
          tma_load_2d(a + kABytes, &map_b, kb * BK, nt * BN, full + stage);
          if (++stage == kStages) {
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
        mbar_wait(acc_empty + acc, acc_phase ^ 1, p.status, 2);
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(full + stage, phase, p.status, 3);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
          const uint64_t da = smem_desc_sw128(a_addr), db = smem_desc_sw128(a_addr + kABytes);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k)
            umma_f16(d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), p.idesc, (uint32_t)((kb | k) != 0));
          umma_commit(empty + stage);
          if (++stage == kStages) {
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
  } else {
    // ================================ epilogue: 8 warps ================================
    const int ew = warp - 2;                       // 0..7
    const int quarter = warp & 3;                  // TMEM lanes 32 * (warp % 4) .. + 31 are this warp's
    const int g = ew >> 2;                         // which 32 of a box's 64 columns
    const int row_in_tile = quarter * 32 + lane;
    const int et = threadIdx.x - 64;               // 0..255
    const bool elected = et == 0;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    // box jj of this CTA's sequence: `pre` box -> staging buffer jj % 4
    auto issue_aux = [&](int jj) {
      const int t = blockIdx.x + (jj / kBoxes) * gridDim.x;
      if (t >= total) return;
      const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles, b = jj % kBoxes, buf = jj % kStaging;
      mbar_expect_tx(aux_full + buf, kBoxBytes);
      tma_load_2d(staging + buf * kBoxBytes, &map_aux, nt * BN + b * kBoxCols, mt * BM, aux_full + buf);
    };
    if (elected) {
      issue_aux(0);
      issue_aux(1);
      issue_aux(2);
    }
    int acc = 0, j = 0, parity = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x, parity ^= 1) {
      const int mt = t / p.n_tiles, nt = t - mt * p.n_tiles;
      float* cp = colpart + parity * (4 * BN);
      mbar_wait(acc_full + acc, acc_phase, p.status, 4);
      tc_fence_after();
#pragma unroll 1
      for (int b = 0; b < kBoxes; ++b, ++j) {
        const int buf = j % kStaging;
        mbar_wait(aux_full + buf, (uint32_t)((j / kStaging) & 1), p.status, 5);
        float v[32];
        tmem_ld32(lane_base + (uint32_t)(acc * BN + b * kBoxCols + g * 32), v);
        if (b == kBoxes - 1) {                     // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_empty + acc);
        }
        uint8_t* rowp = staging + buf * kBoxBytes + row_in_tile * 128;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4* ptr = reinterpret_cast<uint4*>(rowp + (((g * 4 + c) ^ (row_in_tile & 7)) << 4));
          const uint4 x = *ptr;
          const uint32_t xin[4] = {x.x, x.y, x.z, x.w};
          uint32_t xo[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = unpack2<T>(xin[e]);
            float d0 = v[c * 8 + 2 * e] * dgelu(f.x), d1 = v[c * 8 + 2 * e + 1] * dgelu(f.y);
            xo[e] = pack2<T>(d0, d1);
            v[c * 8 + 2 * e] = d0;
            v[c * 8 + 2 * e + 1] = d1;
          }
          *ptr = make_uint4(xo[0], xo[1], xo[2], xo[3]);
        }
        cp[quarter * BN + b * kBoxCols + g * 32 + lane] = colsum32(v, lane);
        fence_async_smem();
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
        if (elected) {
          tma_store_2d(&map_out, staging + buf * kBoxBytes, nt * BN + b * kBoxCols, mt * BM);
          tma_store_commit();
          tma_store_wait_read<1>();                // the store before this one has drained its buffer ...
          issue_aux(j + 3);                        // ... which is the one box j + 3 lands in
        }
      }
      // all four row quarters of this tile's column sums are in shared memory (barrier of the last box)
      const int col = nt * BN + et;
      if (col < p.n) p.part[(int64_t)mt * p.n + col] = (cp[et] + cp[BN + et]) + (cp[2 * BN + et] + cp[3 * BN + et]);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (elected) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// bias gradient = sum over the row tiles of the per-tile column sums (fixed order: deterministic)
template <typename T>
__global__ void __launch_bounds__(256) dgelu_bias_finalize_kernel(const float* __restrict__ part, T* __restrict__ out, int m_tiles,
                                                                  int n, int accumulate) {
  __shared__ float red[4][64];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int col = blockIdx.x * 64 + tx;
  float s = 0.f;
  if (col < n)
    for (int mt = ty; mt < m_tiles; mt += 4) s += part[(int64_t)mt * n + col];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && col < n) {
    float r = (red[0][tx] + red[1][tx]) + (red[2][tx] + red[3][tx]);
    if (accumulate) r += DT<T>::to_f(out[col]);
    out[col] = DT<T>::from_f(r);
  }
}

}  // namespace gd
}  // namespace dlion

using namespace dlion;

extern "C" int dlion_dgrad_dgelu_parts(int64_t m) { return (int)((m + gd::BM - 1) / gd::BM); }

template <typename T>
static int launch_dgrad_dgelu(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mx, const CUtensorMap& mo,
                              const gd::Params& p, cudaStream_t s) {
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(gd::dgrad_dgelu_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, gd::kSmemBytes);
  });
  if (attr_err != cudaSuccess) return check_cuda(attr_err, "dlion_dgrad_dgelu: shared-memory attribute");
  const int total = p.m_tiles * p.n_tiles;
  const int grid = total < kNumSMs ? total : kNumSMs;
  gd::dgrad_dgelu_kernel<T><<<grid, gd::kThreads, gd::kSmemBytes, s>>>(ma, mb, mx, mo, p);
  return DLION_LAUNCH_CHECK("dlion_dgrad_dgelu");
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
  p.m = m;
  p.n = n;
  p.k = k;
  p.m_tiles = (m + gd::BM - 1) / gd::BM;
  p.n_tiles = (n + gd::BN - 1) / gd::BN;
  p.idesc = tc::instr_desc(dtype == DLION_BF16 ? 1u : 0u, gd::BM, gd::BN);
  CUtensorMap ma, mb, mx, mo;
  if (!tc::make_map(&ma, dy, k, m, ld_dy, gd::BM, dtype) || !tc::make_map(&mb, w, k, n, ld_w, gd::BN, dtype) ||
      !tc::make_map(&mx, pre, n, m, ld_pre, gd::BM, dtype) || !tc::make_map(&mo, dpre, n, m, ld_dpre, gd::BM, dtype)) {
    set_error("dlion_dgrad_dgelu: cuTensorMapEncodeTiled unavailable or refused the layout");
    return DLION_E_UNSUPPORTED;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == DLION_BF16) return launch_dgrad_dgelu<__nv_bfloat16>(ma, mb, mx, mo, p, s);
  return launch_dgrad_dgelu<__half>(ma, mb, mx, mo, p, s);
}

extern "C" int dlion_dgrad_dgelu_bias_finalize(const float* part, void* out, int m_tiles, int n, int accumulate, int dtype,
                                               void* stream) {
  DLION_CHECK_ARG(part != nullptr && out != nullptr && m_tiles > 0 && n > 0, "dlion_dgrad_dgelu_bias_finalize: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, gd::dgelu_bias_finalize_kernel<T><<<(n + 63) / 64, 256, 0, s>>>(part, (T*)out, m_tiles, n, accumulate));
  return DLION_LAUNCH_CHECK("dlion_dgrad_dgelu_bias_finalize");
}
