// norm.cu — N3: fused residual-add + LayerNorm / RMSNorm, forward and backward.
//
// One CTA owns one row at a time and keeps it in registers (VPT 16-byte vectors per thread),
// so the forward reads x (+ residual) once and writes y (+ the new residual stream) once, and
// the backward reads dy, the saved residual stream and the incoming residual gradient once
// and writes dx once.  dweight / dbias are accumulated per CTA in registers across the rows
// it visits and written as fp32 partial rows; a second tiny kernel folds the partials.
//
// reference: the HF GPT2Block (ln_1 / ln_2 around residual adds) and LlamaDecoderLayer
// (input_layernorm / post_attention_layernorm) that the reference reaches through
// Trainer.compute_loss (async_trainer.py:24,52,80) as separate add + layer_norm ATen kernels.
#include <stdlib.h>

#include "common.cuh"

namespace dlion {

template <int BLOCK>
__device__ __forceinline__ float block_allreduce_sum(float v, float* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if constexpr (BLOCK > 32) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();   // protect smem reuse across consecutive reductions
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    v = lane < (BLOCK >> 5) ? smem[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  }
  return v;
}

// keep_mask8 (the dropout keep-mask of one 8-element vector) lives in common.cuh: norm_tma.cu shares it

// Thread t owns vectors t, t+BLOCK, ... (VPT of them); vector j covers columns [8j, 8j+8).
template <typename T, int VPT, int BLOCK, bool RMS, bool DROP>
__global__ void __launch_bounds__(BLOCK) norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                         const T* __restrict__ w, const T* __restrict__ b,
                                                         T* __restrict__ y, T* __restrict__ sum_out,
                                                         float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                         int64_t rows, int cols, float eps, float inv_keep,
                                                         uint32_t thr16, uint64_t seed) {
  __shared__ float red[32];
  const int nvec = cols >> 3;
  float wv[VPT][8], bv[VPT][8];
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int j = threadIdx.x + k * BLOCK;
    if (j < nvec) {
      load8<T>(w + j * 8, wv[k]);
      if (!RMS && b != nullptr) load8<T>(b + j * 8, bv[k]);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) bv[k][i] = 0.f;
      }
    }
  }
  const float inv_cols = 1.0f / (float)cols;
  // Software pipeline over the rows this CTA visits: the raw vectors of the NEXT row are requested before
  // the current row is reduced, so a thread keeps two rows of loads in flight (the kernel is bound by load
  // latency, not bandwidth: ncu showed 48 % of stalls on the long scoreboard at 30 % of DRAM peak).
  Raw8<T> xcur[VPT], rcur[VPT], xnext[VPT], rnext[VPT];
  auto request = [&](int64_t row, Raw8<T>(&xo)[VPT], Raw8<T>(&ro)[VPT]) {
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int j = threadIdx.x + k * BLOCK;
      if (j < nvec) {
        xo[k] = load_raw8<T>(x + row * cols + j * 8);
        if (res != nullptr) ro[k] = load_raw8<T>(res + row * cols + j * 8);
      }
    }
  };
  if ((int64_t)blockIdx.x < rows) request(blockIdx.x, xcur, rcur);
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    if (row + gridDim.x < rows) request(row + gridDim.x, xnext, rnext);
    float s[VPT][8];
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int j = threadIdx.x + k * BLOCK;
      if (j < nvec) {
        unpack_raw8<T>(xcur[k], s[k]);
        if (DROP) {                                                         // y = dropout(x) in T, like unfused PyTorch
          const uint32_t keep = keep_mask8(seed, row * nvec + j, thr16);
#pragma unroll
          for (int i = 0; i < 8; ++i) s[k][i] = ((keep >> i) & 1u) ? s[k][i] * inv_keep : 0.f;
          round8<T>(s[k]);
        }
        if (res != nullptr) {
          float r[8];
          unpack_raw8<T>(rcur[k], r);
#pragma unroll
          for (int i = 0; i < 8; ++i) s[k][i] += r[i];
          round8<T>(s[k]);                                                    // the residual stream lives in T
        }
        if (sum_out != nullptr) store8<T>(sum_out + row * cols + j * 8, s[k]);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc += RMS ? s[k][i] * s[k][i] : s[k][i];
      }
    }
    float mean = 0.f, rstd;
    if (RMS) {
      const float ss = block_allreduce_sum<BLOCK>(acc, red);
      rstd = rsqrtf(ss * inv_cols + eps);
    } else {
      mean = block_allreduce_sum<BLOCK>(acc, red) * inv_cols;
      float var = 0.f;
#pragma unroll
      for (int k = 0; k < VPT; ++k) {
        const int j = threadIdx.x + k * BLOCK;
        if (j < nvec) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float d = s[k][i] - mean;
            var = fmaf(d, d, var);
          }
        }
      }
      var = block_allreduce_sum<BLOCK>(var, red) * inv_cols;
      rstd = rsqrtf(var + eps);
    }
    if (threadIdx.x == 0) {
      if (mean_out != nullptr) mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int j = threadIdx.x + k * BLOCK;
      if (j < nvec) {
        float o[8];
        if (RMS) {                                                        // HF: weight * h.to(input_dtype)
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = s[k][i] * rstd;
          round8<T>(o);
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] *= wv[k][i];
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = fmaf((s[k][i] - mean) * rstd, wv[k][i], bv[k][i]);
        }
        store8<T>(y + row * cols + j * 8, o);
      }
    }
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      xcur[k] = xnext[k];
      rcur[k] = rnext[k];
    }
  }
}

// Sum NV values per thread over the block; every thread gets all NV totals.
template <int BLOCK, int NV>
__device__ __forceinline__ void block_allreduce_sum_n(float (&v)[NV], float* smem /* [NV][32] */) {
#pragma unroll
  for (int q = 0; q < NV; ++q)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
  if constexpr (BLOCK > 32) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) {
#pragma unroll
      for (int q = 0; q < NV; ++q) smem[q * 32 + warp] = v[q];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NV; ++q) {
      v[q] = lane < (BLOCK >> 5) ? smem[q * 32 + lane] : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    }
  }
}

// ROWS rows are processed per trip so that ROWS x (dy, xs, dsum) loads are in flight per thread
// before the first block reduction: the kernel is latency-bound otherwise (cols = 768 leaves
// only 96 threads per row).
// Resident CTAs per SM the backward is compiled for: the kernel is latency-bound (ncu: 53 % of
// stall samples on long scoreboard at 21 % warp occupancy when it was register-limited to 7
// CTAs/SM and launched as 1.14 waves), so registers are capped for ~30 warps/SM and the grid is
// exactly one wave.
constexpr int bwd_ctas_per_sm(int vpt, int block) {
  return vpt == 1 ? (1024 / block > 16 ? 16 : 1024 / block) : (vpt == 2 ? 3 : 2);
}
// with the dx column sums riding along (8 more accumulators per thread) the narrow-row kernel gets the registers
// of one CTA in five back, so that nothing spills
constexpr int bwd_ctas_per_sm_dxs(int vpt, int block) {
  return vpt == 1 ? (bwd_ctas_per_sm(vpt, block) * 4 + 4) / 5 : bwd_ctas_per_sm(vpt, block);
}

template <typename T, int VPT, int BLOCK, bool RMS, int ROWS, bool DROP, bool DXS>
__global__ void __launch_bounds__(BLOCK, DXS ? bwd_ctas_per_sm_dxs(VPT, BLOCK) : bwd_ctas_per_sm(VPT, BLOCK)) norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ dsum_in,
                                                         const T* __restrict__ xs, const T* __restrict__ w,
                                                         const float* __restrict__ mean_in,
                                                         const float* __restrict__ rstd_in, T* __restrict__ dx,
                                                         T* __restrict__ dres, float* __restrict__ dw_part,
                                                         float* __restrict__ db_part, float* __restrict__ dxs_part,
                                                         int64_t rows, int cols,
                                                         float inv_keep, uint32_t thr16, uint64_t seed) {
  __shared__ float red[2 * ROWS * 32];
  const int nvec = cols >> 3;
  float wv[VPT][8], dwv[VPT][8], dbv[VPT][8];
  // column sums of dx, the gradient handed to the layer that produced x: when that layer is a biased linear
  // (GPT-2's attn.c_proj / mlp.c_proj) they ARE its bias gradient, so its own pass over dx disappears
  float dxv[VPT][8];
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int j = threadIdx.x + k * BLOCK;
    if (j < nvec) load8<T>(w + j * 8, wv[k]);
#pragma unroll
    for (int i = 0; i < 8; ++i) dwv[k][i] = 0.f, dbv[k][i] = 0.f, dxv[k][i] = 0.f;
  }
  const float inv_cols = 1.0f / (float)cols;
  for (int64_t row0 = (int64_t)blockIdx.x * ROWS; row0 < rows; row0 += (int64_t)gridDim.x * ROWS) {
    float g[ROWS][VPT][8], xh[ROWS][VPT][8], rs[ROWS][VPT][8];
    float sums[2 * ROWS];
    float rstd[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int64_t row = row0 + r;
      const bool live = row < rows;
      const float mean = (RMS || !live) ? 0.f : mean_in[row];
      rstd[r] = live ? rstd_in[row] : 0.f;
      sums[2 * r] = 0.f;
      sums[2 * r + 1] = 0.f;
#pragma unroll
      for (int k = 0; k < VPT; ++k) {
        const int j = threadIdx.x + k * BLOCK;
        if (j < nvec && live) {
          float d[8], s[8];
          load8<T>(dy + row * cols + j * 8, d);
          load8<T>(xs + row * cols + j * 8, s);
          if (dsum_in != nullptr) load8<T>(dsum_in + row * cols + j * 8, rs[r][k]);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            xh[r][k][i] = (s[i] - mean) * rstd[r];
            g[r][k][i] = d[i] * wv[k][i];
            dwv[k][i] = fmaf(d[i], RMS ? DT<T>::rnd(xh[r][k][i]) : xh[r][k][i], dwv[k][i]);
            dbv[k][i] += d[i];
            sums[2 * r] += g[r][k][i];
            sums[2 * r + 1] = fmaf(g[r][k][i], xh[r][k][i], sums[2 * r + 1]);
          }
        }
      }
    }
    block_allreduce_sum_n<BLOCK, 2 * ROWS>(sums, red);
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int64_t row = row0 + r;
      if (row >= rows) break;
      const float mg = RMS ? 0.f : sums[2 * r] * inv_cols;
      const float mgx = sums[2 * r + 1] * inv_cols;
#pragma unroll
      for (int k = 0; k < VPT; ++k) {
        const int j = threadIdx.x + k * BLOCK;
        if (j < nvec) {
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            o[i] = rstd[r] * (g[r][k][i] - mg - xh[r][k][i] * mgx);
            if (dsum_in != nullptr) o[i] += rs[r][k][i];
          }
          if (DROP) {                       // d residual = o ; d x = o * keep / (1 - p)
            if (dres != nullptr) store8<T>(dres + row * cols + j * 8, o);
            const uint32_t keep = keep_mask8(seed, row * nvec + j, thr16);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = ((keep >> i) & 1u) ? o[i] * inv_keep : 0.f;
          }
          store8<T>(dx + row * cols + j * 8, o);
          if (DXS) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dxv[k][i] += o[i];
          }
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int j = threadIdx.x + k * BLOCK;
    if (j < nvec) {
      if (DXS) {
        float* dxp = dxs_part + (int64_t)blockIdx.x * cols + j * 8;
        reinterpret_cast<float4*>(dxp)[0] = make_float4(dxv[k][0], dxv[k][1], dxv[k][2], dxv[k][3]);
        reinterpret_cast<float4*>(dxp)[1] = make_float4(dxv[k][4], dxv[k][5], dxv[k][6], dxv[k][7]);
      }
      float* dwp = dw_part + (int64_t)blockIdx.x * cols + j * 8;
      reinterpret_cast<float4*>(dwp)[0] = make_float4(dwv[k][0], dwv[k][1], dwv[k][2], dwv[k][3]);
      reinterpret_cast<float4*>(dwp)[1] = make_float4(dwv[k][4], dwv[k][5], dwv[k][6], dwv[k][7]);
      if (!RMS && db_part != nullptr) {
        float* dbp = db_part + (int64_t)blockIdx.x * cols + j * 8;
        reinterpret_cast<float4*>(dbp)[0] = make_float4(dbv[k][0], dbv[k][1], dbv[k][2], dbv[k][3]);
        reinterpret_cast<float4*>(dbp)[1] = make_float4(dbv[k][4], dbv[k][5], dbv[k][6], dbv[k][7]);
      }
    }
  }
}

// partials [parts, cols] -> dweight/dbias (T).  blockDim (8, 128): 8 threads x float4 = 32 columns,
// 128 row lanes; every thread folds parts/128 rows with independent 16-byte loads, then the CTA
// folds its 128 lanes through shared memory.  (The first version used 32 row lanes and scalar
// loads: 46 dependent trips per thread, 10 us per call, 146 calls per micro-step.)
template <typename T>
__global__ void __launch_bounds__(1024) norm_bwd_finalize_kernel(const float* __restrict__ dw_part,
                                                                 const float* __restrict__ db_part,
                                                                 const float* __restrict__ dxs_part,
                                                                 T* __restrict__ dw, T* __restrict__ db,
                                                                 T* __restrict__ dxs, int parts,
                                                                 int cols, int acc_mask) {
  __shared__ float4 sw[128][8], sb[128][8], sx[128][8];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c0 = blockIdx.x * 32 + tx * 4;
  float4 aw = make_float4(0.f, 0.f, 0.f, 0.f), ab = aw, ax = aw;
  if (c0 < cols) {
    for (int r = ty; r < parts; r += 128) {
      const float4 v = *reinterpret_cast<const float4*>(dw_part + (int64_t)r * cols + c0);
      aw.x += v.x; aw.y += v.y; aw.z += v.z; aw.w += v.w;
      if (db_part != nullptr) {
        const float4 u = *reinterpret_cast<const float4*>(db_part + (int64_t)r * cols + c0);
        ab.x += u.x; ab.y += u.y; ab.z += u.z; ab.w += u.w;
      }
      if (dxs_part != nullptr) {
        const float4 u = *reinterpret_cast<const float4*>(dxs_part + (int64_t)r * cols + c0);
        ax.x += u.x; ax.y += u.y; ax.z += u.z; ax.w += u.w;
      }
    }
  }
  sw[ty][tx] = aw;
  sb[ty][tx] = ab;
  sx[ty][tx] = ax;
  __syncthreads();
  for (int half = 64; half > 0; half >>= 1) {      // tree over the 128 row lanes
    if (ty < half) {
      float4 a = sw[ty][tx], b2 = sw[ty + half][tx];
      sw[ty][tx] = make_float4(a.x + b2.x, a.y + b2.y, a.z + b2.z, a.w + b2.w);
      a = sb[ty][tx]; b2 = sb[ty + half][tx];
      sb[ty][tx] = make_float4(a.x + b2.x, a.y + b2.y, a.z + b2.z, a.w + b2.w);
      if (dxs_part != nullptr) {
        a = sx[ty][tx]; b2 = sx[ty + half][tx];
        sx[ty][tx] = make_float4(a.x + b2.x, a.y + b2.y, a.z + b2.z, a.w + b2.w);
      }
    }
    __syncthreads();
  }
  if (ty == 0 && c0 < cols) {
    if (dxs != nullptr && dxs_part != nullptr) {
      float4 x4 = sx[0][tx];
      if (acc_mask & 4) {
        x4.x += DT<T>::to_f(dxs[c0]); x4.y += DT<T>::to_f(dxs[c0 + 1]);
        x4.z += DT<T>::to_f(dxs[c0 + 2]); x4.w += DT<T>::to_f(dxs[c0 + 3]);
      }
      dxs[c0] = DT<T>::from_f(x4.x); dxs[c0 + 1] = DT<T>::from_f(x4.y);
      dxs[c0 + 2] = DT<T>::from_f(x4.z); dxs[c0 + 3] = DT<T>::from_f(x4.w);
    }
    float4 a = sw[0][tx], b2 = sb[0][tx];
    if (acc_mask & 1) {     // accumulate into the existing gradient: one fp32 add, one rounding (no add kernel later)
      a.x += DT<T>::to_f(dw[c0]); a.y += DT<T>::to_f(dw[c0 + 1]);
      a.z += DT<T>::to_f(dw[c0 + 2]); a.w += DT<T>::to_f(dw[c0 + 3]);
    }
    dw[c0] = DT<T>::from_f(a.x); dw[c0 + 1] = DT<T>::from_f(a.y);
    dw[c0 + 2] = DT<T>::from_f(a.z); dw[c0 + 3] = DT<T>::from_f(a.w);
    if (db != nullptr && db_part != nullptr) {
      if (acc_mask & 2) {
        b2.x += DT<T>::to_f(db[c0]); b2.y += DT<T>::to_f(db[c0 + 1]);
        b2.z += DT<T>::to_f(db[c0 + 2]); b2.w += DT<T>::to_f(db[c0 + 3]);
      }
      db[c0] = DT<T>::from_f(b2.x); db[c0 + 1] = DT<T>::from_f(b2.y);
      db[c0 + 2] = DT<T>::from_f(b2.z); db[c0 + 3] = DT<T>::from_f(b2.w);
    }
  }
}

struct DropArgs {
  float inv_keep;
  uint32_t thr16;
  uint64_t seed;
  bool on;
};

template <typename T, int VPT, int BLOCK>
static int launch_fwd(const void* x, const void* res, const void* w, const void* b, void* y, void* sum_out,
                      float* mean, float* rstd, int64_t rows, int cols, float eps, int rms, DropArgs d,
                      cudaStream_t s) {
  const int64_t cap = (int64_t)kNumSMs * (2048 / BLOCK);
  const int grid = (int)(rows < cap ? rows : cap);
#define DLION_NORM_FWD(RMS_, DROP_)                                                                              \
  norm_fwd_kernel<T, VPT, BLOCK, RMS_, DROP_><<<grid, BLOCK, 0, s>>>((const T*)x, (const T*)res, (const T*)w,      \
                                                                     RMS_ ? nullptr : (const T*)b, (T*)y,          \
                                                                     (T*)sum_out, mean, rstd, rows, cols, eps,     \
                                                                     d.inv_keep, d.thr16, d.seed)
  if (rms) { if (d.on) DLION_NORM_FWD(true, true); else DLION_NORM_FWD(true, false); }
  else     { if (d.on) DLION_NORM_FWD(false, true); else DLION_NORM_FWD(false, false); }
#undef DLION_NORM_FWD
  return DLION_LAUNCH_CHECK("dlion_norm_fwd");
}

static int tuning(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}
static int bwd_rows_per_trip(int cols) {           // 1 by default; 2 trades occupancy for loads in flight per thread
  static const int forced = tuning("DLION_NORM_BWD_ROWS", 1);
  if ((cols >> 3) > 256) return 1;
  return forced == 2 ? 2 : 1;
}
// mirrors NORM_SHAPE_DISPATCH: (VPT, BLOCK) for a column count
static void norm_shape(int cols, int* vpt, int* block) {
  const int nvec = cols >> 3;
  *vpt = nvec <= 256 ? 1 : (nvec <= 512 ? 2 : 4);
  *block = nvec <= 32 ? 32 : nvec <= 64 ? 64 : nvec <= 96 ? 96 : nvec <= 128 ? 128 : 256;
}
static int bwd_max_parts(int cols) {
  int vpt, block;
  norm_shape(cols, &vpt, &block);
  return kNumSMs * bwd_ctas_per_sm(vpt, block);
}

template <typename T, int VPT, int BLOCK>
static int launch_bwd(const void* dy, const void* dsum_in, const void* xs, const void* w, const float* mean,
                      const float* rstd, void* dx, void* dres, float* dwp, float* dbp, float* dxsp, int64_t rows, int cols,
                      int rms, DropArgs d, cudaStream_t s) {
  const bool dxs = dxsp != nullptr;
  const int rows_per_trip = dxs ? 1 : bwd_rows_per_trip(cols);
  const int64_t trips = (rows + rows_per_trip - 1) / rows_per_trip;
  const int cap = kNumSMs * (dxs ? bwd_ctas_per_sm_dxs(VPT, BLOCK) : bwd_ctas_per_sm(VPT, BLOCK));
  const int grid = (int)(trips < cap ? trips : cap);
  if (dxs) {          // LayerNorm with narrow rows only (dlion_norm_bwd_dxsum_supported): GPT-2's shapes
    if constexpr (VPT == 1) {
      if (!rms) {
        if (d.on)
          norm_bwd_kernel<T, VPT, BLOCK, false, 1, true, true><<<grid, BLOCK, 0, s>>>(
              (const T*)dy, (const T*)dsum_in, (const T*)xs, (const T*)w, mean, rstd, (T*)dx, (T*)dres, dwp, dbp, dxsp,
              rows, cols, d.inv_keep, d.thr16, d.seed);
        else
          norm_bwd_kernel<T, VPT, BLOCK, false, 1, false, true><<<grid, BLOCK, 0, s>>>(
              (const T*)dy, (const T*)dsum_in, (const T*)xs, (const T*)w, mean, rstd, (T*)dx, (T*)dres, dwp, dbp, dxsp,
              rows, cols, d.inv_keep, d.thr16, d.seed);
        return DLION_LAUNCH_CHECK("dlion_norm_bwd");
      }
    }
    set_error("dlion_norm_bwd_dxsum: dx column sums are available for LayerNorm with cols <= 2048 only");
    return DLION_E_UNSUPPORTED;
  }
#define DLION_NORM_BWD(RMS_, DROP_, ROWS_)                                                                        \
  norm_bwd_kernel<T, VPT, BLOCK, RMS_, ROWS_, DROP_, false><<<grid, BLOCK, 0, s>>>(                                 \
      (const T*)dy, (const T*)dsum_in, (const T*)xs, (const T*)w, mean, rstd, (T*)dx, (T*)dres, dwp, dbp, dxsp,    \
      rows, cols, d.inv_keep, d.thr16, d.seed)
#define DLION_NORM_BWD_R(RMS_, DROP_)                                                                            \
  do {                                                                                                           \
    if (VPT == 1 && rows_per_trip == 2) DLION_NORM_BWD(RMS_, DROP_, (VPT == 1 ? 2 : 1));                         \
    else DLION_NORM_BWD(RMS_, DROP_, 1);                                                                         \
  } while (0)
  if (rms) { if (d.on) DLION_NORM_BWD_R(true, true); else DLION_NORM_BWD_R(true, false); }
  else     { if (d.on) DLION_NORM_BWD_R(false, true); else DLION_NORM_BWD_R(false, false); }
#undef DLION_NORM_BWD_R
#undef DLION_NORM_BWD
  return DLION_LAUNCH_CHECK("dlion_norm_bwd");
}

// choose (VPT, BLOCK) so that BLOCK * VPT * 8 >= cols with few idle threads
#define NORM_SHAPE_DISPATCH(FN, ...)                                              \
  do {                                                                            \
    const int nvec = cols >> 3;                                                   \
    if (nvec <= 32) return FN<T, 1, 32>(__VA_ARGS__);                             \
    if (nvec <= 64) return FN<T, 1, 64>(__VA_ARGS__);                             \
    if (nvec <= 96) return FN<T, 1, 96>(__VA_ARGS__);                             \
    if (nvec <= 128) return FN<T, 1, 128>(__VA_ARGS__);                           \
    if (nvec <= 256) return FN<T, 1, 256>(__VA_ARGS__);                           \
    if (nvec <= 512) return FN<T, 2, 256>(__VA_ARGS__);                           \
    if (nvec <= 1024) return FN<T, 4, 256>(__VA_ARGS__);                          \
    ::dlion::set_error("norm: cols %d > 8192 unsupported", cols);                 \
    return DLION_E_UNSUPPORTED;                                                   \
  } while (0)

template <typename T>
static int fwd_t(const void* x, const void* res, const void* w, const void* b, void* y, void* sum_out, float* mean,
                 float* rstd, int64_t rows, int cols, float eps, int rms, DropArgs d, cudaStream_t s) {
  NORM_SHAPE_DISPATCH(launch_fwd, x, res, w, b, y, sum_out, mean, rstd, rows, cols, eps, rms, d, s);
}
template <typename T>
static int bwd_t(const void* dy, const void* dsum_in, const void* xs, const void* w, const float* mean,
                 const float* rstd, void* dx, void* dres, float* dwp, float* dbp, float* dxsp, int64_t rows, int cols,
                 int rms, DropArgs d, cudaStream_t s) {
  NORM_SHAPE_DISPATCH(launch_bwd, dy, dsum_in, xs, w, mean, rstd, dx, dres, dwp, dbp, dxsp, rows, cols, rms, d, s);
}

static DropArgs make_drop(float p, uint64_t seed) {
  DropArgs d;
  d.on = p > 0.f;
  d.inv_keep = d.on ? 1.0f / (1.0f - p) : 1.0f;
  float t = p * 65536.0f;
  d.thr16 = t >= 65535.f ? 65535u : (uint32_t)(t + 0.5f);
  d.seed = seed;
  return d;
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_norm_fwd(const void* x, const void* residual, const void* weight, const void* bias, void* y,
                              void* sum_out, float* mean, float* rstd, int64_t rows, int cols, float eps, int rms,
                              float drop_p, uint64_t drop_seed, int dtype, void* stream) {
  DLION_CHECK_ARG(rows >= 0 && cols > 0 && (cols & 7) == 0, "dlion_norm_fwd: cols (%d) must be a positive multiple of 8",
                  cols);
  DLION_CHECK_ARG(aligned16(x) && aligned16(y) && aligned16(weight) && aligned16(residual) && aligned16(sum_out) &&
                      aligned16(bias),
                  "dlion_norm_fwd: pointers must be 16-byte aligned");
  DLION_CHECK_ARG(rstd != nullptr && (rms || mean != nullptr), "dlion_norm_fwd: mean/rstd outputs required");
  DLION_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "dlion_norm_fwd: dropout probability %f outside [0, 1)", drop_p);
  DLION_CHECK_ARG(drop_p == 0.f || sum_out != nullptr, "dlion_norm_fwd: dropout needs sum_out (the pre-norm stream)");
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const DropArgs d = make_drop(drop_p, drop_seed);
  if (norm_tma_ok(cols)) {      // narrow rows: TMA tile ring + warp-per-row kernel (norm_tma.cu)
    DISPATCH_DTYPE(dtype, return launch_norm_fwd_tma<T>(x, residual, weight, bias, y, sum_out, mean, rstd, rows, cols, eps,
                                                        rms != 0, d.on, d.inv_keep, d.thr16, d.seed, s));
  }
  switch (dtype) {
    case DLION_F32: return fwd_t<float>(x, residual, weight, bias, y, sum_out, mean, rstd, rows, cols, eps, rms, d, s);
    case DLION_BF16:
      return fwd_t<__nv_bfloat16>(x, residual, weight, bias, y, sum_out, mean, rstd, rows, cols, eps, rms, d, s);
    case DLION_F16: return fwd_t<__half>(x, residual, weight, bias, y, sum_out, mean, rstd, rows, cols, eps, rms, d, s);
  }
  set_error("dlion_norm_fwd: unsupported dtype %d", dtype);
  return DLION_E_UNSUPPORTED;
}

extern "C" int dlion_norm_bwd_partial_rows(void) { return kNumSMs * 16; }

extern "C" int dlion_norm_bwd_parts(int64_t rows, int cols) {
  const int rows_per_trip = bwd_rows_per_trip(cols);        // must mirror launch_bwd
  const int64_t trips = (rows + rows_per_trip - 1) / rows_per_trip;
  const int cap = bwd_max_parts(cols);
  return (int)(trips < cap ? trips : cap);
}

extern "C" int dlion_norm_bwd_dxsum_supported(int cols, int rms) { return !rms && (cols >> 3) <= 256 ? 1 : 0; }

extern "C" int dlion_norm_bwd_dxsum_parts(int64_t rows, int cols) {      // mirrors launch_bwd's dxs branch
  int vpt, block;
  norm_shape(cols, &vpt, &block);
  const int cap = kNumSMs * bwd_ctas_per_sm_dxs(vpt, block);
  return (int)(rows < cap ? rows : cap);
}

extern "C" int dlion_norm_bwd(const void* dy, const void* dsum_in, const void* xsum, const void* weight,
                              const float* mean, const float* rstd, void* dx, void* dres, float* dw_partial,
                              float* db_partial, int64_t rows, int cols, int rms, float drop_p, uint64_t drop_seed,
                              int dtype, void* stream) {
  return dlion_norm_bwd_dxsum(dy, dsum_in, xsum, weight, mean, rstd, dx, dres, dw_partial, db_partial, nullptr, rows,
                              cols, rms, drop_p, drop_seed, dtype, stream);
}

extern "C" int dlion_norm_bwd_dxsum(const void* dy, const void* dsum_in, const void* xsum, const void* weight,
                                    const float* mean, const float* rstd, void* dx, void* dres, float* dw_partial,
                                    float* db_partial, float* dxs_partial, int64_t rows, int cols, int rms, float drop_p,
                                    uint64_t drop_seed, int dtype, void* stream) {
  DLION_CHECK_ARG(rows >= 0 && cols > 0 && (cols & 7) == 0, "dlion_norm_bwd: cols (%d) must be a positive multiple of 8",
                  cols);
  DLION_CHECK_ARG(aligned16(dy) && aligned16(dsum_in) && aligned16(xsum) && aligned16(weight) && aligned16(dx) &&
                      aligned16(dres) && aligned16(dw_partial) && aligned16(db_partial) && aligned16(dxs_partial),
                  "dlion_norm_bwd: pointers must be 16-byte aligned");
  DLION_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "dlion_norm_bwd: dropout probability %f outside [0, 1)", drop_p);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const DropArgs d = make_drop(drop_p, drop_seed);
  switch (dtype) {
    case DLION_F32:
      return bwd_t<float>(dy, dsum_in, xsum, weight, mean, rstd, dx, dres, dw_partial, db_partial, dxs_partial, rows,
                          cols, rms, d, s);
    case DLION_BF16:
      return bwd_t<__nv_bfloat16>(dy, dsum_in, xsum, weight, mean, rstd, dx, dres, dw_partial, db_partial, dxs_partial,
                                  rows, cols, rms, d, s);
    case DLION_F16:
      return bwd_t<__half>(dy, dsum_in, xsum, weight, mean, rstd, dx, dres, dw_partial, db_partial, dxs_partial, rows,
                           cols, rms, d, s);
  }
  set_error("dlion_norm_bwd: unsupported dtype %d", dtype);
  return DLION_E_UNSUPPORTED;
}

extern "C" int dlion_norm_bwd_finalize_dxsum(const float* dw_partial, const float* db_partial,
                                             const float* dxs_partial, void* dweight, void* dbias, void* dxsum,
                                             int parts, int cols, int acc_mask, int dtype, void* stream) {
  DLION_CHECK_ARG(parts > 0 && cols > 0 && (cols & 3) == 0 && dw_partial && dweight,
                  "dlion_norm_bwd_finalize: bad arguments");
  DLION_CHECK_ARG((dxs_partial == nullptr) == (dxsum == nullptr), "dlion_norm_bwd_finalize: dxs_partial and dxsum go together");
  cudaStream_t s = (cudaStream_t)stream;
  const dim3 block(8, 128);
  const int grid = (cols + 31) / 32;
  DISPATCH_DTYPE(dtype, norm_bwd_finalize_kernel<T><<<grid, block, 0, s>>>(dw_partial, db_partial, dxs_partial,
                                                                             (T*)dweight, (T*)dbias, (T*)dxsum, parts,
                                                                             cols, acc_mask));
  return DLION_LAUNCH_CHECK("dlion_norm_bwd_finalize");
}

extern "C" int dlion_norm_bwd_finalize_acc(const float* dw_partial, const float* db_partial, void* dweight,
                                           void* dbias, int parts, int cols, int acc_mask, int dtype, void* stream) {
  return dlion_norm_bwd_finalize_dxsum(dw_partial, db_partial, nullptr, dweight, dbias, nullptr, parts, cols, acc_mask,
                                       dtype, stream);
}

extern "C" int dlion_norm_bwd_finalize(const float* dw_partial, const float* db_partial, void* dweight, void* dbias,
                                       int parts, int cols, int dtype, void* stream) {
  return dlion_norm_bwd_finalize_acc(dw_partial, db_partial, dweight, dbias, parts, cols, 0, dtype, stream);
}
