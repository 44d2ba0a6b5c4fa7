// embedding.cu — token + position embedding, forward and (deterministic) backward.
//
//   dlion_embed_fwd        out[row] = wte[ids[row]] + wpe[row % seq]      one pass: gather, add, store
//   dlion_embed_bwd_sorted grad_wte[id] (+)= sum of dh rows whose token is id
//
// The backward takes the token ids SORTED (stable) together with the permutation that sorted them.  One CTA
// per sorted position; the CTA sitting on the first position of a run of equal ids walks the run, sums the
// corresponding dh rows in fp32 in sorted order (deterministic) and adds the sum into that id's gradient row —
// exactly one CTA touches a given row, so there are no atomics and no zero-filled [vocab, hidden] temporary.
// With `accumulate` the row is added into the existing gradient (the flat gradient bucket in steady state), which
// removes autograd's dense add of a [vocab, hidden] tensor as well.
// reference: HF GPT2Model's wte / wpe lookups and their backward under autograd (embedding_dense_backward:
// sort + two segment kernels + a zero-filled dense gradient + AccumulateGrad add), reached from
// accelerator.backward (async_trainer.py:33).
#include "common.cuh"

namespace dlion {

template <typename T>
__global__ void __launch_bounds__(256) embed_fwd_kernel(const int64_t* __restrict__ ids, const T* __restrict__ wte,
                                                        const T* __restrict__ wpe, T* __restrict__ out, int64_t rows,
                                                        int seq, int hidden, int64_t vocab) {
  const int nvec = hidden >> 3;
  const int vecs_per_block = blockDim.x;
  // a block covers blockDim.x 8-element vectors; rows are walked by blocks in y
  for (int64_t row = blockIdx.y; row < rows; row += gridDim.y) {
    int64_t id = ids[row];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);          // out-of-range ids never fault (torch raises earlier)
    const T* w = wte + id * hidden;
    const T* p = wpe != nullptr ? wpe + (int64_t)(row % seq) * hidden : nullptr;
    for (int j = blockIdx.x * vecs_per_block + threadIdx.x; j < nvec; j += gridDim.x * vecs_per_block) {
      float a[8];
      load8<T>(w + j * 8, a);
      if (p != nullptr) {
        float b[8];
        load8<T>(p + j * 8, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] += b[i];
      }
      store8<T>(out + row * hidden + j * 8, a);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(128) embed_bwd_sorted_kernel(const T* __restrict__ dh,
                                                               const int64_t* __restrict__ sorted_ids,
                                                               const int64_t* __restrict__ perm, T* __restrict__ grad,
                                                               int64_t n, int hidden, int64_t vocab, int accumulate) {
  const int64_t i = blockIdx.x;
  const int64_t id = sorted_ids[i];
  if (id < 0 || id >= vocab) return;
  if (i > 0 && sorted_ids[i - 1] == id) return;                // not the head of its run (block-uniform)
  const int nvec = hidden >> 3;
  for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int64_t k = i; k < n && sorted_ids[k] == id; ++k) {
      float v[8];
      load8<T>(dh + perm[k] * hidden + j * 8, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += v[e];
    }
    T* g = grad + id * hidden + j * 8;
    if (accumulate) {
      float old[8];
      load8<T>(g, old);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += old[e];
    }
    store8<T>(g, acc);
  }
}

}  // namespace dlion

using namespace dlion;

extern "C" int dlion_embed_fwd(const int64_t* ids, const void* wte, const void* wpe, void* out, int64_t rows, int seq,
                               int hidden, int64_t vocab, int dtype, void* stream) {
  DLION_CHECK_ARG(rows >= 0 && seq > 0 && hidden > 0 && (hidden & 7) == 0 && vocab > 0,
                  "dlion_embed_fwd: hidden (%d) must be a positive multiple of 8", hidden);
  DLION_CHECK_ARG(ids != nullptr && aligned16(wte) && aligned16(wpe) && aligned16(out), "dlion_embed_fwd: bad pointers");
  if (rows == 0) return 0;
  const int nvec = hidden >> 3;
  const int block = nvec >= 256 ? 256 : ((nvec + 31) / 32 * 32);
  const int gx = (nvec + block - 1) / block;
  int64_t gy = rows < (int64_t)kNumSMs * 16 ? rows : (int64_t)kNumSMs * 16;
  const dim3 grid(gx, (unsigned)gy);
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, embed_fwd_kernel<T><<<grid, block, 0, s>>>(ids, (const T*)wte, (const T*)wpe, (T*)out, rows, seq,
                                                                    hidden, vocab));
  return DLION_LAUNCH_CHECK("dlion_embed_fwd");
}

extern "C" int dlion_embed_bwd_sorted(const void* dh, const int64_t* sorted_ids, const int64_t* perm, void* grad,
                                      int64_t n, int hidden, int64_t vocab, int accumulate, int dtype, void* stream) {
  DLION_CHECK_ARG(n >= 0 && hidden > 0 && (hidden & 7) == 0 && vocab > 0,
                  "dlion_embed_bwd_sorted: hidden (%d) must be a positive multiple of 8", hidden);
  DLION_CHECK_ARG(n < (1ll << 31), "dlion_embed_bwd_sorted: at most 2^31 - 1 tokens per call");
  DLION_CHECK_ARG(aligned16(dh) && aligned16(grad) && sorted_ids != nullptr && perm != nullptr,
                  "dlion_embed_bwd_sorted: bad pointers");
  if (n == 0) return 0;
  const int nvec = hidden >> 3;
  const int block = nvec >= 128 ? 128 : ((nvec + 31) / 32 * 32);
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_DTYPE(dtype, embed_bwd_sorted_kernel<T><<<(unsigned)n, block, 0, s>>>((const T*)dh, sorted_ids, perm, (T*)grad, n,
                                                                                  hidden, vocab, accumulate));
  return DLION_LAUNCH_CHECK("dlion_embed_bwd_sorted");
}
