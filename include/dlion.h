/*
 * dlion.h — C ABI of libdlion.so, the sm_100a kernel library of distributed_lion_pytorch_b200.
 *
 * Conventions
 *   - every entry point returns 0 on success, a cudaError_t (> 0) when the CUDA runtime
 *     reported a failure, or a negative DLION_E_* code for argument errors;
 *     dlion_last_error() returns a human-readable description of the last failure on the
 *     calling thread;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - tensors are raw device pointers, 16-byte aligned unless stated otherwise; `dtype` is a
 *     dlion_dtype; element counts are int64_t;
 *   - nothing in this library allocates device memory unless the comment says so: callers
 *     (PyTorch's caching allocator, in this repository) own every buffer.
 *
 * "reference:" comments cite kyleliang919/distributed-lion-pytorch (file:line) — the Python
 * op sequence each kernel replaces.
 */
#ifndef DLION_H_
#define DLION_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLION_VERSION 103

typedef enum { DLION_F32 = 0, DLION_BF16 = 1, DLION_F16 = 2 } dlion_dtype;

/* vote rule applied to cnt (# of live +1 ballots) out of L live workers */
typedef enum {
  DLION_VOTE_MAJORITY_NEG = 0,  /* +1 iff 2cnt > L else -1      (reference: distributed_lion.py:33-43,91) */
  DLION_VOTE_MAJORITY_ZERO = 1, /* as above, 0 on 2cnt == L     (paper: sign(sum delta_i))               */
  DLION_VOTE_AVERAGE = 2        /* (2cnt - L) / L               (paper: Averaging)                       */
} dlion_vote;

#define DLION_E_BADARG (-1)
#define DLION_E_UNSUPPORTED (-2)
#define DLION_E_TOOBIG (-3)
#define DLION_E_TIMEOUT (-4)

#define DLION_MAX_RANKS 16

const char* dlion_last_error(void);
int dlion_version(void);
/* number of SMs / max co-resident CTAs of the fused kernel on the current device */
int dlion_device_sm_count(int* out);

/* ---------------------------------------------------------------- optimizer kernels */

/* N6 — single-worker Lion, one pass: p*=decay; p-=lr*sign(b1*m+(1-b1)*g); m=b2*m+(1-b2)*g.
 * `decay` = 1 - lr*wd, `omb1` = 1-beta1, `omb2` = 1-beta2 are computed by the caller in double
 * and rounded once, like ATen does with Python scalars.
 * reference: distributed_lion.py:47-59 (update_fn: 7 ATen kernels, 2 temporaries). */
int dlion_lion_local(void* p, const void* g, void* m, int64_t n, int dtype, float lr, float decay,
                     float beta1, float omb1, float beta2, float omb2, void* stream);
/* Same with opt-in stochastic rounding of the parameter write (bf16 only; other dtypes ignore it): sr_key != 0
 * selects it, the random bits are a counter hash of (sr_key, elem_base + index), so every rank and every
 * back-end draws the same bits for the same bucket element.  Not in the reference (at lr ~ 1e-4
 * a Lion update is below half a bf16 ulp of a typical weight). */
int dlion_lion_local_sr(void* p, const void* g, void* m, int64_t n, int dtype, float lr, float decay, float beta1,
                        float omb1, float beta2, float omb2, uint64_t sr_key, int64_t elem_base, void* stream);

/* The same three kernels with the step's neighbours folded in (reference: HF Trainer's clip_grad_norm_ before
 * optimizer.step() and zero_grad() after it, which upstream inherits — run_clm.py:588-622):
 *   clip_sumsq != NULL  device pointer to sum(g^2) over all gradients of the model; the kernel works on
 *                       g' = round_T(g * min(1, clip_max_norm / (sqrt(*clip_sumsq) + 1e-6))) — bit for bit what
 *                       dlion_clip_scale followed by the plain kernel computes — without rewriting g;
 *   zero_g != 0         g is overwritten with zeros once consumed (lion_local, vote_update). */
int dlion_lion_local_clip(void* p, void* g, void* m, int64_t n, int dtype, float lr, float decay, float beta1,
                          float omb1, float beta2, float omb2, uint64_t sr_key, int64_t elem_base,
                          const float* clip_sumsq, float clip_max_norm, int zero_g, void* stream);
int dlion_sign_pack_clip(const void* m, const void* g, uint32_t* words, int64_t n, int dtype, float beta1, float omb1,
                         float stoch_r, uint64_t seed, uint32_t step, uint32_t rank, const float* clip_sumsq,
                         float clip_max_norm, void* stream);
int dlion_vote_update_clip(void* p, void* g, void* m, const uint32_t* gathered, int64_t row_words, int world,
                           uint32_t live_mask, int64_t n, int dtype, float lr, float decay, float beta2, float omb2,
                           int vote, uint64_t sr_key, const float* clip_sumsq, float clip_max_norm, int zero_g,
                           void* stream);

/* Sign + pack: words[e/32] bit e%32 = (b1*m + (1-b1)*g)[e] > 0; tail bits of the last word are 0.
 * stoch_r > 0 selects stochastic binarisation, bit ~ Bernoulli((c + r) / 2r) with a Philox
 * stream keyed by (seed, step, rank, element).
 * reference: distributed_lion.py:68-77 (and :106-117 for the stochastic variant). */
int dlion_sign_pack(const void* m, const void* g, uint32_t* words, int64_t n, int dtype, float beta1,
                    float omb1, float stoch_r, uint64_t seed, uint32_t step, uint32_t rank, void* stream);

/* Vote + update from gathered words (`gathered` = world rows of `row_words` uint32 each).
 * live_mask bit r = rank r's ballot counts.
 * reference: distributed_lion.py:64,84-96 (unpack ×W, torch.mode, add_, momentum EMA). */
int dlion_vote_update(void* p, const void* g, void* m, const uint32_t* gathered, int64_t row_words,
                      int world, uint32_t live_mask, int64_t n, int dtype, float lr, float decay,
                      float beta2, float omb2, int vote, void* stream);
int dlion_vote_update_sr(void* p, const void* g, void* m, const uint32_t* gathered, int64_t row_words, int world,
                         uint32_t live_mask, int64_t n, int dtype, float lr, float decay, float beta2, float omb2,
                         int vote, uint64_t sr_key, void* stream);

/* N1 — the fused sign-pack-exchange-vote-update kernel (one persistent cooperative launch).
 *
 * Every rank launches the same grid over the same n (multiple of the tile).  For tile t the
 * CTA that owns it TMA-loads g/m/p into shared memory, packs the signs, stores the words into
 * arena slot [parity][my rank] of every peer (`arena[r]` is rank r's arena as mapped into THIS
 * process; arena[rank] is the local one), releases a per-(source rank, CTA) progress flag on
 * each peer, and `lookahead` tiles later acquires the peers' flags, popcounts the live ballots
 * and writes p and m.  No other kernel and no NCCL call takes part.
 *
 *   arena layout (uint32): [2 parities][world][arena_words]      arena_words >= n/32
 *   flags layout (uint64): [world][DLION_FUSED_MAX_CTAS]         monotone, never reset
 *   epoch: strictly increasing launch counter shared by all ranks (parity = epoch & 1)
 *
 * emulate_ranks > 0: single-GPU protocol emulation — the grid is split into `emulate_ranks`
 * groups of CTAs, group v acting as rank v on p/g/m + v*emu_stride elements, all arenas /
 * flags living on this device (arena[v], flags[v]).  Used by the one-GPU tests.
 *
 * status (device int32[4], may be NULL): [0]=0 ok / 1 flag-wait timeout, [1]=CTA, [2]=src rank; tiles whose wait
 * expired are not applied (see wait_timeout_us).
 * reference: distributed_lion.py:61-96 + the NCCL all_gather it calls at :81. */
typedef struct {
  void* p;
  const void* g;
  void* m;
  int64_t n;           /* elements in this launch (multiple of dlion_fused_tile(dtype)) */
  int64_t elem_base;   /* index of p[0] within its bucket: keeps the Philox stream chunk-invariant */
  int dtype;
  int world;
  int rank;
  uint32_t live_mask;  /* ballots that count */
  uint32_t wait_mask;  /* ranks whose flags are awaited (normally all) */
  float lr, decay, beta1, omb1, beta2, omb2;
  int vote;
  float stoch_r;
  uint64_t seed;
  uint32_t step;
  uint64_t epoch;
  uint32_t* arena[DLION_MAX_RANKS];
  uint64_t* flags[DLION_MAX_RANKS];
  uint32_t* mc_arena;  /* optional NVLS multicast alias of the arenas (same layout on every rank): the packed
                          words then leave the GPU once (multimem.st) and the switch replicates them; NULL = unicast */
  int64_t arena_words;
  int grid;            /* CTAs per (virtual) rank; identical on every rank */
  int emulate_ranks;
  int64_t emu_stride;
  int32_t* status;
  int64_t spin_limit;  /* flag polls before giving up (0 = default) */
  uint64_t sr_key;     /* != 0: stochastic rounding of the bf16 parameter write, keyed by (sr_key, element index) —
                          identical on every rank; 0 = round to nearest (the reference's arithmetic) */
  /* Folded neighbours of the step (reference: HF Trainer calls clip_grad_norm_ before optimizer.step() and
   * model.zero_grad() after it; upstream inherits both through transformers.Trainer, run_clm.py:588-622).
   * clip_sumsq != NULL: device pointer to sum(g^2) over ALL gradients of the model (dlion_sumsq); the kernel uses
   * g' = round_T(g * min(1, clip_max_norm / (sqrt(*clip_sumsq) + 1e-6))) for the sign and the momentum — bit for
   * bit what dlion_clip_scale followed by this kernel would compute — and g itself is not rewritten.
   * zero_g != 0: g is overwritten with zeros once consumed (the zero_grad that would follow the step). */
  const float* clip_sumsq;
  float clip_max_norm;
  int zero_g;
  /* Upper bound, in microseconds of GPU wall clock (%globaltimer), on the time a voter waits for one peer's
   * progress flag; 0 = only the poll-count limit `spin_limit` applies.  When a wait expires the kernel records it
   * in `status`, leaves the parameter tiles whose ballots are incomplete UNTOUCHED (momenta are local and still
   * advance) and runs to completion — a stalled peer can neither hang the GPU nor inject an update voted from
   * stale words.  (reference: dist.all_gather at distributed_lion.py:81 blocks forever on a missing rank.) */
  uint64_t wait_timeout_us;
} dlion_fused_args;

#define DLION_FUSED_MAX_CTAS 1024
int dlion_fused_step(const dlion_fused_args* args, void* stream);
/* tile size (elements) for a dtype, and the largest grid the device can co-schedule */
int dlion_fused_tile(int dtype);
int dlion_fused_max_grid(int dtype, int* out);

/* ---------------------------------------------------------------- bucket kernels (N2) */

/* Copy/cast up to `count` tensors into their slots of a flat bucket with one launch per 96
 * tensors.  srcs/dsts/numels are HOST arrays.  reference: the implicit per-tensor loop of
 * distributed_lion.py:179-198 and flatten_and_pad :14-24. */
int dlion_multi_gather(const void* const* srcs, void* const* dsts, const int64_t* numels, int count,
                       int src_dtype, int dst_dtype, void* stream);

/* Sum of squares of x into out[0] (fp32).  `scratch` is caller-provided device memory of at
 * least dlion_sumsq_scratch_bytes() bytes whose first 8 bytes are zero before the first use;
 * the kernel leaves them zero again, so one zero-initialised buffer serves every later call
 * on the same stream.  Used for local gradient-norm clipping (HF Trainer's max_grad_norm,
 * which the reference inherits: SURVEY §3.1). */
int64_t dlion_sumsq_scratch_bytes(void);
int dlion_sumsq(const void* x, int64_t n, int dtype, void* scratch, float* out, void* stream);
/* Same result as dlion_sumsq, for callers that cannot hand over a scratch buffer because they
 * cannot allocate device memory: a plugin inside a host application that gave its whole
 * device-memory budget out at start-up and forbids cudaMalloc afterwards, a monitoring hook
 * that must not disturb the framework's caching allocator, a plain C caller without an
 * allocator layer.  The scratch is a 16,640-byte array built into the library image on every
 * GPU that loads it.  Limit: n <= dlion_sumsq_static_max_n() = 268,435,456 elements, larger n is
 * refused with DLION_E_TOOBIG.  The first call on each device looks the array's device address
 * up (cudaGetSymbolAddress) and must therefore be made outside stream capture.  Calls share the
 * array: serialise them — two calls overlapping on the GPU race.  Never allocates.
 * (No counterpart in the reference: it has no native code and clips through HF's Trainer.) */
int64_t dlion_sumsq_static_max_n(void);
int dlion_sumsq_static(const void* x, int64_t n, int dtype, float* out, void* stream);
/* x *= min(1, max_norm / (sqrt(*sumsq) + 1e-6)) — torch.nn.utils.clip_grad_norm_'s coefficient */
int dlion_clip_scale(void* x, int64_t n, int dtype, const float* sumsq, float max_norm, void* stream);

/* ---------------------------------------------------------------- model kernels */

/* N3 — y = LayerNorm(dropout(x) + residual) (residual may be NULL; drop_p = 0 disables dropout); also
 * writes the pre-norm sum to `sum_out` (may be NULL only when residual is NULL and drop_p = 0) and
 * saves mean/rstd (fp32, one per row).  RMSNorm when `rms != 0` (bias ignored).  The dropout
 * keep-mask is a pure function of (drop_seed, element index): it is regenerated, not stored.
 * reference: HF GPT2Block / LlamaDecoderLayer reached via async_trainer.py:24 — resid_dropout,
 * the residual add and ln_1 / ln_2 are three ATen kernels and a mask tensor there. */
int dlion_norm_fwd(const void* x, const void* residual, const void* weight, const void* bias, void* y,
                   void* sum_out, float* mean, float* rstd, int64_t rows, int cols, float eps, int rms,
                   float drop_p, uint64_t drop_seed, int dtype, void* stream);
/* backward: d(sum) = norm-backward(dy) + dsum_in (may be NULL).  Without dropout dx = d(sum) (it is
 * also the residual's gradient).  With dropout dres (may be NULL) receives d(sum) and
 * dx = d(sum) * keep / (1 - p).  dweight/dbias partials are accumulated in fp32
 * `dw_partial`/`db_partial` of shape [dlion_norm_bwd_parts(rows, cols), cols] (one row per CTA
 * launched, at most dlion_norm_bwd_partial_rows()); reduce them with dlion_norm_bwd_finalize. */
int dlion_norm_bwd_partial_rows(void);
/* number of partial rows dlion_norm_bwd writes for this shape (= CTAs it launches) */
int dlion_norm_bwd_parts(int64_t rows, int cols);
int dlion_norm_bwd(const void* dy, const void* dsum_in, const void* xsum, const void* weight,
                   const float* mean, const float* rstd, void* dx, void* dres, float* dw_partial,
                   float* db_partial, int64_t rows, int cols, int rms, float drop_p, uint64_t drop_seed,
                   int dtype, void* stream);
int dlion_norm_bwd_finalize(const float* dw_partial, const float* db_partial, void* dweight, void* dbias,
                            int parts, int cols, int dtype, void* stream);
/* same, but bit 0 / bit 1 of acc_mask make dweight / dbias ACCUMULATE (out = T(float(out) + sum)): the
 * gradient lands in an existing .grad buffer with one rounding and no separate add kernel */
int dlion_norm_bwd_finalize_acc(const float* dw_partial, const float* db_partial, void* dweight, void* dbias,
                                int parts, int cols, int acc_mask, int dtype, void* stream);

/* The same two with one more reduction riding along: dxs_partial[parts, cols] receives per-CTA column sums of dx,
 * folded by the finalize into dxsum[cols] (bit 2 of acc_mask = accumulate).  dx is the gradient handed to whatever
 * produced x; when that is a biased linear (GPT-2's attn.c_proj / mlp.c_proj in front of ln_2 / the next ln_1)
 * dxsum IS its bias gradient and the layer's own pass over dx (`grad_output.sum(0)` in HF's Conv1D backward,
 * reached from accelerator.backward, async_trainer.py:33) is not needed. */
int dlion_norm_bwd_dxsum_supported(int cols, int rms);      /* LayerNorm with cols <= 2048 */
int dlion_norm_bwd_dxsum_parts(int64_t rows, int cols);     /* partial rows written when dxs_partial != NULL */
int dlion_norm_bwd_dxsum(const void* dy, const void* dsum_in, const void* xsum, const void* weight, const float* mean,
                         const float* rstd, void* dx, void* dres, float* dw_partial, float* db_partial,
                         float* dxs_partial, int64_t rows, int cols, int rms, float drop_p, uint64_t drop_seed,
                         int dtype, void* stream);
int dlion_norm_bwd_finalize_dxsum(const float* dw_partial, const float* db_partial, const float* dxs_partial,
                                  void* dweight, void* dbias, void* dxsum, int parts, int cols, int acc_mask, int dtype,
                                  void* stream);

/* Bias gradient: fp32 partial column sums of x [rows, cols] (cols % 8 == 0) into
 * partial[dlion_colsum_parts(rows, cols)][cols]; fold them with dlion_norm_bwd_finalize
 * (db_partial = NULL).  reference: `grad_output.sum(0)` of every biased Conv1D in HF GPT-2's
 * backward, reached from accelerator.backward (async_trainer.py:33). */
int dlion_colsum_parts(int64_t rows, int cols);
int dlion_colsum(const void* x, int64_t rows, int cols, float* partial, int dtype, void* stream);
/* One-launch form: the last CTA to finish on each 256-column tile folds the tile's partial rows and writes
 * (accumulate = 0) or adds into (accumulate = 1) out[cols] (same dtype as x).  `tickets` is
 * uint32[dlion_colsum_tiles(cols)], zero-initialised once by the caller; the kernel leaves it zeroed, so one
 * buffer serves every call ISSUED ON ONE STREAM (calls that overlap on the GPU must not share it). */
int dlion_colsum_tiles(int cols);
int dlion_colsum_fused(const void* x, int64_t rows, int cols, float* partial, unsigned int* tickets, void* out,
                       int accumulate, int dtype, void* stream);

/* Producer + bias-gradient reduction in one pass (same tiling, tickets and partial buffer as dlion_colsum_fused):
 *   dlion_gelu_tanh_bwd_colsum  dpre = dy * gelu_tanh'(pre) is written AND its column sums land in out[cols] —
 *                               GPT-2's c_fc bias gradient without a second pass over the [tokens, 4*hidden] tensor;
 *   dlion_qkv_pack_colsum       d(qkv)[B, S, 3, H, D] is packed from attention's dq / dk / dv ([B, H, S, D] with the
 *                               element strides strides_bhs[9] = {b, h, s} x {q, k, v}, d contiguous; HOST array) AND
 *                               its column sums (c_attn's bias gradient) land in out[3*H*D].
 * reference: the NewGELUActivation / unbind backward kernels followed by `grad_output.sum(0)` of the biased
 * Conv1D in HF GPT-2's backward, reached from accelerator.backward (async_trainer.py:33). */
int dlion_gelu_tanh_bwd_colsum(const void* pre, const void* dy, void* dpre, int64_t rows, int cols, float* partial,
                               unsigned int* tickets, void* out, int accumulate, int dtype, void* stream);
int dlion_qkv_pack_colsum(const void* dq, const void* dk, const void* dv, const int64_t* strides_bhs, void* dqkv,
                          int batch, int seq, int heads, int head_dim, float* partial, unsigned int* tickets, void* out,
                          int accumulate, int dtype, void* stream);

/* Token + position embedding.  Forward: out[row] = wte[ids[row]] + wpe[row % seq] in one pass (wpe may be NULL).
 * Backward of the token table: `sorted_ids` are the n token ids sorted ascending (stable) and `perm[k]` is the row of
 * dh that sorted position k came from; the gradient rows are summed in fp32 in sorted order (deterministic, no
 * atomics) and written (accumulate = 0: rows of absent tokens are NOT touched — pass a zeroed buffer) or added into
 * grad (accumulate = 1: grad is the live .grad buffer).
 * reference: HF GPT2Model wte / wpe and autograd's embedding_dense_backward + AccumulateGrad (async_trainer.py:33). */
int dlion_embed_fwd(const int64_t* ids, const void* wte, const void* wpe, void* out, int64_t rows, int seq, int hidden,
                    int64_t vocab, int dtype, void* stream);
int dlion_embed_bwd_sorted(const void* dh, const int64_t* sorted_ids, const int64_t* perm, void* grad, int64_t n,
                           int hidden, int64_t vocab, int accumulate, int dtype, void* stream);

/* MLP activations as single passes over n elements (n % 8 == 0): GPT-2's tanh-form GELU and
 * Llama's silu(gate) * up, forward and backward.
 * reference: HF GPT2MLP / LlamaMLP via Trainer.compute_loss (async_trainer.py:24). */
int dlion_gelu_tanh_fwd(const void* x, void* y, int64_t n, int dtype, void* stream);
int dlion_gelu_tanh_bwd(const void* x, const void* dy, void* dx, int64_t n, int dtype, void* stream);
int dlion_swiglu_fwd(const void* gate, const void* up, void* y, int64_t n, int dtype, void* stream);
int dlion_swiglu_bwd(const void* gate, const void* up, const void* dy, void* dgate, void* dup, int64_t n,
                     int dtype, void* stream);

/* N4 — rotary embedding applied in place to q [tokens, q_heads, head_dim] and k
 * [tokens, k_heads, head_dim] (HF "rotate_half" pairing: i <-> i + head_dim/2); cos/sin are
 * fp32 [seq, head_dim/2]; pos = token index % seq_len.  inverse != 0 applies the transpose
 * rotation (the backward).  reference: HF LlamaAttention via sft_llama2.py:147. */
int dlion_rope(void* q, void* k, const float* cos, const float* sin, int64_t tokens, int seq_len,
               int q_heads, int k_heads, int head_dim, int64_t q_token_stride, int64_t k_token_stride,
               int inverse, int dtype, void* stream);

/* N5 — cross-entropy over one chunk of logits rows, in place: on return logits holds
 * d loss / d logits * grad_scale (softmax - onehot, zero rows where label == ignore_index),
 * loss[row] = -log softmax[label] (0 if ignored), logp optional same thing negated.
 * reference: HF ForCausalLM loss in compute_loss (async_trainer.py:24); DPO log-probs in TRL. */
int dlion_ce_fwd_bwd(void* logits, const int64_t* labels, float* loss, int64_t rows, int vocab,
                     int64_t row_stride, int64_t ignore_index, float grad_scale, int write_grad, int dtype,
                     void* stream);
/* per-row backward scale variant used by DPO: logits <- (softmax - onehot) * row_scale[row] */
int dlion_ce_bwd_rowscale(void* logits, const int64_t* labels, const float* row_scale, int64_t rows,
                          int vocab, int64_t row_stride, int64_t ignore_index, int dtype, void* stream);

/* N5, no-grad side — vocabulary projection fused with log-sum-exp on the tcgen05 tensor cores (TMA-fed, TMEM
 * accumulators): lse[r] = log sum_v exp(h[r] . w[v]) and logp[r] = h[r] . w[labels[r]] - lse[r] (0 where the label is
 * outside [0, vocab), e.g. -100) without a [rows, vocab] logits tensor in HBM.  h [rows, hidden] and w [vocab, hidden]
 * are bf16 / fp16, row-major with row strides ld_h / ld_w elements (multiples of 8), 16-byte aligned.  logp or lse may
 * be NULL; labels may be NULL when logp is.  `workspace` holds dlion_lmhead_plan()'s number of floats; `status`
 * (device int, may be NULL) receives a code before the kernel traps if one of its bounded pipeline waits expires.
 * reference: HF ForCausalLM evaluation loss (run_clm.py:626-639); TRL DPOTrainer reference-model log-probs
 * (dpo_llama2.py:219-231). */
int dlion_lmhead_plan(int rows, int vocab, int* splits, int64_t* workspace_floats);
int dlion_lmhead_logprobs(const void* h, int64_t ld_h, const void* w, int64_t ld_w, const int64_t* labels, float* logp,
                          float* lse, float* workspace, int64_t workspace_floats, int* status, int rows, int vocab,
                          int hidden, int dtype, void* stream);

/* Data-gradient GEMM of an MLP down-projection with the tanh-GELU backward and the up-projection's bias gradient in
 * its epilogue (tcgen05 / TMEM / 2-D TMA kernel): dpre[m, n] = (dy[m, k] . w[n, k]^T) * gelu_tanh'(pre[m, n]);
 * part[m_tile, n] (fp32, dlion_dgrad_dgelu_parts(m) x n) receives the column sums of dpre per 128-row tile, which
 * dlion_dgrad_dgelu_bias_finalize folds (fixed order) into `out[n]` (+= with accumulate).  bf16 / fp16, row-major,
 * row strides in elements (multiples of 8), 16-byte aligned; dpre may not alias pre.
 * reference: HF GPT2MLP backward under autograd (Conv1D -> NewGELUActivation -> Conv1D), async_trainer.py:33. */
int dlion_dgrad_dgelu_parts(int64_t m);
int dlion_dgrad_dgelu_set_profile(long long* buf);   /* diagnostics: 148 x 16 int64 device counters, NULL = off */
int dlion_dgrad_dgelu(const void* dy, int64_t ld_dy, const void* w, int64_t ld_w, const void* pre, int64_t ld_pre, void* dpre,
                      int64_t ld_dpre, float* part, int* status, int m, int n, int k, int dtype, void* stream);
int dlion_dgrad_dgelu_bias_finalize(const float* part, void* out, int m_tiles, int n, int accumulate, int dtype, void* stream);

/* Blockwise 4-bit weight codec for frozen LoRA base weights (bitsandbytes layout: element 2i in
 * the high nibble of byte i, one fp32 absmax per `blocksize` elements, 16-entry code book given
 * as a HOST pointer — NF4 or FP4).  n must be a multiple of blocksize; quantize needs
 * blocksize % 64 == 0, dequantize a power of two >= 8; `packed` holds n/2 bytes and `absmax`
 * n/blocksize floats.  dtype is the type of w / out.
 * reference: BitsAndBytesConfig(load_in_4bit, nf4) at sft_llama2.py:141-145, dpo_llama2.py:137,151. */
int dlion_q4_quantize(const void* w, void* packed, float* absmax, int64_t n, int blocksize,
                      const float* code16, int dtype, void* stream);
int dlion_q4_dequantize(const void* packed, const float* absmax, void* out, int64_t n, int blocksize,
                        const float* code16, int dtype, void* stream);

/* Library GEMMs with cuBLASLt's training epilogues (the ones PyTorch does not expose).  Row-major:
 * d[m,n] (ld ldd) = alpha * op(a) @ op(b) + beta * d, a stored [m,k] (trans_a = 0) or [k,m], b stored
 * [k,n] (trans_b = 0) or [n,k].  Epilogues:
 *   DLION_LT_BIAS           d = a@b + bias[n]
 *   DLION_LT_GELU_AUX_BIAS  aux[m,n] = a@b + bias;  d = gelu_tanh(aux)
 *   DLION_LT_DGELU_BGRAD    d = gelu_tanh'(aux) * (a@b);  bias_grad[n] = sum_m d[m,n]
 *   DLION_LT_BGRAD_B        d = a@b (+ beta*d);  bias_grad[n] = sum_k b[k,n]   (weight + bias gradient in one pass)
 *   DLION_LT_DGELU          d = gelu_tanh'(aux) * (a@b)
 * `algo_rank` picks the r-th candidate of the library's heuristic (0 = its first choice); callers that tune time
 * a few ranks on scratch buffers and keep the fastest.
 * cuBLASLt is looked up at run time (dlopen), not linked: dlion_lt_available() says whether it was found;
 * without it, or when the library has no kernel for the request, dlion_lt_gemm returns DLION_E_UNSUPPORTED
 * and the caller keeps its unfused path.  `workspace` is caller-owned scratch (32 MiB is plenty).
 * reference: HF GPT-2 MLP forward/backward under autograd (run_clm.py:588-639). */
enum { DLION_LT_NONE = 0, DLION_LT_BIAS = 1, DLION_LT_GELU_AUX_BIAS = 2, DLION_LT_DGELU_BGRAD = 3, DLION_LT_BGRAD_B = 4,
       DLION_LT_DGELU = 5 };
typedef struct {
  const void* a;
  const void* b;
  void* d;
  int64_t m, n, k, lda, ldb, ldd;
  int trans_a, trans_b;
  float alpha, beta;
  int epilogue;
  const void* bias;
  void* bias_grad;
  void* aux;
  int64_t ld_aux;
  void* workspace;
  int64_t workspace_bytes;
  int dtype;
  int algo_rank;
} dlion_lt_args;
int dlion_lt_available(void);
int dlion_lt_gemm(const dlion_lt_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLION_H_ */
