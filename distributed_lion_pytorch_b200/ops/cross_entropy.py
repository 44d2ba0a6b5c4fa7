"""N5 — chunked linear + cross-entropy: the [tokens, vocab] logits are never materialised.

``linear_cross_entropy(h, weight, labels)`` == ``F.cross_entropy(h @ weight.T, labels)`` (mean
over non-ignored tokens).  The token dimension is tiled; per tile: one cuBLAS GEMM makes a
bf16 logits block, the CE kernel (csrc/cross_entropy.cu) turns it into per-row losses and, in
place, into d loss / d logits; two more GEMMs fold that block into dh and dW straight away
(the gradient is produced during the forward, as the loss is always the last node).

``linear_log_probs(h, weight, labels)`` returns per-token log-probabilities (for DPO); its
backward recomputes each logits block and applies the per-row upstream gradient in the kernel.

reference: HF ``ForCausalLM`` loss inside ``compute_loss`` (async_trainer.py:24) — full logits
+ fp32 upcast + log_softmax; TRL's ``get_batch_logps`` for DPO.
"""
from __future__ import annotations

import contextlib
import os
from typing import Optional

import torch
import torch.nn.functional as F

from . import _cabi, lm_head, require_library

IGNORE_INDEX = -100
_DEFAULT_CHUNK = int(os.environ.get("DLION_CE_CHUNK", "4096"))   # rows of logits alive at a time


def _pad_rows(weight: torch.Tensor, mult: int = 64) -> torch.Tensor:
    """[V, H] -> [Vp, H] with zero rows so every logits row is 16-byte aligned and the GEMMs see
    friendly dimensions; ``wp[:V]`` is a contiguous prefix view.

    Measured alternative, rejected (``scripts/ab_ce_gemm_shapes.py``, B200, GPT-2 shape): leaving the weight
    alone and padding only the logits *stride* (GEMMs on ``buf[:, :V]`` views) makes cuBLAS fall off its TMA
    kernels — 2,895 us instead of 226 us for the logits GEMM, 2,072 instead of 204 us for dh — while this copy
    costs 36 us per micro-batch."""
    V = weight.shape[0]
    Vp = (V + mult - 1) // mult * mult
    if Vp == V:
        return weight
    wp = weight.new_empty(Vp, weight.shape[1])
    wp[:V].copy_(weight)
    wp[V:].zero_()
    return wp


def _compute_dtype(h: torch.Tensor) -> torch.dtype:
    """The dtype the logits GEMMs run in: the autocast dtype when autocast is on (mixed-precision training of fp32
    master weights: ``Trainer.autocast`` for ``--bf16`` / ``--fp16`` without ``--torch_dtype``), else h's own."""
    if h.is_cuda and torch.is_autocast_enabled("cuda"):
        return torch.get_autocast_dtype("cuda")
    if not h.is_cuda and _EMULATE and torch.is_autocast_enabled("cpu"):
        return torch.get_autocast_dtype("cpu")
    return h.dtype


_FUSED_FWD = os.environ.get("DLION_LMHEAD_FUSED_FWD", "1") == "1"   # forward of linear_log_probs WITH grad on the tcgen05 kernel
_EMULATE = os.environ.get("DLION_EMULATE_KERNELS") == "1"   # CPU stand-ins, to test the autograd plumbing


def _kernel_ok(h: torch.Tensor) -> bool:
    if not h.is_cuda:
        return _EMULATE
    return require_library() and h.dtype in (torch.float32, torch.bfloat16, torch.float16)


def _ce_kernel(logits, labels, losses, V: int, ignore_index: int, write_grad: bool, row_scale=None) -> None:
    """logits [c, Vp] (modified in place when write_grad / row_scale), losses fp32 [c] or None."""
    if logits.is_cuda:
        L = _cabi.lib()
        dt = _cabi.dtype_code(logits.dtype)
        st = _cabi.stream_ptr(logits.device)
        rows, Vp = logits.shape
        if row_scale is None:
            _cabi.check(L.dlion_ce_fwd_bwd(logits.data_ptr(), labels.data_ptr(), losses.data_ptr(), rows, V, Vp,
                                           ignore_index, 1.0, int(write_grad), dt, st), "ce_fwd_bwd")
        else:
            _cabi.check(L.dlion_ce_bwd_rowscale(logits.data_ptr(), labels.data_ptr(), row_scale.data_ptr(), rows, V,
                                                Vp, ignore_index, dt, st), "ce_bwd_rowscale")
        return
    # torch stand-in with the kernel's exact contract
    x = logits[:, :V].float()
    valid = (labels != ignore_index) & (labels >= 0) & (labels < V)
    safe = labels.clamp(0, V - 1)
    lse = torch.logsumexp(x, dim=-1)
    if losses is not None:
        losses.copy_(torch.where(valid, lse - x.gather(1, safe[:, None]).squeeze(1), torch.zeros_like(lse)))
    if write_grad or row_scale is not None:
        g = torch.softmax(x, dim=-1)
        g.scatter_add_(1, safe[:, None], -torch.ones_like(lse)[:, None])
        scale = valid.float() if row_scale is None else valid.float() * row_scale
        logits.zero_()
        logits[:, :V].copy_(g * scale[:, None])


def _device_guard(t: torch.Tensor):
    return torch.cuda.device(t.device) if t.is_cuda else contextlib.nullcontext()


def _no_autocast(t: torch.Tensor):
    return torch.autocast(t.device.type, enabled=False)


class _LinearCE(torch.autograd.Function):
    """The body runs with autocast OFF on operands cast once to the compute dtype: ``out=`` / in-place GEMMs are not
    autocast-aware, so under autocast bf16 logits would otherwise meet fp32 weights."""

    @staticmethod
    def forward(ctx, h, weight, labels, ignore_index: int, chunk: int):
        N, H = h.shape
        V = weight.shape[0]
        cd = _compute_dtype(h)
        hq = h if h.dtype == cd else h.to(cd)
        wp = _pad_rows(weight if weight.dtype == cd else weight.to(cd))
        losses = torch.empty(N, dtype=torch.float32, device=h.device)
        n_valid = (labels != ignore_index).sum().clamp_(min=1).to(torch.float32)
        need_grad = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        dh = torch.empty_like(hq) if need_grad else None
        dwp = torch.empty_like(wp) if need_grad else None              # written with beta = 0 by the first chunk
        with _device_guard(h), _no_autocast(h):
            for s in range(0, N, chunk):
                e = min(N, s + chunk)
                hc = hq[s:e]
                logits = hc @ wp.t()                                   # [c, Vp]
                _ce_kernel(logits, labels[s:e], losses[s:e], V, ignore_index, need_grad)
                if need_grad:
                    torch.mm(logits, wp, out=dh[s:e])                  # d h   (unscaled)
                    if s == 0:
                        torch.mm(logits.t(), hc, out=dwp)              # d W   (unscaled)
                    else:
                        dwp.addmm_(logits.t(), hc)                     # beta = 1 epilogue
        loss = losses.sum() / n_valid
        if need_grad:
            if N == 0:
                dwp.zero_()
            ctx.save_for_backward(dh, dwp, n_valid)
            ctx.V, ctx.h_dtype, ctx.w_dtype = V, h.dtype, weight.dtype
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        dh, dwp, n_valid = ctx.saved_tensors
        scale = (grad_out.float() / n_valid)
        dh.mul_(scale.to(dh.dtype))
        dw = dwp[:ctx.V]
        if dw.dtype == ctx.w_dtype:
            dw.mul_(scale.to(dw.dtype))
        else:                                                          # fp32 master weights under autocast
            dw = dw.to(ctx.w_dtype).mul_(scale.to(ctx.w_dtype))
        return (dh if dh.dtype == ctx.h_dtype else dh.to(ctx.h_dtype)), dw, None, None, None


class _LinearLogProbs(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, weight, labels, ignore_index: int, chunk: int):
        N, H = h.shape
        V = weight.shape[0]
        cd = _compute_dtype(h)
        hq = h if h.dtype == cd else h.to(cd)
        ctx.save_for_backward(h, weight, labels)
        ctx.ignore_index, ctx.chunk, ctx.cd = ignore_index, chunk, cd
        if _FUSED_FWD and ignore_index < 0 and N > 0 and lm_head.supported(hq, weight, cd):
            # forward half on the tcgen05 kernel (no logits block, no padded head); backward recomputes blocks below
            with _no_autocast(h):
                return lm_head.lm_head_log_probs(hq, weight if weight.dtype == cd else weight.to(cd), labels, want_lse=False)[0]
        wp = _pad_rows(weight if weight.dtype == cd else weight.to(cd))
        losses = torch.empty(N, dtype=torch.float32, device=h.device)
        with _device_guard(h), _no_autocast(h):
            for s in range(0, N, chunk):
                e = min(N, s + chunk)
                logits = hq[s:e] @ wp.t()
                _ce_kernel(logits, labels[s:e], losses[s:e], V, ignore_index, False)
        return -losses

    @staticmethod
    def backward(ctx, grad_logp):
        h, weight, labels = ctx.saved_tensors
        N, H = h.shape
        V = weight.shape[0]
        cd = ctx.cd
        hq = h if h.dtype == cd else h.to(cd)
        wp = _pad_rows(weight if weight.dtype == cd else weight.to(cd))
        dh = torch.empty_like(hq)
        dwp = torch.empty_like(wp)
        # logp = -ce  =>  d logits = (softmax - onehot) * (-grad_logp)
        row_scale = (-grad_logp).float().contiguous()
        with _device_guard(h), _no_autocast(h):
            for s in range(0, N, ctx.chunk):
                e = min(N, s + ctx.chunk)
                hc = hq[s:e]
                logits = hc @ wp.t()
                _ce_kernel(logits, labels[s:e], None, V, ctx.ignore_index, False, row_scale=row_scale[s:e])
                torch.mm(logits, wp, out=dh[s:e])
                if s == 0:
                    torch.mm(logits.t(), hc, out=dwp)
                else:
                    dwp.addmm_(logits.t(), hc)
            if N == 0:
                dwp.zero_()
        dw = dwp[:V]
        return (dh if dh.dtype == h.dtype else dh.to(h.dtype)), (dw if dw.dtype == weight.dtype else dw.to(weight.dtype)), \
            None, None, None


def _fused_no_grad(h: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int):
    """(hq, wq) when this call needs no gradient and the tcgen05 LM-head kernel (``ops/lm_head.py``) applies:
    log-probabilities straight out of TMEM, no logits block, no padded copy of the head.  Else None."""
    if torch.is_grad_enabled() and (h.requires_grad or weight.requires_grad):
        return None
    if ignore_index >= 0 or h.shape[0] == 0 or not (h.is_cuda or lm_head._EMULATE):
        return None
    cd = _compute_dtype(h)
    if not lm_head.supported(h, weight, cd):
        return None
    return (h if h.dtype == cd else h.to(cd)), (weight if weight.dtype == cd else weight.to(cd))


def linear_cross_entropy(h: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                         ignore_index: int = IGNORE_INDEX, chunk: Optional[int] = None) -> torch.Tensor:
    """h: [N, H], weight: [V, H], labels: [N] int64 → scalar mean CE over non-ignored rows."""
    fused = _fused_no_grad(h, weight, labels, ignore_index)
    if fused is not None:
        logp, _ = lm_head.lm_head_log_probs(fused[0], fused[1], labels, want_lse=False)
        return -logp.sum() / (labels != ignore_index).sum().clamp_(min=1).to(torch.float32)
    if _kernel_ok(h):
        return _LinearCE.apply(h.contiguous(), weight, labels.contiguous(), ignore_index, chunk or _DEFAULT_CHUNK)
    # CPU / fallback: still chunked so memory stays bounded
    chunk = chunk or _DEFAULT_CHUNK
    total = h.new_zeros((), dtype=torch.float32)
    for s in range(0, h.shape[0], chunk):
        logits = (h[s:s + chunk] @ weight.t()).float()
        total = total + F.cross_entropy(logits, labels[s:s + chunk], ignore_index=ignore_index, reduction="sum")
    return total / (labels != ignore_index).sum().clamp(min=1)


def linear_log_probs(h: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                     ignore_index: int = IGNORE_INDEX, chunk: Optional[int] = None) -> torch.Tensor:
    """Per-token log p(label) as fp32 [N]; ignored positions give 0."""
    fused = _fused_no_grad(h, weight, labels, ignore_index)
    if fused is not None:
        return lm_head.lm_head_log_probs(fused[0], fused[1], labels, want_lse=False)[0]
    if _kernel_ok(h):
        return _LinearLogProbs.apply(h.contiguous(), weight, labels.contiguous(), ignore_index,
                                     chunk or _DEFAULT_CHUNK)
    chunk = chunk or _DEFAULT_CHUNK
    outs = []
    for s in range(0, h.shape[0], chunk):
        logits = (h[s:s + chunk] @ weight.t()).float()
        outs.append(-F.cross_entropy(logits, labels[s:s + chunk], ignore_index=ignore_index, reduction="none"))
    return torch.cat(outs)


def cross_entropy_rows(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = IGNORE_INDEX) -> torch.Tensor:
    """Per-row CE of an existing [rows, vocab] logits tensor (not modified); test/eval helper."""
    if logits.is_cuda and _kernel_ok(logits) and logits.stride(-1) == 1 and logits.stride(0) % 8 == 0 and logits.data_ptr() % 16 == 0:
        out = torch.empty(logits.shape[0], dtype=torch.float32, device=logits.device)
        with torch.cuda.device(logits.device):
            _cabi.check(_cabi.lib().dlion_ce_fwd_bwd(logits.data_ptr(), labels.contiguous().data_ptr(), out.data_ptr(),
                                                     logits.shape[0], logits.shape[1], logits.stride(0), ignore_index,
                                                     1.0, 0, _cabi.dtype_code(logits.dtype),
                                                     _cabi.stream_ptr(logits.device)), "ce_fwd")
        return out
    return F.cross_entropy(logits.float(), labels, ignore_index=ignore_index, reduction="none")
