"""N3 — fused residual-add + LayerNorm / RMSNorm (forward and backward kernels).

``add_layer_norm(x, residual, weight, bias, eps)`` returns ``(y, new_residual)`` where
``new_residual = x + residual`` (``x`` itself when ``residual is None``) and
``y = LayerNorm(new_residual)``; ``add_rms_norm`` is the Llama flavour.  On CUDA both passes
run in ``csrc/norm.cu``; on CPU the same math runs as plain PyTorch so the models work
under gloo.  reference: HF GPT2Block / LlamaDecoderLayer via async_trainer.py:24.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _cabi, require_library


def _use_kernel(x: torch.Tensor) -> bool:
    return x.is_cuda and require_library() and x.shape[-1] % 8 == 0 and x.shape[-1] <= 8192 \
        and x.dtype in (torch.float32, torch.bfloat16, torch.float16)


class _FusedNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, eps: float, rms: bool, drop_p: float, seed: int, x_bias=None):
        cols = x.shape[-1]
        xc = x.contiguous()
        rc = residual.contiguous() if residual is not None else None
        rows = xc.numel() // cols
        y = torch.empty_like(xc)
        summed = torch.empty_like(xc) if (rc is not None or drop_p > 0) else xc
        mean = torch.empty(rows, dtype=torch.float32, device=x.device) if not rms else None
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().dlion_norm_fwd(
                xc.data_ptr(), _cabi.ptr(rc), weight.data_ptr(), _cabi.ptr(bias), y.data_ptr(),
                summed.data_ptr() if summed is not xc else None, _cabi.ptr(mean), rstd.data_ptr(), rows, cols,
                float(eps), int(rms), float(drop_p), int(seed), _cabi.dtype_code(x.dtype),
                _cabi.stream_ptr(x.device)), "norm_fwd")
        ctx.save_for_backward(summed, weight, mean, rstd, bias)
        ctx.rms, ctx.has_res, ctx.has_bias = rms, residual is not None, bias is not None
        ctx.drop_p, ctx.seed = float(drop_p), int(seed)
        # the bias of the linear layer that produced x (if any): its gradient is the column sum of OUR dx
        ctx.x_bias = x_bias if (x_bias is not None and x_bias.requires_grad and x_bias.shape == (cols,)
                                and x_bias.dtype == x.dtype
                                and _cabi.lib().dlion_norm_bwd_dxsum_supported(cols, int(rms))) else None
        ctx.set_materialize_grads(False)      # an unused residual output costs nothing in backward
        return y, summed

    @staticmethod
    def backward(ctx, dy, dsum):
        summed, weight, mean, rstd, bias = ctx.saved_tensors
        cols = summed.shape[-1]
        rows = summed.numel() // cols
        if dy is None:      # only the residual stream was consumed downstream
            dy = torch.zeros_like(summed)
        dyc = dy.contiguous()
        dsc = dsum.contiguous() if dsum is not None else None
        L = _cabi.lib()
        xb = ctx.x_bias if ctx.needs_input_grad[0] else None
        parts = int(L.dlion_norm_bwd_dxsum_parts(rows, cols) if xb is not None else L.dlion_norm_bwd_parts(rows, cols))
        dx = torch.empty_like(summed)
        drop = ctx.drop_p > 0
        dres = torch.empty_like(summed) if (drop and ctx.has_res) else None
        dwp = torch.empty(parts, cols, dtype=torch.float32, device=dy.device)
        dbp = torch.empty(parts, cols, dtype=torch.float32, device=dy.device) if ctx.has_bias else None
        # weight / bias gradients go straight into existing .grad buffers (the flat-bucket views once Lion has
        # aliased them): the fold kernel adds in fp32 and rounds once, and autograd has no add kernel left to run
        from .linear import _can_accumulate
        acc_w = ctx.needs_input_grad[2] and _can_accumulate(weight, weight)
        acc_b = ctx.has_bias and ctx.needs_input_grad[3] and _can_accumulate(bias, bias)
        from . import mark_grad_written
        if acc_w:
            mark_grad_written(weight)
        if acc_b:
            mark_grad_written(bias)
        dw = weight.grad if acc_w else torch.empty_like(weight)
        db = (bias.grad if acc_b else torch.empty_like(weight)) if ctx.has_bias else None
        dxsp = torch.empty(parts, cols, dtype=torch.float32, device=dy.device) if xb is not None else None
        dxsum = torch.empty(cols, dtype=weight.dtype, device=dy.device) if xb is not None else None
        with torch.cuda.device(dy.device):
            st = _cabi.stream_ptr(dy.device)
            _cabi.check(L.dlion_norm_bwd_dxsum(dyc.data_ptr(), _cabi.ptr(dsc), summed.data_ptr(), weight.data_ptr(),
                                               _cabi.ptr(mean), rstd.data_ptr(), dx.data_ptr(), _cabi.ptr(dres),
                                               dwp.data_ptr(), _cabi.ptr(dbp), _cabi.ptr(dxsp), rows, cols, int(ctx.rms),
                                               ctx.drop_p, ctx.seed, _cabi.dtype_code(dy.dtype), st), "norm_bwd")
            _cabi.check(L.dlion_norm_bwd_finalize_dxsum(dwp.data_ptr(), _cabi.ptr(dbp), _cabi.ptr(dxsp), dw.data_ptr(),
                                                        _cabi.ptr(db), _cabi.ptr(dxsum), parts, cols,
                                                        int(acc_w) | (int(acc_b) << 1), _cabi.dtype_code(weight.dtype),
                                                        st), "norm_bwd_finalize")
        if xb is not None:
            from .linear import attach_column_sums
            attach_column_sums(dx, xb, dxsum)        # rides with dx to the backward of the linear that produced x
        if ctx.has_res:
            g_res = dres if drop else dx
        else:
            g_res = None
        return dx, g_res, (None if acc_w else dw), (None if acc_b else db), None, None, None, None, None


def _draw_seed() -> int:
    """Seed for the in-kernel dropout mask, drawn from torch's CPU generator: cheap, reproducible
    under ``set_seed`` / checkpointed RNG state, and replayed identically by activation
    checkpointing (which restores the generator before recomputing)."""
    return int(torch.randint(0, 2 ** 62, (1,)).item())


def _ref_norm(x, residual, weight, bias, eps, rms, drop_p=0.0):
    if drop_p > 0:
        x = torch.nn.functional.dropout(x, drop_p, True)
    s = x if residual is None else x + residual
    if rms:
        h = s.float()
        h = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps)
        return weight * h.to(s.dtype), s
    return torch.nn.functional.layer_norm(s, (s.shape[-1],), weight, bias, eps), s


def add_layer_norm(x: torch.Tensor, residual: Optional[torch.Tensor], weight: torch.Tensor,
                   bias: Optional[torch.Tensor], eps: float = 1e-5, dropout_p: float = 0.0,
                   training: bool = False, x_bias: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``s = dropout(x) + residual``; returns ``(LayerNorm(s), s)``.  One kernel on CUDA.

    ``x_bias``: the bias parameter of the linear layer whose output ``x`` is (GPT-2: ``attn.c_proj`` in front of
    ``ln_2``, ``mlp.c_proj`` in front of the next ``ln_1`` / ``ln_f``).  The backward kernel then also reduces dx
    over rows — that IS that bias's gradient — and hands it over with dx, so the linear's backward makes no pass
    of its own over dx."""
    p = float(dropout_p) if training else 0.0
    if _use_kernel(x):
        return _FusedNorm.apply(x, residual, weight, bias, eps, False, p, _draw_seed() if p > 0 else 0, x_bias)
    return _ref_norm(x, residual, weight, bias, eps, False, p)


def add_rms_norm(x: torch.Tensor, residual: Optional[torch.Tensor], weight: torch.Tensor,
                 eps: float = 1e-5, dropout_p: float = 0.0, training: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    p = float(dropout_p) if training else 0.0
    if _use_kernel(x):
        return _FusedNorm.apply(x, residual, weight, None, eps, True, p, _draw_seed() if p > 0 else 0)
    return _ref_norm(x, residual, weight, None, eps, True, p)
