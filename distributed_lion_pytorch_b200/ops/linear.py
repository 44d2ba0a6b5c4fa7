"""Linear layers whose backward (a) accumulates the weight gradient inside the GEMM
(``grad.addmm_``: beta = 1 epilogue, no separate add kernel per parameter per micro-batch) when
``weight.grad`` already exists — which it always does once ``Lion`` has aliased the gradients to
its flat bucket — and (b) computes the bias gradient with the column-sum kernel
(csrc/colsum.cu) instead of ATen's generic reduction.

reference: HF GPT-2 ``Conv1D`` / ``nn.Linear`` backward + autograd's AccumulateGrad, reached
from ``accelerator.backward(loss)`` (async_trainer.py:33): one extra read-modify-write pass over
every parameter per micro-batch (8 per optimizer step in the README recipe).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _cabi, require_library

_FUSE_WGRAD = os.environ.get("DLION_FUSE_WGRAD", "1") == "1"


def column_sum(dy2: torch.Tensor, out_dtype: torch.dtype) -> torch.Tensor:
    """[rows, cols] -> [cols] (fp32 accumulation)."""
    rows, cols = dy2.shape
    if dy2.is_cuda and require_library() and cols % 8 == 0 and dy2.is_contiguous() and dy2.data_ptr() % 16 == 0 \
            and dy2.dtype in (torch.float32, torch.bfloat16, torch.float16):
        L = _cabi.lib()
        parts = int(L.dlion_colsum_parts(rows, cols))
        partial = torch.empty(parts, cols, dtype=torch.float32, device=dy2.device)
        out = torch.empty(cols, dtype=out_dtype, device=dy2.device)
        with torch.cuda.device(dy2.device):
            st = _cabi.stream_ptr(dy2.device)
            _cabi.check(L.dlion_colsum(dy2.data_ptr(), rows, cols, partial.data_ptr(), _cabi.dtype_code(dy2.dtype), st),
                        "colsum")
            _cabi.check(L.dlion_norm_bwd_finalize(partial.data_ptr(), None, out.data_ptr(), None, parts, cols,
                                                  _cabi.dtype_code(out_dtype), st), "colsum_finalize")
        return out
    return dy2.sum(0).to(out_dtype)


def _can_accumulate(param: torch.Tensor, like: torch.Tensor) -> bool:
    g = param.grad
    return _FUSE_WGRAD and g is not None and g.dtype == like.dtype and g.is_contiguous() and g.device == like.device \
        and not g.requires_grad


class _FusedLinear(torch.autograd.Function):
    """y = x @ W (+ b) with W stored [in, out] (``in_out=True``, HF Conv1D) or [out, in] (nn.Linear)."""

    @staticmethod
    def forward(ctx, x, weight, bias, in_out: bool):
        x2 = x.reshape(-1, x.shape[-1])
        w = weight if in_out else weight.t()
        # the output is allocated in its final shape (not returned as a view of a 2-D result) so
        # that downstream in-place kernels (RoPE) may own it
        out = x.new_empty(*x.shape[:-1], w.shape[1])
        out2 = out.view(-1, w.shape[1])
        if bias is not None:
            torch.addmm(bias, x2, w, out=out2)
        else:
            torch.mm(x2, w, out=out2)
        ctx.save_for_backward(x2, weight, bias)
        ctx.in_out = in_out
        ctx.x_shape = x.shape
        return out

    @staticmethod
    def backward(ctx, dy):
        x2, weight, bias = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = (dy2 @ (weight.t() if ctx.in_out else weight)).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            if _can_accumulate(weight, dy2):
                if ctx.in_out:
                    weight.grad.addmm_(x2.t(), dy2)          # dW += x^T dy, accumulated in the GEMM epilogue
                else:
                    weight.grad.addmm_(dy2.t(), x2)
            else:
                dw = x2.t() @ dy2 if ctx.in_out else dy2.t() @ x2
        if bias is not None and ctx.needs_input_grad[2]:
            db = column_sum(dy2, bias.dtype)
            if _can_accumulate(bias, db):
                bias.grad.add_(db)
                db = None
        return dx, dw, db, None


def fused_linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], in_out: bool) -> torch.Tensor:
    if x.is_cuda and torch.is_grad_enabled() and (weight.requires_grad or x.requires_grad):
        return _FusedLinear.apply(x, weight, bias, in_out)
    if not in_out:
        return torch.nn.functional.linear(x, weight, bias)
    x2 = x.reshape(-1, x.shape[-1])
    w = weight if in_out else weight.t()
    y = torch.addmm(bias, x2, w) if bias is not None else x2 @ w
    return y.view(*x.shape[:-1], y.shape[-1])
