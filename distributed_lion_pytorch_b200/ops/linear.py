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

from . import _cabi, mark_grad_written, require_library

_FUSE_WGRAD = os.environ.get("DLION_FUSE_WGRAD", "1") == "1"
_FUSE_COLSUM = os.environ.get("DLION_FUSE_COLSUM", "1") == "1"


_TICKETS: dict = {}
_MAX_TILES = 64


def _tickets(device: torch.device) -> torch.Tensor:
    """Zeroed uint32 counters for the one-launch column sum; the kernel leaves them zeroed, and every call on a
    device goes through the one stream autograd uses there, so a single cached buffer is enough."""
    t = _TICKETS.get(device)
    if t is None:
        t = _TICKETS[device] = torch.zeros(_MAX_TILES, dtype=torch.int32, device=device)
    return t


def _kernel_ok(dy2: torch.Tensor) -> bool:
    return dy2.is_cuda and require_library() and dy2.shape[1] % 8 == 0 and dy2.is_contiguous() \
        and dy2.data_ptr() % 16 == 0 and dy2.dtype in (torch.float32, torch.bfloat16, torch.float16) and dy2.shape[0] > 0


def column_sum(dy2: torch.Tensor, out_dtype: torch.dtype, accumulate_into: Optional[torch.Tensor] = None):
    """[rows, cols] -> [cols] (fp32 accumulation).  With ``accumulate_into`` (a [cols] gradient buffer) the sums are
    added to it — inside the same kernel when dtypes allow — and ``None`` is returned."""
    rows, cols = dy2.shape
    if _kernel_ok(dy2):
        L = _cabi.lib()
        parts = int(L.dlion_colsum_parts(rows, cols))
        partial = torch.empty(parts, cols, dtype=torch.float32, device=dy2.device)
        acc = accumulate_into
        one_launch = out_dtype == dy2.dtype and int(L.dlion_colsum_tiles(cols)) <= _MAX_TILES and _FUSE_COLSUM \
            and (acc is None or (acc.dtype == out_dtype and acc.is_contiguous() and acc.shape == (cols,)))
        with torch.cuda.device(dy2.device):
            st = _cabi.stream_ptr(dy2.device)
            if one_launch:
                out = acc if acc is not None else torch.empty(cols, dtype=out_dtype, device=dy2.device)
                _cabi.check(L.dlion_colsum_fused(dy2.data_ptr(), rows, cols, partial.data_ptr(),
                                                 _tickets(dy2.device).data_ptr(), out.data_ptr(), int(acc is not None),
                                                 _cabi.dtype_code(dy2.dtype), st), "colsum_fused")
                return None if acc is not None else out
            out = torch.empty(cols, dtype=out_dtype, device=dy2.device)
            _cabi.check(L.dlion_colsum(dy2.data_ptr(), rows, cols, partial.data_ptr(), _cabi.dtype_code(dy2.dtype), st),
                        "colsum")
            _cabi.check(L.dlion_norm_bwd_finalize(partial.data_ptr(), None, out.data_ptr(), None, parts, cols,
                                                  _cabi.dtype_code(out_dtype), st), "colsum_finalize")
    else:
        out = dy2.sum(0).to(out_dtype)
    if accumulate_into is not None:
        accumulate_into.add_(out)
        return None
    return out


HANDOFFS = 0      # bias gradients that arrived ready-made with the incoming gradient (tests read this)


def attach_column_sums(grad: torch.Tensor, bias: torch.Tensor, sums: torch.Tensor) -> None:
    """Producer side of the bias-gradient hand-off: a backward kernel that already has ``grad`` in registers (the
    norm backward for the linear in front of it, the qkv packer for ``c_attn``) also reduced it over rows; the sums
    travel WITH the gradient tensor, as a Python attribute, to the backward of the linear that owns ``bias``.
    If autograd hands that linear a different tensor object (it had to add several gradients, say), the attribute
    is simply not there and the linear reduces ``grad`` itself — never both."""
    grad._dlion_colsum = (id(bias), sums)


def _take_column_sums(dy: torch.Tensor, bias: torch.Tensor) -> Optional[torch.Tensor]:
    tag = getattr(dy, "_dlion_colsum", None)
    if tag is None or tag[0] != id(bias) or tag[1].shape != bias.shape or tag[1].dtype != bias.dtype:
        return None
    global HANDOFFS
    HANDOFFS += 1
    return tag[1]


def bias_grad(dy2: torch.Tensor, bias: torch.Tensor, dy: Optional[torch.Tensor] = None):
    """Gradient of a bias added to every row of the tensor whose gradient is ``dy2``: accumulated straight into
    ``bias.grad`` when that buffer exists (returns ``None``), returned otherwise.  ``dy`` is the gradient tensor as
    autograd delivered it (``dy2`` may be a reshaped view): if its producer attached the column sums, no pass over
    ``dy2`` is made at all."""
    ready = _take_column_sums(dy, bias) if dy is not None else None
    if ready is not None:
        if _can_accumulate(bias, dy2) and bias.grad.shape == ready.shape:
            mark_grad_written(bias)
            bias.grad.add_(ready)
            return None
        return ready
    if _can_accumulate(bias, dy2) and bias.grad.shape == (dy2.shape[1],):
        mark_grad_written(bias)
        return column_sum(dy2, bias.dtype, accumulate_into=bias.grad)
    return column_sum(dy2, bias.dtype)


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
                mark_grad_written(weight)
                if ctx.in_out:
                    weight.grad.addmm_(x2.t(), dy2)          # dW += x^T dy, accumulated in the GEMM epilogue
                else:
                    weight.grad.addmm_(dy2.t(), x2)
            else:
                dw = x2.t() @ dy2 if ctx.in_out else dy2.t() @ x2
        if bias is not None and ctx.needs_input_grad[2]:
            db = bias_grad(dy2, bias, dy)                    # ready-made, or one launch folded and added into bias.grad
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
