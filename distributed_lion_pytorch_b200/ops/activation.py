"""MLP activations as single-pass kernels (csrc/activation.cu): GPT-2's tanh-form GELU and
Llama's SwiGLU gate.  reference: HF GPT2MLP / LlamaMLP reached via async_trainer.py:24."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _cabi, require_library


def _ok(*ts: torch.Tensor) -> bool:
    t0 = ts[0]
    return t0.is_cuda and require_library() and t0.dtype in (torch.float32, torch.bfloat16, torch.float16) \
        and t0.numel() % 8 == 0 and all(t.is_contiguous() and t.data_ptr() % 16 == 0 and t.dtype == t0.dtype
                                        and t.shape == t0.shape for t in ts)


class _GeluTanh(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().dlion_gelu_tanh_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _cabi.dtype_code(x.dtype),
                                                        _cabi.stream_ptr(x.device)), "gelu_fwd")
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().dlion_gelu_tanh_bwd(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), x.numel(),
                                                        _cabi.dtype_code(x.dtype), _cabi.stream_ptr(x.device)), "gelu_bwd")
        return dx


class _SwiGLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gate, up):
        y = torch.empty_like(gate)
        with torch.cuda.device(gate.device):
            _cabi.check(_cabi.lib().dlion_swiglu_fwd(gate.data_ptr(), up.data_ptr(), y.data_ptr(), gate.numel(),
                                                     _cabi.dtype_code(gate.dtype), _cabi.stream_ptr(gate.device)),
                        "swiglu_fwd")
        ctx.save_for_backward(gate, up)
        return y

    @staticmethod
    def backward(ctx, dy):
        gate, up = ctx.saved_tensors
        dy = dy.contiguous()
        dgate, dup = torch.empty_like(gate), torch.empty_like(up)
        with torch.cuda.device(gate.device):
            _cabi.check(_cabi.lib().dlion_swiglu_bwd(gate.data_ptr(), up.data_ptr(), dy.data_ptr(), dgate.data_ptr(),
                                                     dup.data_ptr(), gate.numel(), _cabi.dtype_code(gate.dtype),
                                                     _cabi.stream_ptr(gate.device)), "swiglu_bwd")
        return dgate, dup


def gelu_tanh(x: torch.Tensor) -> torch.Tensor:
    if _ok(x):
        return _GeluTanh.apply(x)
    return F.gelu(x, approximate="tanh")


def swiglu(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """silu(gate) * up"""
    if _ok(gate, up):
        return _SwiGLU.apply(gate, up)
    return F.silu(gate) * up
