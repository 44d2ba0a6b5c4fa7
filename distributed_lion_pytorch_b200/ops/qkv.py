"""Splitting a packed qkv projection into attention's q, k, v — and, in backward, packing dq, dk, dv back into
d(qkv) while reducing it over rows, so that the stack/cat kernel of ``unbind``'s backward and the bias-gradient
pass of ``c_attn`` are one kernel (csrc/colsum.cu: ``QkvPackSource``).

reference: HF ``GPT2Attention`` splits with ``.split(self.split_size, dim=2)`` and permutes; its backward is three
strided copies, two adds and — for the Conv1D bias — a separate ``grad_output.sum(0)``, reached from
accelerator.backward (async_trainer.py:33).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi, require_library


def _kernel_ok(t: torch.Tensor, head_dim: int) -> bool:
    return t.is_cuda and require_library() and head_dim % 8 == 0 \
        and t.dtype in (torch.float32, torch.bfloat16, torch.float16)


class _SplitQKV(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, n_head: int, bias):
        B, S, E3 = qkv.shape
        D = E3 // 3 // n_head
        v5 = qkv.view(B, S, 3, n_head, D)
        ctx.shape = (B, S, n_head, D)
        ctx.bias = bias if (bias is not None and bias.requires_grad) else None
        return tuple(v5[:, :, i].transpose(1, 2) for i in range(3))       # three [B, H, S, D] views, no copy

    @staticmethod
    def backward(ctx, dq, dk, dv):
        B, S, H, D = ctx.shape
        ref = next(g for g in (dq, dk, dv) if g is not None)
        grads = []
        for g in (dq, dk, dv):
            if g is None:
                g = torch.zeros(B, H, S, D, dtype=ref.dtype, device=ref.device)
            if g.stride(-1) != 1 or any(st % 8 for st in g.stride()[:3]) or g.data_ptr() % 16:
                g = g.contiguous()
            grads.append(g)
        dqkv = torch.empty(B, S, 3 * H * D, dtype=ref.dtype, device=ref.device)
        cols = 3 * H * D
        L = _cabi.lib()
        from . import linear as _lin
        if int(L.dlion_colsum_tiles(cols)) > _lin._MAX_TILES:
            torch.stack([g.transpose(1, 2) for g in grads], dim=2, out=dqkv.view(B, S, 3, H, D))
            return dqkv, None, None
        strides = (C.c_int64 * 9)(*[g.stride(i) for g in grads for i in range(3)])
        sums = torch.empty(cols, dtype=ref.dtype, device=ref.device)
        partial = torch.empty(int(L.dlion_colsum_parts(B * S, cols)), cols, dtype=torch.float32, device=ref.device)
        with torch.cuda.device(ref.device):
            _cabi.check(L.dlion_qkv_pack_colsum(grads[0].data_ptr(), grads[1].data_ptr(), grads[2].data_ptr(), strides,
                                                dqkv.data_ptr(), B, S, H, D, partial.data_ptr(),
                                                _lin._tickets(ref.device).data_ptr(), sums.data_ptr(), 0,
                                                _cabi.dtype_code(ref.dtype), _cabi.stream_ptr(ref.device)),
                        "qkv_pack_colsum")
        if ctx.bias is not None and ctx.bias.dtype == sums.dtype:
            _lin.attach_column_sums(dqkv, ctx.bias, sums)       # c_attn's bias gradient rides with d(qkv)
        return dqkv, None, None


def split_qkv(qkv: torch.Tensor, n_head: int, bias: Optional[torch.Tensor] = None):
    """[B, S, 3*E] -> q, k, v as [B, H, S, D] views.  ``bias`` is the bias of the projection that produced ``qkv``
    (its gradient comes out of the backward's packing pass)."""
    B, S, E3 = qkv.shape
    D = E3 // 3 // n_head
    if torch.is_grad_enabled() and qkv.requires_grad and _kernel_ok(qkv, D) and qkv.is_contiguous():
        return _SplitQKV.apply(qkv, n_head, bias)
    # unbind (backward = one stack) rather than three selects (backward = 3 zero-fills + 3 copies + 2 adds)
    return tuple(t.transpose(1, 2) for t in qkv.view(B, S, 3, n_head, D).unbind(2))
