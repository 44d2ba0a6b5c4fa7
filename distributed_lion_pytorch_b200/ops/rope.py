"""N4 — rotary embedding, in place on q and k (forward) and on dq, dk (backward).
reference: HF LlamaAttention.apply_rotary_pos_emb via sft_llama2.py:147 / dpo_llama2.py:133."""
from __future__ import annotations

import os
from typing import Tuple

import torch

from . import _cabi, require_library

_EMULATE = os.environ.get("DLION_EMULATE_KERNELS") == "1"   # CPU stand-ins, to test the autograd plumbing


def rope_tables(seq_len: int, head_dim: int, theta: float = 10000.0, device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """fp32 cos/sin tables of shape [seq_len, head_dim/2]."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32, device=device) / head_dim))
    ang = torch.outer(torch.arange(seq_len, dtype=torch.float32, device=device), inv)
    return ang.cos().contiguous(), ang.sin().contiguous()


def _rotate_(x: torch.Tensor, cos, sin, seq_len: int, head_dim: int, inverse: bool) -> None:
    """torch stand-in for the kernel: x [..., S, H*D] rotated in place."""
    half = head_dim // 2
    xv = x.view(-1, seq_len, x.shape[-1] // head_dim, head_dim)
    c = cos[:seq_len].view(1, seq_len, 1, half)
    s = sin[:seq_len].view(1, seq_len, 1, half) * (-1.0 if inverse else 1.0)
    x1, x2 = xv[..., :half].float(), xv[..., half:].float()
    o1, o2 = x1 * c - x2 * s, x2 * c + x1 * s
    xv[..., :half].copy_(o1)
    xv[..., half:].copy_(o2)


def _launch(q, k, cos, sin, seq_len, head_dim, inverse):
    if not q.is_cuda:
        _rotate_(q, cos, sin, seq_len, head_dim, inverse)
        _rotate_(k, cos, sin, seq_len, head_dim, inverse)
        return
    tokens = q.numel() // q.shape[-1]
    with torch.cuda.device(q.device):
        _cabi.check(_cabi.lib().dlion_rope(q.data_ptr(), k.data_ptr(), cos.data_ptr(), sin.data_ptr(), tokens, seq_len,
                                           q.shape[-1] // head_dim, k.shape[-1] // head_dim, head_dim, q.shape[-1],
                                           k.shape[-1], int(inverse), _cabi.dtype_code(q.dtype),
                                           _cabi.stream_ptr(q.device)), "rope")


class _Rope(torch.autograd.Function):
    """Takes the projection outputs themselves ([B, S, H*D], not views of them) so marking
    them dirty never runs into autograd's in-place-on-a-view restrictions."""

    @staticmethod
    def forward(ctx, q, k, cos, sin, seq_len: int, head_dim: int):
        _launch(q, k, cos, sin, seq_len, head_dim, False)
        ctx.mark_dirty(q, k)
        ctx.save_for_backward(cos, sin)
        ctx.seq_len, ctx.head_dim = seq_len, head_dim
        return q, k

    @staticmethod
    def backward(ctx, dq, dk):
        cos, sin = ctx.saved_tensors
        dq = dq.contiguous().clone()
        dk = dk.contiguous().clone()
        _launch(dq, dk, cos, sin, ctx.seq_len, ctx.head_dim, True)
        return dq, dk, None, None, None, None


def apply_rope(q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, head_dim: int):
    """q: [B, S, Hq*D], k: [B, S, Hk*D] — the contiguous outputs of the q/k projections.
    Rotated IN PLACE (the projection's backward never needs its own output) and returned."""
    seq_len = q.shape[-2]
    use = (q.is_cuda and require_library()) or (_EMULATE and not q.is_cuda)
    if use and q.is_contiguous() and k.is_contiguous() and head_dim % 16 == 0 and q._base is None and k._base is None:
        return _Rope.apply(q, k, cos, sin, seq_len, head_dim)
    return _rope_functional(q, cos, sin, seq_len, head_dim), _rope_functional(k, cos, sin, seq_len, head_dim)


def _rope_functional(x, cos, sin, seq_len, head_dim):
    half = head_dim // 2
    xv = x.reshape(-1, seq_len, x.shape[-1] // head_dim, head_dim)
    c = cos[:seq_len].view(1, seq_len, 1, half)
    s = sin[:seq_len].view(1, seq_len, 1, half)
    x1, x2 = xv[..., :half].float(), xv[..., half:].float()
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1).to(x.dtype).reshape(x.shape)
