"""Token + position embedding with a one-pass forward and a deterministic, atomics-free backward that adds straight
into the existing gradient buffer (csrc/embedding.cu).

reference: HF ``GPT2Model`` does ``wte(input_ids) + wpe(position_ids)`` (two gathers and an add); autograd's
``embedding_dense_backward`` sorts the ids, runs two segment kernels into a zero-filled ``[vocab, hidden]`` tensor
and AccumulateGrad adds that dense tensor into ``.grad`` — reached from accelerator.backward (async_trainer.py:33).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _cabi, mark_grad_written, require_library


def _kernel_ok(w: torch.Tensor) -> bool:
    return w.is_cuda and require_library() and w.shape[1] % 8 == 0 and w.is_contiguous() \
        and w.dtype in (torch.float32, torch.bfloat16, torch.float16)


class _Embed(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, wte, wpe):
        B, S = ids.shape
        V, H = wte.shape
        idc = ids.contiguous()
        out = torch.empty(B, S, H, dtype=wte.dtype, device=wte.device)
        with torch.cuda.device(wte.device):
            _cabi.check(_cabi.lib().dlion_embed_fwd(idc.data_ptr(), wte.data_ptr(), _cabi.ptr(wpe), out.data_ptr(), B * S, S,
                                                    H, V, _cabi.dtype_code(wte.dtype), _cabi.stream_ptr(wte.device)),
                        "embed_fwd")
        ctx.save_for_backward(idc)
        ctx.wte, ctx.wpe = wte, wpe          # parameters (not saved tensors: the backward adds into their .grad)
        return out

    @staticmethod
    def backward(ctx, dh):
        from .linear import _can_accumulate
        (ids,) = ctx.saved_tensors
        wte, wpe = ctx.wte, ctx.wpe
        B, S = ids.shape
        V, H = wte.shape
        dhc = dh.contiguous()
        d_wte = d_wpe = None
        if ctx.needs_input_grad[1]:
            sorted_ids, perm = torch.sort(ids.view(-1), stable=True)
            acc = _can_accumulate(wte, dhc) and wte.grad.shape == wte.shape
            target = wte.grad if acc else torch.zeros_like(wte)
            if acc:
                mark_grad_written(wte)
            with torch.cuda.device(dh.device):
                _cabi.check(_cabi.lib().dlion_embed_bwd_sorted(dhc.data_ptr(), sorted_ids.data_ptr(), perm.data_ptr(),
                                                               target.data_ptr(), B * S, H, V, int(acc),
                                                               _cabi.dtype_code(dh.dtype), _cabi.stream_ptr(dh.device)),
                            "embed_bwd_sorted")
            d_wte = None if acc else target
        if wpe is not None and ctx.needs_input_grad[2]:
            pos = dhc.sum(0, dtype=torch.float32).to(dh.dtype)            # [S, H]: every batch row uses positions 0..S-1
            if _can_accumulate(wpe, dhc) and wpe.grad.shape == wpe.shape:
                mark_grad_written(wpe)
                wpe.grad[:S].add_(pos)
            else:
                d_wpe = torch.zeros_like(wpe)
                d_wpe[:S] = pos
        return None, d_wte, d_wpe


def embed(ids: torch.Tensor, wte: torch.Tensor, wpe: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``wte[ids] + wpe[arange(S)]`` for ``ids`` [B, S] (``wpe`` optional)."""
    if ids.dim() == 2 and _kernel_ok(wte) and (wpe is None or (_kernel_ok(wpe) and wpe.dtype == wte.dtype
                                                                and wpe.shape[1] == wte.shape[1]
                                                                and wpe.shape[0] >= ids.shape[1])):
        if torch.is_grad_enabled() and (wte.requires_grad or (wpe is not None and wpe.requires_grad)):
            return _Embed.apply(ids, wte, wpe)
        return _Embed.forward(_NoCtx(), ids, wte, wpe)
    h = torch.nn.functional.embedding(ids, wte)
    if wpe is not None:
        h = h + wpe[: ids.shape[-1]]
    return h


class _NoCtx:
    def save_for_backward(self, *a):
        pass
