"""Flat parameter / momentum / gradient buckets.

The reference walks parameters one tensor at a time in Python (distributed_lion.py:179-198:
148 tensors for GPT-2 small, 291 for Llama-2-7B, one blocking all_gather each).  Here every
``(param_group, dtype, device)`` owns three flat tensors ``p``, ``m``, ``g`` and each
parameter / ``exp_avg`` / ``.grad`` is a *view* into them, so one kernel launch covers the
whole group and "bucket flatten" degenerates to nothing in steady state.

Layout rules
------------
* every parameter starts at a multiple of ``ALIGN`` (=128) elements: 256 B for bf16, 512 B
  for fp32.  That keeps cuBLAS on its aligned kernels, makes every view a legal 16-byte
  vector / ``cp.async.bulk`` source, and keeps a 32-sign word from straddling two tensors;
* the bucket length is a multiple of ``TILE_PAD`` (=8192) elements so the fused kernel
  never sees a ragged tile.  Pad elements hold ``p = m = g = 0``; their sign bit is 0 and
  whatever the vote does to the pad is never read.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

ALIGN = 128        # elements
TILE_PAD = 8192    # elements; lcm of the fused kernel's tile sizes


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass
class Bucket:
    group_index: int
    dtype: torch.dtype
    device: torch.device
    params: List[torch.nn.Parameter]
    offsets: List[int]
    numel: int                      # padded length
    flat_p: torch.Tensor
    flat_m: torch.Tensor
    flat_g: torch.Tensor
    p_views: List[torch.Tensor] = field(default_factory=list)
    m_views: List[torch.Tensor] = field(default_factory=list)
    g_views: List[torch.Tensor] = field(default_factory=list)
    scratch: Dict[str, torch.Tensor] = field(default_factory=dict)   # wire buffers etc., owned by backends

    @property
    def payload(self) -> int:
        return sum(p.numel() for p in self.params)

    @property
    def wire_bytes(self) -> int:
        return self.numel // 8


def plan_offsets(numels: List[int]) -> (List[int], int):
    offs, cur = [], 0
    for n in numels:
        offs.append(cur)
        cur = _round_up(cur + n, ALIGN)
    return offs, max(_round_up(cur, TILE_PAD), TILE_PAD)


def build_bucket(group_index: int, params: List[torch.nn.Parameter], state: dict, *,
                 alias_params: bool = True) -> Bucket:
    """Allocate flat storage for ``params`` (same dtype/device), move the live values in and
    re-point ``p.data`` (and the ``exp_avg`` state, if present) at the views."""
    assert params
    dtype, device = params[0].dtype, params[0].device
    offs, total = plan_offsets([p.numel() for p in params])
    flat_p = torch.zeros(total, dtype=dtype, device=device)
    flat_m = torch.zeros(total, dtype=dtype, device=device)
    flat_g = torch.zeros(total, dtype=dtype, device=device)
    b = Bucket(group_index, dtype, device, list(params), offs, total, flat_p, flat_m, flat_g)
    for p, o in zip(params, offs):
        n = p.numel()
        pv = flat_p[o:o + n].view(p.shape)
        mv = flat_m[o:o + n].view(p.shape)
        gv = flat_g[o:o + n].view(p.shape)
        pv.copy_(p.data)
        st = state.get(p)
        if st is not None and "exp_avg" in st:
            mv.copy_(st["exp_avg"])
            st["exp_avg"] = mv
        if alias_params:
            p.data = pv
        b.p_views.append(pv)
        b.m_views.append(mv)
        b.g_views.append(gv)
    return b


def gather_grads(b: Bucket, *, use_kernel: bool = True, untouched_is_missing: bool = False):
    """Make ``b.flat_g`` hold this step's gradients.  Views that already alias the bucket
    cost nothing; foreign ``.grad`` tensors are copied in (one multi-tensor kernel on CUDA);
    missing grads contribute zeros.  Returns ``(number of tensors copied, indices of the
    parameters without a gradient this step)``.

    ``untouched_is_missing``: a parameter whose ``.grad`` aliases the bucket but which no backward pass wrote to
    since the last step (``p._dlion_touched`` is False — see ``Lion._install_grad_tracking``) counts as missing
    too: the reference skips ``p.grad is None`` parameters (distributed_lion.py:180), and once gradients alias
    the bucket "None" can only be told apart from "zero" this way.  The marks are consumed here."""
    src, dst, missing = [], [], []
    for i, (p, gv) in enumerate(zip(b.params, b.g_views)):
        g = p.grad
        touched = getattr(p, "_dlion_touched", None)
        if touched:
            p._dlion_touched = False
        if g is None:
            gv.zero_()
            missing.append(i)
        elif g.is_sparse:
            raise RuntimeError("Lion does not support sparse gradients")
        elif g.data_ptr() == gv.data_ptr() and g.dtype == gv.dtype:
            if untouched_is_missing and touched is False and not bool(gv.any()):
                # aliased, no backward pass wrote to it since the last step AND it still holds the zeros of
                # zero_grad(): "grad is None" upstream.  (A caller who fills .grad by hand — copy_ into the aliased
                # view — leaves no mark but leaves non-zero values: that parameter HAS a gradient.  The check syncs,
                # and only runs for parameters without a mark: none in an ordinary training step.)
                missing.append(i)
            continue
        else:
            src.append(g)
            dst.append(gv)
    if src:
        done = False
        if use_kernel and b.device.type == "cuda":
            from ..ops import bucket_ops
            done = bucket_ops.multi_tensor_gather(src, dst)
        if not done:
            torch._foreach_copy_(dst, src)
    return len(src), missing


def alias_grads(b: Bucket) -> None:
    for p, gv in zip(b.params, b.g_views):
        p.grad = gv
