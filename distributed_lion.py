"""Public entry point kept from the reference: ``from distributed_lion import Lion``
(reference: README.md:8-15, distributed_lion.py:138).  The optimizer lives in
``distributed_lion_pytorch_b200.optim.lion``; the module-level helpers of the reference
(``exists``, ``flatten_and_pad``, ``restore_flattened_tensor``, ``majority_vote``, ``update_fn``,
``update_fn_distributed``, ``update_fn_distributed_stoc`` — distributed_lion.py:8-136) are
re-provided with the same signatures on top of this package's semantics, for code that calls
them directly."""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist

from distributed_lion_pytorch_b200.optim import semantics as _sem
from distributed_lion_pytorch_b200.optim.lion import Lion  # noqa: F401
from distributed_lion_pytorch_b200.optim.semantics import (  # noqa: F401
    count_votes, pack_bits, unpack_bits, vote_direction)

__all__ = ["Lion", "exists", "flatten_and_pad", "restore_flattened_tensor", "majority_vote", "update_fn",
           "update_fn_distributed", "update_fn_distributed_stoc", "pack_bits", "unpack_bits", "count_votes",
           "vote_direction"]


def exists(val) -> bool:
    """reference: distributed_lion.py:8-9"""
    return val is not None


def flatten_and_pad(tensor: torch.Tensor) -> Tuple[torch.Tensor, torch.Size]:
    """Flatten and zero-pad to a multiple of 8 elements (reference: distributed_lion.py:14-24)."""
    flat = tensor.reshape(-1)
    want = _sem.padded_len(flat.numel(), 8)
    if want != flat.numel():
        flat = torch.cat([flat, flat.new_zeros(want - flat.numel())])
    return flat, tensor.shape


def restore_flattened_tensor(padded_tensor: torch.Tensor, original_shape: torch.Size) -> torch.Tensor:
    """reference: distributed_lion.py:27-31"""
    return padded_tensor[: original_shape.numel()].reshape(original_shape)


def majority_vote(boolean_tensors: List[torch.Tensor]) -> torch.Tensor:
    """Strict majority of True; a tie gives False (reference: distributed_lion.py:33-43 via torch.mode)."""
    votes = sum(t.to(torch.int16) for t in boolean_tensors)
    return votes * 2 > len(boolean_tensors)


def update_fn(p, grad, exp_avg, lr, wd, beta1, beta2) -> None:
    """Single-worker Lion on one tensor (reference: distributed_lion.py:47-59)."""
    if p.is_cuda and p.is_contiguous() and grad.is_contiguous() and exp_avg.is_contiguous() \
            and p.data_ptr() % 16 == 0 and grad.data_ptr() % 16 == 0 and exp_avg.data_ptr() % 16 == 0:
        from distributed_lion_pytorch_b200 import ops
        if ops.require_library():
            from distributed_lion_pytorch_b200.ops import lion_ops
            lion_ops.local_step(p.data, grad, exp_avg, lr, wd, beta1, beta2)
            return
    _sem.local_lion_step_(p.data, grad, exp_avg, lr, wd, beta1, beta2)


def _exchange_and_apply(p, grad, exp_avg, lr, wd, beta1, beta2, max_grad_norm=None) -> None:
    W = dist.get_world_size()
    bits = _sem.local_bits(exp_avg.reshape(-1), grad.reshape(-1), beta1, max_grad_norm=max_grad_norm)
    mine = _sem.pack_bits(bits)                       # true 1 bit / element (upstream sends an int64 per 8: D1)
    gathered = torch.empty(W, mine.numel(), dtype=torch.uint8, device=mine.device)
    try:
        dist.all_gather_into_tensor(gathered.view(-1), mine)
    except (RuntimeError, NotImplementedError):
        dist.all_gather(list(gathered.unbind(0)), mine)
    cnt = _sem.count_votes(gathered, p.numel())
    _sem.apply_vote_(p.data, grad, exp_avg, cnt, W, lr, wd, beta2, "majority", "neg")


def update_fn_distributed(p, grad, exp_avg, lr, wd, beta1, beta2) -> None:
    """Majority-vote Distributed Lion on one tensor (reference: distributed_lion.py:61-96).  The
    per-tensor functional form; ``Lion.step`` uses flat buckets and the fused kernel instead."""
    _exchange_and_apply(p, grad, exp_avg, lr, wd, beta1, beta2)


def update_fn_distributed_stoc(p, grad, exp_avg, lr, wd, beta1, beta2, max_grad_norm) -> None:
    """Stochastic binarisation over workers (reference: distributed_lion.py:98-136 — unreachable
    upstream; this one runs)."""
    _exchange_and_apply(p, grad, exp_avg, lr, wd, beta1, beta2, max_grad_norm=max_grad_norm)
