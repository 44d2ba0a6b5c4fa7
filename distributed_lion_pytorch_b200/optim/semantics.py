"""Frozen semantics of Distributed Lion, as plain PyTorch.

This module is the *oracle*: every other implementation in the repository (the CPU/gloo
path, the ``nccl_1bit`` split kernels, the fused NVLink kernel) is tested against it, and
it is itself tested bit-for-bit against the reference's ``update_fn`` /
``update_fn_distributed`` (reference: distributed_lion.py:47-96).

Wire format
-----------
One bit per element, ``bit = (beta1*m + (1-beta1)*g) > 0``, packed LSB-first: element
``e`` lives in byte ``e // 8`` at bit ``e % 8`` (reference: distributed_lion.py:75-77, where
the packed word is accidentally int64).  On a little-endian machine this is
identical to 32 signs per ``uint32`` word with element ``e`` at word ``e // 32`` bit
``e % 32``, which is the layout the CUDA kernels use.  Pad bits are 0 and never vote.

Vote rules (``L`` = number of live workers, ``cnt`` = number of live workers whose bit is 1)
----------------------------------------------------------------------------------------
* ``majority`` / ``tie="neg"``  : ``u = +1 if 2*cnt > L else -1``  (reference-exact, D4)
* ``majority`` / ``tie="zero"`` : as above but ``u = 0`` when ``2*cnt == L`` (paper's sign(Σδ))
* ``average``                   : ``u = (2*cnt - L) / L``  (paper, Algorithm 1 "Averaging")
* ``stochastic``                : the local bit is ``Bernoulli((c + r) / 2r)``,
  ``r = (1 + 1/beta1) * max_grad_norm`` (intended behaviour of the reference's dead
  ``update_fn_distributed_stoc``, distributed_lion.py:98-136), then majority vote.
"""
from __future__ import annotations

from typing import Iterable, Optional, Sequence

import torch

VOTE_MODES = ("majority", "average")
TIE_RULES = ("neg", "zero")

_BIT_WEIGHTS = {}


def _weights(device) -> torch.Tensor:
    w = _BIT_WEIGHTS.get(device)
    if w is None:
        w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=device)
        _BIT_WEIGHTS[device] = w
    return w


def padded_len(n: int, multiple: int = 8) -> int:
    return (n + multiple - 1) // multiple * multiple


def pack_bits(bits: torch.Tensor) -> torch.Tensor:
    """bool[n] -> uint8[ceil(n/8)], LSB-first, pad bits zero (reference: distributed_lion.py:14-24,75-77)."""
    flat = bits.reshape(-1)
    n = flat.numel()
    n8 = padded_len(n, 8)
    if n8 != n:
        flat = torch.nn.functional.pad(flat, (0, n8 - n))
    # (x * w).sum over 8 lanes never exceeds 255
    return (flat.view(-1, 8).to(torch.uint8) * _weights(flat.device)).sum(dim=-1, dtype=torch.uint8)


def unpack_bits(packed: torch.Tensor, n: int) -> torch.Tensor:
    """uint8[...] -> bool[n] (reference: distributed_lion.py:84-88)."""
    p = packed.reshape(-1, 1)
    bits = (p & _weights(p.device)) != 0
    return bits.reshape(-1)[:n]


def live_ranks(world_size: int, live_mask: Optional[int]) -> list:
    if live_mask is None:
        return list(range(world_size))
    return [r for r in range(world_size) if (live_mask >> r) & 1]


def count_votes(gathered: torch.Tensor, n: int, live_mask: Optional[int] = None) -> torch.Tensor:
    """gathered: uint8[W, ceil(n/8)] -> int16[n] count of live +1 bits."""
    W = gathered.shape[0]
    cnt = torch.zeros(n, dtype=torch.int16, device=gathered.device)
    for r in live_ranks(W, live_mask):
        cnt += unpack_bits(gathered[r], n)
    return cnt


def vote_direction(cnt: torch.Tensor, L: int, vote: str = "majority", tie: str = "neg",
                   dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """Map a per-element +1 count among ``L`` live workers to the update direction."""
    if L <= 0:
        raise ValueError("live worker set is empty")
    if vote == "majority":
        twice = cnt.to(torch.int32) * 2
        u = torch.where(twice > L, 1.0, -1.0)
        if tie == "zero":
            u = torch.where(twice == L, 0.0, u)
        elif tie != "neg":
            raise ValueError(f"tie must be one of {TIE_RULES}")
        return u.to(dtype)
    if vote == "average":
        return ((cnt.to(torch.float32) * 2 - L) / L).to(dtype)
    raise ValueError(f"vote must be one of {VOTE_MODES}")


def lion_blend(m: torch.Tensor, g: torch.Tensor, beta1: float) -> torch.Tensor:
    """``beta1*m + (1-beta1)*g`` with the reference's op order and per-op rounding
    (reference: distributed_lion.py:54,68)."""
    return m.clone().mul_(beta1).add(g, alpha=1 - beta1)


def local_bits(m: torch.Tensor, g: torch.Tensor, beta1: float, *, max_grad_norm: Optional[float] = None,
               generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """The bool tensor a worker puts on the wire."""
    c = lion_blend(m, g, beta1)
    if max_grad_norm is None:
        return c > 0
    r = (1.0 + 1.0 / beta1) * float(max_grad_norm)
    prob = ((c.float() + r) / (2.0 * r)).clamp_(0.0, 1.0)
    return torch.bernoulli(prob, generator=generator) > 0


def momentum_update_(m: torch.Tensor, g: torch.Tensor, beta2: float) -> None:
    """reference: distributed_lion.py:59,96"""
    m.mul_(beta2).add_(g, alpha=1 - beta2)


# ----------------------------------------------------------------------------- stochastic rounding (opt-in)
_M32 = 0xFFFFFFFF


def sr_key(seed: int, step: int) -> int:
    """Non-zero 64-bit key of one optimizer step's rounding noise; identical on every rank."""
    return ((int(seed) * 0x9E3779B97F4A7C15 + (int(step) + 1) * 0xD1B54A32D192ED03) & (2 ** 64 - 1)) | 1


def _mix32(x: torch.Tensor) -> torch.Tensor:
    """The kernels' counter hash ("lowbias32") on int64 tensors holding 32-bit values."""
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M32
    return x ^ (x >> 16)


def stochastic_round_bf16(x: torch.Tensor, key: int, elem_base: int = 0) -> torch.Tensor:
    """fp32 ``x`` (flat) → bf16, rounding up in magnitude with probability (discarded bits) / 2^16; the random bits
    are the kernels' hash of ``(key, elem_base + index)`` — same bits on CPU, in the split kernels and in the
    fused kernel (``csrc/common.cuh: store8_sr``)."""
    assert x.dtype == torch.float32 and x.dim() == 1
    idx = torch.arange(x.numel(), dtype=torch.int64, device=x.device) + int(elem_base)
    vec, lane = idx >> 3, idx & 7
    lo, hi = key & _M32, (key >> 32) & _M32
    base = _mix32(((vec & _M32) * 0x9E3779B9 + lo) & _M32) ^ _mix32(((vec >> 32) + hi) & _M32)
    word = _mix32((base + (lane >> 1) * 0x85EBCA6B + hi) & _M32)
    r16 = torch.where((lane & 1) == 0, word & 0xFFFF, word >> 16)
    bits = x.view(torch.int32).to(torch.int64) & _M32
    finite = (bits & 0x7F800000) != 0x7F800000
    bits = torch.where(finite, (bits + r16) & _M32, bits)
    out16 = (bits >> 16).to(torch.int32)
    out16 = torch.where(out16 >= 0x8000, out16 - 0x10000, out16).to(torch.int16)      # reinterpret as signed
    return out16.view(torch.bfloat16)


def _apply_sr_(p: torch.Tensor, u: torch.Tensor, lr: float, wd: float, key: int, elem_base: int) -> None:
    """p ← SR(p·(1−lr·wd) − lr·u), all in fp32 with one stochastic rounding (csrc/common.cuh: apply8_fp32)."""
    decay = torch.tensor(1 - lr * wd, dtype=torch.float32, device=p.device)
    neg_lr = torch.tensor(-lr, dtype=torch.float32, device=p.device)
    x = p.float().reshape(-1) * decay
    x = torch.addcmul(x, u.float().reshape(-1), neg_lr)
    p.copy_(stochastic_round_bf16(x, key, elem_base).view_as(p))


def local_lion_step_(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, lr: float, wd: float,
                     beta1: float, beta2: float, sr_key: int = 0, elem_base: int = 0) -> None:
    """Single-worker Lion (reference: distributed_lion.py:47-59); ``sign(0) == 0``."""
    if sr_key and p.dtype == torch.bfloat16:
        _apply_sr_(p, lion_blend(m, g, beta1).sign_(), lr, wd, sr_key, elem_base)
        momentum_update_(m, g, beta2)
        return
    p.mul_(1 - lr * wd)
    u = lion_blend(m, g, beta1).sign_()
    p.add_(u, alpha=-lr)
    momentum_update_(m, g, beta2)


def apply_vote_(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, cnt: torch.Tensor, L: int, lr: float,
                wd: float, beta2: float, vote: str = "majority", tie: str = "neg", sr_key: int = 0) -> None:
    """Decay, voted update, momentum write-back (reference: distributed_lion.py:64,91-96)."""
    if sr_key and p.dtype == torch.bfloat16:
        _apply_sr_(p, vote_direction(cnt.view_as(p), L, vote, tie, dtype=torch.float32), lr, wd, sr_key, 0)
        momentum_update_(m, g, beta2)
        return
    p.mul_(1 - lr * wd)
    u = vote_direction(cnt.view_as(p), L, vote, tie, dtype=torch.float32)
    if vote == "majority":
        p.add_(u.to(p.dtype), alpha=-lr)
    else:  # fractional direction: keep it in fp32 until the single rounding into p
        p.copy_(torch.addcmul(p.float(), u, torch.tensor(-lr, dtype=torch.float32, device=p.device)))
    momentum_update_(m, g, beta2)


def simulate_workers_step_(ps: Sequence[torch.Tensor], gs: Sequence[torch.Tensor], ms: Sequence[torch.Tensor],
                           lr: float, wd: float, beta1: float, beta2: float, vote: str = "majority",
                           tie: str = "neg", live_mask: Optional[int] = None,
                           max_grad_norm: Optional[float] = None,
                           generators: Optional[Iterable[torch.Generator]] = None) -> torch.Tensor:
    """Run one Distributed-Lion step for ``W`` workers held in one process.

    This is the single-process restatement used by the tests and by the one-GPU rank
    emulation checks; returns the gathered packed bits ``uint8[W, ceil(n/8)]``.
    """
    W = len(ps)
    n = ps[0].numel()
    gens = list(generators) if generators is not None else [None] * W
    gathered = torch.stack([
        pack_bits(local_bits(ms[r].reshape(-1), gs[r].reshape(-1), beta1, max_grad_norm=max_grad_norm,
                             generator=gens[r]))
        for r in range(W)
    ])
    cnt = count_votes(gathered, n, live_mask)
    L = len(live_ranks(W, live_mask))
    for r in range(W):
        apply_vote_(ps[r], gs[r], ms[r], cnt, L, lr, wd, beta2, vote, tie)
    return gathered
