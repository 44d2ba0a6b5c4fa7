"""Optimizer kernels: N6 local Lion, sign-pack, vote-update and the fused N1 step.
reference: distributed_lion.py:47-96 (update_fn / update_fn_distributed)."""
from __future__ import annotations

from typing import Optional

import torch

from . import _cabi


def _scalars(lr: float, wd: float, beta1: float, beta2: float):
    # computed in double exactly like the reference's Python expressions, rounded to fp32 once
    return float(lr), float(1 - lr * wd), float(beta1), float(1 - beta1), float(beta2), float(1 - beta2)


def _check_triplet(p, g, m):
    assert p.is_cuda and p.is_contiguous() and g.is_contiguous() and m.is_contiguous()
    assert p.dtype == g.dtype == m.dtype and p.numel() == g.numel() == m.numel()


def _clip_args(clip):
    """``clip`` = None or ``(sumsq_dev fp32[1], max_norm)``: the gradient clip folded into the kernel."""
    if clip is None:
        return None, 0.0
    sumsq_dev, max_norm = clip
    assert sumsq_dev.is_cuda and sumsq_dev.dtype == torch.float32 and sumsq_dev.numel() >= 1
    return sumsq_dev.data_ptr(), float(max_norm)


def local_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, lr: float, wd: float, beta1: float,
               beta2: float, sr_key: int = 0, elem_base: int = 0, clip=None, zero_g: bool = False) -> None:
    """``sr_key != 0``: stochastic rounding of the bf16 parameter write (``optim.semantics.sr_key``);
    ``elem_base`` = index of ``p[0]`` within its bucket (keeps the random stream layout-invariant).
    ``clip=(sumsq_dev, max_norm)`` folds ``clip_grad_norm_`` in (g itself is not rewritten);
    ``zero_g`` clears g once consumed (the ``zero_grad`` that would follow)."""
    _check_triplet(p, g, m)
    lr_, decay, b1, omb1, b2, omb2 = _scalars(lr, wd, beta1, beta2)
    cptr, cmax = _clip_args(clip)
    with torch.cuda.device(p.device):
        _cabi.check(_cabi.lib().dlion_lion_local_clip(p.data_ptr(), g.data_ptr(), m.data_ptr(), p.numel(),
                                                      _cabi.dtype_code(p.dtype), lr_, decay, b1, omb1, b2, omb2,
                                                      int(sr_key) & (2 ** 64 - 1), int(elem_base), cptr, cmax,
                                                      int(bool(zero_g)), _cabi.stream_ptr(p.device)), "lion_local")


def stoch_radius(beta1: float, max_grad_norm: Optional[float]) -> float:
    """r = (1 + 1/beta1) * max_grad_norm (reference: distributed_lion.py:106); 0 disables."""
    if max_grad_norm is None:
        return 0.0
    return float((1.0 + 1.0 / beta1) * max_grad_norm)


def sign_pack(m: torch.Tensor, g: torch.Tensor, beta1: float, out: torch.Tensor, *,
              max_grad_norm: Optional[float] = None, seed: int = 0, step: int = 0, rank: int = 0,
              clip=None) -> torch.Tensor:
    """out: uint8[>= ceil(n/32)*4] (or int32 words); bit e%8 of byte e//8 = sign bit of element e."""
    assert m.is_cuda and m.is_contiguous() and g.is_contiguous() and out.is_contiguous()
    n = m.numel()
    assert out.numel() * out.element_size() >= (n + 31) // 32 * 4, "wire buffer too small"
    cptr, cmax = _clip_args(clip)
    with torch.cuda.device(m.device):
        _cabi.check(_cabi.lib().dlion_sign_pack_clip(m.data_ptr(), g.data_ptr(), out.data_ptr(), n,
                                                     _cabi.dtype_code(m.dtype), float(beta1), float(1 - beta1),
                                                     stoch_radius(beta1, max_grad_norm), int(seed) & (2**64 - 1),
                                                     int(step) & 0xffffffff, int(rank), cptr, cmax,
                                                     _cabi.stream_ptr(m.device)), "sign_pack")
    return out


def vote_update(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, gathered: torch.Tensor, world: int,
                live_mask: Optional[int], lr: float, wd: float, beta2: float, vote: str = "majority",
                tie: str = "neg", sr_key: int = 0, clip=None, zero_g: bool = False) -> None:
    """gathered: [world, row_bytes] uint8 (row_bytes % 4 == 0), row r = rank r's packed signs."""
    _check_triplet(p, g, m)
    assert gathered.is_contiguous() and gathered.dim() == 2 and gathered.shape[0] == world
    row_bytes = gathered.shape[1] * gathered.element_size()
    assert row_bytes % 4 == 0
    mask = (1 << world) - 1 if live_mask is None else int(live_mask) & ((1 << world) - 1)
    lr_, decay, _, _, b2, omb2 = _scalars(lr, wd, 0.0, beta2)
    cptr, cmax = _clip_args(clip)
    with torch.cuda.device(p.device):
        _cabi.check(_cabi.lib().dlion_vote_update_clip(p.data_ptr(), g.data_ptr(), m.data_ptr(), gathered.data_ptr(),
                                                       row_bytes // 4, world, mask, p.numel(),
                                                       _cabi.dtype_code(p.dtype), lr_, decay, b2, omb2,
                                                       _cabi.VOTE_CODES[(vote, tie)], int(sr_key) & (2 ** 64 - 1),
                                                       cptr, cmax, int(bool(zero_g)), _cabi.stream_ptr(p.device)),
                    "vote_update")


# --------------------------------------------------------------------------- N1 fused step
def fused_tile(dtype: torch.dtype) -> int:
    return int(_cabi.lib().dlion_fused_tile(_cabi.dtype_code(dtype)))


def fused_max_grid(dtype: torch.dtype, device=None) -> int:
    import ctypes as C
    out = C.c_int(0)
    with torch.cuda.device(device):
        _cabi.check(_cabi.lib().dlion_fused_max_grid(_cabi.dtype_code(dtype), C.byref(out)), "fused_max_grid")
    return out.value


def fused_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, *, world: int, rank: int, arena_ptrs, flag_ptrs,
               arena_words: int, epoch: int, grid: int, lr: float, wd: float, beta1: float, beta2: float,
               vote: str = "majority", tie: str = "neg", live_mask: Optional[int] = None,
               wait_mask: Optional[int] = None, max_grad_norm: Optional[float] = None, seed: int = 0, step: int = 0,
               elem_base: int = 0, emulate_ranks: int = 0, emu_stride: int = 0, status: Optional[torch.Tensor] = None,
               spin_limit: int = 0, n: Optional[int] = None, mc_arena_ptr: int = 0, sr_key: int = 0,
               clip=None, zero_g: bool = False, wait_timeout_us: int = 0) -> None:
    """Launch the fused sign-pack-exchange-vote-update kernel on ``n`` elements of p/g/m.

    ``arena_ptrs[r]`` / ``flag_ptrs[r]`` are device addresses of rank r's receive arena and
    progress flags as mapped into THIS process (``parallel.symm``).  With ``emulate_ranks=W``
    the W ranks are CTA groups of one launch on one GPU working on p/g/m + v*emu_stride."""
    a = _cabi.FusedArgs()
    a.p, a.g, a.m = p.data_ptr(), g.data_ptr(), m.data_ptr()
    a.n = int(p.numel() if n is None else n)
    a.elem_base = int(elem_base)
    a.dtype = _cabi.dtype_code(p.dtype)
    a.world, a.rank = int(world), int(rank)
    full = (1 << world) - 1
    a.live_mask = full if live_mask is None else int(live_mask) & full
    a.wait_mask = full if wait_mask is None else int(wait_mask) & full
    a.lr, a.decay, a.beta1, a.omb1, a.beta2, a.omb2 = _scalars(lr, wd, beta1, beta2)
    a.vote = _cabi.VOTE_CODES[(vote, tie)]
    a.stoch_r = stoch_radius(beta1, max_grad_norm)
    a.seed, a.step, a.epoch = int(seed) & (2 ** 64 - 1), int(step) & 0xffffffff, int(epoch)
    for r in range(world):
        a.arena[r] = int(arena_ptrs[r])
        a.flags[r] = int(flag_ptrs[r])
    a.mc_arena = int(mc_arena_ptr) if mc_arena_ptr else None
    a.arena_words, a.grid = int(arena_words), int(grid)
    a.emulate_ranks, a.emu_stride = int(emulate_ranks), int(emu_stride)
    a.status = status.data_ptr() if status is not None else None
    a.spin_limit = int(spin_limit)
    a.sr_key = int(sr_key) & (2 ** 64 - 1)
    a.clip_sumsq, a.clip_max_norm = _clip_args(clip)
    a.zero_g = int(bool(zero_g))
    a.wait_timeout_us = int(wait_timeout_us)
    import ctypes as C
    with torch.cuda.device(p.device):
        _cabi.check(_cabi.lib().dlion_fused_step(C.byref(a), _cabi.stream_ptr(p.device)), "fused_step")
