"""Step back-ends for :class:`Lion` — where the sign exchange happens.

reference: distributed_lion.py:61-96 is the only strategy upstream (encode → list-API
``all_gather`` of an int64 tensor per parameter → decode → ``torch.mode``).  Here the
strategy is pluggable per bucket; see ``optim/lion.py`` for the user-facing description.
"""
from __future__ import annotations

import os
import weakref
from typing import List, Optional

import torch
import torch.distributed as dist

from . import semantics as sem
from .bucket import Bucket


class _Backend:
    name = "base"

    def __init__(self, opt):
        self.world_size = opt.world_size
        # weak: the optimizer owns its back-end; a strong back-reference would form a cycle and keep
        # tens of GB of flat buckets alive until the cyclic GC happens to run
        self._opt_ref = weakref.ref(opt)

    @property
    def opt(self):
        o = self._opt_ref()
        if o is None:
            raise ReferenceError("the Lion optimizer that owned this back-end is gone")
        return o

    def on_rebucket(self, opt) -> None:
        pass

    def _use_kernels(self, b: Bucket) -> bool:
        return self.opt._kernels_ok(b.device)

    def step(self, b: Bucket, lr: float, wd: float, beta1: float, beta2: float, missing: List[int],
             clip=None, zero_g: bool = False) -> bool:
        """One optimizer step on bucket ``b``.  ``clip=(sumsq_dev, max_norm)`` is the gradient clip to fold into the
        kernels, ``zero_g`` asks them to clear the gradient bucket once consumed.  Returns True when the gradient
        bucket was left zeroed (only kernel paths do that; the caller then skips its memset)."""
        raise NotImplementedError


class LocalBackend(_Backend):
    """Single-worker Lion (reference: distributed_lion.py:47-59)."""
    name = "local"

    def step(self, b, lr, wd, beta1, beta2, missing, clip=None, zero_g=False):
        if missing:
            # reference semantics: a parameter without a gradient is left untouched this step
            # (distributed_lion.py:180) — no decay, no momentum decay, no sign(beta1*m) move
            skip = set(missing)
            spans = [(pv, gv, mv, off) for i, (pv, gv, mv, off) in
                     enumerate(zip(b.p_views, b.g_views, b.m_views, b.offsets)) if i not in skip]
        else:
            spans = [(b.flat_p, b.flat_g, b.flat_m, 0)]
        key = self.opt.sr_key()
        if self._use_kernels(b):
            from ..ops import lion_ops
            for p, g, m, off in spans:
                lion_ops.local_step(p.view(-1), g.view(-1), m.view(-1), lr, wd, beta1, beta2, sr_key=key, elem_base=off,
                                    clip=clip, zero_g=zero_g)
            return bool(zero_g)      # skipped spans hold no gradient: the whole bucket is clean
        assert clip is None, "the eager path clips eagerly (Lion.clip_grad_norm_)"
        for p, g, m, off in spans:
            sem.local_lion_step_(p.view(-1), g.view(-1), m.view(-1), lr, wd, beta1, beta2, sr_key=key, elem_base=off)
        return False


class AllGatherBackend(_Backend):
    """``nccl_1bit``: split pack / all-gather / vote+update.  The true 1-bit payload goes
    through ``torch.distributed`` (NCCL on GPUs, gloo on CPU) — the baseline the fused
    kernel is measured against, and the fallback for multi-node jobs."""
    name = "nccl_1bit"

    def _wire(self, b: Bucket):
        W = self.world_size
        mine = b.scratch.get("wire_out")
        if mine is None or mine.numel() != b.wire_bytes:
            mine = torch.zeros(b.wire_bytes, dtype=torch.uint8, device=b.device)
            b.scratch["wire_out"] = mine
            b.scratch["wire_in"] = torch.zeros(W, b.wire_bytes, dtype=torch.uint8, device=b.device)
        return mine, b.scratch["wire_in"]

    def _all_gather(self, out: torch.Tensor, inp: torch.Tensor) -> None:
        pg = self.opt.process_group
        try:
            dist.all_gather_into_tensor(out.view(-1), inp, group=pg)
        except (RuntimeError, NotImplementedError):
            chunks = list(out.unbind(0))
            dist.all_gather(chunks, inp, group=pg)

    def step(self, b, lr, wd, beta1, beta2, missing, clip=None, zero_g=False):
        opt = self.opt
        W = self.world_size
        mine, gathered = self._wire(b)
        # the vote kernels count ballots in 5 bit planes (<= DLION_MAX_RANKS workers); larger jobs — which the
        # reference's all_gather path supports — take the torch / cpu_ops vote path below
        from ..ops import _cabi
        kern = self._use_kernels(b) and W <= _cabi.MAX_RANKS
        if clip is not None and not kern:
            raise RuntimeError("a folded gradient clip needs the kernel path (Lion.clip_grad_norm_ decides this)")
        if kern:
            from ..ops import lion_ops
            lion_ops.sign_pack(b.flat_m, b.flat_g, beta1, mine, max_grad_norm=opt.max_grad_norm,
                               seed=opt.seed, step=opt._steps_done, rank=opt.rank, clip=clip)
        else:
            gen = opt.cpu_generator() if opt.max_grad_norm is not None else None
            from ..ops import cpu_ops
            bits = sem.local_bits(b.flat_m, b.flat_g, beta1, max_grad_norm=opt.max_grad_norm, generator=gen)
            mine.copy_(cpu_ops.pack_bits(bits))
        self._all_gather(gathered, mine)
        L = opt.live_count()
        if kern:
            lion_ops.vote_update(b.flat_p, b.flat_g, b.flat_m, gathered, W, opt.live_mask, lr, wd, beta2,
                                 opt.vote, opt.tie, sr_key=opt.sr_key(), clip=clip, zero_g=zero_g)
            return bool(zero_g)
        if b.device.type == "cuda":
            cnt = sem.count_votes(gathered, b.numel, opt.live_mask)
        else:
            from ..ops import cpu_ops
            cnt = cpu_ops.count_votes(gathered, b.numel, opt.live_mask)
        sem.apply_vote_(b.flat_p, b.flat_g, b.flat_m, cnt, L, lr, wd, beta2, opt.vote, opt.tie, sr_key=opt.sr_key())
        return False


class FusedBackend(_Backend):
    """One persistent sm_100a kernel per bucket chunk; the sign words travel as peer stores
    over NVLink into symmetric arenas (``parallel.symm``); no NCCL call on the hot path."""
    name = "fused"

    def __init__(self, opt):
        super().__init__(opt)
        from ..parallel import symm
        self._symm = symm
        self._ctx = None

    def on_rebucket(self, opt) -> None:
        self._ctx = None

    def _try_create(self, device) -> bool:
        """Set the exchange up and agree collectively that it worked everywhere (a rank-local failure
        must not leave the others spinning on flags that will never be written)."""
        import logging
        err = None
        try:
            self._ctx = self._symm.FusedExchange(self.opt.process_group, device)
        except Exception as e:      # noqa: BLE001 — any failure means "no peer mapping here"
            err = e
            self._ctx = None
        ok = torch.tensor([0 if err is not None else 1], dtype=torch.int32, device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.opt.process_group)
        if int(ok.item()) == 1:
            return True
        self._ctx = None
        if self.opt.comm == "fused":
            raise RuntimeError(f"comm='fused' was requested but peer-mapped memory could not be set up: {err}")
        logging.getLogger("dlion.lion").warning("fused NVLink exchange unavailable (%s); using nccl_1bit", err)
        return False

    def step(self, b, lr, wd, beta1, beta2, missing, clip=None, zero_g=False):
        opt = self.opt
        if b.device.type != "cuda":
            raise RuntimeError("comm='fused' needs CUDA tensors; use comm='nccl_1bit' on CPU")
        if self._ctx is None and not self._try_create(b.device):
            # peer mapping is unavailable on at least one rank: every rank switches to the split path
            fallback = AllGatherBackend(opt)
            opt._backend = fallback
            return fallback.step(b, lr, wd, beta1, beta2, missing, clip=clip, zero_g=zero_g)
        self._ctx.step(b, lr, wd, beta1, beta2, vote=opt.vote, tie=opt.tie, live_mask=opt.live_mask,
                       max_grad_norm=opt.max_grad_norm, seed=opt.seed, step_index=opt._steps_done, sr_key=opt.sr_key(),
                       clip=clip, zero_g=zero_g, wait_mask=opt.wait_mask)
        return bool(zero_g)


def pick_auto(opt) -> _Backend:
    """CUDA + every rank on one NVLink domain + peer-mapped memory available → fused; else nccl_1bit."""
    dev = None
    for g in opt.param_groups:
        for p in g["params"]:
            dev = p.device
            break
        if dev is not None:
            break
    if dev is not None and dev.type == "cuda" and opt._kernels_ok(dev):
        from ..parallel import symm
        if symm.fused_supported(opt.process_group):
            return FusedBackend(opt)
    return AllGatherBackend(opt)
