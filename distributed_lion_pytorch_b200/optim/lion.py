"""``Lion`` — local and Distributed Lion over flat buckets.

API contract (reference: distributed_lion.py:138-200):
``Lion(params, lr=1e-4, betas=(0.9, 0.99), weight_decay=0.0, max_grad_norm=None)``;
``step(closure=None)`` returns the closure's loss; the only per-parameter state key is
``exp_avg`` (same shape/dtype as the parameter, created on the first step that sees a
gradient); ``param_groups`` carry ``lr / betas / weight_decay`` only, so a reference
``optimizer.pt`` loads here and ours loads there.

What is different from the reference (on purpose):

* the update runs on flat buckets — one launch per group instead of one Python iteration,
  ~50 ATen ops and one blocking ``all_gather`` per tensor (distributed_lion.py:179-198);
* the world size / backend is resolved lazily at the first ``step`` (fixes SURVEY D7: the
  reference freezes it in ``__init__`` and silently degrades to local Lion);
* ``max_grad_norm`` really selects stochastic binarisation (D2/D3: dead code upstream);
* a parameter with ``grad is None`` on *this* rank but active elsewhere contributes a zero
  gradient instead of desynchronising the collective sequence (D6);
* new keyword-only options: ``vote`` (majority | average), ``tie`` (neg | zero), ``comm``
  (auto | fused | nccl_1bit | local), ``live_mask`` / :meth:`set_live_mask`, ``seed``,
  ``stochastic_rounding`` (bf16 parameter writes; off by default = the reference's arithmetic);
* the two neighbours of ``step()`` in every training loop can ride in the same kernel pass:
  :meth:`clip_grad_norm_` measures the gradient norm and hands the *scaling* to the next ``step()``
  (the kernels work on ``round(g * coef)`` in registers), and ``zero_grads_in_step=True`` lets the
  kernels clear the gradient bucket once consumed, so ``zero_grad()`` after ``step()`` is free.
  Both are off unless asked for (``Trainer`` asks); results are bit-identical to the separate passes.

Communication back-ends (``comm``):

``fused``      one persistent sm_100a kernel per bucket that packs signs, stores them into
               every peer's receive arena over NVLink, waits on per-CTA flags, votes and
               applies the update (``ops.lion_ops.fused_step``).  No NCCL on the hot path.
``nccl_1bit``  pack kernel → ``all_gather_into_tensor`` of the true 1-bit payload →
               vote/update kernel.  Works on any backend (gloo on CPU, multi-node NCCL).
``local``      single worker; fused local Lion kernel.
"""
from __future__ import annotations

import os
from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch.optim.optimizer import Optimizer

from . import semantics as sem
from .bucket import Bucket, alias_grads, build_bucket, gather_grads

_COMM_CHOICES = ("auto", "fused", "nccl_1bit", "local")


def _dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized()


class Lion(Optimizer):
    def __init__(self, params, lr: float = 1e-4, betas: Tuple[float, float] = (0.9, 0.99),
                 weight_decay: float = 0.0, max_grad_norm: Optional[float] = None, *,
                 vote: str = "majority", tie: str = "neg", comm: str = "auto",
                 process_group=None, live_mask: Optional[int] = None, seed: int = 0,
                 flat_grads: bool = True, use_kernels: Optional[bool] = None, stochastic_rounding: bool = False,
                 zero_grads_in_step: bool = False, wait_mask: Optional[int] = None):
        if not lr > 0.0:
            raise ValueError(f"lr must be positive, got {lr}")
        if not all(0.0 <= b <= 1.0 for b in betas):
            raise ValueError(f"betas must lie in [0, 1], got {betas}")
        if vote not in sem.VOTE_MODES:
            raise ValueError(f"vote must be one of {sem.VOTE_MODES}")
        if tie not in sem.TIE_RULES:
            raise ValueError(f"tie must be one of {sem.TIE_RULES}")
        if comm not in _COMM_CHOICES:
            raise ValueError(f"comm must be one of {_COMM_CHOICES}")
        if max_grad_norm is not None and not max_grad_norm > 0:
            raise ValueError("max_grad_norm must be positive when given")
        super().__init__(params, dict(lr=lr, betas=betas, weight_decay=weight_decay))
        self.max_grad_norm = max_grad_norm
        self.vote, self.tie, self.comm = vote, tie, comm
        self.process_group = process_group
        self.live_mask = live_mask
        self.seed = int(seed)
        self.flat_grads = flat_grads
        self.use_kernels = use_kernels
        # opt-in: bf16 parameters are written with stochastic instead of nearest rounding (not in the reference;
        # SURVEY §7.3 — at lr ~ 1e-4 a sign update is below half a bf16 ulp of a typical weight).  The noise is a
        # counter hash of (seed, step, element), identical on every rank, so replicas stay bit-identical.
        self.stochastic_rounding = bool(stochastic_rounding)
        # opt-in: step() leaves the flat gradient buckets zeroed (the kernels clear g once consumed), which turns
        # the zero_grad() that follows into a no-op.  Off by default: torch optimizers leave .grad alone in step().
        self.zero_grads_in_step = bool(zero_grads_in_step)
        # ranks whose progress flags the fused kernel waits for (None = every rank).  A rank the caller KNOWS to be
        # gone is taken out of both masks so that nobody spins on flags that will never arrive.
        self.wait_mask = wait_mask
        self._pending_clip = None       # (sumsq_dev, max_norm) measured by clip_grad_norm_, consumed by step()
        self._g_zeroed = False          # the last step() left every flat gradient bucket zeroed
        self._track_hooks = []
        self._buckets: Optional[List[Bucket]] = None
        self._bucket_sig = None
        self._frozen_actives: List[list] = []
        self._backend = None            # resolved at first step
        self._steps_done = 0
        self._cpu_gen: Optional[torch.Generator] = None
        self.last_step_stats: Dict[str, float] = {}
        self._install_grad_tracking()

    # ------------------------------------------------------------------ which parameters got a gradient this step
    def _install_grad_tracking(self) -> None:
        """With flat buckets ``p.grad`` aliases the bucket from the first step on and is never ``None`` again, so
        "did this parameter receive a gradient in this step" (reference: ``if p.grad is not None`` at
        distributed_lion.py:180) needs its own record: a post-accumulate hook (and ``ops.mark_grad_written`` in the
        custom backward functions that add into ``.grad`` themselves) sets ``p._dlion_touched``; ``step()`` reads
        and clears it.  Only the single-worker path consults it — with W > 1 a parameter without a gradient
        contributes zeros (see the module docstring, D6)."""
        from ..ops import mark_grad_written
        seen = set()
        for group in self.param_groups:
            for p in group["params"]:
                if id(p) in seen or not p.requires_grad or not p.is_leaf:
                    continue
                seen.add(id(p))
                if hasattr(p, "register_post_accumulate_grad_hook"):
                    self._track_hooks.append(p.register_post_accumulate_grad_hook(mark_grad_written))

    # ------------------------------------------------------------------ public extras
    def set_live_mask(self, mask: Optional[int]) -> None:
        """Bit ``r`` set ⇔ worker ``r``'s signs count in the vote for the following steps.
        Must be identical on every rank (it changes the majority threshold / the averaging
        denominator).  ``None`` = everybody.  A masked-out worker still *receives* the voted
        update, so replicas stay identical; only its ballot is discarded."""
        if mask is not None:
            W = self.world_size
            if mask & ((1 << W) - 1) == 0:
                raise ValueError("live_mask excludes every worker")
        self.live_mask = mask

    @property
    def world_size(self) -> int:
        if _dist_ready():
            return dist.get_world_size(self.process_group)
        return 1

    @property
    def rank(self) -> int:
        return dist.get_rank(self.process_group) if _dist_ready() else 0

    @property
    def backend_name(self) -> str:
        self._resolve_backend()
        return self._backend.name

    @property
    def buckets(self) -> List[Bucket]:
        return self._buckets or []

    @property
    def steps_done(self) -> int:
        """Optimizer steps taken so far.  It indexes the Philox stream of the stochastic binarisation, and
        ``optimizer.pt`` keeps the reference's schema (no room for it), so whoever resumes a run sets it
        back (``Trainer`` does, from ``trainer_state.json``) — otherwise the noise sequence restarts."""
        return self._steps_done

    @steps_done.setter
    def steps_done(self, n: int) -> None:
        if int(n) < 0:
            raise ValueError("steps_done must be non-negative")
        self._steps_done = int(n)

    def wire_bytes_per_step(self) -> int:
        """Bytes this rank *sends* per optimizer step (true 1-bit payload × peers)."""
        W = self.world_size
        return sum(b.wire_bytes for b in self.buckets) * max(W - 1, 0)

    # ------------------------------------------------------------------ bucket plumbing
    def _active_params(self, group) -> List[torch.nn.Parameter]:
        return [p for p in group["params"] if p.grad is not None or "exp_avg" in self.state.get(p, {})]

    def _ensure_buckets(self) -> None:
        W = self.world_size
        actives = []
        for gi, group in enumerate(self.param_groups):
            actives.append(self._active_params(group))
        if W > 1 and self._buckets is None:
            actives = self._sync_active_sets(actives)
        elif W > 1:
            # distributed: the active set is frozen after the first step (see class docstring)
            actives = [list(ps) for ps in self._frozen_actives]
        sig = tuple(tuple(id(p) for p in ps) for ps in actives)
        if self._buckets is not None and sig == self._bucket_sig:
            return
        buckets: List[Bucket] = []
        for gi, ps in enumerate(actives):
            by_kind: Dict[tuple, list] = {}
            for p in ps:
                by_kind.setdefault((p.dtype, p.device), []).append(p)
            for (_dt, _dev), plist in by_kind.items():
                buckets.append(build_bucket(gi, plist, self.state))
        for b in buckets:
            for p, mv in zip(b.params, b.m_views):
                self.state[p]["exp_avg"] = mv      # lazily created state (distributed_lion.py:185-186)
        self._buckets, self._bucket_sig = buckets, sig
        self._frozen_actives = actives
        if self._backend is not None:
            self._backend.on_rebucket(self)

    def _sync_active_sets(self, actives):
        """One-off agreement on which parameters take part (union over ranks)."""
        flat = [p for g in self.param_groups for p in g["params"]]
        if not flat:
            return actives
        mine = {id(p) for ps in actives for p in ps}
        dev = flat[0].device
        flags = torch.tensor([1 if id(p) in mine else 0 for p in flat], dtype=torch.int32, device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MAX, group=self.process_group)
        on = flags.cpu().tolist()
        keep = {id(p) for p, f in zip(flat, on) if f}
        return [[p for p in g["params"] if id(p) in keep] for g in self.param_groups]

    def refresh_active_set(self) -> None:
        """Collective: re-derive which parameters participate (call on every rank)."""
        self._buckets = None
        self._bucket_sig = None

    def add_param_group(self, param_group) -> None:
        """Adding a group after the first step re-plans the buckets at the next step (with W > 1 that
        step re-runs the one-off agreement on the active set, so add the group on every rank)."""
        super().add_param_group(param_group)
        if getattr(self, "_buckets", None) is not None:
            self.refresh_active_set()
        if hasattr(self, "_track_hooks"):           # not during Optimizer.__init__
            from ..ops import mark_grad_written
            for p in self.param_groups[-1]["params"]:
                if p.requires_grad and p.is_leaf and hasattr(p, "register_post_accumulate_grad_hook"):
                    self._track_hooks.append(p.register_post_accumulate_grad_hook(mark_grad_written))

    # ------------------------------------------------------------------ backend choice
    def _kernels_ok(self, device: torch.device) -> bool:
        if device.type != "cuda":
            return False
        if self.use_kernels is False:
            return False
        from .. import ops
        return ops.require_library()

    def _resolve_backend(self) -> None:
        if self._backend is not None and self._backend.world_size == self.world_size:
            return
        from . import backends
        W = self.world_size
        choice = self.comm
        if choice == "auto":
            env = os.environ.get("DLION_COMM")
            if env in _COMM_CHOICES and env != "auto":
                choice = env
        if W == 1 or choice == "local":
            self._backend = backends.LocalBackend(self)
        elif choice == "fused":
            self._backend = backends.FusedBackend(self)
        elif choice == "nccl_1bit":
            self._backend = backends.AllGatherBackend(self)
        else:  # auto
            self._backend = backends.pick_auto(self)

    # ------------------------------------------------------------------ Optimizer API
    @torch.no_grad()
    def step(self, closure: Optional[Callable] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self._resolve_backend()
        self._ensure_buckets()
        clip, self._pending_clip = self._pending_clip, None
        zero_g = self.zero_grads_in_step and self.flat_grads
        local = self._backend.name == "local"
        cleaned = bool(self._buckets)
        for b in self._buckets:
            _, missing = gather_grads(b, use_kernel=self._kernels_ok(b.device), untouched_is_missing=local)
            group = self.param_groups[b.group_index]
            beta1, beta2 = group["betas"]
            cleaned &= bool(self._backend.step(b, float(group["lr"]), float(group["weight_decay"]), float(beta1),
                                               float(beta2), missing, clip=clip, zero_g=zero_g))
            if self.flat_grads:
                alias_grads(b)
        self._g_zeroed = cleaned
        self._steps_done += 1
        return loss

    # ------------------------------------------------------------------ gradient clipping folded into step()
    def _grads_aliased(self) -> bool:
        """Every bucketed parameter's ``.grad`` IS its slice of the flat gradient bucket, and no parameter of this
        optimizer holds a gradient outside the buckets."""
        if not (self._buckets and self.flat_grads):
            return False
        owned = set()
        for b in self._buckets:
            for p, gv in zip(b.params, b.g_views):
                g = p.grad
                if g is None or g.data_ptr() != gv.data_ptr() or g.dtype != gv.dtype:
                    return False
                owned.add(id(p))
        return all(p.grad is None or id(p) in owned for grp in self.param_groups for p in grp["params"])

    @torch.no_grad()
    def clip_grad_norm_(self, max_norm: float, extra_sumsq: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``torch.nn.utils.clip_grad_norm_`` over this optimizer's parameters (local gradients only — HF Trainer's
        ``max_grad_norm``, which the reference inherits: SURVEY §3.1).  Returns the total norm (a tensor; no host
        sync).

        In steady state (gradients aliased to the flat buckets, CUDA, library loaded) only the *norm* is computed
        here — one ``sumsq`` launch per bucket — and the scaling is handed to the next :meth:`step`, whose kernels
        form ``round(g * coef)`` in registers: the gradient bucket is read twice per optimizer step instead of
        three times and written zero times instead of once.  ``p.grad`` therefore still holds the *unclipped*
        values afterwards; the parameters and momenta after ``step()`` are bit-identical to clip-then-step.
        Otherwise (first step, CPU, foreign ``.grad`` tensors) the gradients are scaled in place, eagerly.

        ``extra_sumsq`` (fp32[1] on the same device): sum of squares of gradients that are clipped together with
        ours but live elsewhere (another optimizer's parameters)."""
        max_norm = float(max_norm)
        if self._grads_aliased() and self._buckets[0].device.type == "cuda" and self._kernels_ok(self._buckets[0].device) \
                and len({b.device for b in self._buckets}) == 1:
            from ..ops import bucket_ops
            parts = [bucket_ops.sumsq(b.flat_g) for b in self._buckets]
            if extra_sumsq is not None:
                parts.append(extra_sumsq.view(1).float())
            total = parts[0] if len(parts) == 1 else torch.stack([x.view(()) for x in parts]).sum().view(1)
            self._pending_clip = (total, max_norm)
            return total.sqrt().view(())
        self._pending_clip = None
        params = [p for grp in self.param_groups for p in grp["params"] if p.grad is not None]
        if not params:
            return torch.zeros(())
        if extra_sumsq is not None:
            raise RuntimeError("extra_sumsq needs the flat-bucket CUDA path")
        return torch.nn.utils.clip_grad_norm_(params, max_norm)

    def zero_grad(self, set_to_none: bool = True) -> None:
        """With flat gradient buckets the gradients are cleared by one memset per bucket and
        ``p.grad`` keeps aliasing the bucket, so the next backward accumulates straight into
        the flat tensor the optimizer kernel reads."""
        if self._buckets and self.flat_grads:
            owned = set()
            clean, self._g_zeroed = self._g_zeroed, False       # the step's kernels already cleared the buckets
            self._pending_clip = None                           # a norm measured on these gradients is void now
            for b in self._buckets:
                if not clean:
                    b.flat_g.zero_()
                alias_grads(b)
                owned.update(id(p) for p in b.params)
            for group in self.param_groups:
                for p in group["params"]:
                    if id(p) not in owned and p.grad is not None:
                        if set_to_none:
                            p.grad = None
                        else:
                            p.grad.detach_().zero_()
        else:
            super().zero_grad(set_to_none=set_to_none)

    def state_dict(self):
        """Reference-compatible: independent ``exp_avg`` tensors, not views of the bucket."""
        sd = super().state_dict()
        sd["state"] = {k: {kk: (vv.detach().clone() if torch.is_tensor(vv) else vv) for kk, vv in st.items()}
                       for k, st in sd["state"].items()}
        return sd

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        # the base class installed fresh tensors; fold them back into the flat momentum
        if self._buckets:
            owned = set()
            for b in self._buckets:
                for p, mv in zip(b.params, b.m_views):
                    owned.add(id(p))
                    st = self.state.get(p)
                    if st is not None and "exp_avg" in st and st["exp_avg"].data_ptr() != mv.data_ptr():
                        mv.copy_(st["exp_avg"])
                        st["exp_avg"] = mv
            stray = any("exp_avg" in self.state.get(p, {}) and id(p) not in owned
                        for g in self.param_groups for p in g["params"])
            if stray:   # the checkpoint knows parameters we have not bucketed yet: re-plan at next step
                self._buckets, self._bucket_sig = None, None

    def declare_dead(self, rank: int) -> None:
        """Call on every surviving rank (same argument) once a peer is known to be gone: its ballots no longer count
        (the majority threshold / averaging denominator shrink) and — on the fused path — nobody waits for its
        progress flags any more, so a step costs no spin on flags that will never arrive.  The reference's README
        advertises robustness to worker drop-out (README.md:2) but ``dist.all_gather`` simply hangs on a missing
        rank (distributed_lion.py:81).  The ``nccl_1bit`` path is a collective and cannot shed a dead rank."""
        W = self.world_size
        if not 0 <= rank < W:
            raise ValueError(f"rank {rank} outside [0, {W})")
        full = (1 << W) - 1
        live = (full if self.live_mask is None else self.live_mask) & ~(1 << rank)
        wait = (full if self.wait_mask is None else self.wait_mask) & ~(1 << rank)
        if live == 0:
            raise ValueError("no live worker left")
        self.live_mask, self.wait_mask = live, wait
        ctx = getattr(self._backend, "_ctx", None)
        if ctx is not None:
            ctx.clear_status()

    def check_health(self) -> None:
        """Raise if the fused exchange kernel recorded a stalled flag wait since the last check
        (synchronises the device; the trainer calls it whenever it logs).  No-op for other back-ends."""
        ctx = getattr(self._backend, "_ctx", None)
        if ctx is not None:
            ctx.check_status()

    @torch.no_grad()
    def resync_replicas(self, src: int = 0) -> None:
        """Collective over ``process_group`` (``src`` is a rank *within that group*): overwrite every replica's
        parameters with rank ``src``'s.  This is how a
        worker that really missed steps rejoins (rejoin = parameter re-broadcast);
        momenta stay per worker, as the algorithm prescribes."""
        from ..parallel import runtime as rt
        self._ensure_buckets()
        if self._buckets:
            rt.broadcast_parameters([b.flat_p for b in self._buckets], src, group=self.process_group)
        else:
            rt.broadcast_parameters([p for g in self.param_groups for p in g["params"]], src, group=self.process_group)

    # ------------------------------------------------------------------ diagnostics
    @torch.no_grad()
    def step_crosscheck(self) -> Dict[str, bool]:
        """Collective diagnostic (call on every rank in place of one ``step()``): take this optimizer step twice —
        through the active back-end on the real buckets, and through the split ``nccl_1bit`` path (pack kernel →
        ``all_gather`` → vote/update kernel) on clones of the same parameters, gradients and momenta — and compare
        the results bit for bit.  Returns ``{"equal": …, "replicas_identical": …}``, both reduced over ranks.

        This is how ``bench.py`` and the multi-GPU tests prove that the fused NVLink kernel computed *the* update
        (reference: distributed_lion.py:61-96) at full size on real hardware, not merely a finite loss.  Costs one
        extra copy of p, g, m."""
        import dataclasses
        from . import backends
        self._resolve_backend()
        self._ensure_buckets()
        kern = bool(self._buckets) and self._kernels_ok(self._buckets[0].device)
        clones = []
        for b in self._buckets:
            gather_grads(b, use_kernel=kern)
            clones.append((b.flat_p.clone(), b.flat_g.clone(), b.flat_m.clone()))
        clip, steps_done = self._pending_clip, self._steps_done
        hp = [(float(self.param_groups[b.group_index]["lr"]), float(self.param_groups[b.group_index]["weight_decay"]),
               *map(float, self.param_groups[b.group_index]["betas"])) for b in self._buckets]
        self.step()
        equal = True
        if self.world_size > 1:
            split = backends.AllGatherBackend(self)
            self._steps_done = steps_done               # same noise streams as the step just taken
            try:
                for b, (p, g, m), (lr, wd, b1, b2) in zip(self._buckets, clones, hp):
                    shadow = dataclasses.replace(b, flat_p=p, flat_g=g, flat_m=m, scratch={})
                    split.step(shadow, lr, wd, b1, b2, [], clip=clip if kern else None, zero_g=False)
                    equal = equal and bool(torch.equal(p, b.flat_p)) and bool(torch.equal(m, b.flat_m))
            finally:
                self._steps_done = steps_done + 1
            dev = self._buckets[0].device if self._buckets else torch.device("cpu")
            flag = torch.tensor([1 if equal else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.process_group)
            equal = bool(int(flag.item()))
        return {"equal": equal, "replicas_identical": self.replicas_identical()}

    @torch.no_grad()
    def replica_checksum(self) -> torch.Tensor:
        """One int64 per bucket: the wrap-around sum of the parameter bits.  Bit-identical replicas — the
        invariant the whole algorithm rests on (every worker applies the same voted update to the same
        parameters) — have identical checksums."""
        self._ensure_buckets()
        sums = []
        for b in self._buckets or []:
            bits = b.flat_p.view({1: torch.int8, 2: torch.int16, 4: torch.int32, 8: torch.int64}[b.flat_p.element_size()])
            sums.append(bits.sum(dtype=torch.int64))
        dev = self._buckets[0].device if self._buckets else torch.device("cpu")
        return torch.stack(sums) if sums else torch.zeros(0, dtype=torch.int64, device=dev)

    @torch.no_grad()
    def replicas_identical(self) -> bool:
        """Collective (call on every rank): do all replicas hold bit-identical parameters?  One small reduction
        per bucket plus two all-reduces of a few integers; cheap enough for every logging step.  The reference
        has no such check: a diverged replica there trains on silently."""
        cs = self.replica_checksum()
        if self.world_size == 1 or cs.numel() == 0:
            return True
        lo, hi = cs.clone(), cs.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.process_group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.process_group)
        return bool(torch.equal(lo, hi))

    @torch.no_grad()
    def vote_statistics(self, chunk_elems: int = 1 << 24) -> Dict[str, float]:
        """Collective, off the hot path: what would the ballot of the CURRENT gradients look like?
        Returns the fraction of coordinates that tie, that are unanimous, and the mean agreement
        ``|2*cnt - L| / L`` (1 = every worker agrees, 0 = perfect split) — the observability the
        reference lacks.  Uses the split path (pack → all_gather → count) on the
        gradients currently held in ``p.grad``; does not modify parameters or momenta."""
        self._resolve_backend()
        self._ensure_buckets()
        W = self.world_size
        L = self.live_count()
        live = self.live_mask
        ties = unanimous = agree = total = 0.0
        for b in self._buckets or []:
            gather_grads(b, use_kernel=self._kernels_ok(b.device))
            beta1 = float(self.param_groups[b.group_index]["betas"][0])
            for off in range(0, b.numel, chunk_elems):
                n = min(chunk_elems, b.numel - off)
                bits = sem.local_bits(b.flat_m[off:off + n], b.flat_g[off:off + n], beta1)
                mine = sem.pack_bits(bits)
                if W > 1:
                    gathered = torch.empty(W, mine.numel(), dtype=torch.uint8, device=mine.device)
                    try:
                        dist.all_gather_into_tensor(gathered.view(-1), mine, group=self.process_group)
                    except (RuntimeError, NotImplementedError):
                        dist.all_gather(list(gathered.unbind(0)), mine, group=self.process_group)
                else:
                    gathered = mine.unsqueeze(0)
                cnt = sem.count_votes(gathered, n, live).to(torch.int32)
                ties += float((cnt * 2 == L).sum())
                unanimous += float(((cnt == 0) | (cnt == L)).sum())
                agree += float((cnt * 2 - L).abs().sum()) / L
                total += n
            # alignment padding holds m = g = 0 on every rank: bit 0 everywhere, i.e. "unanimous" with
            # agreement 1 and never a tie — take it out again so only real coordinates are reported
            pad = b.numel - b.payload
            unanimous -= pad
            agree -= pad
            total -= pad
        total = max(total, 1.0)
        return {"tie_fraction": ties / total, "unanimous_fraction": unanimous / total, "mean_agreement": agree / total,
                "live_workers": float(L), "world_size": float(W)}

    # ------------------------------------------------------------------ helpers for backends
    def cpu_generator(self) -> torch.Generator:
        if self._cpu_gen is None:
            self._cpu_gen = torch.Generator()
            self._cpu_gen.manual_seed(self.seed * 1000003 + self.rank)
        return self._cpu_gen

    def live_count(self) -> int:
        return len(sem.live_ranks(self.world_size, self.live_mask))

    def sr_key(self) -> int:
        """Key of this step's rounding noise (0 = round to nearest)."""
        return sem.sr_key(self.seed, self._steps_done) if self.stochastic_rounding else 0
