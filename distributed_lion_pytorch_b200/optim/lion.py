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
  ``stochastic_rounding`` (bf16 parameter writes; off by default = the reference's arithmetic).

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
                 flat_grads: bool = True, use_kernels: Optional[bool] = None, stochastic_rounding: bool = False):
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
        self._buckets: Optional[List[Bucket]] = None
        self._bucket_sig = None
        self._frozen_actives: List[list] = []
        self._backend = None            # resolved at first step
        self._steps_done = 0
        self._cpu_gen: Optional[torch.Generator] = None
        self.last_step_stats: Dict[str, float] = {}

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
        for b in self._buckets:
            _, missing = gather_grads(b, use_kernel=self._kernels_ok(b.device))
            group = self.param_groups[b.group_index]
            beta1, beta2 = group["betas"]
            self._backend.step(b, float(group["lr"]), float(group["weight_decay"]), float(beta1), float(beta2),
                               missing)
            if self.flat_grads:
                alias_grads(b)
        self._steps_done += 1
        return loss

    def zero_grad(self, set_to_none: bool = True) -> None:
        """With flat gradient buckets the gradients are cleared by one memset per bucket and
        ``p.grad`` keeps aliasing the bucket, so the next backward accumulates straight into
        the flat tensor the optimizer kernel reads."""
        if self._buckets and self.flat_grads:
            owned = set()
            for b in self._buckets:
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

    def check_health(self) -> None:
        """Raise if the fused exchange kernel recorded a stalled flag wait since the last check
        (synchronises the device; the trainer calls it whenever it logs).  No-op for other back-ends."""
        ctx = getattr(self._backend, "_ctx", None)
        if ctx is not None:
            ctx.check_status()

    @torch.no_grad()
    def resync_replicas(self, src: int = 0) -> None:
        """Collective: overwrite every replica's parameters with rank ``src``'s.  This is how a
        worker that really missed steps rejoins (rejoin = parameter re-broadcast);
        momenta stay per worker, as the algorithm prescribes."""
        from ..parallel import runtime as rt
        self._ensure_buckets()
        if self._buckets:
            rt.broadcast_parameters([b.flat_p for b in self._buckets], src)
        else:
            rt.broadcast_parameters([p for g in self.param_groups for p in g["params"]], src)

    # ------------------------------------------------------------------ diagnostics
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
