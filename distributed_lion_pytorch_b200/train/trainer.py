"""Self-contained training loop with the HF ``Trainer`` surface the reference uses.

reference: run_clm.py:588-639 drives ``transformers.Trainer`` (gradient accumulation, local
gradient clipping at ``max_grad_norm=1.0``, LR schedule, logging, evaluation, checkpoint
rotation and resume); async_trainer.py:8-90 overrides ``training_step`` with a
``model.no_sync()`` body so DDP never all-reduces.  ``accelerate`` is not installed on the box,
so HF's Trainer cannot even be constructed; this module re-provides the loop and writes the
same checkpoint directory layout:

    checkpoint-<step>/ model.safetensors config.json optimizer.pt scheduler.pt
                       rng_state[_<rank>].pth trainer_state.json training_args.bin
                       [optimizer_rank<r>.pt]      (opt-in: per-worker momenta, fixes D14)

Gradient synchronisation policy
    ``Trainer``       : synchronous baseline — flat all-reduce(mean) of the gradient buckets
                        before the optimizer step (what DDP does);
    ``AsyncTrainer``  : never touches gradients across ranks; the only cross-rank traffic of a
                        step is the 1-bit sign exchange inside ``Lion.step`` (async_trainer.py:15).
There is no DDP wrapper: replicas are made identical by one flat broadcast at start (X4).
"""
from __future__ import annotations

import json
import logging
import math
import os
import random
import re
import shutil
import time
from dataclasses import asdict, dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..optim.lion import Lion
from ..parallel import runtime as rt
from ..parallel.fault import DropoutSchedule
from ..utils.profiling import StepTimer, nvtx_range
from .args import TrainingArguments
from .data import build_loader, default_data_collator
from .schedules import get_scheduler

logger = logging.getLogger("dlion.trainer")
PREFIX_CHECKPOINT_DIR = "checkpoint"
_re_ckpt = re.compile(r"^" + PREFIX_CHECKPOINT_DIR + r"\-(\d+)$")


def get_last_checkpoint(folder: str) -> Optional[str]:
    """reference: transformers.trainer_utils.get_last_checkpoint used at run_clm.py:289-302"""
    if not os.path.isdir(folder):
        return None
    cks = [p for p in os.listdir(folder) if _re_ckpt.search(p) and os.path.isdir(os.path.join(folder, p))]
    if not cks:
        return None
    return os.path.join(folder, max(cks, key=lambda x: int(_re_ckpt.search(x).group(1))))


def lion_comm_for(args, sync_gradients: bool) -> str:
    """Map ``--comm`` onto a Lion back-end.  ``nccl_allreduce`` is the synchronous-gradient
    baseline: gradients are averaged first, every rank then holds the same gradient, so the
    plain local Lion step keeps replicas identical (SURVEY X5 / BASELINE B2)."""
    if getattr(args, "lion_bucket_mb", None):
        os.environ["DLION_CHUNK_ELEMS"] = str(int(args.lion_bucket_mb) * (1 << 20) // 2)
    if args.comm == "nccl_allreduce" or (sync_gradients and rt.world_size() > 1):
        return "local"
    return args.comm if args.comm in ("auto", "fused", "nccl_1bit", "local") else "auto"


def set_seed(seed: int) -> None:
    random.seed(seed)
    np.random.seed(seed % (2 ** 32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


@dataclass
class TrainerState:
    global_step: int = 0
    epoch: float = 0.0
    max_steps: int = 0
    log_history: List[Dict[str, float]] = field(default_factory=list)
    best_metric: Optional[float] = None
    best_model_checkpoint: Optional[str] = None
    total_flos: float = 0.0
    num_input_tokens_seen: int = 0

    def save_to_json(self, path: str):
        with open(path, "w") as f:
            json.dump(asdict(self), f, indent=2, sort_keys=True)

    @classmethod
    def load_from_json(cls, path: str):
        with open(path) as f:
            return cls(**json.load(f))


@dataclass
class TrainOutput:
    global_step: int
    training_loss: float
    metrics: Dict[str, float]


class LossScaler:
    """Dynamic loss scaling for fp16 training (``--fp16``, or a model held in float16 as dpo_llama2.py:133-139 loads
    it): the loss is multiplied by ``scale`` before backward, the gradients are divided by it before clipping; a step
    whose gradients contain inf / nan is skipped and the scale halved, and after ``growth_interval`` clean steps the
    scale doubles — ``torch.amp.GradScaler``'s policy, restated because GradScaler refuses fp16 *parameters* and
    knows nothing about flat gradient buckets.  One host sync per optimizer step (the skip decision)."""

    def __init__(self, init_scale: float = 2.0 ** 16, growth_interval: int = 2000, min_scale: float = 1.0):
        self.scale, self.growth_interval, self.min_scale = float(init_scale), int(growth_interval), float(min_scale)
        self._good = 0
        self.skipped = 0

    def unscale_and_check(self, grads: List[torch.Tensor]) -> bool:
        """Divide ``grads`` by the scale in place; False when any of them is not finite."""
        if not grads:
            return True
        inv = 1.0 / self.scale
        torch._foreach_mul_(grads, inv)
        norms = torch._foreach_norm(grads)
        return bool(torch.isfinite(torch.stack([n.float() for n in norms])).all().item())

    def update(self, ok: bool) -> None:
        if ok:
            self._good += 1
            if self._good >= self.growth_interval:
                self.scale, self._good = self.scale * 2.0, 0
        else:
            self.scale, self._good = max(self.scale * 0.5, self.min_scale), 0
            self.skipped += 1


class Trainer:
    """Synchronous-gradient trainer (the comparison baseline).  See module docstring."""
    sync_gradients = True

    def __init__(self, model=None, args: Optional[TrainingArguments] = None, train_dataset=None, eval_dataset=None,
                 tokenizer=None, data_collator=None, compute_metrics: Optional[Callable] = None,
                 preprocess_logits_for_metrics: Optional[Callable] = None,
                 optimizers: Tuple[Optional[torch.optim.Optimizer], Optional[Any]] = (None, None),
                 callbacks: Optional[List[Callable]] = None, processing_class=None, **_unused):
        self.args = args or TrainingArguments()
        self.device = self.args.torch_device
        self.model = model.to(self.device) if model is not None else None
        self.train_dataset, self.eval_dataset = train_dataset, eval_dataset
        self.tokenizer = tokenizer or processing_class
        self.data_collator = data_collator or default_data_collator
        self.compute_metrics = compute_metrics
        self.preprocess_logits_for_metrics = preprocess_logits_for_metrics
        self.optimizer, self.lr_scheduler = optimizers
        self.callbacks = callbacks or []
        self.state = TrainerState()
        self.is_world_process_zero = rt.rank() == 0
        self._dropout_schedule: Optional[DropoutSchedule] = None
        self.timer = StepTimer(enabled=bool(getattr(self.args, "profile_steps", False)))
        self._wandb = None
        self._last_grad_norm = None
        self.optimizer_step_events = None       # set to a list to collect (start, end) CUDA events per step
        self.loss_scaler: Optional[LossScaler] = None
        if self.model is not None and self.device.type == "cuda":
            p0 = next(self.model.parameters(), None)
            if self.args.fp16 or (p0 is not None and p0.dtype == torch.float16):
                self.loss_scaler = LossScaler()
        if self.args.gradient_checkpointing and hasattr(self.model, "gradient_checkpointing_enable"):
            self.model.gradient_checkpointing_enable()

    # ------------------------------------------------------------------ optimizer / schedule
    def create_optimizer_and_scheduler(self, num_training_steps: int):
        a = self.args
        if self.optimizer is None:
            params = [p for p in self.model.parameters() if p.requires_grad]
            if a.lion:
                vote = "majority" if a.lion_vote == "stochastic" else a.lion_vote
                mgn = a.lion_max_grad_norm if a.lion_vote != "stochastic" else (a.lion_max_grad_norm or 1.0)
                comm = lion_comm_for(a, self.sync_gradients)
                self.optimizer = Lion(params, lr=a.learning_rate, weight_decay=a.weight_decay, max_grad_norm=mgn,
                                      vote=vote, tie=a.lion_tie, comm=comm, seed=a.seed,
                                      stochastic_rounding=a.lion_stochastic_rounding)
            else:   # reference: run_clm.py:584 (hard-codes weight_decay=0.1 there — D13; we honour the flag)
                self.optimizer = torch.optim.AdamW(params, lr=a.learning_rate, weight_decay=a.weight_decay)
        if self.lr_scheduler is None:
            warm = a.warmup_steps if a.warmup_steps > 0 else math.ceil(num_training_steps * a.warmup_ratio)
            self.lr_scheduler = get_scheduler(a.lr_scheduler_type, self.optimizer, warm, num_training_steps)

    # ------------------------------------------------------------------ data
    def get_train_dataloader(self):
        a = self.args
        return build_loader(self.train_dataset, a.per_device_train_batch_size, rank=rt.rank(), world=rt.world_size(),
                            seed=a.seed, shuffle=True, collate_fn=self.data_collator,
                            num_workers=a.dataloader_num_workers,
                            pin_memory=a.dataloader_pin_memory and self.device.type == "cuda")

    def get_eval_dataloader(self, eval_dataset=None):
        a = self.args
        ds = eval_dataset if eval_dataset is not None else self.eval_dataset
        return build_loader(ds, a.per_device_eval_batch_size, rank=rt.rank(), world=rt.world_size(), seed=a.seed,
                            shuffle=False, collate_fn=self.data_collator, num_workers=0,
                            pin_memory=a.dataloader_pin_memory and self.device.type == "cuda", drop_last=False)

    def _prepare_inputs(self, inputs: Dict[str, Any]) -> Dict[str, Any]:
        return {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v) for k, v in inputs.items()}

    # ------------------------------------------------------------------ the step
    def autocast(self):
        a = self.args
        p = next(self.model.parameters())
        if self.device.type == "cuda" and p.dtype == torch.float32 and (a.bf16 or a.fp16):
            return torch.autocast("cuda", dtype=torch.bfloat16 if a.bf16 else torch.float16)
        import contextlib
        return contextlib.nullcontext()

    def compute_loss(self, model, inputs, return_outputs: bool = False):
        out = model(**inputs)
        loss = out.loss if hasattr(out, "loss") else out[0]
        return (loss, out) if return_outputs else loss

    def training_step(self, model, inputs, num_items_in_batch=None) -> torch.Tensor:
        """One micro-batch: forward, backward (scaled by 1/GA).  Gradients stay local; the base
        class synchronises them once per optimizer step in :meth:`_sync_gradients`."""
        model.train()
        inputs = self._prepare_inputs(inputs)
        with self.timer.phase("forward"), self.autocast():
            loss = self.compute_loss(model, inputs)
        ga = self.args.gradient_accumulation_steps
        with self.timer.phase("backward"):
            scaled = loss / ga if ga > 1 else loss
            if self.loss_scaler is not None:
                scaled = scaled * self.loss_scaler.scale
            scaled.backward()
        return loss.detach() / ga

    def _flat_grads(self) -> List[torch.Tensor]:
        if isinstance(self.optimizer, Lion) and self.optimizer.buckets and self.optimizer.flat_grads:
            owned = {id(p) for b in self.optimizer.buckets for p in b.params}
            ok = all(p.grad is None or id(p) in owned for p in self.model.parameters() if p.requires_grad)
            aliased = all(p.grad is not None and p.grad.data_ptr() == gv.data_ptr()
                          for b in self.optimizer.buckets for p, gv in zip(b.params, b.g_views))
            if ok and aliased:
                return [b.flat_g for b in self.optimizer.buckets]
        return []

    def _sync_gradients(self) -> None:
        if not self.sync_gradients or not rt.is_dist():
            return
        flats = self._flat_grads()
        rt.all_reduce_mean_(flats if flats else [p.grad for p in self.model.parameters() if p.grad is not None])

    def _clip_gradients(self) -> Optional[torch.Tensor]:
        """Local gradient-norm clipping (HF Trainer's ``max_grad_norm``, default 1.0, which the reference inherits).
        With ``Lion`` in steady state this only *measures* the norm (one ``sumsq`` launch per bucket); the scaling
        rides inside the optimizer kernel (``Lion.clip_grad_norm_``), so no clip kernel runs."""
        mg = self.args.max_grad_norm
        if mg is None or mg <= 0:
            return None
        if isinstance(self.optimizer, Lion):
            mine = self._lion_param_ids()
            if all(p.grad is None or id(p) in mine for p in self.model.parameters()):
                return self.optimizer.clip_grad_norm_(mg)
        flats = self._flat_grads()
        if flats and flats[0].is_cuda:
            from ..ops import bucket_ops, require_library
            if require_library():
                return bucket_ops.clip_grad_norm_flat_(flats, mg)
        params = [p for p in self.model.parameters() if p.grad is not None]
        return torch.nn.utils.clip_grad_norm_(params, mg)

    def _unscale_or_skip(self) -> bool:
        """fp16: bring the gradients back to true scale; on overflow drop them, shrink the scale and skip the step.
        With W > 1 every rank must take the same decision (a rank that skipped alone would desynchronise the sign
        exchange), so the verdict is reduced over ranks."""
        flats = self._flat_grads()
        grads = flats if flats else [p.grad for p in self.model.parameters() if p.grad is not None]
        ok = self.loss_scaler.unscale_and_check(grads)
        if rt.is_dist():
            ok = rt.all_reduce_scalar(1.0 if ok else 0.0, "min", self.device) > 0.5
        self.loss_scaler.update(ok)
        if not ok:
            if isinstance(self.optimizer, Lion):
                self.optimizer.zero_grad()
            else:
                self.optimizer.zero_grad(set_to_none=True)
            logger.warning("step %d: non-finite gradients, step skipped, loss scale -> %g", self.state.global_step,
                           self.loss_scaler.scale)
        return ok

    def _lion_param_ids(self) -> set:
        ids = getattr(self, "_lion_ids", None)
        n = sum(len(g["params"]) for g in self.optimizer.param_groups)
        if ids is None or len(ids) != n:
            ids = self._lion_ids = {id(p) for g in self.optimizer.param_groups for p in g["params"]}
        return ids

    def _optimizer_step(self) -> None:
        if isinstance(self.optimizer, Lion):
            if self._dropout_schedule is not None:
                self.optimizer.set_live_mask(self._dropout_schedule.mask(self.state.global_step))
            # this loop always calls zero_grad() right after step(): let the optimizer kernel clear the gradient
            # bucket while it has it in hand, so that zero_grad() has nothing left to do
            self.optimizer.zero_grads_in_step = True
        self.optimizer.step()
        self.lr_scheduler.step()
        if isinstance(self.optimizer, Lion):
            self.optimizer.zero_grad()
        else:
            self.optimizer.zero_grad(set_to_none=True)

    def _check_replicas(self, logs: Dict[str, float]) -> None:
        """Distributed Lion only works while every replica holds the same parameters; verify it whenever we log
        (collective; one checksum per bucket).  ``--check_replicas resync`` repairs a divergence by re-broadcasting
        rank 0's parameters instead of stopping."""
        mode = getattr(self.args, "check_replicas", "off")
        if mode == "off" or rt.world_size() == 1 or self.sync_gradients:
            return
        same = self.optimizer.replicas_identical()
        logs["replicas_identical"] = int(same)
        if same:
            return
        if mode == "resync":
            logger.warning("replicas diverged at step %d: re-broadcasting rank 0's parameters", self.state.global_step)
            self.optimizer.resync_replicas(0)
            return
        raise RuntimeError(f"parameter replicas diverged by step {self.state.global_step} (checksums differ across "
                           "ranks); rerun with --check_replicas resync to repair in place, or off to ignore")

    def train_step(self, micro_batches) -> torch.Tensor:
        """One optimizer step from ``gradient_accumulation_steps`` micro-batches (host or device
        tensors): forward/backward each, then gradient sync (baseline trainers only), local
        clipping, optimizer, schedule, zero_grad.  Returns the step's mean loss as a device tensor
        (no host sync).  ``train()`` is this in a loop plus logging/eval/checkpoints; it is public
        so that custom loops and ``bench.py`` drive exactly the code path ``train()`` runs."""
        if self.optimizer is None or self.lr_scheduler is None:
            raise RuntimeError("call create_optimizer_and_scheduler(num_training_steps) or pass optimizers=(opt, sched)")
        total = None
        for inputs in micro_batches:
            loss = self.training_step(self.model, inputs)
            total = loss if total is None else total + loss
        with self.timer.phase("grad_sync"):
            self._sync_gradients()
        if self.loss_scaler is not None and not self._unscale_or_skip():
            self.state.global_step += 1         # a skipped step still counts (HF semantics); nothing was applied
            return total
        ev = self.optimizer_step_events
        if ev is not None:                      # bench.py: device time of clip + optimizer + zero_grad
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
        with self.timer.phase("clip"):
            self._last_grad_norm = self._clip_gradients()
        with self.timer.phase("optimizer"):
            self._optimizer_step()
        if ev is not None:
            b.record()
            ev.append((a, b))
        self.state.global_step += 1
        return total

    # ------------------------------------------------------------------ train
    def train(self, resume_from_checkpoint: Optional[str] = None, **_) -> TrainOutput:
        a = self.args
        set_seed(a.seed)
        loader, sampler = self.get_train_dataloader()
        ga = a.gradient_accumulation_steps
        try:
            steps_per_epoch = max(len(loader) // ga, 1)
        except TypeError:
            steps_per_epoch = None
        if a.max_steps > 0:
            max_steps = a.max_steps
        elif steps_per_epoch is not None:
            max_steps = math.ceil(a.num_train_epochs * steps_per_epoch)
        else:
            raise ValueError("an iterable dataset needs --max_steps")
        self.state.max_steps = max_steps
        # replicas start identical (DDP constructor broadcast upstream)
        rt.broadcast_parameters(list(self.model.parameters()) + list(self.model.buffers()))
        self.create_optimizer_and_scheduler(max_steps)
        if a.lion and rt.world_size() > 1 and (a.lion_drop_prob > 0 or a.lion_live_mask is not None):
            self._dropout_schedule = DropoutSchedule(rt.world_size(), a.lion_drop_prob, a.seed, a.live_mask_int())
        if resume_from_checkpoint is True:
            resume_from_checkpoint = get_last_checkpoint(a.output_dir)
        if resume_from_checkpoint:
            self._load_checkpoint(resume_from_checkpoint)
        self._init_reporting()
        start_step = self.state.global_step
        tokens_per_micro = None
        world = rt.world_size()
        if self.is_world_process_zero:
            logger.info("***** Running training *****  steps=%d  per-device batch=%d  GA=%d  world=%d  sync_grads=%s",
                        max_steps, a.per_device_train_batch_size, ga, world,
                        self.sync_gradients and world > 1)
        tr_loss = torch.zeros((), device=self.device)
        logged_loss, logged_step = 0.0, start_step
        logged_time = time.time()
        t_start = time.time()
        epoch = start_step // steps_per_epoch if steps_per_epoch else 0
        skip_micro = (start_step % steps_per_epoch) * ga if steps_per_epoch else start_step * ga
        done = self.state.global_step >= max_steps
        grad_norm = None
        while not done:
            if sampler is not None:
                sampler.set_epoch(epoch)
            micro = 0
            pending = []
            for inputs in loader:
                if skip_micro > 0:          # resume: drop the batches the checkpointed run consumed
                    skip_micro -= 1
                    continue
                if tokens_per_micro is None and "input_ids" in inputs:
                    tokens_per_micro = inputs["input_ids"].numel()
                pending.append(inputs)
                micro += 1
                if micro % ga:
                    continue
                tr_loss += self.train_step(pending)
                pending = []
                grad_norm = self._last_grad_norm
                self.state.epoch = epoch + (micro / ga) / steps_per_epoch if steps_per_epoch else 0.0
                if tokens_per_micro:
                    self.state.num_input_tokens_seen += tokens_per_micro * ga * world
                step = self.state.global_step
                if (a.logging_steps and step % a.logging_steps == 0) or (a.logging_first_step and step == 1):
                    cur = float(tr_loss.item())
                    logs = {"loss": round((cur - logged_loss) / max(step - logged_step, 1), 6),
                            "learning_rate": self.lr_scheduler.get_last_lr()[0], "epoch": round(self.state.epoch, 4)}
                    if grad_norm is not None:
                        logs["grad_norm"] = float(grad_norm)
                    if isinstance(self.optimizer, Lion) and self._dropout_schedule is not None:
                        logs["live_workers"] = self._dropout_schedule.live_count(step - 1)
                    now = time.time()                     # tr_loss.item() above synchronised the device
                    if tokens_per_micro and step > logged_step:
                        logs["tokens_per_second"] = round((step - logged_step) * tokens_per_micro * ga * world
                                                          / max(now - logged_time, 1e-9), 1)
                    logged_time = now
                    logged_loss, logged_step = cur, step
                    logs.update(self.timer.summary())          # device-timed phase means (only with --profile_steps)
                    if isinstance(self.optimizer, Lion):
                        self.optimizer.check_health()          # a stalled sign exchange surfaces here, not as a hang
                        self._check_replicas(logs)
                    self.log(logs)
                if a.evaluation_strategy == "steps" and a.eval_steps and step % a.eval_steps == 0 \
                        and self.eval_dataset is not None:
                    self.evaluate()
                if a.save_strategy == "steps" and a.save_steps and step % a.save_steps == 0:
                    self._save_checkpoint()
                if step >= max_steps:
                    done = True
                    break
            else:
                if micro == 0 and skip_micro == 0:
                    raise RuntimeError("the training dataloader produced no batches")
            epoch += 1
            if a.evaluation_strategy == "epoch" and self.eval_dataset is not None and not done:
                self.evaluate()
        if self.device.type == "cuda":
            torch.cuda.synchronize()
        runtime = time.time() - t_start
        steps_run = max(self.state.global_step - start_step, 1)
        train_loss = float(tr_loss.item()) / steps_run
        samples = steps_run * a.per_device_train_batch_size * ga * world
        metrics = {"train_runtime": round(runtime, 4), "train_samples_per_second": round(samples / runtime, 3),
                   "train_steps_per_second": round(steps_run / runtime, 3), "train_loss": train_loss,
                   "epoch": self.state.epoch}
        if tokens_per_micro:
            metrics["train_tokens_per_second"] = round(steps_run * tokens_per_micro * ga * world / runtime, 1)
        if isinstance(self.optimizer, Lion):
            metrics["lion_backend"] = self.optimizer.backend_name
            metrics["lion_wire_bytes_per_step"] = self.optimizer.wire_bytes_per_step()
        self.log(metrics)
        return TrainOutput(self.state.global_step, train_loss, metrics)

    # ------------------------------------------------------------------ evaluate
    @torch.no_grad()
    def evaluate(self, eval_dataset=None, metric_key_prefix: str = "eval") -> Dict[str, float]:
        """Mean loss over the eval set (+ ``compute_metrics`` on argmax predictions, the
        reference's accuracy metric: run_clm.py:562-577).  Results are averaged over ranks."""
        loader, _ = self.get_eval_dataloader(eval_dataset)
        self.model.eval()
        tot_loss, tot_n = 0.0, 0
        preds_all, labels_all = [], []
        t0 = time.time()
        for i, inputs in enumerate(loader):
            if self.args.max_eval_batches is not None and i >= self.args.max_eval_batches:
                break
            inputs = self._prepare_inputs(inputs)
            with self.autocast():
                want_logits = self.compute_metrics is not None
                try:
                    out = self.model(**inputs, return_logits=want_logits)
                except TypeError:
                    out = self.model(**inputs)
            bs = next(iter(inputs.values())).shape[0]
            tot_loss += float(out.loss) * bs
            tot_n += bs
            if want_logits and out.logits is not None:
                lg = out.logits
                if self.preprocess_logits_for_metrics is not None:
                    lg = self.preprocess_logits_for_metrics(lg, inputs.get("labels"))
                preds_all.append(lg.cpu())
                labels_all.append(inputs["labels"].cpu())
        loss = rt.all_reduce_scalar(tot_loss / max(tot_n, 1), "mean", self.device)
        metrics = {f"{metric_key_prefix}_loss": loss, f"{metric_key_prefix}_runtime": round(time.time() - t0, 4),
                   f"{metric_key_prefix}_samples": tot_n * rt.world_size()}
        if self.compute_metrics is not None and preds_all:
            extra = self.compute_metrics((torch.cat(preds_all).numpy(), torch.cat(labels_all).numpy()))
            metrics.update({f"{metric_key_prefix}_{k}": rt.all_reduce_scalar(float(v), "mean", self.device)
                            for k, v in extra.items()})
        self.log(metrics)
        self.model.train()
        return metrics

    # ------------------------------------------------------------------ logging / metrics files
    def _init_reporting(self):
        if "wandb" in (self.args.report_to or "") and self.is_world_process_zero and self._wandb is None:
            try:   # never log in with a baked-in key (the reference does: run_clm.py:59)
                import wandb
                os.environ.setdefault("WANDB_MODE", "offline")
                self._wandb = wandb.init(project=os.environ.get("WANDB_PROJECT", "distributed-lion-b200"),
                                         name=self.args.run_name, config=self.args.to_dict(), reinit=True)
            except Exception as e:   # reporting must never kill training
                logger.warning("wandb disabled: %s", e)

    def log(self, logs: Dict[str, float]) -> None:
        logs = dict(logs)
        logs["step"] = self.state.global_step
        self.state.log_history.append(logs)
        if self.is_world_process_zero:
            logger.info("%s", logs)
            if self._wandb is not None:
                self._wandb.log({k: v for k, v in logs.items() if isinstance(v, (int, float))},
                                step=self.state.global_step)
        for cb in self.callbacks:
            cb(self, logs)

    def log_metrics(self, split: str, metrics: Dict[str, float]) -> None:
        if not self.is_world_process_zero:
            return
        print(f"***** {split} metrics *****")
        width = max((len(k) for k in metrics), default=0)
        for k in sorted(metrics):
            print(f"  {k:<{width}} = {metrics[k]}")

    def save_metrics(self, split: str, metrics: Dict[str, float], combined: bool = True) -> None:
        if not self.is_world_process_zero:
            return
        os.makedirs(self.args.output_dir, exist_ok=True)
        with open(os.path.join(self.args.output_dir, f"{split}_results.json"), "w") as f:
            json.dump(metrics, f, indent=4, sort_keys=True, default=str)
        if combined:
            path = os.path.join(self.args.output_dir, "all_results.json")
            allm = {}
            if os.path.exists(path):
                with open(path) as f:
                    allm = json.load(f)
            allm.update(metrics)
            with open(path, "w") as f:
                json.dump(allm, f, indent=4, sort_keys=True, default=str)

    def create_model_card(self, **kwargs) -> None:
        if not self.is_world_process_zero:
            return
        os.makedirs(self.args.output_dir, exist_ok=True)
        with open(os.path.join(self.args.output_dir, "README.md"), "w") as f:
            f.write("# Model card\n\nTrained with distributed_lion_pytorch_b200.\n\n```json\n"
                    + json.dumps({k: str(v) for k, v in kwargs.items()}, indent=2) + "\n```\n")

    def push_to_hub(self, **kwargs):
        raise RuntimeError("no network access on this box; push_to_hub is unavailable")

    # ------------------------------------------------------------------ checkpoints
    def save_model(self, output_dir: Optional[str] = None) -> None:
        output_dir = output_dir or self.args.output_dir
        if self.is_world_process_zero:
            os.makedirs(output_dir, exist_ok=True)
            if hasattr(self.model, "save_pretrained"):
                self.model.save_pretrained(output_dir)
            else:
                torch.save(self.model.state_dict(), os.path.join(output_dir, "pytorch_model.bin"))
            torch.save(self.args.to_dict(), os.path.join(output_dir, "training_args.bin"))
            if self.tokenizer is not None and hasattr(self.tokenizer, "save_pretrained"):
                self.tokenizer.save_pretrained(output_dir)
        rt.barrier()

    def save_state(self) -> None:
        if self.is_world_process_zero:
            os.makedirs(self.args.output_dir, exist_ok=True)
            self.state.save_to_json(os.path.join(self.args.output_dir, "trainer_state.json"))

    def _rng_file(self, d: str) -> str:
        return os.path.join(d, "rng_state.pth" if rt.world_size() == 1 else f"rng_state_{rt.rank()}.pth")

    def _save_checkpoint(self) -> str:
        a = self.args
        if isinstance(self.optimizer, Lion):
            # never persist parameters that came out of a stalled sign exchange: raises here, before anything is written
            self.optimizer.check_health()
        d = os.path.join(a.output_dir, f"{PREFIX_CHECKPOINT_DIR}-{self.state.global_step}")
        if self.is_world_process_zero:
            os.makedirs(d, exist_ok=True)
        rt.barrier()
        self.save_model(d)
        if self.is_world_process_zero:        # rank 0's momenta only — HF behaviour the reference inherits (D14)
            torch.save(self.optimizer.state_dict(), os.path.join(d, "optimizer.pt"))
            torch.save(self.lr_scheduler.state_dict(), os.path.join(d, "scheduler.pt"))
            self.state.save_to_json(os.path.join(d, "trainer_state.json"))
        if a.per_rank_optimizer_state and rt.world_size() > 1:
            torch.save(self.optimizer.state_dict(), os.path.join(d, f"optimizer_rank{rt.rank()}.pt"))
        rng = {"python": random.getstate(), "numpy": np.random.get_state(), "cpu": torch.random.get_rng_state()}
        if self.device.type == "cuda":
            rng["cuda"] = torch.cuda.get_rng_state(self.device)
        torch.save(rng, self._rng_file(d))
        rt.barrier()
        if self.is_world_process_zero:
            self._rotate_checkpoints()
        return d

    def _rotate_checkpoints(self) -> None:
        lim = self.args.save_total_limit
        if not lim or lim <= 0:
            return
        out = self.args.output_dir
        cks = sorted((int(_re_ckpt.search(p).group(1)), p) for p in os.listdir(out) if _re_ckpt.search(p))
        for _, p in cks[:-lim]:
            shutil.rmtree(os.path.join(out, p), ignore_errors=True)

    def _load_checkpoint(self, d: str) -> None:
        if not os.path.isdir(d):
            raise FileNotFoundError(d)
        from ..models.common import load_checkpoint_into
        load_checkpoint_into(self.model, d)          # single-file or sharded, safetensors or .bin
        if hasattr(self.model, "tie_weights"):
            self.model.tie_weights()
        mine = os.path.join(d, f"optimizer_rank{rt.rank()}.pt")
        opt_file = mine if os.path.exists(mine) else os.path.join(d, "optimizer.pt")
        if os.path.exists(opt_file):
            self.optimizer.load_state_dict(torch.load(opt_file, map_location=self.device, weights_only=False))
        sch = os.path.join(d, "scheduler.pt")
        if os.path.exists(sch):
            self.lr_scheduler.load_state_dict(torch.load(sch, weights_only=False))
        ts = os.path.join(d, "trainer_state.json")
        if os.path.exists(ts):
            self.state = TrainerState.load_from_json(ts)
            if isinstance(self.optimizer, Lion):     # continue the stochastic-sign noise stream where it stopped
                self.optimizer.steps_done = self.state.global_step
        rf = self._rng_file(d)
        if os.path.exists(rf):
            rng = torch.load(rf, weights_only=False)
            random.setstate(rng["python"])
            np.random.set_state(rng["numpy"])
            torch.random.set_rng_state(rng["cpu"])
            if "cuda" in rng and self.device.type == "cuda":
                torch.cuda.set_rng_state(rng["cuda"], self.device)
        if self.is_world_process_zero:
            logger.info("resumed from %s at step %d", d, self.state.global_step)


class NoSyncMixin:
    """``training_step`` that never lets gradients cross ranks: if the model is DDP-wrapped the
    reducer is silenced with ``no_sync`` (exactly the reference's body, async_trainer.py:15);
    otherwise there is nothing to silence.  ``sync_gradients = False`` also switches off this
    package's own flat all-reduce."""
    sync_gradients = False

    def training_step(self, model, inputs, num_items_in_batch=None) -> torch.Tensor:
        no_sync = getattr(model, "no_sync", None)
        if callable(no_sync):
            with no_sync():
                return Trainer.training_step(self, model, inputs, num_items_in_batch)
        return Trainer.training_step(self, model, inputs, num_items_in_batch)


class AsyncTrainer(NoSyncMixin, Trainer):
    """Never synchronises gradients (reference: async_trainer.py:8-34).  Upstream this is a
    ``with model.no_sync():`` around HF's training_step and therefore needs a DDP-wrapped
    model (it crashes on one process); here there is no reducer to silence, so
    the same class works for any world size, and a DDP-wrapped model is still honoured."""
