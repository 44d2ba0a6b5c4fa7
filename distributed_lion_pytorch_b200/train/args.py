"""Argument dataclasses and a small ``HfArgumentParser``-compatible parser.

reference: ``HfArgumentParser((ModelArguments, DataTrainingArguments, AsyncTrainingArguments))``
in run_clm.py:252-258 (including the single-``.json``-file mode), ``AsyncTrainingArguments``
run_clm.py:73-86 (``--lion``, ``--async_grad``).  ``accelerate`` is not installed on the box, so
HF's own ``TrainingArguments`` cannot be instantiated; this one keeps its field names.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
import sys
from dataclasses import dataclass, field, fields
from typing import Any, List, Optional, Sequence, Tuple, Union, get_args, get_origin, get_type_hints


@dataclass
class TrainingArguments:
    output_dir: str = field(default="output", metadata={"help": "where checkpoints and metrics go"})
    overwrite_output_dir: bool = False
    do_train: bool = False
    do_eval: bool = False
    per_device_train_batch_size: int = 8
    per_device_eval_batch_size: int = 8
    gradient_accumulation_steps: int = 1
    learning_rate: float = 5e-5
    weight_decay: float = 0.0
    max_grad_norm: float = 1.0
    num_train_epochs: float = 3.0
    max_steps: int = -1
    lr_scheduler_type: str = "linear"
    warmup_steps: int = 0
    warmup_ratio: float = 0.0
    logging_steps: int = 500
    logging_first_step: bool = False
    save_steps: int = 500
    save_strategy: str = "steps"
    save_total_limit: Optional[int] = None
    eval_steps: Optional[int] = None
    evaluation_strategy: str = "no"
    eval_strategy: Optional[str] = None
    seed: int = 42
    bf16: bool = False
    fp16: bool = False
    gradient_checkpointing: bool = False
    group_by_length: bool = False
    remove_unused_columns: bool = True
    dataloader_num_workers: int = 0
    dataloader_pin_memory: bool = True
    optim: str = "adamw_torch"
    report_to: str = "none"
    run_name: Optional[str] = None
    resume_from_checkpoint: Optional[str] = None
    push_to_hub: bool = False
    ddp_find_unused_parameters: Optional[bool] = None
    log_level: str = "passive"
    max_eval_batches: Optional[int] = None
    # ---- Distributed Lion (reference flags, run_clm.py:75-86, + the new knobs of SURVEY §5.6)
    lion: bool = field(default=False, metadata={"help": "use the Lion optimizer"})
    async_grad: bool = field(default=False, metadata={"help": "never synchronise gradients; exchange signs"})
    lion_vote: str = field(default="majority", metadata={"help": "majority | average | stochastic"})
    lion_tie: str = field(default="neg", metadata={"help": "neg (reference) | zero (paper)"})
    lion_stochastic_rounding: bool = field(default=False, metadata={
        "help": "write bf16 parameters with stochastic rounding (same noise on every rank; not in the reference)"})
    lion_max_grad_norm: Optional[float] = field(default=None, metadata={"help": "range for stochastic binarisation"})
    lion_drop_prob: float = field(default=0.0, metadata={"help": "simulated per-step worker drop-out probability"})
    lion_live_mask: Optional[str] = field(default=None, metadata={"help": "fixed live-worker mask, e.g. 0b1011"})
    comm: str = field(default="auto", metadata={"help": "auto | fused | nccl_1bit | nccl_allreduce"})
    lion_bucket_mb: Optional[int] = field(default=None, metadata={"help": "MiB of bf16 parameters per fused launch"})
    profile_steps: bool = field(default=False, metadata={"help": "log device-timed forward/backward/optimizer ms"})
    per_rank_optimizer_state: bool = field(default=False, metadata={"help": "also save optimizer_rank<r>.pt"})
    check_replicas: str = field(default="raise", metadata={
        "help": "at every logging step compare parameter checksums across ranks: raise | resync | off"})
    synthetic_data: bool = field(default=False, metadata={"help": "random tokens of the right shape"})
    synthetic_samples: int = 4096
    device: Optional[str] = None

    def __post_init__(self):
        if self.eval_strategy is not None:
            self.evaluation_strategy = self.eval_strategy
        if isinstance(self.report_to, (list, tuple)):
            self.report_to = ",".join(self.report_to)

    # -- process / device topology (HF: touching ``.device`` brings the process group up, run_clm.py:284)
    @property
    def world_size(self) -> int:
        return int(os.environ.get("WORLD_SIZE", "1"))

    @property
    def process_index(self) -> int:
        return int(os.environ.get("RANK", "0"))

    @property
    def local_process_index(self) -> int:
        return int(os.environ.get("LOCAL_RANK", "0"))

    @property
    def local_rank(self) -> int:
        return self.local_process_index if self.world_size > 1 else -1

    @property
    def should_save(self) -> bool:
        return self.process_index == 0

    @property
    def should_log(self) -> bool:
        return self.process_index == 0

    @property
    def n_gpu(self) -> int:
        return 1

    @property
    def torch_device(self):
        import torch
        from ..parallel.runtime import init_distributed
        return init_distributed(self.device)

    def to_dict(self):
        return {f.name: getattr(self, f.name) for f in fields(self)}

    def to_json_string(self):
        return json.dumps(self.to_dict(), indent=2, default=str)

    def get_process_log_level(self):
        import logging
        return logging.INFO if self.should_log else logging.WARNING

    def live_mask_int(self) -> Optional[int]:
        if self.lion_live_mask is None:
            return None
        return int(str(self.lion_live_mask), 0)


@dataclass
class AsyncTrainingArguments(TrainingArguments):
    """Name kept from the reference (run_clm.py:73-86); the flags live on the base class here."""


# --------------------------------------------------------------------------- parser
def _str2bool(v):
    if isinstance(v, bool):
        return v
    s = str(v).lower()
    if s in ("1", "true", "t", "yes", "y"):
        return True
    if s in ("0", "false", "f", "no", "n"):
        return False
    raise argparse.ArgumentTypeError(f"expected a boolean, got {v!r}")


def _base_type(tp):
    origin = get_origin(tp)
    if origin is Union:
        args = [a for a in get_args(tp) if a is not type(None)]
        return _base_type(args[0]) if args else str
    if origin in (list, List):
        return ("list", _base_type(get_args(tp)[0]) if get_args(tp) else str)
    return tp


class HfArgumentParser(argparse.ArgumentParser):
    """``--field value`` / ``--field=value`` / bare ``--flag`` for bools, one group per dataclass."""

    def __init__(self, dataclass_types, **kw):
        kw.setdefault("formatter_class", argparse.ArgumentDefaultsHelpFormatter)
        kw.setdefault("allow_abbrev", False)
        super().__init__(**kw)
        if dataclasses.is_dataclass(dataclass_types):
            dataclass_types = [dataclass_types]
        self.dataclass_types = list(dataclass_types)
        seen = set()
        for dt in self.dataclass_types:
            hints = get_type_hints(dt)
            for f in fields(dt):
                if not f.init or f.name in seen:
                    continue
                seen.add(f.name)
                tp = _base_type(hints.get(f.name, str))
                default = f.default if f.default is not dataclasses.MISSING else (
                    f.default_factory() if f.default_factory is not dataclasses.MISSING else None)
                kwargs = dict(help=f.metadata.get("help", ""), dest=f.name, default=default)
                if tp is bool:
                    kwargs.update(type=_str2bool, nargs="?", const=True)
                elif isinstance(tp, tuple) and tp[0] == "list":
                    kwargs.update(type=tp[1], nargs="+")
                else:
                    kwargs.update(type=tp if callable(tp) else str)
                self.add_argument(f"--{f.name}", **kwargs)

    def _build(self, ns: dict, allow_extra=False):
        outs = []
        used = set()
        for dt in self.dataclass_types:
            names = {f.name for f in fields(dt) if f.init}
            outs.append(dt(**{k: v for k, v in ns.items() if k in names}))
            used |= names
        extra = {k: v for k, v in ns.items() if k not in used}
        if extra and not allow_extra:
            raise ValueError(f"unknown arguments: {sorted(extra)}")
        return tuple(outs)

    def parse_args_into_dataclasses(self, args: Optional[Sequence[str]] = None, return_remaining_strings=False):
        ns, remaining = self.parse_known_args(args=args)
        outs = self._build(vars(ns))
        if return_remaining_strings:
            return (*outs, remaining)
        if remaining:
            raise ValueError(f"Some specified arguments are not used by the HfArgumentParser: {remaining}")
        return outs

    def parse_json_file(self, json_file: str, allow_extra_keys: bool = False):
        with open(json_file) as f:
            data = json.load(f)
        defaults = vars(self.parse_args([]))
        defaults.update(data)
        return self._build(defaults, allow_extra=allow_extra_keys)

    def parse_dict(self, d: dict, allow_extra_keys: bool = False):
        defaults = vars(self.parse_args([]))
        defaults.update(d)
        return self._build(defaults, allow_extra=allow_extra_keys)


def parse_cli(dataclass_types, argv: Optional[Sequence[str]] = None):
    """CLI or a single ``.json`` path (reference: run_clm.py:253-258)."""
    argv = list(sys.argv[1:] if argv is None else argv)
    parser = HfArgumentParser(dataclass_types)
    if len(argv) == 1 and argv[0].endswith(".json"):
        return parser.parse_json_file(os.path.abspath(argv[0]))
    return parser.parse_args_into_dataclasses(argv)
