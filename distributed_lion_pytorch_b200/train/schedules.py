"""LR schedules with HF's formulas (reference: get_cosine_schedule_with_warmup at
run_clm.py:582, sft_llama2.py:165, dpo_llama2.py:211).  Pure ``LambdaLR``: the optimizer only
sees ``group['lr']`` change between steps, which the Lion kernels take as a launch scalar."""
from __future__ import annotations

import math
from functools import partial

from torch.optim.lr_scheduler import LambdaLR


def _cosine(step, *, warmup, total, cycles):
    if step < warmup:
        return step / max(1, warmup)
    prog = (step - warmup) / max(1, total - warmup)
    return max(0.0, 0.5 * (1.0 + math.cos(math.pi * cycles * 2.0 * prog)))


def _linear(step, *, warmup, total):
    if step < warmup:
        return step / max(1, warmup)
    return max(0.0, (total - step) / max(1, total - warmup))


def _constant(step, *, warmup):
    return step / max(1, warmup) if step < warmup else 1.0


def get_cosine_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps, num_cycles=0.5, last_epoch=-1):
    return LambdaLR(optimizer, partial(_cosine, warmup=num_warmup_steps, total=num_training_steps, cycles=num_cycles),
                    last_epoch)


def get_linear_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps, last_epoch=-1):
    return LambdaLR(optimizer, partial(_linear, warmup=num_warmup_steps, total=num_training_steps), last_epoch)


def get_constant_schedule_with_warmup(optimizer, num_warmup_steps, last_epoch=-1):
    return LambdaLR(optimizer, partial(_constant, warmup=num_warmup_steps), last_epoch)


def get_scheduler(name, optimizer, num_warmup_steps, num_training_steps):
    name = str(name).lower()
    if name == "cosine":
        return get_cosine_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps)
    if name == "linear":
        return get_linear_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps)
    if name in ("constant", "constant_with_warmup"):
        return get_constant_schedule_with_warmup(optimizer, num_warmup_steps if name != "constant" else 0)
    raise ValueError(f"unknown lr_scheduler_type {name}")
