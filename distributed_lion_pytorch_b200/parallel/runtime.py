"""Process-group plumbing: one process per GPU, ``torch.distributed`` (NCCL / gloo).

reference: the reference never calls ``init_process_group`` itself — HF ``TrainingArguments``
/ ``accelerate.PartialState`` does it as a side effect (run_clm.py:284, sft_llama2.py:150,
dpo_llama2.py:171) — and relies on DDP's constructor broadcast to make replicas
start identical.  Here both are explicit and there is no DDP wrapper at all: in async mode
nothing ever all-reduces gradients, so a one-off flat parameter broadcast is all that is left.
"""
from __future__ import annotations

import datetime
import os
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist

_DEVICE: Optional[torch.device] = None


def env_world_size() -> int:
    return int(os.environ.get("WORLD_SIZE", "1"))


def init_distributed(device: Optional[str] = None, timeout_s: int = 1800) -> torch.device:
    """Pick this process's device and bring the default process group up when launched by
    torchrun (RANK/WORLD_SIZE/MASTER_* in the environment).  Idempotent."""
    global _DEVICE
    if _DEVICE is not None and (device is None or str(_DEVICE).startswith(device)):
        return _DEVICE
    want_cuda = torch.cuda.is_available() and (device is None or device.startswith("cuda"))
    if want_cuda:
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local % torch.cuda.device_count())
        dev = torch.device("cuda", torch.cuda.current_device())
    else:
        dev = torch.device("cpu")
    if env_world_size() > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "nccl" if dev.type == "cuda" else "gloo"
        kwargs = dict(backend=backend, timeout=datetime.timedelta(seconds=timeout_s))
        if dev.type == "cuda":
            kwargs["device_id"] = dev
        dist.init_process_group(**kwargs)
    _DEVICE = dev
    return dev


def is_dist() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def world_size() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def barrier() -> None:
    if is_dist():
        dist.barrier()


import contextlib


@contextlib.contextmanager
def main_process_first():
    """Rank 0 runs the block first, the others after it (dataset preparation that fills a cache);
    reference: ``training_args.main_process_first`` at run_clm.py:474,531."""
    if is_dist() and rank() != 0:
        dist.barrier()
    try:
        yield
    finally:
        if is_dist() and rank() == 0:
            dist.barrier()


def shutdown() -> None:
    global _DEVICE
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
    _DEVICE = None


@torch.no_grad()
def broadcast_parameters(tensors: Iterable[torch.Tensor], src: int = 0, bucket_bytes: int = 256 << 20,
                         group=None) -> None:
    """Make every replica start from rank ``src``'s values (what DDP's constructor does, X4).  With ``group``,
    ``src`` is a rank *within that group* and only its members take part."""
    if not is_dist():
        return
    pg = group
    gsrc = dist.get_global_rank(pg, src) if pg is not None else src
    group: List[torch.Tensor] = []
    size = 0

    def flush():
        nonlocal group, size
        if not group:
            return
        if len(group) == 1:
            dist.broadcast(group[0], gsrc, group=pg)
        else:
            flat = torch.cat([t.reshape(-1) for t in group])
            dist.broadcast(flat, gsrc, group=pg)
            off = 0
            for t in group:
                t.copy_(flat[off:off + t.numel()].view_as(t))
                off += t.numel()
        group, size = [], 0

    last_kind = None
    for t in tensors:
        kind = (t.dtype, t.device)
        nbytes = t.numel() * t.element_size()
        if last_kind is not None and (kind != last_kind or size + nbytes > bucket_bytes):
            flush()
        group.append(t.data)
        size += nbytes
        last_kind = kind
    flush()


@torch.no_grad()
def all_reduce_mean_(tensors: Iterable[torch.Tensor]) -> None:
    """The synchronous-gradient baseline (what DDP's reducer does): average over ranks."""
    if not is_dist():
        return
    W = dist.get_world_size()
    for t in tensors:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t.div_(W)


def all_reduce_scalar(x: float, op: str = "mean", device: Optional[torch.device] = None) -> float:
    if not is_dist():
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64 if (device or _DEVICE).type == "cpu" else torch.float32,
                     device=device or _DEVICE)
    dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}.get(op, dist.ReduceOp.SUM))
    v = float(t.item())
    return v / dist.get_world_size() if op == "mean" else v
