"""Fake-cluster harness: W processes on 127.0.0.1 over gloo (CPU) or nccl (one GPU each)."""
import os
import socket
import sys
import traceback
from pathlib import Path

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _portable(x):
    """Tensors cannot ride an mp.Queue once their producer has exited: ship numpy instead."""
    if torch.is_tensor(x):
        t = x.detach().cpu()
        if t.dtype in (torch.bfloat16, torch.float16):
            return ("__tensor__", t.float().numpy(), str(t.dtype))
        return ("__tensor__", t.numpy(), str(t.dtype))
    if isinstance(x, dict):
        return {k: _portable(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_portable(v) for v in x)
    if isinstance(x, torch.dtype):
        return ("__dtype__", str(x))
    return x


def _unportable(x):
    if isinstance(x, tuple) and len(x) == 3 and x[0] == "__tensor__":
        return torch.from_numpy(x[1]).to(getattr(torch, x[2].split(".")[-1]))
    if isinstance(x, tuple) and len(x) == 2 and x[0] == "__dtype__":
        return getattr(torch, x[1].split(".")[-1])
    if isinstance(x, dict):
        return {k: _unportable(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_unportable(v) for v in x)
    return x


def _entry(rank, world, port, backend, fn, args, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    if str(ROOT) not in sys.path:
        sys.path.insert(0, str(ROOT))
    torch.set_num_threads(1)
    try:
        if backend == "nccl":
            torch.cuda.set_device(rank)
        dist.init_process_group(backend, rank=rank, world_size=world)
        out = fn(rank, world, *args)
        q.put((rank, "ok", _portable(out)))
    except Exception:
        q.put((rank, "err", traceback.format_exc()))
    finally:
        if dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass


def run_distributed(fn, world: int, *args, backend: str = "gloo", timeout: float = 180.0, _retries: int = 2):
    """Run ``fn(rank, world, *args)`` in ``world`` spawned processes; returns results by rank.
    A rendezvous that loses the race for its TCP port (seen once on an 8-GPU box) is retried."""
    try:
        return _run_once(fn, world, args, backend, timeout)
    except RuntimeError as e:
        msg = str(e)
        if _retries > 0 and ("TCPStore" in msg or "EADDRINUSE" in msg or "address already in use" in msg.lower()):
            return run_distributed(fn, world, *args, backend=backend, timeout=timeout, _retries=_retries - 1)
        raise


def _run_once(fn, world, args, backend, timeout):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, backend, fn, args, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world):
            rank, status, out = q.get(timeout=timeout)
            if status != "ok":
                raise RuntimeError(f"rank {rank} failed:\n{out}")
            results[rank] = _unportable(out)
    finally:
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world)]
