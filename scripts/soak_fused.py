#!/usr/bin/env python
"""Randomised-skew soak of the fused exchange kernel on real GPUs — the stand-in for `compute-sanitizer --tool
racecheck`, which is closed on this pool.  One process per GPU under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 scripts/soak_fused.py \
        [--launches 2000] [--out profiles/soak_w8.json]

For each configuration (grid of 3 / 37 / all CTAs, unicast and NVLS multicast send path, 1 Mi / 2 Mi-element chunks)
an optimizer with two parameter groups and bf16 + fp32 buckets of different sizes takes steps until the requested
number of kernel launches is reached; before every step each rank sleeps a random 0-3 ms (ranks enter the kernel at
different times, in a different order every step); gradient clip and zero-fill ride inside the kernel.  Every 50
steps the step is taken through `Lion.step_crosscheck()`: the fused result must equal the split pack / all_gather /
vote path bit for bit on every rank and the replica checksums must agree; the kernel's status word must stay clean.
"""
import argparse
import json
import os
import random
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion import Lion  # noqa: E402
from distributed_lion_pytorch_b200.parallel import runtime as rt  # noqa: E402


def make_params(dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    mk = lambda shape, dt: torch.nn.Parameter((torch.randn(*shape, device=dev, generator=g) * 0.05).to(dt))  # noqa: E731
    g0 = [mk((1024, 2048), torch.bfloat16), mk((301, 1003), torch.float32), mk((513, 2047), torch.bfloat16), mk((77,), torch.bfloat16)]
    g1 = [mk((256, 4096), torch.bfloat16), mk((555, 33), torch.float32)]
    return g0, g1


def run_config(dev, rank, world, grid, chunk, multicast, launches, seed):
    os.environ["DLION_MULTICAST"] = "1" if multicast else "0"
    os.environ["DLION_CHUNK_ELEMS"] = str(chunk)
    if grid:
        os.environ["DLION_FUSED_GRID"] = str(grid)
    else:
        os.environ.pop("DLION_FUSED_GRID", None)
    g0, g1 = make_params(dev, seed=0)
    opt = Lion([dict(params=g0, lr=1e-3, weight_decay=0.1), dict(params=g1, lr=3e-4, weight_decay=0.0)], comm="fused",
               zero_grads_in_step=True)
    params = g0 + g1
    gen = torch.Generator(device=dev).manual_seed(1000 + rank + 17 * seed)
    rnd = random.Random(seed * 101 + rank)
    steps = checks = 0
    ok = True
    t0 = time.time()
    while True:
        for p in params:
            gr = (torch.randn(p.shape, device=dev, generator=gen) * (1e-3 if steps % 3 else 1e-5)).to(p.dtype)
            if p.grad is None:
                p.grad = gr
            else:
                p.grad.copy_(gr)
        torch.cuda.synchronize()
        time.sleep(rnd.random() * 0.003)
        opt.clip_grad_norm_(1.0)
        if steps % 50 == 49:
            chk = opt.step_crosscheck()
            ok = ok and chk["equal"] and chk["replicas_identical"]
            checks += 1
        else:
            opt.step()
        opt.zero_grad()
        steps += 1
        ctx = opt._backend._ctx
        if ctx.epoch >= launches and steps % 50 == 0:
            break
    torch.cuda.synchronize()
    ctx.check_status()
    res = {"grid": ctx.grid_for(torch.bfloat16), "chunk_elems": chunk, "multicast": bool(ctx.uses_multicast),
           "steps": steps, "launches": ctx.epoch, "crosschecks": checks, "all_equal": bool(ok),
           "buckets": [(str(b.dtype), b.numel) for b in opt.buckets], "seconds": round(time.time() - t0, 1)}
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    res["all_equal"] = bool(int(flag))
    del opt, params, g0, g1
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches", type=int, default=2000)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dev = rt.init_distributed()
    rank, world = rt.rank(), rt.world_size()
    configs = [(3, 1 << 20, False), (37, 1 << 20, True), (0, 1 << 21, False), (0, 1 << 21, True)]
    per = max(a.launches // len(configs), 100)
    rows = []
    for i, (grid, chunk, mc) in enumerate(configs):
        r = run_config(dev, rank, world, grid, chunk, mc, per, seed=i + 1)
        rows.append(r)
        if rank == 0:
            print(json.dumps(r), flush=True)
    ok = all(r["all_equal"] for r in rows)
    if rank == 0:
        summary = {"world": world, "total_launches": sum(r["launches"] for r in rows), "ok": ok, "configs": rows}
        print(json.dumps({"soak_ok": ok, "world": world, "total_launches": summary["total_launches"]}))
        if a.out:
            Path(a.out).parent.mkdir(parents=True, exist_ok=True)
            Path(a.out).write_text(json.dumps(summary, indent=1))
    rt.shutdown()
    return 0 if ok else 4


if __name__ == "__main__":
    sys.exit(main())
