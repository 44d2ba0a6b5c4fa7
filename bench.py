#!/usr/bin/env python
"""Headline benchmark: tokens/s of GPT-2 small Lion training (BASELINE.json config 2).

    python bench.py --gpus N --steps K --warmup W [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

One "step" is one optimizer step of the README recipe (reference: README.md:19-37): per-device
batch 20 x 1024 tokens, gradient accumulation 8, bf16 parameters, local grad clipping at 1.0,
majority-vote Distributed Lion (lr 1e-4, wd 0.1, cosine schedule), gradients never synchronised.
Synthetic tokens, random-init weights.  Rank 0 prints ONE JSON line (see the task contract).

  value      device-timed (CUDA events, max over ranks) with the step's batches already on the GPU;
  e2e.value  the same steps through the public trainer API with every micro-batch copied from
             pinned host memory and the step loss read back to the host.

--impl reference runs the UNMODIFIED reference optimizer (baseline/_ref/distributed_lion.py) on
HF transformers' GPT2LMHeadModel with the no_sync loop of the reference's AsyncTrainer.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "tokens/sec/box GPT-2 small Lion training (per-device batch 20x1024, grad-accum 8, bf16)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="gpt2")
    ap.add_argument("--batch", type=int, default=20)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--ga", type=int, default=8)
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the post-run fused == split / replica check")
    ap.add_argument("--lora", action="store_true", help="non-default: LoRA on q_proj/v_proj (the sft_llama2 arrangement)")
    ap.add_argument("--load-in-4bit", action="store_true", help="non-default: NF4 frozen base under --lora")
    return ap.parse_args()


from distributed_lion_pytorch_b200.utils.profiling import ClockSampler  # noqa: E402


def make_batches(n_micro, batch, seq, vocab, seed, pin):
    import torch
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_micro):
        ids = torch.randint(0, vocab, (batch, seq), generator=g, dtype=torch.int64)
        lab = ids.clone()
        if pin:
            ids, lab = ids.pin_memory(), lab.pin_memory()
        out.append({"input_ids": ids, "labels": lab})
    return out


def timed_region(step_fn, steps, warmup, device, e2e: bool):
    """Returns (seconds for `steps` steps, per-step ms list).  Device mode: CUDA events around
    every step, an L2-sized scratch write between steps; e2e mode: wall clock between syncs."""
    import torch
    import torch.distributed as dist
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    for i in range(warmup):
        step_fn(i)
    torch.cuda.synchronize(device)
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize(device)
    per_step = []
    t0 = time.perf_counter()
    if e2e:
        for i in range(steps):
            step_fn(warmup + i)
        torch.cuda.synchronize(device)
        total = time.perf_counter() - t0
    else:
        evs = []
        for i in range(steps):
            flush.fill_(i & 0xff)                       # evict L2 between timed steps (outside the event pair)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step_fn(warmup + i)
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize(device)
        per_step = [a.elapsed_time(b) for a, b in evs]
        total = sum(per_step) / 1e3
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize(device)
    return total, per_step


def load_peaks():
    """Roofline denominators: the driver-measured MEASURED_PEAKS.json, else the profiling recipe's fallback."""
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            d = json.loads(f.read_text())
            return {"hbm_gbs": float(d["hbm_gbs"]), "source": "measured"}
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "source": "fallback"}


NVLINK_GBS = 770.0      # measured peer-copy rate per direction per GPU (B200_PROFILING.md)


def optimizer_step_report(events, n_params, elem_size, world, device):
    """Device time of clip + optimizer + zero_grad per step (CUDA events inside the timed steps, max over ranks)
    against its roofline: the slower of the HBM bytes at the measured copy peak and the packed sign bits over
    NVLink.  HBM bytes per parameter: g is read twice (the norm must be known before any element is scaled) and
    cleared once, m and p are read and written once: 7 * elem_size, plus (W + 1) / 8 bytes of sign words."""
    if not events:
        return None
    us = sum(a.elapsed_time(b) for a, b in events) / len(events) * 1e3
    us = max_over_ranks(us, device)
    peaks = load_peaks()
    hbm_bytes = n_params * (7 * elem_size + ((world + 1) / 8 if world > 1 else 0))
    nvl_bytes = n_params / 8 * (world - 1)
    roof_us = max(hbm_bytes / (peaks["hbm_gbs"] * 1e9), nvl_bytes / (NVLINK_GBS * 1e9)) * 1e6
    return {"optimizer_step_us": round(us, 1), "optimizer_step_roofline_us": round(roof_us, 1),
            "optimizer_step_roofline_frac": round(roof_us / us, 3), "optimizer_step_includes": "grad-norm + clip + "
            "sign exchange + vote + update + momentum + gradient zero-fill", "roofline_peaks": peaks["source"],
            "optimizer_params": int(n_params)}


def max_over_ranks(x, device):
    import torch
    import torch.distributed as dist
    if not dist.is_initialized():
        return x
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ----------------------------------------------------------------------------- our arm
def build_ours(a, device):
    import torch
    from async_trainer import AsyncTrainer
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.parallel import runtime as rt
    from distributed_lion_pytorch_b200.train import TrainingArguments, get_cosine_schedule_with_warmup
    torch.manual_seed(1234)
    model = build_model(a.model, dtype=torch.bfloat16, device=device, pretrained=False)
    if a.load_in_4bit:
        from distributed_lion_pytorch_b200.models import quantize_base_
        quantize_base_(model)
    if a.lora:
        from distributed_lion_pytorch_b200.models import LoraConfig, apply_lora
        apply_lora(model, LoraConfig(target_modules=["q_proj", "v_proj"]))
    targs = TrainingArguments(output_dir="/tmp/dlion_bench", per_device_train_batch_size=a.batch,
                              gradient_accumulation_steps=a.ga, learning_rate=1e-4, weight_decay=0.1,
                              max_grad_norm=1.0, lion=True, async_grad=True, comm=a.comm, max_steps=100000,
                              warmup_steps=2000)
    comm = a.comm if a.comm in ("auto", "fused", "nccl_1bit", "local") else "auto"
    opt = Lion([p for p in model.parameters() if p.requires_grad], lr=1e-4, weight_decay=0.1, comm=comm)
    sched = get_cosine_schedule_with_warmup(opt, 2000, 100000)
    trainer = AsyncTrainer(model=model, args=targs, optimizers=(opt, sched))
    rt.broadcast_parameters(list(model.parameters()) + list(model.buffers()))
    return model, opt, trainer


def run_ours(a, device, rank, world):
    import torch
    from distributed_lion_pytorch_b200.ops import _cabi
    model, opt, trainer = build_ours(a, device)
    vocab = model.config.vocab_size
    n_steps = a.warmup + a.steps
    host = make_batches(a.ga * min(n_steps, 4), a.batch, a.seq, vocab, seed=100 + rank, pin=True)
    dev = [{k: v.to(device) for k, v in b.items()} for b in host]
    last = {}

    def step_device(i):      # batches already resident on the GPU, no read-back
        trainer.train_step([dev[(i * a.ga + j) % len(dev)] for j in range(a.ga)])

    def step_e2e(i):         # the public API as a user calls it: pinned host batches in, the step's loss out
        loss = trainer.train_step([host[(i * a.ga + j) % len(host)] for j in range(a.ga)])   # H2D copies inside
        last["loss"] = float(loss.item())                                                    # D2H read of the loss

    sampler = ClockSampler(torch.cuda.current_device())
    n0 = _cabi.LAUNCHES
    sampler.start()
    opt_events = []

    def step_device_timed(i):
        trainer.optimizer_step_events = opt_events if i >= a.warmup else None
        step_device(i)

    total, per_step = timed_region(step_device_timed, a.steps, a.warmup, device, e2e=False)
    trainer.optimizer_step_events = None
    clocks = sampler.stop()
    launches = (_cabi.LAUNCHES - n0) * a.steps // (a.steps + a.warmup)
    e2e_total = None
    if not a.no_e2e:
        e2e_total, _ = timed_region(step_e2e, a.steps, 1, device, e2e=True)
    tokens_per_step = a.batch * a.seq * a.ga * world
    total = max_over_ranks(total, device)
    res = {"ms_per_step": total / a.steps * 1e3, "value": tokens_per_step * a.steps / total,
           "gpu_launches": launches, "clocks": clocks, "backend": opt.backend_name, "loss": last.get("loss"),
           "peak_mem_gb": round(torch.cuda.max_memory_allocated(device) / 2 ** 30, 2)}
    if e2e_total is not None:
        e2e_total = max_over_ranks(e2e_total, device)
        h2d = a.ga * 2 * a.batch * a.seq * 8
        res["e2e"] = {"value": tokens_per_step * a.steps / e2e_total, "unit": "tokens/s",
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "ms_per_step": e2e_total / a.steps * 1e3}
    n_params = sum(b.numel for b in opt.buckets)
    esize = opt.buckets[0].flat_p.element_size() if opt.buckets else 2
    rep = optimizer_step_report(opt_events, n_params, esize, world, device)
    if rep:
        res.update(rep)
    if hasattr(opt._backend, "_ctx") and opt._backend._ctx is not None:
        opt._backend._ctx.check_status()
        res["nvls_multicast"] = bool(opt._backend._ctx.uses_multicast)
    # Proof, outside the timed region, that the step the numbers above come from is THE update: one more optimizer
    # step taken twice — through the active back-end and through the split pack / all_gather / vote path on clones —
    # compared bit for bit on every rank, plus the replica checksums (reference: distributed_lion.py:61-96).
    if world > 1 and not a.no_verify:
        for j in range(a.ga):
            trainer.training_step(trainer.model, dev[j % len(dev)])
        trainer._clip_gradients()
        chk = opt.step_crosscheck()
        opt.zero_grad()
        res["fused_equals_split"] = chk["equal"]
        res["replicas_identical"] = chk["replicas_identical"]
    return res


# ----------------------------------------------------------------------------- reference arm
def run_reference(a, device, rank, world):
    """Stock path: HF GPT2LMHeadModel + the reference's Lion, driven by the loop of the reference's
    AsyncTrainer.training_step (`with model.no_sync(): loss = model(**inputs).loss; backward`) —
    HF's Trainer itself cannot be constructed here (accelerate / trl are not installed)."""
    import importlib.util
    import torch
    import torch.distributed as dist
    ref_dir = ROOT / "baseline" / "_ref"
    f = ref_dir / "distributed_lion.py"
    if not f.exists():
        return {"unavailable": "baseline/_ref/distributed_lion.py missing (reference not installed)"}
    spec = importlib.util.spec_from_file_location("ref_distributed_lion", f)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    try:
        from transformers import GPT2Config, GPT2LMHeadModel, get_cosine_schedule_with_warmup
    except Exception as e:
        return {"unavailable": f"transformers not importable: {e}"}
    if a.lora or a.load_in_4bit:
        return {"unavailable": "reference arm: peft / bitsandbytes are not installed on this image"}
    torch.manual_seed(1234)
    if a.model == "gpt2":
        cfg = GPT2Config()
        model = GPT2LMHeadModel(cfg).to(device=device, dtype=torch.bfloat16)
    elif a.model == "llama-2-7b":      # non-default: the second model BASELINE names (reference: sft_llama2.py:147)
        from transformers import LlamaConfig, LlamaForCausalLM
        cfg = LlamaConfig()            # transformers' defaults are the Llama-2-7B shape
        old_dtype = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            with torch.device(device):
                model = LlamaForCausalLM(cfg)
        finally:
            torch.set_default_dtype(old_dtype)
    else:
        return {"unavailable": f"reference arm implements gpt2 and llama-2-7b, not {a.model}"}
    model.train()
    net = model
    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[device.index])
    opt = ref.Lion(model.parameters(), lr=1e-4, weight_decay=0.1)      # picks update_fn_distributed when W > 1
    sched = get_cosine_schedule_with_warmup(opt, 2000, 100000)
    n_steps = a.warmup + a.steps
    host = make_batches(a.ga * min(n_steps, 4), a.batch, a.seq, cfg.vocab_size, seed=100 + rank, pin=True)
    dev = [{k: v.to(device) for k, v in b.items()} for b in host]
    last = {}

    def micro(batch):
        if world > 1:
            with net.no_sync():                                         # reference: async_trainer.py:15
                loss = net(**batch).loss
                (loss / a.ga).backward()
        else:
            loss = net(**batch).loss
            (loss / a.ga).backward()
        return loss.detach() / a.ga

    opt_events = []
    rec = {"on": False}

    def finish():
        if rec["on"]:
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)         # HF Trainer default max_grad_norm
        opt.step()
        sched.step()
        model.zero_grad()
        if rec["on"]:
            eb.record()
            opt_events.append((ea, eb))

    def step_device(i):
        for j in range(a.ga):
            micro(dev[(i * a.ga + j) % len(dev)])
        rec["on"] = i >= a.warmup
        finish()
        rec["on"] = False

    def step_e2e(i):
        tot = None
        for j in range(a.ga):
            b = {k: v.to(device, non_blocking=True) for k, v in host[(i * a.ga + j) % len(host)].items()}
            l = micro(b)
            tot = l if tot is None else tot + l
        finish()
        last["loss"] = float(tot.item())

    sampler = ClockSampler(torch.cuda.current_device())
    sampler.start()
    total, _ = timed_region(step_device, a.steps, a.warmup, device, e2e=False)
    clocks = sampler.stop()
    e2e_total = None
    if not a.no_e2e:
        e2e_total, _ = timed_region(step_e2e, a.steps, 1, device, e2e=True)
    tokens_per_step = a.batch * a.seq * a.ga * world
    total = max_over_ranks(total, device)
    res = {"ms_per_step": total / a.steps * 1e3, "value": tokens_per_step * a.steps / total, "gpu_launches": 0,
           "clocks": clocks, "backend": "reference:" + opt.update_fn.__name__, "loss": last.get("loss")}
    n_params = sum(p.numel() for p in model.parameters())
    rep = optimizer_step_report(opt_events, n_params, 2, world, device)
    if rep:
        res.update(rep)
    if e2e_total is not None:
        e2e_total = max_over_ranks(e2e_total, device)
        res["e2e"] = {"value": tokens_per_step * a.steps / e2e_total, "unit": "tokens/s",
                      "h2d_bytes_per_step": a.ga * 2 * a.batch * a.seq * 8, "d2h_bytes_per_step": 4,
                      "ms_per_step": e2e_total / a.steps * 1e3}
    return res


def main():
    a = parse()
    import torch
    if not torch.cuda.is_available():
        line = {"impl": a.impl, "unavailable": "no CUDA device visible"} if a.impl == "reference" else \
            {"metric": METRIC, "value": None, "error": "no CUDA device visible"}
        print(json.dumps(line))
        return 0 if a.impl == "reference" else 1
    from distributed_lion_pytorch_b200.parallel import runtime as rt
    device = rt.init_distributed()
    rank, world = rt.rank(), rt.world_size()
    if world != a.gpus and rank == 0:
        print(f"[bench] warning: --gpus {a.gpus} but WORLD_SIZE={world}; launch with torchrun for N>1", file=sys.stderr)
    res = run_reference(a, device, rank, world) if a.impl == "reference" else run_ours(a, device, rank, world)
    if rank == 0:
        if "unavailable" in res:
            print(json.dumps({"impl": "reference", "unavailable": res["unavailable"]}))
        else:
            default_cfg = (a.model, a.batch, a.seq, a.ga, a.lora, a.load_in_4bit) == ("gpt2", 20, 1024, 8, False, False)
            metric = METRIC if default_cfg else (
                f"tokens/sec/box {a.model} Lion training (per-device batch {a.batch}x{a.seq}, grad-accum {a.ga}, bf16"
                + (", LoRA" if a.lora else "") + (", NF4 base" if a.load_in_4bit else "") + ")")
            line = {"metric": metric, "value": res["value"], "unit": "tokens/s", "n_gpus": world, "steps": a.steps,
                    "warmup": a.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
                    "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                    "impl": a.impl,
                    "config": {"model": "gpt2-small-124M" if a.model == "gpt2" else a.model,
                               "global_batch": a.batch * world * a.ga, "per_device_batch": a.batch, "seq_len": a.seq,
                               "grad_accum": a.ga, "parallelism": f"dp{world}", "optimizer": "lion-majority-vote",
                               "sign_exchange": res["backend"], "named_config": default_cfg,
                               "lora": bool(a.lora), "load_in_4bit": bool(a.load_in_4bit),
                               "l2": "working set >> 126 MB L2 and a 256 MB scratch write between timed steps"},
                    "clocks": res["clocks"], "gpu_launches": res["gpu_launches"], "loss": res.get("loss")}
            for k in ("optimizer_step_us", "optimizer_step_roofline_us", "optimizer_step_roofline_frac",
                      "optimizer_step_includes", "roofline_peaks", "optimizer_params", "fused_equals_split",
                      "replicas_identical", "nvls_multicast", "peak_mem_gb"):
                if k in res:
                    line[k] = res[k]
            if "e2e" in res:
                line["e2e"] = res["e2e"]
            print(json.dumps(line))
    rt.shutdown()
    if res.get("fused_equals_split") is False or res.get("replicas_identical") is False:
        print("[bench] the fused step disagrees with the split path or replicas diverged", file=sys.stderr)
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
