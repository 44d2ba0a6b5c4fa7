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
import statistics
import sys
import threading
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
    ap.add_argument("--lora", action="store_true", help="non-default: LoRA on q_proj/v_proj (the sft_llama2 arrangement)")
    ap.add_argument("--load-in-4bit", action="store_true", help="non-default: NF4 frozen base under --lora")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of this rank's GPU every 200 ms during the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80)}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, v in names.items():
                    if bits & v:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


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
    total, per_step = timed_region(step_device, a.steps, a.warmup, device, e2e=False)
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
    if hasattr(opt._backend, "_ctx") and opt._backend._ctx is not None:
        opt._backend._ctx.check_status()
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

    def finish():
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)         # HF Trainer default max_grad_norm
        opt.step()
        sched.step()
        model.zero_grad()

    def step_device(i):
        for j in range(a.ga):
            micro(dev[(i * a.ga + j) % len(dev)])
        finish()

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
            if "peak_mem_gb" in res:
                line["peak_mem_gb"] = res["peak_mem_gb"]
            if "e2e" in res:
                line["e2e"] = res["e2e"]
            print(json.dumps(line))
    rt.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
