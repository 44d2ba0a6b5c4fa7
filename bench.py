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
    ap.add_argument("--task", default="clm", choices=["clm", "dpo"],
                    help="clm: causal-LM training (run_clm / sft_llama2 shape).  dpo: BASELINE config 4 — policy + "
                         "frozen reference model, averaged-sign Lion, simulated worker drop-out (defaults: "
                         "llama-2-7b, per-device batch 4 pairs, GA 4, gradient checkpointing: dpo_llama2.py:37-53)")
    ap.add_argument("--model", default=None)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--ga", type=int, default=None)
    ap.add_argument("--drop-prob", type=float, default=0.2, help="dpo: per-rank per-step ballot drop probability")
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the post-run fused == split / replica check")
    ap.add_argument("--lora", action="store_true", help="non-default: LoRA on q_proj/v_proj (the sft_llama2 arrangement)")
    ap.add_argument("--load-in-4bit", action="store_true", help="non-default: NF4 frozen base under --lora")
    a = ap.parse_args()
    dflt = {"clm": ("gpt2", 20, 8), "dpo": ("llama-2-7b", 4, 4)}[a.task]
    a.model = a.model or dflt[0]
    a.batch = a.batch or dflt[1]
    a.ga = a.ga or dflt[2]
    return a


from distributed_lion_pytorch_b200.utils.profiling import ClockSampler  # noqa: E402


def make_batches(n_micro, batch, seq, vocab, seed, pin, task="clm"):
    """clm: [batch, seq] ids, labels = ids.  dpo: [2*batch, seq] — rows 0..batch-1 prompt+chosen, the rest
    prompt+rejected (same prompt), labels masked (-100) over the 512-token prompt (dpo_llama2.py:51-52)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_micro):
        ids = torch.randint(0, vocab, (batch, seq), generator=g, dtype=torch.int64)
        lab = ids.clone()
        if task == "dpo":
            rej = torch.randint(0, vocab, (batch, seq), generator=g, dtype=torch.int64)
            plen = seq // 2
            rej[:, :plen] = ids[:, :plen]
            ids = torch.cat([ids, rej])
            lab = ids.clone()
            lab[:, :plen] = -100
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


def optimizer_step_report(events, in_situ_events, n_params, elem_size, world, device):
    """Device time of grad-norm + clip + optimizer + zero_grad per step against its roofline: the slower of the HBM
    bytes at the measured copy peak and the packed sign bits over NVLink.  HBM bytes per parameter: g is read
    twice (the norm must be known before any element is scaled) and cleared once, m and p are read and written
    once: 7 * elem_size, plus (W + 1) / 8 bytes of sign words.

    ``events``: steps taken after the timed region with the ranks aligned by a barrier right before the optimizer
    phase — the cost of the phase itself, max over ranks.  ``in_situ_events``: the same phase inside the timed
    steps, where a rank that finishes its backward passes early waits INSIDE the fused kernel for the slowest
    rank's sign words; max over ranks, so it mostly measures how far apart the ranks' backward passes finish."""
    if not events:
        return None
    us = sorted(a.elapsed_time(b) for a, b in events)[len(events) // 2] * 1e3
    us = max_over_ranks(us, device)
    peaks = load_peaks()
    hbm_bytes = n_params * (7 * elem_size + ((world + 1) / 8 if world > 1 else 0))
    nvl_bytes = n_params / 8 * (world - 1)
    roof_us = max(hbm_bytes / (peaks["hbm_gbs"] * 1e9), nvl_bytes / (NVLINK_GBS * 1e9)) * 1e6
    out = {"optimizer_step_us": round(us, 1), "optimizer_step_roofline_us": round(roof_us, 1),
           "optimizer_step_roofline_frac": round(roof_us / us, 3), "optimizer_step_includes": "grad-norm + clip + "
           "sign exchange + vote + update + momentum + gradient zero-fill; ranks aligned by a barrier before the phase, "
           "median of the post-run steps, max over ranks", "roofline_peaks": peaks["source"],
           "optimizer_params": int(n_params)}
    if in_situ_events:
        t = sum(a.elapsed_time(b) for a, b in in_situ_events) / len(in_situ_events) * 1e3
        out["optimizer_step_us_in_situ"] = round(max_over_ranks(t, device), 1)
    return out


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
    if a.task == "dpo":
        # BASELINE config 4 (reference: dpo_llama2.py:133-152, 219-231): policy + frozen reference model from the
        # same weights, averaged-sign vote (the paper's Averaging), Bernoulli ballot drop-out from a shared seed
        from async_trainer import AsyncDPOTrainer
        from distributed_lion_pytorch_b200.parallel.fault import DropoutSchedule
        ref_model = build_model(a.model, dtype=torch.bfloat16, device=device, pretrained=False)
        ref_model.load_state_dict(model.state_dict())
        targs.gradient_checkpointing = True
        opt = Lion([p for p in model.parameters() if p.requires_grad], lr=5e-4, weight_decay=0.05, comm=comm,
                   vote="average")
        sched = get_cosine_schedule_with_warmup(opt, 100, 1000)
        trainer = AsyncDPOTrainer(model, ref_model, args=targs, beta=0.1, optimizers=(opt, sched),
                                  max_prompt_length=a.seq // 2, max_length=a.seq)
        if rt.world_size() > 1 and a.drop_prob > 0:
            trainer._dropout_schedule = DropoutSchedule(rt.world_size(), a.drop_prob, seed=1234)
        rt.broadcast_parameters(list(model.parameters()) + list(model.buffers()))
        return model, opt, trainer
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
    host = make_batches(a.ga * min(n_steps, 4), a.batch, a.seq, vocab, seed=100 + rank, pin=True, task=a.task)
    dev = [{k: v.to(device) for k, v in b.items()} for b in host]
    last = {}
    seqs = 2 if a.task == "dpo" else 1          # chosen + rejected rows per pair

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
    tokens_per_step = seqs * a.batch * a.seq * a.ga * world
    total = max_over_ranks(total, device)
    res = {"ms_per_step": total / a.steps * 1e3, "value": tokens_per_step * a.steps / total,
           "gpu_launches": launches, "clocks": clocks, "backend": opt.backend_name, "loss": last.get("loss"),
           "peak_mem_gb": round(torch.cuda.max_memory_allocated(device) / 2 ** 30, 2)}
    if e2e_total is not None:
        e2e_total = max_over_ranks(e2e_total, device)
        h2d = a.ga * 2 * seqs * a.batch * a.seq * 8
        res["e2e"] = {"value": tokens_per_step * a.steps / e2e_total, "unit": "tokens/s",
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "ms_per_step": e2e_total / a.steps * 1e3}
    # the optimizer phase on its own: a few more steps, outside the timed region, with the ranks aligned by a
    # barrier right before the phase (inside the timed steps an early rank waits in the kernel for the slowest one)
    import torch.distributed as dist
    aligned = []
    for i in range(5):
        for j in range(a.ga):
            trainer.training_step(trainer.model, dev[(i * a.ga + j) % len(dev)])
        torch.cuda.synchronize(device)
        if dist.is_initialized():
            dist.barrier()
            torch.cuda.synchronize(device)
        torch.cuda._sleep(6_000_000)            # ~3 ms of GPU idle spin: the host enqueues the whole phase behind it, so
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)   # the events see GPU time only
        ea.record()
        trainer._last_grad_norm = trainer._clip_gradients()
        trainer._optimizer_step()
        eb.record()
        aligned.append((ea, eb))
    torch.cuda.synchronize(device)
    n_params = sum(b.numel for b in opt.buckets)
    esize = opt.buckets[0].flat_p.element_size() if opt.buckets else 2
    rep = optimizer_step_report(aligned[1:], opt_events, n_params, esize, world, device)
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
    ref_model = None
    if a.task == "dpo":
        # reference: dpo_llama2.py:133-152,219-231 — TRL's DPOTrainer is not installed on this image, so its step is
        # restated with stock parts: HF policy (gradient checkpointing, as the script's default) + frozen HF reference
        # model, per-sequence log-prob sums over the answer tokens, loss -log sigmoid(beta * ((pi_c - pi_r) -
        # (ref_c - ref_r))), and the reference's own Lion (majority vote: it has no averaging variant and no drop-out).
        import copy
        ref_model = copy.deepcopy(model).eval().requires_grad_(False)
        model.gradient_checkpointing_enable()
        model.config.use_cache = False
    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[device.index])
    lr, wd = (5e-4, 0.0) if a.task == "dpo" else (1e-4, 0.1)           # D11: the DPO script never forwards weight_decay
    opt = ref.Lion(model.parameters(), lr=lr, weight_decay=wd)         # picks update_fn_distributed when W > 1
    sched = get_cosine_schedule_with_warmup(opt, 2000, 100000)
    n_steps = a.warmup + a.steps
    host = make_batches(a.ga * min(n_steps, 4), a.batch, a.seq, cfg.vocab_size, seed=100 + rank, pin=True, task=a.task)
    dev = [{k: v.to(device) for k, v in b.items()} for b in host]
    last = {}
    seqs = 2 if a.task == "dpo" else 1

    def seq_logps(m, ids, labels):
        logits = m(input_ids=ids).logits[:, :-1].float()
        tgt = labels[:, 1:]
        mask = tgt != -100
        lp = torch.gather(logits.log_softmax(-1), 2, tgt.clamp(min=0).unsqueeze(2)).squeeze(2)
        return (lp * mask).sum(-1)

    def dpo_forward(batch):
        ids, labels = batch["input_ids"], batch["labels"]
        B = ids.shape[0] // 2
        pi = seq_logps(net, ids, labels)
        with torch.no_grad():
            rf = seq_logps(ref_model, ids, labels)
        return -torch.nn.functional.logsigmoid(0.1 * ((pi[:B] - pi[B:]) - (rf[:B] - rf[B:]))).mean()

    def forward(batch):
        return dpo_forward(batch) if a.task == "dpo" else net(**batch).loss

    def micro(batch):
        if world > 1:
            with net.no_sync():                                         # reference: async_trainer.py:15
                loss = forward(batch)
                (loss / a.ga).backward()
        else:
            loss = forward(batch)
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
    tokens_per_step = seqs * a.batch * a.seq * a.ga * world
    total = max_over_ranks(total, device)
    res = {"ms_per_step": total / a.steps * 1e3, "value": tokens_per_step * a.steps / total, "gpu_launches": 0,
           "clocks": clocks, "backend": "reference:" + opt.update_fn.__name__, "loss": last.get("loss")}
    n_params = sum(p.numel() for p in model.parameters())
    aligned = []
    for i in range(2):      # the reference's distributed step takes 0.6 s (GPT-2) to ~30 s (Llama-2-7B): two samples
        for j in range(a.ga):
            micro(dev[(i * a.ga + j) % len(dev)])
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(device)
        n_before = len(opt_events)
        torch.cuda._sleep(6_000_000)
        rec["on"] = True
        finish()
        rec["on"] = False
        aligned.append(opt_events.pop())
        assert len(opt_events) == n_before
    torch.cuda.synchronize(device)
    rep = optimizer_step_report(aligned[1:], opt_events, n_params, 2, world, device)
    if rep:
        res.update(rep)
    if e2e_total is not None:
        e2e_total = max_over_ranks(e2e_total, device)
        res["e2e"] = {"value": tokens_per_step * a.steps / e2e_total, "unit": "tokens/s",
                      "h2d_bytes_per_step": a.ga * 2 * seqs * a.batch * a.seq * 8, "d2h_bytes_per_step": 4,
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
            default_cfg = (a.task, a.model, a.batch, a.seq, a.ga, a.lora, a.load_in_4bit) == \
                ("clm", "gpt2", 20, 1024, 8, False, False)
            if a.task == "dpo":
                metric = (f"tokens/sec/box {a.model} DPO (policy + frozen reference model, per-device batch {a.batch} pairs "
                          f"x {a.seq}, grad-accum {a.ga}, bf16, gradient checkpointing; chosen + rejected tokens)")
            else:
                metric = METRIC if default_cfg else (
                    f"tokens/sec/box {a.model} Lion training (per-device batch {a.batch}x{a.seq}, grad-accum {a.ga}, bf16"
                    + (", LoRA" if a.lora else "") + (", NF4 base" if a.load_in_4bit else "") + ")")
            line = {"metric": metric, "value": res["value"], "unit": "tokens/s", "n_gpus": world, "steps": a.steps,
                    "warmup": a.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
                    "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                    "impl": a.impl,
                    "config": {"model": "gpt2-small-124M" if a.model == "gpt2" else a.model,
                               "global_batch": a.batch * world * a.ga, "per_device_batch": a.batch, "seq_len": a.seq,
                               "grad_accum": a.ga, "parallelism": f"dp{world}", "task": a.task,
                               "optimizer": ("lion-majority-vote" if a.task == "clm" or a.impl == "reference" else
                                             f"lion-averaged-sign, ballot drop-out p={a.drop_prob}"),
                               "sign_exchange": res["backend"], "named_config": default_cfg,
                               "lora": bool(a.lora), "load_in_4bit": bool(a.load_in_4bit),
                               "l2": "working set >> 126 MB L2 and a 256 MB scratch write between timed steps"},
                    "clocks": res["clocks"], "gpu_launches": res["gpu_launches"], "loss": res.get("loss")}
            for k in ("optimizer_step_us", "optimizer_step_us_in_situ", "optimizer_step_roofline_us", "optimizer_step_roofline_frac",
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
