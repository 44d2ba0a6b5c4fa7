"""Trainer tier (CPU, tiny GPT-2, synthetic tokens): gradient accumulation, zero gradient
all-reduces in async mode, HF checkpoint layout + rotation, bit-exact resume, entry points."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

from distutil import run_distributed

ROOT = Path(__file__).resolve().parent.parent


def _mk(tmp, **kw):
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.train import AsyncTrainer, SyntheticTokenDataset, TrainingArguments
    torch.manual_seed(0)
    model = build_model("gpt2-tiny", config_overrides="resid_pdrop=0.0,embd_pdrop=0.0,attn_pdrop=0.0")
    base = dict(output_dir=str(tmp), per_device_train_batch_size=2, gradient_accumulation_steps=2, max_steps=6,
                learning_rate=1e-3, weight_decay=0.1, lion=True, async_grad=True, logging_steps=1, save_steps=3,
                lr_scheduler_type="cosine", warmup_steps=1, seed=7)
    base.update(kw)
    args = TrainingArguments(**base)
    ds = SyntheticTokenDataset(512, 32, 64, seed=1)
    return AsyncTrainer(model=model, args=args, train_dataset=ds, eval_dataset=SyntheticTokenDataset(512, 32, 8, seed=2))


def test_checkpoint_layout_rotation_and_metrics(tmp_path):
    tr = _mk(tmp_path, save_total_limit=1)
    out = tr.train()
    assert out.global_step == 6
    cks = sorted(p.name for p in tmp_path.iterdir() if p.name.startswith("checkpoint-"))
    assert cks == ["checkpoint-6"]                                  # save_total_limit rotated checkpoint-3 away
    files = {p.name for p in (tmp_path / "checkpoint-6").iterdir()}
    assert {"model.safetensors", "config.json", "optimizer.pt", "scheduler.pt", "rng_state.pth",
            "trainer_state.json", "training_args.bin"} <= files
    sd = torch.load(tmp_path / "checkpoint-6" / "optimizer.pt", weights_only=False)
    assert set(sd.keys()) == {"state", "param_groups"}
    n_params = len(list(tr.model.parameters()))
    assert len(sd["state"]) == n_params and all(list(v.keys()) == ["exp_avg"] for v in sd["state"].values())
    assert {"lr", "betas", "weight_decay", "initial_lr", "params"} <= set(sd["param_groups"][0].keys())
    st = json.loads((tmp_path / "checkpoint-6" / "trainer_state.json").read_text())
    assert st["global_step"] == 6 and len(st["log_history"]) >= 6
    tr.save_model()
    tr.save_state()
    m = tr.evaluate()
    tr.save_metrics("eval", m)
    assert (tmp_path / "eval_results.json").exists() and (tmp_path / "all_results.json").exists()
    # the saved model reloads with HF-style names
    from distributed_lion_pytorch_b200.models import GPT2LMHeadModel
    again = GPT2LMHeadModel.from_pretrained(str(tmp_path))
    for (k, a), (_, b) in zip(again.state_dict().items(), tr.model.state_dict().items()):
        assert torch.equal(a, b.cpu()), k


def test_resume_reproduces_the_run_bit_exactly(tmp_path):
    a = _mk(tmp_path / "a")
    a.train()
    ref = [p.detach().clone() for p in a.model.parameters()]
    ref_losses = [h["loss"] for h in a.state.log_history if "loss" in h]
    b = _mk(tmp_path / "b")            # same schedule; its checkpoint-3 is the "interrupted" state
    b.train()
    c = _mk(tmp_path / "b")
    c.train(resume_from_checkpoint=str(tmp_path / "b" / "checkpoint-3"))
    assert c.state.global_step == 6
    for x, y in zip(ref, c.model.parameters()):
        assert torch.equal(x, y)
    tail = [h["loss"] for h in c.state.log_history if "loss" in h][-3:]
    assert tail == pytest.approx(ref_losses[-3:], abs=1e-5)     # logged from a running sum: rounding only


def test_gradient_accumulation_equals_big_batch(tmp_path):
    t1 = _mk(tmp_path / "x", per_device_train_batch_size=4, gradient_accumulation_steps=1, max_steps=2, save_steps=100,
             max_grad_norm=0.0)
    t2 = _mk(tmp_path / "y", per_device_train_batch_size=2, gradient_accumulation_steps=2, max_steps=2, save_steps=100,
             max_grad_norm=0.0)
    t1.train()
    t2.train()
    for a, b in zip(t1.model.parameters(), t2.model.parameters()):
        # same samples per optimizer step; the sign update makes tiny fp reassociation differences invisible
        assert (a - b).abs().max() <= 2.1e-3 and (a != b).float().mean() < 0.02


def _worker_async_counts(rank, world, outdir, async_grad):
    import torch.distributed as dist
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.train import AsyncTrainer, SyntheticTokenDataset, Trainer, TrainingArguments
    calls = {"all_reduce": 0, "all_gather": 0}
    orig_ar, orig_ag = dist.all_reduce, dist.all_gather_into_tensor

    def ar(t, *a, **k):
        if t.numel() > 64:                      # ignore the one-off active-set / scalar reductions
            calls["all_reduce"] += 1
        return orig_ar(t, *a, **k)

    def ag(o, i, *a, **k):
        calls["all_gather"] += 1
        return orig_ag(o, i, *a, **k)
    dist.all_reduce, dist.all_gather_into_tensor = ar, ag
    torch.manual_seed(rank)                      # replicas start DIFFERENT on purpose: the broadcast must fix it
    model = build_model("gpt2-tiny")
    args = TrainingArguments(output_dir=outdir, per_device_train_batch_size=2, gradient_accumulation_steps=2, max_steps=3,
                             learning_rate=1e-3, weight_decay=0.1, lion=True, async_grad=async_grad, logging_steps=1,
                             save_steps=100, lion_drop_prob=0.3 if async_grad else 0.0, seed=3)
    cls = AsyncTrainer if async_grad else Trainer
    tr = cls(model=model, args=args, train_dataset=SyntheticTokenDataset(512, 32, 64, seed=1))
    tr.train()
    dist.all_reduce, dist.all_gather_into_tensor = orig_ar, orig_ag
    flat = torch.cat([p.detach().flatten() for p in model.parameters()])
    return dict(calls=calls, params=flat, backend=tr.optimizer.backend_name)


@pytest.mark.parametrize("async_grad", [True, False])
def test_async_trainer_never_allreduces_gradients(tmp_path, async_grad):
    outs = run_distributed(_worker_async_counts, 2, str(tmp_path), async_grad)
    assert torch.equal(outs[0]["params"], outs[1]["params"])          # replicas identical after training
    # async: signs are exchanged; sync baseline: gradients are averaged, so the local step suffices
    assert outs[0]["backend"] == ("nccl_1bit" if async_grad else "local")
    if async_grad:
        assert outs[0]["calls"]["all_reduce"] == 0                     # no gradient ever crossed ranks
        assert outs[0]["calls"]["all_gather"] == 3                     # one 1-bit exchange per optimizer step
    else:
        assert outs[0]["calls"]["all_reduce"] >= 3                     # the synchronous baseline does all-reduce


def test_dropout_schedule_is_deterministic_and_never_empty():
    from distributed_lion_pytorch_b200.parallel.fault import DropoutSchedule
    s1, s2 = DropoutSchedule(8, 0.5, seed=3), DropoutSchedule(8, 0.5, seed=3)
    masks = [s1.mask(i) for i in range(200)]
    assert masks == [s2.mask(i) for i in range(200)]
    assert all(m and 0 < m < 256 for m in masks)
    assert len(set(masks)) > 20
    assert DropoutSchedule(4, 0.0).mask(5) is None
    assert DropoutSchedule(4, 0.0, fixed_mask=0b1011).mask(9) == 0b1011


def test_arg_parser_cli_and_json(tmp_path):
    from distributed_lion_pytorch_b200.train import AsyncTrainingArguments, HfArgumentParser, parse_cli
    (a,) = HfArgumentParser(AsyncTrainingArguments).parse_args_into_dataclasses(
        ["--lion", "--async_grad", "--learning_rate", "0.0001", "--bf16=True", "--output_dir", "o", "--save_total_limit", "2"])
    assert a.lion and a.async_grad and a.bf16 and a.learning_rate == 1e-4 and a.save_total_limit == 2
    f = tmp_path / "cfg.json"
    f.write_text(json.dumps({"output_dir": "x", "lion": True, "max_steps": 5}))
    (b,) = parse_cli((AsyncTrainingArguments,), [str(f)])
    assert b.lion and b.max_steps == 5 and not b.async_grad


def test_entry_points_import_and_readme_contract():
    import async_trainer
    import distributed_lion
    import dpo_llama2
    import run_clm
    import sft_llama2
    assert hasattr(distributed_lion, "Lion")
    for n in ("AsyncTrainer", "AsyncSFTTrainer", "AsyncDPOTrainer"):
        assert hasattr(async_trainer, n)
    for mod in (run_clm, sft_llama2, dpo_llama2):
        assert callable(mod.main)


@pytest.mark.timeout(300)
def test_run_clm_torchrun_gloo_world2(tmp_path):
    """BASELINE config 1: run_clm --lion --async_grad on CPU/gloo, world_size 2, synthetic tokens."""
    from distutil import free_port
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(free_port()), "-m", "run_clm", "--model_type", "gpt2-tiny",
           "--synthetic_data", "--do_train", "--do_eval", "--output_dir", str(tmp_path / "o"), "--block_size", "64",
           "--per_device_train_batch_size", "2", "--gradient_accumulation_steps", "2", "--max_steps", "3", "--lion",
           "--async_grad", "--learning_rate", "1e-3", "--weight_decay", "0.1", "--logging_steps", "1", "--save_steps",
           "3", "--max_eval_samples", "4", "--per_device_eval_batch_size", "2", "--per_rank_optimizer_state"]
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=280)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    ck = tmp_path / "o" / "checkpoint-3"
    for f in ("optimizer.pt", "optimizer_rank0.pt", "optimizer_rank1.pt", "rng_state_0.pth", "rng_state_1.pth"):
        assert (ck / f).exists(), f
    res = json.loads((tmp_path / "o" / "all_results.json").read_text())
    assert res["lion_backend"] == "nccl_1bit" and "perplexity" in res
    hist = json.loads((ck / "trainer_state.json").read_text())["log_history"]
    checked = [h["replicas_identical"] for h in hist if "replicas_identical" in h]
    assert checked and all(v == 1 for v in checked)              # the per-log-step replica checksum ran and agreed


def test_reference_readme_command_lines_parse():
    """The three launch recipes of the reference README (README.md:19-72) parse unchanged."""
    import dpo_llama2
    import run_clm
    import sft_llama2
    from distributed_lion_pytorch_b200.train import AsyncTrainingArguments, TrainingArguments, parse_cli
    gpt2 = ("--config_name gpt2 --tokenizer_name gpt2 --dataset_name openwebtext --per_device_train_batch_size 20 "
            "--per_device_eval_batch_size 24 --do_train --do_eval --output_dir gpt2_lion_wd_0.1 --report_to wandb "
            "--torch_dtype bfloat16 --gradient_accumulation_steps 8 --max_steps 100000 --warmup_steps 2000 --lion "
            "--save_total_limit 2 --learning_rate 0.0001 --weight_decay 0.1 --async_grad").split()
    m, d, t = parse_cli((run_clm.ModelArguments, run_clm.DataTrainingArguments, AsyncTrainingArguments), gpt2)
    assert m.config_name == "gpt2" and m.torch_dtype == "bfloat16" and d.dataset_name == "openwebtext"
    assert t.lion and t.async_grad and t.gradient_accumulation_steps == 8 and t.save_total_limit == 2
    sft = ['--output_dir=./sft', '--max_steps=500', '--logging_steps=10', '--save_steps=10',
           '--per_device_train_batch_size=4', '--per_device_eval_batch_size=1', '--gradient_accumulation_steps=2',
           '--gradient_checkpointing=False', '--group_by_length=False', '--learning_rate=1e-4',
           '--lr_scheduler_type=cosine', '--warmup_steps=100', '--weight_decay=0.05', '--optim=paged_adamw_32bit',
           '--bf16=True', '--remove_unused_columns=False', '--run_name=sft_llama2', '--report_to=wandb', '--lion',
           '--async_grad']
    s, ta = parse_cli((sft_llama2.ScriptArguments, TrainingArguments), sft)
    assert ta.lion and ta.async_grad and ta.bf16 and not ta.gradient_checkpointing and ta.weight_decay == 0.05
    assert s.model_name == "meta-llama/Llama-2-7b-hf" and s.seq_length == 1024 and s.lora_r == 8
    dpo = ['--model_name_or_path=sft/final_checkpoint', '--output_dir=dpo', '--lion', '--async_grad']
    (ds,) = parse_cli((dpo_llama2.ScriptArguments,), dpo)
    assert ds.lion and ds.async_grad and ds.beta == 0.1 and ds.max_length == 1024 and ds.max_prompt_length == 512


def _worker_per_rank_resume(rank, world, outdir):
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.train import AsyncTrainer, SyntheticTokenDataset, TrainingArguments

    def mk(sub, **kw):
        torch.manual_seed(0)
        model = build_model("gpt2-tiny", config_overrides="resid_pdrop=0.0,embd_pdrop=0.0,attn_pdrop=0.0")
        base = dict(output_dir=f"{outdir}/{sub}", per_device_train_batch_size=2, gradient_accumulation_steps=1,
                    max_steps=4, learning_rate=1e-3, weight_decay=0.1, lion=True, async_grad=True, logging_steps=1,
                    save_steps=2, lr_scheduler_type="cosine", warmup_steps=1, seed=11, per_rank_optimizer_state=True)
        base.update(kw)
        return AsyncTrainer(model=model, args=TrainingArguments(**base),
                            train_dataset=SyntheticTokenDataset(512, 32, 64, seed=1))
    a = mk("a")
    a.train()
    b = mk("a")                                       # resume from the mid-run checkpoint of the same directory
    b.train(resume_from_checkpoint=f"{outdir}/a/checkpoint-2")
    pa = torch.cat([p.detach().flatten() for p in a.model.parameters()])
    pb = torch.cat([p.detach().flatten() for p in b.model.parameters()])
    ma = torch.cat([a.optimizer.state[p]["exp_avg"].flatten() for p in a.model.parameters()])
    mb = torch.cat([b.optimizer.state[p]["exp_avg"].flatten() for p in b.model.parameters()])
    return dict(params_equal=torch.equal(pa, pb), mom_equal=torch.equal(ma, mb), mom=ma)


def test_per_rank_momentum_survives_resume(tmp_path):
    """--per_rank_optimizer_state: every worker gets ITS momentum back (upstream only rank 0's is
    saved and all workers collapse onto it: SURVEY D14), so the resumed run is bit-identical."""
    outs = run_distributed(_worker_per_rank_resume, 2, str(tmp_path))
    assert all(o["params_equal"] and o["mom_equal"] for o in outs)
    assert not torch.equal(outs[0]["mom"], outs[1]["mom"])


def test_run_clm_on_text_files_with_eval_metrics(tmp_path):
    """reference: run_clm.py:352-381 (text files), :562-577 (accuracy), :628-639 (perplexity)."""
    import json
    import run_clm
    text = "\n".join(f"the quick brown fox number {i} jumps over the lazy dog" for i in range(400))
    (tmp_path / "train.txt").write_text(text)
    (tmp_path / "valid.txt").write_text(text[:4000])
    out = tmp_path / "out"
    res = run_clm.main([
        "--model_type", "gpt2-tiny", "--train_file", str(tmp_path / "train.txt"), "--validation_file",
        str(tmp_path / "valid.txt"), "--block_size", "64", "--do_train", "--do_eval", "--max_steps", "12",
        "--per_device_train_batch_size", "4", "--per_device_eval_batch_size", "4", "--learning_rate", "3e-3", "--lion",
        "--async_grad", "--logging_steps", "4", "--save_strategy", "no", "--output_dir", str(out)])
    assert res["train_samples"] > 0 and 0.0 <= res["eval_accuracy"] <= 1.0
    assert abs(res["perplexity"] - __import__("math").exp(res["eval_loss"])) < 1e-6
    hist = json.loads((out / "trainer_state.json").read_text())["log_history"]
    losses = [h["loss"] for h in hist if "loss" in h]
    assert losses[-1] < losses[0]                              # byte-level LM on repetitive text learns quickly
    assert {"all_results.json", "train_results.json", "eval_results.json", "README.md"} <= set(p.name for p in out.iterdir())


def _worker_divergence(rank, world, outdir, mode):
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.train import AsyncTrainer, SyntheticTokenDataset, TrainingArguments
    torch.manual_seed(0)
    model = build_model("gpt2-tiny")
    args = TrainingArguments(output_dir=outdir, per_device_train_batch_size=2, max_steps=5, learning_rate=1e-3, lion=True,
                             async_grad=True, logging_steps=1, save_steps=100, seed=3, check_replicas=mode)

    def corrupt(trainer, logs):                    # a worker's memory is silently corrupted after step 2
        if trainer.state.global_step == 2 and rank == 1 and "loss" in logs:
            with torch.no_grad():
                next(model.parameters()).view(-1)[7] += 1.0
    tr = AsyncTrainer(model=model, args=args, train_dataset=SyntheticTokenDataset(512, 32, 64, seed=1), callbacks=[corrupt])
    try:
        tr.train()
        err = None
    except RuntimeError as e:
        err = str(e)
    checks = [h["replicas_identical"] for h in tr.state.log_history if "replicas_identical" in h]
    flat = torch.cat([p.detach().flatten() for p in model.parameters()])
    return dict(err=err, checks=checks, params=flat, step=tr.state.global_step)


def test_trainer_detects_and_repairs_a_diverged_replica(tmp_path):
    outs = run_distributed(_worker_divergence, 2, str(tmp_path / "a"), "resync")
    assert all(o["err"] is None and o["step"] == 5 for o in outs)
    assert outs[0]["checks"] == [1, 1, 0, 1, 1]                           # seen at step 3, repaired, clean afterwards
    assert torch.equal(outs[0]["params"], outs[1]["params"])
    outs = run_distributed(_worker_divergence, 2, str(tmp_path / "b"), "raise")
    assert all(o["err"] is not None and "diverged" in o["err"] and o["step"] == 3 for o in outs)
    outs = run_distributed(_worker_divergence, 2, str(tmp_path / "c"), "off")
    assert all(o["err"] is None and not o["checks"] for o in outs) and not torch.equal(outs[0]["params"], outs[1]["params"])


def test_hf_trainer_subclass_wraps_training_step_in_no_sync(monkeypatch):
    """reference: async_trainer.py:8-34 — the HF ``Trainer`` subclass whose ``training_step`` runs under
    ``model.no_sync()``.  HF's Trainer cannot be *constructed* on this image (no accelerate), so the class is built
    and its ``training_step`` is driven on an un-initialised instance with HF's own step stubbed out: the override must
    enter ``no_sync`` exactly once around the parent call, pass ``num_items_in_batch`` through, and also work for a
    model that has no ``no_sync`` (one process)."""
    import transformers
    from async_trainer import hf_async_trainer_class
    calls = []
    monkeypatch.setattr(transformers.Trainer, "training_step",
                        lambda self, model, inputs, num_items_in_batch=None: calls.append(("step", model.depth, num_items_in_batch)) or 7)
    cls = hf_async_trainer_class()
    assert issubclass(cls, transformers.Trainer)
    tr = object.__new__(cls)

    class _DDPLike:
        depth = 0

        def no_sync(self):
            import contextlib

            @contextlib.contextmanager
            def cm():
                self.depth += 1
                calls.append(("enter", self.depth))
                try:
                    yield
                finally:
                    self.depth -= 1
                    calls.append(("exit", self.depth))
            return cm()

    m = _DDPLike()
    assert tr.training_step(m, {"x": 1}, num_items_in_batch=5) == 7
    assert calls == [("enter", 1), ("step", 1, 5), ("exit", 0)]          # HF's step ran INSIDE no_sync
    calls.clear()

    class _Plain:
        depth = 0
    assert tr.training_step(_Plain(), {}) == 7 and calls == [("step", 0, None)]
