"""GPU tier: the three reference-compatible entry points end to end on one B200 with tiny models — every
kernel-backed path (fused norms, RoPE, chunked CE, GELU-epilogue GEMM, one-launch bias gradients, 4-bit base,
Lion kernels with stochastic rounding, HF-layout checkpoints) in the arrangement a user launches."""
import json

import pytest
import torch

pytestmark = pytest.mark.gpu


def _launches():
    from distributed_lion_pytorch_b200.ops import _cabi
    return _cabi.LAUNCHES


def test_run_clm_gpt2_on_gpu(tmp_path):
    import run_clm
    out = tmp_path / "clm"
    n0 = _launches()
    res = run_clm.main([
        "--model_type", "gpt2-tiny", "--synthetic_data", "--torch_dtype", "bfloat16", "--block_size", "64", "--do_train",
        "--do_eval", "--max_steps", "6", "--per_device_train_batch_size", "4", "--gradient_accumulation_steps", "2",
        "--learning_rate", "1e-3", "--weight_decay", "0.1", "--lion", "--async_grad", "--lion_stochastic_rounding",
        "--logging_steps", "2", "--save_steps", "3", "--save_total_limit", "2", "--max_eval_samples", "8",
        "--output_dir", str(out)])
    assert _launches() - n0 > 100                                   # our kernels ran, not an eager fallback
    assert res["lion_backend"] == "local" and res["train_loss"] > 0 and res["perplexity"] > 1
    ck = out / "checkpoint-6"
    for f in ("model.safetensors", "config.json", "optimizer.pt", "scheduler.pt", "rng_state.pth", "trainer_state.json"):
        assert (ck / f).exists(), f
    sd = torch.load(ck / "optimizer.pt", weights_only=False)
    assert set(sd.keys()) == {"state", "param_groups"} and list(sd["state"][0].keys()) == ["exp_avg"]
    # resume continues from the checkpoint (and restores the rounding-noise stream position)
    res2 = run_clm.main([
        "--model_type", "gpt2-tiny", "--synthetic_data", "--torch_dtype", "bfloat16", "--block_size", "64", "--do_train",
        "--max_steps", "8", "--per_device_train_batch_size", "4", "--gradient_accumulation_steps", "2",
        "--learning_rate", "1e-3", "--weight_decay", "0.1", "--lion", "--async_grad", "--lion_stochastic_rounding",
        "--logging_steps", "1", "--save_strategy", "no", "--output_dir", str(out)])
    hist = json.loads((out / "trainer_state.json").read_text())["log_history"]
    assert min(h["step"] for h in hist if "loss" in h and h["step"] > 6) == 7 and res2["train_loss"] > 0


def test_sft_llama_lora_on_4bit_base_on_gpu(tmp_path):
    import sft_llama2
    out = tmp_path / "sft"
    n0 = _launches()
    metrics = sft_llama2.main([
        "--model_name", "llama-tiny", "--synthetic_data", "--seq_length", "64", "--max_steps", "4",
        "--per_device_train_batch_size", "4", "--gradient_accumulation_steps", "2", "--learning_rate", "1e-3", "--lion",
        "--async_grad", "--load_in_4bit", "--logging_steps", "1", "--save_strategy", "no", "--output_dir", str(out)])
    assert _launches() - n0 > 100 and metrics["train_loss"] > 0
    from distributed_lion_pytorch_b200.models import build_model
    merged = build_model(str(out / "final_merged_checkpoint"), device="cuda")          # plain 16-bit HF layout again
    assert next(merged.parameters()).dtype == torch.bfloat16
    ids = torch.randint(0, 512, (2, 32), device="cuda")
    assert torch.isfinite(merged(ids, labels=ids).loss)


def test_dpo_llama_on_gpu(tmp_path):
    import dpo_llama2
    out = tmp_path / "dpo"
    n0 = _launches()
    dpo_llama2.main([
        "--model_name_or_path", "llama-tiny", "--synthetic_data", "--synthetic_samples", "32", "--max_prompt_length", "16",
        "--max_length", "48", "--max_steps", "3", "--per_device_train_batch_size", "2", "--gradient_accumulation_steps", "2",
        "--learning_rate", "1e-3", "--lion", "--async_grad", "--lion_vote", "average", "--logging_steps", "1",
        "--eval_steps", "3", "--save_steps", "100", "--output_dir", str(out)])
    assert _launches() - n0 > 50
    assert (out / "final_checkpoint").is_dir()


def test_fp16_training_uses_dynamic_loss_scaling(tmp_path):
    """``--fp16`` / a float16 model (reference: dpo_llama2.py:133-139 loads its models in float16): the trainer scales
    the loss, unscales the gradients before clipping, skips a step whose gradients overflow and halves the scale."""
    import torch
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.train import AsyncTrainer, TrainingArguments, get_cosine_schedule_with_warmup
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = build_model("gpt2-tiny", dtype=torch.float16, device=dev, pretrained=False)
    opt = Lion(model.parameters(), lr=1e-3, weight_decay=0.0)
    tr = AsyncTrainer(model=model, args=TrainingArguments(output_dir=str(tmp_path), gradient_accumulation_steps=2, lion=True,
                                                          async_grad=True, max_grad_norm=1.0),
                      optimizers=(opt, get_cosine_schedule_with_warmup(opt, 1, 100)))
    assert tr.loss_scaler is not None and tr.loss_scaler.scale == 2.0 ** 16
    tr.loss_scaler.scale = 2.0 ** 30          # certain to overflow fp16 gradients at first
    ids = torch.randint(0, 512, (2, 32), device=dev)
    batch = {"input_ids": ids, "labels": ids}
    before = [p.detach().clone() for p in model.parameters()]
    scales, moved = [], []
    for _ in range(24):
        loss = tr.train_step([batch, batch])
        assert torch.isfinite(loss)
        scales.append(tr.loss_scaler.scale)
        moved.append(not all(torch.equal(a, p.detach()) for a, p in zip(before, model.parameters())))
    # a scale of 2^30 overflows fp16 gradients: those steps are skipped and the scale halves until the gradients fit
    assert tr.loss_scaler.skipped >= 1 and scales[-1] < 2.0 ** 30
    assert not moved[0] and moved[-1]                       # skipped steps apply nothing; later steps train
    assert all(torch.isfinite(p).all() for p in model.parameters())
