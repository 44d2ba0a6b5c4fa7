"""CPU tier: model parity with HF transformers (same checkpoint, same logits), LoRA, data
packing, DPO collator / loss, schedules."""
import math

import pytest
import torch


@pytest.mark.parametrize("family", ["gpt2", "llama"])
def test_checkpoint_interchanges_with_hf_transformers(tmp_path, family):
    """save_pretrained() here -> transformers.from_pretrained() there: identical logits, and back."""
    transformers = pytest.importorskip("transformers")
    from distributed_lion_pytorch_b200.models import GPT2Config, GPT2LMHeadModel, LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    if family == "gpt2":
        ours = GPT2LMHeadModel(GPT2Config(vocab_size=300, n_positions=64, n_embd=48, n_layer=2, n_head=4))
        hf_cls = transformers.GPT2LMHeadModel
    else:
        ours = LlamaForCausalLM(LlamaConfig(vocab_size=300, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                                            num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=64))
        hf_cls = transformers.LlamaForCausalLM
    ours.eval()
    ours.save_pretrained(str(tmp_path / "ours"))
    hf = hf_cls.from_pretrained(str(tmp_path / "ours")).eval()
    ids = torch.randint(0, 300, (2, 33))
    with torch.no_grad():
        a = ours(ids).logits
        b = hf(input_ids=ids).logits
        la = ours(ids, labels=ids).loss
        lb = hf(input_ids=ids, labels=ids).loss
    assert torch.allclose(a, b, atol=2e-5, rtol=1e-4), (a - b).abs().max()
    assert abs(float(la) - float(lb)) < 1e-5
    # and the other direction: an HF-written checkpoint loads here
    hf.save_pretrained(str(tmp_path / "hf"), safe_serialization=True)
    back = type(ours).from_pretrained(str(tmp_path / "hf")).eval()
    with torch.no_grad():
        assert torch.allclose(back(ids).logits, a, atol=1e-6)


def test_lora_wrap_train_merge():
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.models import LoraConfig, apply_lora, build_model, lora_state_dict, merge_and_unload
    torch.manual_seed(0)
    m = build_model("llama-tiny").eval()
    ids = torch.randint(0, 512, (2, 16))
    with torch.no_grad():
        base = m(ids).logits.clone()
    apply_lora(m, LoraConfig(r=4, lora_alpha=8, lora_dropout=0.0, target_modules=["q_proj", "v_proj"]))
    trainable = [n for n, p in m.named_parameters() if p.requires_grad]
    assert trainable and all("lora_" in n for n in trainable)
    with torch.no_grad():
        assert torch.allclose(m(ids).logits, base, atol=1e-6)          # B starts at zero: no change yet
    opt = Lion([p for p in m.parameters() if p.requires_grad], lr=1e-2)       # built AFTER wrapping (D10)
    m.train()
    for _ in range(3):
        opt.zero_grad()
        m(ids, labels=ids).loss.backward()
        opt.step()
    m.eval()
    with torch.no_grad():
        tuned = m(ids).logits.clone()
    assert not torch.allclose(tuned, base, atol=1e-4)
    assert len(lora_state_dict(m)) == 2 * 2 * 2
    merge_and_unload(m)
    assert not any("lora_" in n for n, _ in m.named_parameters())
    with torch.no_grad():
        assert torch.allclose(m(ids).logits, tuned, atol=1e-4)


def test_group_texts_and_constant_length_packing():
    from distributed_lion_pytorch_b200.train.data import (ByteTokenizer, ConstantLengthDataset, chars_token_ratio,
                                                          group_texts, prepare_sample_text)
    blocks = group_texts([[1, 2, 3], [4, 5], [6, 7, 8, 9]], 4)
    assert blocks == [[1, 2, 3, 4], [5, 6, 7, 8]]                      # remainder dropped (run_clm.py:509-522)
    tok = ByteTokenizer()
    data = [{"question": f"q{i} " * 5, "response_j": f"a{i} " * 7} for i in range(20)]
    assert prepare_sample_text(data[0]).startswith("Question: q0") and "\n\nAnswer: a0" in prepare_sample_text(data[0])
    assert 0.9 < chars_token_ratio(data, tok, 10) <= 1.0                # bytes: one token per char
    ds = ConstantLengthDataset(tok, data, infinite=False, seq_length=32, num_of_sequences=4, chars_per_token=1.0)
    items = list(ds)
    assert items and all(it["input_ids"].shape == (32,) and torch.equal(it["input_ids"], it["labels"]) for it in items)
    assert any((it["input_ids"] == tok.eos_token_id).any() for it in items)      # samples separated by EOS
    inf = ConstantLengthDataset(tok, data, infinite=True, seq_length=32, num_of_sequences=4, chars_per_token=1.0)
    it = iter(inf)
    assert len([next(it) for _ in range(len(items) * 3)]) == len(items) * 3       # wraps around


def test_dpo_collator_and_loss():
    from distributed_lion_pytorch_b200.train import DPODataCollator, dpo_loss
    col = DPODataCollator(None, max_prompt_length=4, max_length=10, pad_token_id=0)
    feats = [{"prompt_ids": [5, 6, 7, 8, 9, 10], "chosen_ids": [11, 12, 13], "rejected_ids": [21, 22, 23, 24, 25, 26, 27, 28]},
             {"prompt_ids": [1, 2], "chosen_ids": [3], "rejected_ids": [4, 5]}]
    b = col(feats)
    assert b["input_ids"].shape == (4, 10) and b["labels"].shape == (4, 10)
    assert b["input_ids"][0].tolist() == [7, 8, 9, 10, 11, 12, 13, 0, 0, 0]       # prompt truncated from the left
    assert b["labels"][0].tolist() == [-100] * 4 + [11, 12, 13] + [-100] * 3       # prompt and pad masked
    assert b["input_ids"][2].tolist() == [7, 8, 9, 10, 21, 22, 23, 24, 25, 26]     # rejected row, answer truncated
    pc, pr = torch.tensor([-1.0, -2.0]), torch.tensor([-3.0, -1.0])
    rc, rr = torch.tensor([-1.5, -2.0]), torch.tensor([-2.5, -2.0])
    loss, cr, rj = dpo_loss(pc, pr, rc, rr, beta=0.5)
    want = -torch.nn.functional.logsigmoid(0.5 * ((pc - pr) - (rc - rr))).mean()
    assert torch.allclose(loss, want) and torch.allclose(cr, 0.5 * (pc - rc)) and torch.allclose(rj, 0.5 * (pr - rr))


def test_schedules_match_hf_formulas():
    from distributed_lion_pytorch_b200.train import get_cosine_schedule_with_warmup, get_linear_schedule_with_warmup
    p = [torch.nn.Parameter(torch.zeros(1))]
    opt = torch.optim.SGD(p, lr=1.0)
    s = get_cosine_schedule_with_warmup(opt, 10, 110)
    lrs = []
    for _ in range(110):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        s.step()
    assert lrs[0] == 0 and abs(lrs[5] - 0.5) < 1e-9 and abs(lrs[10] - 1.0) < 1e-9
    assert abs(lrs[60] - 0.5 * (1 + math.cos(math.pi * 0.5))) < 1e-9 and lrs[-1] < 1e-3
    transformers = pytest.importorskip("transformers")
    opt2 = torch.optim.SGD(p, lr=1.0)
    h = transformers.get_cosine_schedule_with_warmup(opt2, 10, 110)
    for i in range(110):
        assert abs(opt2.param_groups[0]["lr"] - lrs[i]) < 1e-12
        opt2.step()
        h.step()
    opt3 = torch.optim.SGD(p, lr=1.0)
    l = get_linear_schedule_with_warmup(opt3, 2, 12)
    vals = []
    for _ in range(12):
        vals.append(opt3.param_groups[0]["lr"])
        opt3.step()
        l.step()
    assert vals[:3] == [0.0, 0.5, 1.0] and abs(vals[7] - 0.5) < 1e-9


def test_dpo_trainer_with_implicit_lora_reference(tmp_path):
    """No ref_model: the LoRA-wrapped policy with adapters disabled is the reference (TRL behaviour)."""
    from distributed_lion_pytorch_b200.models import LoraConfig, build_model
    from distributed_lion_pytorch_b200.train import AsyncDPOTrainer, SyntheticPreferenceDataset, TrainingArguments
    torch.manual_seed(0)
    model = build_model("llama-tiny")
    args = TrainingArguments(output_dir=str(tmp_path), per_device_train_batch_size=2, gradient_accumulation_steps=1,
                             max_steps=3, learning_rate=1e-2, lion=True, async_grad=True, logging_steps=1, save_steps=100)
    tr = AsyncDPOTrainer(model, None, args=args, beta=0.1, train_dataset=SyntheticPreferenceDataset(512, 8, 24, 16),
                         peft_config=LoraConfig(r=4, lora_alpha=8, lora_dropout=0.0), max_prompt_length=8, max_length=24)
    out = tr.train()
    losses = [h["loss"] for h in tr.state.log_history if "loss" in h]
    assert abs(losses[0] - 0.693147) < 1e-3            # adapters start at zero: policy == reference
    assert out.global_step == 3 and losses[-1] != losses[0]
    assert any("rewards/accuracies" in h for h in tr.state.log_history)


def test_print_trainable_parameters_and_paired_dataset_mapping(monkeypatch, capsys):
    """C19 (sft_llama2.py:78-90) and C24 (dpo_llama2.py:84-125), the latter on an in-memory stand-in for the hub set."""
    from distributed_lion_pytorch_b200.models import LoraConfig, apply_lora, build_model
    from distributed_lion_pytorch_b200.models.lora import print_trainable_parameters
    m = apply_lora(build_model("llama-tiny"), LoraConfig(r=4))
    msg = print_trainable_parameters(m)
    trainable = sum(p.numel() for p in m.parameters() if p.requires_grad)
    total = sum(p.numel() for p in m.parameters())
    assert capsys.readouterr().out.strip() == msg
    assert f"trainable params: {trainable} || all params: {total}" in msg and 0 < trainable < total

    import dpo_llama2
    rows = {"qid": [1, 2, 3], "question": ["a?", "b?", "c?"], "date": ["x"] * 3, "metadata": [["m"]] * 3,
            "response_j": ["good a", "good b", "good c"], "response_k": ["bad a", "bad b", "bad c"]}
    assert dpo_llama2.paired_batch_to_prompts(rows) == {
        "prompt": ["Question: a?\n\nAnswer: ", "Question: b?\n\nAnswer: ", "Question: c?\n\nAnswer: "],
        "chosen": rows["response_j"], "rejected": rows["response_k"]}
    datasets = pytest.importorskip("datasets")
    seen = {}

    def fake_load(name, split=None, cache_dir=None, data_dir=None, **kw):
        seen.update(name=name, split=split, data_dir=data_dir)
        return datasets.Dataset.from_dict(rows)
    monkeypatch.setattr(datasets, "load_dataset", fake_load)
    ds = dpo_llama2.get_stack_exchange_paired("data/rl", sanity_check=True, num_proc=None)
    assert seen == {"name": "lvwerra/stack-exchange-paired", "split": "train", "data_dir": "data/rl"}
    assert sorted(ds.column_names) == ["chosen", "prompt", "rejected"] and len(ds) == 3
    assert ds[1] == {"prompt": "Question: b?\n\nAnswer: ", "chosen": "good b", "rejected": "bad b"}


@pytest.mark.parametrize("family", ["gpt2-tiny", "llama-tiny"])
@pytest.mark.parametrize("safe", [True, False])
def test_sharded_checkpoints_round_trip_and_interchange_with_transformers(tmp_path, family, safe):
    """A real Llama-2-7b-hf directory is sharded (model-0000i-of-0000n.safetensors + index); both directions."""
    import os
    from distributed_lion_pytorch_b200.models import build_model, checkpoint_files
    torch.manual_seed(0)
    m = build_model(family).eval()
    ids = torch.randint(0, 512, (2, 16))
    with torch.no_grad():
        ref = m(ids).logits
    d = str(tmp_path / "ours")
    m.save_pretrained(d, safe_serialization=safe, max_shard_size="64KB")
    files = checkpoint_files(d)
    assert len(files) > 2 and all(("-of-" in os.path.basename(f)) for f in files)
    back = build_model(d).eval()
    with torch.no_grad():
        assert torch.equal(back(ids).logits, ref)
    if family == "gpt2-tiny":
        assert back.lm_head.weight is back.transformer.wte.weight           # tying survives the meta-device build
    # bf16 on load: built directly in the target dtype
    assert next(build_model(d, dtype=torch.bfloat16).parameters()).dtype == torch.bfloat16
    # saving again as a single file removes the stale shards and index
    m.save_pretrained(d, safe_serialization=safe)
    assert sorted(os.listdir(d)) == ["config.json", "model.safetensors" if safe else "pytorch_model.bin"]
    if not safe:
        return
    transformers = pytest.importorskip("transformers")
    m.save_pretrained(d, max_shard_size="64KB")
    hf = transformers.AutoModelForCausalLM.from_pretrained(d).eval()       # ours (sharded) -> transformers
    with torch.no_grad():
        assert torch.allclose(hf(ids).logits, ref, atol=1e-5)
    d2 = str(tmp_path / "theirs")
    hf.save_pretrained(d2, max_shard_size="64KB")                          # transformers (sharded) -> ours
    assert len(checkpoint_files(d2)) > 2
    with torch.no_grad():
        assert torch.allclose(build_model(d2).eval()(ids).logits, ref, atol=1e-5)


def test_checkpoint_loader_reports_problems(tmp_path):
    import json
    from distributed_lion_pytorch_b200.models import build_model, checkpoint_files, load_checkpoint_into
    m = build_model("llama-tiny")
    d = str(tmp_path / "ck")
    m.save_pretrained(d, max_shard_size="64KB")
    files = checkpoint_files(d)
    import os
    os.remove(files[1])
    with pytest.raises(FileNotFoundError, match="shards that are not there"):
        checkpoint_files(d)
    with pytest.raises(FileNotFoundError, match="no model weights"):
        checkpoint_files(str(tmp_path))
    m.save_pretrained(d)
    other = build_model("llama-tiny", config_overrides="num_hidden_layers=1")
    missing, unexpected = load_checkpoint_into(other, d)
    assert not missing and unexpected and all(".layers.1." in k for k in unexpected)
    wide = build_model("llama-tiny", config_overrides="intermediate_size=192")
    with pytest.raises(RuntimeError, match="shape"):
        load_checkpoint_into(wide, d)


@pytest.mark.parametrize("family", ["gpt2-tiny", "llama-tiny"])
def test_gradient_checkpointing_gives_the_same_gradients(family):
    """reference: ``--gradient_checkpointing`` of the HF TrainingArguments (README.md:52, dpo_llama2.py:46)."""
    from distributed_lion_pytorch_b200.models import build_model
    torch.manual_seed(0)
    m = build_model(family).train()
    ids = torch.randint(0, 512, (2, 24))
    grads = []
    for ckpt in (False, True):
        m.zero_grad()
        (m.gradient_checkpointing_enable if ckpt else m.gradient_checkpointing_disable)()
        torch.manual_seed(123)                      # same dropout draws in both runs (and in the recomputation)
        loss = m(ids, labels=ids).loss
        loss.backward()
        grads.append((float(loss), {n: p.grad.clone() for n, p in m.named_parameters()}))
    assert grads[0][0] == grads[1][0]
    for n in grads[0][1]:
        assert torch.allclose(grads[0][1][n], grads[1][1][n], atol=1e-6), n
