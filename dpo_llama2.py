#!/usr/bin/env python
"""Llama-2 DPO entry point:  torchrun --nproc_per_node N -m dpo_llama2 --model_name_or_path sft/final_checkpoint ...

Same contract as the reference script as it was *meant* to work (reference: dpo_llama2.py:18-239,
README.md:67-71; upstream the file has a SyntaxError at line 81 and references an undefined
``base_model`` at :210 — SURVEY D11): ``ScriptArguments`` with beta / lengths / LoRA / schedule
fields, a trainable policy and a frozen reference model loaded from the same checkpoint,
prompt/chosen/rejected triples (``"Question: ...\\n\\nAnswer: "`` prompts), ``--lion`` builds
``Lion`` + cosine schedule, ``--async_grad`` picks ``AsyncDPOTrainer``.

Additions for the B200 build: ``--lion_vote average`` (the paper's averaged-sign server rule) and
``--lion_drop_prob`` / ``--lion_live_mask`` (simulated worker drop-out) — BASELINE config 4;
``--weight_decay`` is forwarded (upstream drops it); ``--synthetic_data`` for the offline box.
"""
from __future__ import annotations

import logging
import os
import sys
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from async_trainer import AsyncDPOTrainer
from distributed_lion import Lion
from distributed_lion_pytorch_b200.models import LoraConfig, build_model, quantize_base_
from distributed_lion_pytorch_b200.train import (DPOTrainer, SyntheticPreferenceDataset, TrainingArguments,
                                                 get_cosine_schedule_with_warmup, load_tokenizer, parse_cli, set_seed)

logger = logging.getLogger("dpo_llama2")


@dataclass
class ScriptArguments:
    """reference: dpo_llama2.py:18-81"""
    beta: Optional[float] = field(default=0.1, metadata={"help": "the beta parameter for DPO loss"})
    model_name_or_path: Optional[str] = field(default="../sft/results/final_checkpoint")
    learning_rate: Optional[float] = field(default=5e-4)
    lr_scheduler_type: Optional[str] = field(default="cosine")
    warmup_steps: Optional[int] = field(default=100)
    weight_decay: Optional[float] = field(default=0.05)
    optimizer_type: Optional[str] = field(default="paged_adamw_32bit")
    per_device_train_batch_size: Optional[int] = field(default=4)
    per_device_eval_batch_size: Optional[int] = field(default=1)
    gradient_accumulation_steps: Optional[int] = field(default=4)
    gradient_checkpointing: Optional[bool] = field(default=True)
    lora_alpha: Optional[float] = field(default=16)
    lora_dropout: Optional[float] = field(default=0.05)
    lora_r: Optional[int] = field(default=8)
    use_lora: Optional[bool] = field(default=True)
    max_prompt_length: Optional[int] = field(default=512)
    max_length: Optional[int] = field(default=1024)
    max_steps: Optional[int] = field(default=1000)
    logging_steps: Optional[int] = field(default=10)
    save_steps: Optional[int] = field(default=100)
    eval_steps: Optional[int] = field(default=100)
    output_dir: Optional[str] = field(default="./results")
    log_freq: Optional[int] = field(default=1)
    sanity_check: Optional[bool] = field(default=False, metadata={"help": "only train on 1000 samples"})
    report_to: Optional[str] = field(default="none")      # upstream default "wandb" needs a login; opt in explicitly
    ignore_bias_buffers: Optional[bool] = field(default=False)
    lion: Optional[bool] = field(default=False)
    async_grad: Optional[bool] = field(default=False)
    # ---- additions
    lion_vote: Optional[str] = field(default="majority", metadata={"help": "majority | average | stochastic"})
    lion_tie: Optional[str] = field(default="neg")
    lion_stochastic_rounding: Optional[bool] = field(default=False)
    lion_drop_prob: Optional[float] = field(default=0.0)
    lion_live_mask: Optional[str] = field(default=None)
    comm: Optional[str] = field(default="auto")
    synthetic_data: Optional[bool] = field(default=False)
    synthetic_samples: Optional[int] = field(default=512)
    model_dtype: Optional[str] = field(default="bfloat16")
    config_overrides: Optional[str] = field(default=None)
    load_in_4bit: Optional[bool] = field(default=False, metadata={"help": "4-bit policy base and reference model (reference: :137,:151)"})
    bnb_4bit_quant_type: Optional[str] = field(default="nf4", metadata={"help": "nf4 | fp4"})
    seed: Optional[int] = field(default=42)
    max_eval_batches: Optional[int] = field(default=8)


def paired_batch_to_prompts(samples) -> Dict[str, list]:
    """Column mapping of one batch of ``lvwerra/stack-exchange-paired`` rows (reference: dpo_llama2.py:115-121)."""
    return {"prompt": ["Question: " + q + "\n\nAnswer: " for q in samples["question"]],
            "chosen": samples["response_j"], "rejected": samples["response_k"]}


def get_stack_exchange_paired(data_dir: str = "data/rl", sanity_check: bool = False, cache_dir: Optional[str] = None,
                              num_proc=24):
    """{'prompt': "Question: ...\\n\\nAnswer: ", 'chosen': response_j, 'rejected': response_k}
    (reference: dpo_llama2.py:84-125).  Needs the hub; unavailable on the offline box."""
    from datasets import load_dataset
    dataset = load_dataset("lvwerra/stack-exchange-paired", split="train", cache_dir=cache_dir, data_dir=data_dir)
    original_columns = dataset.column_names
    if sanity_check:
        dataset = dataset.select(range(min(len(dataset), 1000)))
    return dataset.map(paired_batch_to_prompts, batched=True, num_proc=num_proc, remove_columns=original_columns)


def main(argv=None):
    (script_args,) = parse_cli((ScriptArguments,), argv)
    logging.basicConfig(format="%(asctime)s - %(levelname)s - %(name)s - %(message)s", level=logging.INFO,
                        handlers=[logging.StreamHandler(sys.stdout)])
    # reference: dpo_llama2.py:171-190 builds TrainingArguments by hand (and forgets weight_decay)
    training_args = TrainingArguments(
        per_device_train_batch_size=script_args.per_device_train_batch_size,
        per_device_eval_batch_size=script_args.per_device_eval_batch_size, max_steps=script_args.max_steps,
        logging_steps=script_args.logging_steps, save_steps=script_args.save_steps,
        gradient_accumulation_steps=script_args.gradient_accumulation_steps,
        gradient_checkpointing=script_args.gradient_checkpointing, learning_rate=script_args.learning_rate,
        weight_decay=script_args.weight_decay, evaluation_strategy="steps", eval_steps=script_args.eval_steps,
        output_dir=script_args.output_dir, report_to=script_args.report_to,
        lr_scheduler_type=script_args.lr_scheduler_type, warmup_steps=script_args.warmup_steps,
        optim=script_args.optimizer_type, bf16=True, remove_unused_columns=False, run_name="dpo_llama2",
        lion=script_args.lion, async_grad=script_args.async_grad, lion_vote=script_args.lion_vote,
        lion_tie=script_args.lion_tie, lion_drop_prob=script_args.lion_drop_prob,
        lion_live_mask=script_args.lion_live_mask, comm=script_args.comm, seed=script_args.seed,
        max_eval_batches=script_args.max_eval_batches)
    device = training_args.torch_device
    set_seed(script_args.seed)
    dtype = getattr(torch, script_args.model_dtype)

    # 1. the policy and 2. the frozen reference model, from the same (SFT) checkpoint
    model = build_model(script_args.model_name_or_path, dtype=dtype, device=device,
                        config_overrides=script_args.config_overrides)
    model_ref = build_model(script_args.model_name_or_path, dtype=dtype, device=device,
                            config_overrides=script_args.config_overrides)
    model_ref.load_state_dict(model.state_dict())        # identical start even when randomly initialised
    if script_args.load_in_4bit:
        if not script_args.use_lora:
            raise ValueError("--load_in_4bit freezes the base weights; it needs --use_lora True")
        quantize_base_(model, quant_type=script_args.bnb_4bit_quant_type)
        quantize_base_(model_ref, quant_type=script_args.bnb_4bit_quant_type)
    vocab = model.config.vocab_size

    tokenizer = None
    if script_args.synthetic_data:
        train_dataset = SyntheticPreferenceDataset(vocab, script_args.max_prompt_length, script_args.max_length,
                                                   script_args.synthetic_samples, seed=script_args.seed)
        eval_dataset = SyntheticPreferenceDataset(vocab, script_args.max_prompt_length, script_args.max_length, 32,
                                                  seed=script_args.seed + 1)
    else:
        tokenizer = load_tokenizer(script_args.model_name_or_path)
        if getattr(tokenizer, "pad_token", None) is None and hasattr(tokenizer, "eos_token"):
            tokenizer.pad_token = tokenizer.eos_token
        keep = lambda x: len(x["prompt"]) + len(x["chosen"]) <= script_args.max_length \
            and len(x["prompt"]) + len(x["rejected"]) <= script_args.max_length      # char lengths, as upstream :159-168
        train_dataset = get_stack_exchange_paired("data/rl", script_args.sanity_check).filter(keep)
        eval_dataset = get_stack_exchange_paired("data/evaluation", True).filter(keep)

    peft_config = LoraConfig(r=script_args.lora_r, lora_alpha=script_args.lora_alpha, lora_dropout=script_args.lora_dropout,
                             target_modules=["q_proj", "v_proj", "k_proj", "o_proj"], bias="none",
                             task_type="CAUSAL_LM") if script_args.use_lora else None
    trainer_class = AsyncDPOTrainer if script_args.async_grad else DPOTrainer        # reference: dpo_llama2.py:216
    dpo_trainer = trainer_class(model, model_ref, args=training_args, beta=script_args.beta,
                                train_dataset=train_dataset, eval_dataset=eval_dataset, tokenizer=tokenizer,
                                peft_config=peft_config, max_prompt_length=script_args.max_prompt_length,
                                max_length=script_args.max_length)
    if script_args.lion:        # after LoRA wrapping, so the trainable parameters are the ones optimised
        params = [p for p in dpo_trainer.model.parameters() if p.requires_grad]
        comm = script_args.comm if script_args.comm in ("auto", "fused", "nccl_1bit", "local") else "auto"
        vote = "majority" if script_args.lion_vote == "stochastic" else script_args.lion_vote
        optimizer = Lion(params, lr=script_args.learning_rate, weight_decay=script_args.weight_decay, vote=vote,
                         tie=script_args.lion_tie, comm=comm, seed=script_args.seed,
                         max_grad_norm=1.0 if script_args.lion_vote == "stochastic" else None,
                         stochastic_rounding=script_args.lion_stochastic_rounding)
        dpo_trainer.optimizer = optimizer
        dpo_trainer.lr_scheduler = get_cosine_schedule_with_warmup(optimizer, script_args.warmup_steps,
                                                                   script_args.max_steps)
    result = dpo_trainer.train()
    dpo_trainer.save_model(script_args.output_dir)
    dpo_trainer.log_metrics("train", result.metrics)
    dpo_trainer.save_metrics("train", result.metrics)
    output_dir = os.path.join(script_args.output_dir, "final_checkpoint")
    if dpo_trainer.is_world_process_zero:
        dpo_trainer.model.save_pretrained(output_dir)
    return result.metrics


if __name__ == "__main__":
    main()
