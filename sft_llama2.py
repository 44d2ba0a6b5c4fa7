#!/usr/bin/env python
"""Llama-2 supervised fine-tuning entry point:  torchrun --nproc_per_node N -m sft_llama2 ...

Same contract as the reference script (reference: sft_llama2.py:20-199, launch recipe
README.md:42-62): ``ScriptArguments`` + ``TrainingArguments`` from the CLI, LoRA config on
q_proj/v_proj, "Question/Answer" formatting packed into constant-length blocks, ``--lion``
builds ``Lion`` + cosine schedule, ``--async_grad`` picks ``AsyncSFTTrainer``, final checkpoint
and merged weights are written under ``output_dir``.

Deliberate differences: a ``main()`` instead of import-time execution; the optimizer is built
AFTER the adapters exist (upstream optimises nothing with --lion); full-parameter
bf16 training (``--use_lora False``, the BASELINE "Llama-2-7B SFT bf16" config) is supported
next to LoRA; the nf4 4-bit base of upstream (:141-145, bitsandbytes there) is ``--load_in_4bit``
here (``models/quant.py`` + ``csrc/quant4.cu``; off by default because the BASELINE config is bf16); with
``--synthetic_data`` (no network here) packed random-token blocks of the same shape are used.
"""
from __future__ import annotations

import logging
import os
import sys
from dataclasses import dataclass, field
from typing import Optional

import torch

from async_trainer import AsyncSFTTrainer
from distributed_lion import Lion
from distributed_lion_pytorch_b200.models import LoraConfig, build_model, merge_and_unload, quantize_base_
from distributed_lion_pytorch_b200.models.lora import print_trainable_parameters  # noqa: F401  (reference: :78-90)
from distributed_lion_pytorch_b200.train import (ConstantLengthDataset, SFTTrainer, SyntheticTokenDataset,
                                                 TrainingArguments, get_cosine_schedule_with_warmup, load_tokenizer,
                                                 parse_cli, set_seed)
from distributed_lion_pytorch_b200.train.data import chars_token_ratio, prepare_sample_text

logger = logging.getLogger("sft_llama2")


@dataclass
class ScriptArguments:
    """reference: sft_llama2.py:20-40"""
    model_name: Optional[str] = field(default="meta-llama/Llama-2-7b-hf", metadata={"help": "preset or checkpoint dir"})
    dataset_name: Optional[str] = field(default="lvwerra/stack-exchange-paired")
    subset: Optional[str] = field(default="data/finetune")
    split: Optional[str] = field(default="train")
    size_valid_set: Optional[int] = field(default=4000)
    streaming: Optional[bool] = field(default=True)
    shuffle_buffer: Optional[int] = field(default=5000)
    seq_length: Optional[int] = field(default=1024)
    num_workers: Optional[int] = field(default=4)
    packing: Optional[bool] = field(default=True)
    lora_alpha: Optional[float] = field(default=16)
    lora_dropout: Optional[float] = field(default=0.05)
    lora_r: Optional[int] = field(default=8)
    use_lora: Optional[bool] = field(default=True, metadata={"help": "False = full-parameter bf16 SFT"})
    model_dtype: Optional[str] = field(default="bfloat16")
    config_overrides: Optional[str] = field(default=None)
    load_in_4bit: Optional[bool] = field(default=False, metadata={"help": "frozen 4-bit base under LoRA (reference: :141-145)"})
    bnb_4bit_quant_type: Optional[str] = field(default="nf4", metadata={"help": "nf4 | fp4"})


def create_datasets(tokenizer, args: ScriptArguments, training_args: TrainingArguments, vocab_size: int):
    """reference: sft_llama2.py:99-138"""
    if training_args.synthetic_data:
        n_train = max(training_args.synthetic_samples, 64)
        return (SyntheticTokenDataset(vocab_size, args.seq_length, n_train, seed=training_args.seed),
                SyntheticTokenDataset(vocab_size, args.seq_length, 64, seed=training_args.seed + 1))
    from datasets import load_dataset
    dataset = load_dataset(args.dataset_name, data_dir=args.subset, split=args.split,
                           num_proc=args.num_workers if not args.streaming else None, streaming=args.streaming)
    if args.streaming:
        valid_data = dataset.take(args.size_valid_set)
        train_data = dataset.skip(args.size_valid_set).shuffle(buffer_size=args.shuffle_buffer, seed=None)
    else:
        dataset = dataset.train_test_split(test_size=0.005, seed=None)
        train_data, valid_data = dataset["train"], dataset["test"]
        print(f"Size of the train set: {len(train_data)}. Size of the validation set: {len(valid_data)}")
    ratio = chars_token_ratio(train_data, tokenizer)
    print(f"The character to token ratio of the dataset is: {ratio:.2f}")
    train = ConstantLengthDataset(tokenizer, train_data, formatting_func=prepare_sample_text, infinite=True,
                                  seq_length=args.seq_length, chars_per_token=ratio)
    valid = ConstantLengthDataset(tokenizer, valid_data, formatting_func=prepare_sample_text, infinite=False,
                                  seq_length=args.seq_length, chars_per_token=ratio)
    return train, valid


def main(argv=None):
    script_args, training_args = parse_cli((ScriptArguments, TrainingArguments), argv)
    logging.basicConfig(format="%(asctime)s - %(levelname)s - %(name)s - %(message)s", level=logging.INFO,
                        handlers=[logging.StreamHandler(sys.stdout)])
    if training_args.group_by_length and script_args.packing:
        raise ValueError("Cannot use both packing and group by length")
    if training_args.gradient_checkpointing and script_args.use_lora:
        raise ValueError("gradient_checkpointing not supported with LoRA adapters here (reference: sft_llama2.py:57-59)")
    device = training_args.torch_device
    set_seed(training_args.seed)

    peft_config = LoraConfig(r=script_args.lora_r, lora_alpha=script_args.lora_alpha, lora_dropout=script_args.lora_dropout,
                             target_modules=["q_proj", "v_proj"], bias="none", task_type="CAUSAL_LM") \
        if script_args.use_lora else None
    base_model = build_model(script_args.model_name, dtype=getattr(torch, script_args.model_dtype), device=device,
                             config_overrides=script_args.config_overrides)
    if script_args.load_in_4bit:
        if not script_args.use_lora:
            raise ValueError("--load_in_4bit freezes the base weights; it needs --use_lora True")
        quantize_base_(base_model, quant_type=script_args.bnb_4bit_quant_type)
    tokenizer = None
    if not training_args.synthetic_data:
        tokenizer = load_tokenizer(script_args.model_name)
        if getattr(tokenizer, "pad_token", None) is None and hasattr(tokenizer, "eos_token"):
            tokenizer.pad_token = tokenizer.eos_token
        tokenizer.padding_side = "right"
    train_dataset, eval_dataset = create_datasets(tokenizer, script_args, training_args, base_model.config.vocab_size)

    trainer_class = AsyncSFTTrainer if training_args.async_grad else SFTTrainer      # reference: sft_llama2.py:170
    trainer = trainer_class(model=base_model, train_dataset=train_dataset, eval_dataset=eval_dataset,
                            peft_config=peft_config, packing=script_args.packing, max_seq_length=None,
                            tokenizer=tokenizer, args=training_args)
    if training_args.lion:       # built after the adapters exist, so the trainable parameters are in it (D10)
        if training_args.max_steps <= 0:
            raise ValueError("--lion needs --max_steps (the cosine schedule is built up front, as upstream)")
        params = [p for p in trainer.model.parameters() if p.requires_grad]
        comm = training_args.comm if training_args.comm in ("auto", "fused", "nccl_1bit", "local") else "auto"
        vote = "majority" if training_args.lion_vote == "stochastic" else training_args.lion_vote
        optimizer = Lion(params, lr=training_args.learning_rate, weight_decay=training_args.weight_decay, vote=vote,
                         tie=training_args.lion_tie, comm=comm, seed=training_args.seed,
                         max_grad_norm=training_args.lion_max_grad_norm,
                         stochastic_rounding=training_args.lion_stochastic_rounding)
        trainer.optimizer = optimizer
        trainer.lr_scheduler = get_cosine_schedule_with_warmup(optimizer, training_args.warmup_steps,
                                                               training_args.max_steps)
    result = trainer.train(resume_from_checkpoint=training_args.resume_from_checkpoint)
    trainer.save_model(training_args.output_dir)
    trainer.log_metrics("train", result.metrics)
    trainer.save_metrics("train", result.metrics)

    output_dir = os.path.join(training_args.output_dir, "final_checkpoint")
    if trainer.is_world_process_zero:
        trainer.model.save_pretrained(output_dir)
    if script_args.use_lora and trainer.is_world_process_zero:
        # reference: sft_llama2.py:195-199 — merge the adapters and save plain safetensors
        merged = merge_and_unload(trainer.model)
        merged.save_pretrained(os.path.join(training_args.output_dir, "final_merged_checkpoint"), safe_serialization=True)
    return result.metrics


if __name__ == "__main__":
    main()
