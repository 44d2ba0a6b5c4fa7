#!/usr/bin/env python
"""Causal-LM (GPT-2) pre-training entry point:  torchrun --nproc_per_node N -m run_clm ...

Same contract as the reference's run_clm.py (reference: run_clm.py:247-653 and the launch
recipe README.md:19-37): ModelArguments / DataTrainingArguments / AsyncTrainingArguments
parsed from the CLI or a single .json file, last-checkpoint detection, dataset → tokenise →
fixed blocks with labels = input_ids, ``--lion`` picks ``Lion`` + cosine schedule,
``--async_grad`` picks ``AsyncTrainer``, train → save → evaluate → perplexity → metrics files.

Differences that are deliberate: ``--synthetic_data`` (random tokens of the right shape: there
is no network or dataset cache on the B200 box); ``--torch_dtype`` is honoured also with
``--config_name``; the non-Lion branch honours ``--weight_decay`` (D13); no W&B key
is baked in (D15); ``--async_grad`` works on a single process (D8).
"""
from __future__ import annotations

import logging
import math
import os
import sys
from dataclasses import dataclass, field
from typing import Optional

import torch

from async_trainer import AsyncTrainer
from distributed_lion import Lion
from distributed_lion_pytorch_b200.models import build_model, config_for
from distributed_lion_pytorch_b200.train import (AsyncTrainingArguments, SyntheticTokenDataset, Trainer,
                                                 default_data_collator, get_cosine_schedule_with_warmup,
                                                 get_last_checkpoint, load_tokenizer, parse_cli, set_seed)
from distributed_lion_pytorch_b200.train.data import BlockDataset, group_texts, text_file_dataset

logger = logging.getLogger("run_clm")


@dataclass
class ModelArguments:
    """reference: run_clm.py:88-166"""
    model_name_or_path: Optional[str] = field(default=None, metadata={"help": "checkpoint dir or preset for weights"})
    model_type: Optional[str] = field(default=None, metadata={"help": "gpt2 | llama when training from scratch"})
    config_overrides: Optional[str] = field(default=None, metadata={"help": "e.g. n_embd=10,n_layer=2"})
    config_name: Optional[str] = field(default=None, metadata={"help": "preset / config dir (random init)"})
    tokenizer_name: Optional[str] = None
    cache_dir: Optional[str] = None
    use_fast_tokenizer: bool = True
    model_revision: str = "main"
    use_auth_token: bool = False
    torch_dtype: Optional[str] = field(default=None, metadata={"help": "auto | bfloat16 | float16 | float32"})
    low_cpu_mem_usage: bool = False

    def __post_init__(self):
        if self.config_overrides is not None and (self.config_name is not None or self.model_name_or_path is not None):
            raise ValueError("--config_overrides can't be used in combination with --config_name or --model_name_or_path")


@dataclass
class DataTrainingArguments:
    """reference: run_clm.py:169-244"""
    dataset_name: Optional[str] = None
    dataset_config_name: Optional[str] = None
    train_file: Optional[str] = None
    validation_file: Optional[str] = None
    max_train_samples: Optional[int] = None
    max_eval_samples: Optional[int] = None
    streaming: bool = False
    block_size: Optional[int] = None
    overwrite_cache: bool = False
    validation_split_percentage: int = 5
    preprocessing_num_workers: Optional[int] = None
    keep_linebreaks: bool = True

    def validate(self, synthetic: bool):
        if synthetic:
            return
        if self.dataset_name is None and self.train_file is None and self.validation_file is None:
            raise ValueError("Need either a dataset name, a training/validation file, or --synthetic_data.")
        for f in (self.train_file, self.validation_file):
            if f is not None and f.split(".")[-1] not in ("csv", "json", "txt"):
                raise ValueError("`train_file`/`validation_file` should be a csv, a json or a txt file.")


def _text_column(ds) -> str:
    cols = getattr(ds, "column_names", None)
    if not cols:                                   # streaming datasets may not know their columns: peek
        first = next(iter(ds))
        cols = list(first.keys())
    return "text" if "text" in cols else cols[0]


def _hub_dataset(data_args, tokenizer, block_size, load_dataset=None):
    """openwebtext & co. through ``datasets`` when it is reachable (it is not on the B200 box).
    reference: run_clm.py:316-341 (load, carve a validation split when the corpus has none) and :463-544
    (tokenize + group_texts as cached, multi-process ``map`` calls).

    * ``--streaming``: the splits stay iterable; examples are tokenised and packed on the fly, sharded by rank on the
      raw stream (``StreamingBlockDataset``).  Without a validation split the first ``max_eval_samples`` (default
      1000) raw examples become the validation stream and training skips them.
    * otherwise the text column is tokenised once by ``--preprocessing_num_workers`` processes into memory-mapped
      token shards (``train/token_cache.py``), cached under the dataset's fingerprint; ``--overwrite_cache`` rebuilds."""
    if load_dataset is None:
        from datasets import load_dataset
    from distributed_lion_pytorch_b200.train.data import StreamingBlockDataset
    from distributed_lion_pytorch_b200.train.token_cache import MemmapBlockDataset, tokenize_text_iterable
    name, cfgname = data_args.dataset_name, data_args.dataset_config_name
    raw = load_dataset(name, cfgname, streaming=data_args.streaming)
    if data_args.streaming:
        col = _text_column(raw["train"])
        if "validation" in raw:
            train = StreamingBlockDataset(raw["train"], tokenizer, block_size, col, max_blocks=data_args.max_train_samples,
                                          infinite=True)
            valid = StreamingBlockDataset(raw["validation"], tokenizer, block_size, col, max_blocks=data_args.max_eval_samples)
        else:
            n_val = data_args.max_eval_samples or 1000
            train = StreamingBlockDataset(raw["train"], tokenizer, block_size, col, skip=n_val,
                                          max_blocks=data_args.max_train_samples, infinite=True)
            valid = StreamingBlockDataset(raw["train"], tokenizer, block_size, col, take=n_val,
                                          max_blocks=data_args.max_eval_samples)
        return train, valid
    if "validation" not in raw:
        pct = data_args.validation_split_percentage
        raw = {"validation": load_dataset(name, cfgname, split=f"train[:{pct}%]"),
               "train": load_dataset(name, cfgname, split=f"train[{pct}%:]")}
    out = {}
    for split, cap in (("train", data_args.max_train_samples), ("validation", data_args.max_eval_samples)):
        ds = raw[split]
        col = _text_column(ds)
        fp = str(getattr(ds, "_fingerprint", "")) or f"{len(ds)}"
        d = tokenize_text_iterable((ex[col] for ex in ds), tokenizer, f"{name}-{cfgname}-{split}",
                                   num_workers=data_args.preprocessing_num_workers,
                                   overwrite_cache=data_args.overwrite_cache, fingerprint=fp)
        out[split] = MemmapBlockDataset(d, block_size, max_blocks=cap)
    return out["train"], out["validation"]


def main(argv=None):
    model_args, data_args, training_args = parse_cli((ModelArguments, DataTrainingArguments, AsyncTrainingArguments), argv)
    data_args.validate(training_args.synthetic_data)

    logging.basicConfig(format="%(asctime)s - %(levelname)s - %(name)s - %(message)s", datefmt="%m/%d/%Y %H:%M:%S",
                        handlers=[logging.StreamHandler(sys.stdout)], level=training_args.get_process_log_level())
    device = training_args.torch_device      # brings the process group up under torchrun (run_clm.py:284 upstream)
    logger.warning("Process rank: %s, device: %s, world: %s, bf16: %s", training_args.process_index, device,
                   training_args.world_size, training_args.bf16 or model_args.torch_dtype == "bfloat16")

    last_checkpoint = None
    if os.path.isdir(training_args.output_dir) and training_args.do_train and not training_args.overwrite_output_dir:
        last_checkpoint = get_last_checkpoint(training_args.output_dir)
        if last_checkpoint is None and len(os.listdir(training_args.output_dir)) > 0:
            raise ValueError(f"Output directory ({training_args.output_dir}) already exists and is not empty. "
                             "Use --overwrite_output_dir to overcome.")
        if last_checkpoint is not None and training_args.resume_from_checkpoint is None:
            logger.info("Checkpoint detected, resuming training at %s.", last_checkpoint)
    set_seed(training_args.seed)

    # ---- model (random init from a config, or weights from a local checkpoint directory)
    dtype = None
    if model_args.torch_dtype not in (None, "auto"):
        dtype = getattr(torch, model_args.torch_dtype)
    name = model_args.model_name_or_path or model_args.config_name or model_args.model_type or "gpt2"
    model = build_model(name, dtype=dtype, device=device, pretrained=model_args.model_name_or_path is not None,
                        config_overrides=model_args.config_overrides)
    cfg = model.config
    logger.info("model %s: %.1fM parameters, dtype %s", name, model.num_parameters() / 1e6,
                next(model.parameters()).dtype)

    # ---- data
    max_pos = getattr(cfg, "n_positions", None) or getattr(cfg, "max_position_embeddings", 1024)
    block_size = min(data_args.block_size or 1024, max_pos, 1024 if data_args.block_size is None else 1 << 30)
    tokenizer = None
    if training_args.synthetic_data:
        n_train = data_args.max_train_samples or training_args.synthetic_samples
        n_eval = data_args.max_eval_samples or max(training_args.per_device_eval_batch_size * training_args.world_size * 2, 8)
        train_dataset = SyntheticTokenDataset(cfg.vocab_size, block_size, n_train, seed=training_args.seed)
        eval_dataset = SyntheticTokenDataset(cfg.vocab_size, block_size, n_eval, seed=training_args.seed + 1)
    else:
        tokenizer = load_tokenizer(model_args.tokenizer_name or model_args.model_name_or_path)
        if len(tokenizer) > cfg.vocab_size:
            raise ValueError(f"tokenizer has {len(tokenizer)} tokens but the model embeds {cfg.vocab_size}")
        from distributed_lion_pytorch_b200.parallel.runtime import main_process_first
        if data_args.dataset_name is not None:
            with main_process_first():            # reference: run_clm.py:474,531
                train_dataset, eval_dataset = _hub_dataset(data_args, tokenizer, block_size)
        else:
            kw = dict(num_workers=data_args.preprocessing_num_workers, overwrite_cache=data_args.overwrite_cache)
            with main_process_first():            # rank 0 tokenises into the cache, the others open it
                train_dataset = text_file_dataset(data_args.train_file, tokenizer, block_size, data_args.keep_linebreaks,
                                                  max_blocks=data_args.max_train_samples, **kw) \
                    if data_args.train_file else None
                eval_dataset = text_file_dataset(data_args.validation_file, tokenizer, block_size,
                                                 data_args.keep_linebreaks, max_blocks=data_args.max_eval_samples, **kw) \
                    if data_args.validation_file else None

    def preprocess_logits_for_metrics(logits, labels):
        return logits.argmax(dim=-1)

    def compute_metrics(eval_preds):
        preds, labels = eval_preds
        labels, preds = labels[:, 1:].reshape(-1), preds[:, :-1].reshape(-1)      # shift, as run_clm.py:571-576
        return {"accuracy": float((preds == labels).mean())}

    # ---- optimizer + schedule (reference: run_clm.py:580-585)
    optimizers = (None, None)
    if training_args.lion:
        if training_args.max_steps <= 0:
            if not hasattr(train_dataset, "__len__"):
                raise ValueError("a streaming dataset has no length: pass --max_steps (as HF's Trainer requires too)")
            per_epoch = len(train_dataset) // (training_args.per_device_train_batch_size * training_args.world_size
                                               * training_args.gradient_accumulation_steps)
            training_args.max_steps = max(int(training_args.num_train_epochs * per_epoch), 1)
        vote = "majority" if training_args.lion_vote == "stochastic" else training_args.lion_vote
        mgn = training_args.lion_max_grad_norm or (1.0 if training_args.lion_vote == "stochastic" else None)
        from distributed_lion_pytorch_b200.train.trainer import lion_comm_for
        sync = (not training_args.async_grad) or training_args.comm == "nccl_allreduce"
        comm = lion_comm_for(training_args, sync)
        optimizer = Lion(model.parameters(), lr=training_args.learning_rate, weight_decay=training_args.weight_decay,
                         max_grad_norm=mgn, vote=vote, tie=training_args.lion_tie, comm=comm, seed=training_args.seed,
                         stochastic_rounding=training_args.lion_stochastic_rounding)
        optimizers = (optimizer, get_cosine_schedule_with_warmup(optimizer, training_args.warmup_steps,
                                                                 training_args.max_steps))
    # reference: run_clm.py:587; --comm nccl_allreduce forces the synchronous-gradient baseline
    trainer_class = AsyncTrainer if training_args.async_grad and training_args.comm != "nccl_allreduce" else Trainer
    trainer = trainer_class(model=model, args=training_args,
                            train_dataset=train_dataset if training_args.do_train else None,
                            eval_dataset=eval_dataset if training_args.do_eval else None, tokenizer=tokenizer,
                            data_collator=default_data_collator,
                            compute_metrics=compute_metrics if training_args.do_eval else None,
                            preprocess_logits_for_metrics=preprocess_logits_for_metrics if training_args.do_eval else None,
                            optimizers=optimizers)

    results = {}
    if training_args.do_train:
        checkpoint = training_args.resume_from_checkpoint or last_checkpoint
        train_result = trainer.train(resume_from_checkpoint=checkpoint)
        trainer.save_model()
        metrics = train_result.metrics
        metrics["train_samples"] = len(train_dataset)
        trainer.log_metrics("train", metrics)
        trainer.save_metrics("train", metrics)
        trainer.save_state()
        results.update(metrics)
    if training_args.do_eval:
        metrics = trainer.evaluate()
        try:
            metrics["perplexity"] = math.exp(metrics["eval_loss"])
        except OverflowError:
            metrics["perplexity"] = float("inf")
        trainer.log_metrics("eval", metrics)
        trainer.save_metrics("eval", metrics)
        results.update(metrics)
    trainer.create_model_card(finetuned_from=model_args.model_name_or_path, tasks="text-generation",
                              dataset=data_args.dataset_name or ("synthetic" if training_args.synthetic_data else "files"))
    return results


def _mp_fn(index):
    main()


if __name__ == "__main__":
    main()
