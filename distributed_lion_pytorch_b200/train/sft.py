"""Supervised fine-tuning trainers (stands in for TRL's ``SFTTrainer``, which is not installed).

reference: sft_llama2.py:170-186 — ``SFTTrainer(model, train_dataset, eval_dataset,
peft_config, packing=True, max_seq_length=None, tokenizer, args, optimizers)`` over two
``ConstantLengthDataset`` s; async_trainer.py:37-62 is its no-sync twin.  With packing the loss
is plain next-token cross-entropy on constant-length blocks, i.e. the CLM step.
"""
from __future__ import annotations

import logging
from typing import Optional

import torch

from ..models.lora import LoraConfig, apply_lora
from ..optim.lion import Lion
from .data import ConstantLengthDataset
from .trainer import NoSyncMixin, Trainer

logger = logging.getLogger("dlion.sft")


class SFTTrainer(Trainer):
    def __init__(self, model=None, args=None, train_dataset=None, eval_dataset=None, peft_config: Optional[LoraConfig] = None,
                 packing: bool = True, max_seq_length: Optional[int] = None, tokenizer=None, formatting_func=None,
                 optimizers=(None, None), **kw):
        if peft_config is not None:
            model = apply_lora(model, peft_config)
            opt = optimizers[0]
            if opt is not None and not any(p.requires_grad for g in opt.param_groups for p in g["params"]):
                # upstream builds Lion(base_model.parameters()) before the adapters exist,
                # so nothing trainable is in the optimizer.  Re-point the groups at the adapters.
                logger.warning("optimizer was built before LoRA wrapping; re-pointing it at the adapter parameters")
                trainable = [p for p in model.parameters() if p.requires_grad]
                opt.param_groups[0]["params"] = trainable
                for g in opt.param_groups[1:]:
                    g["params"] = []
                opt.state.clear()
                if isinstance(opt, Lion):
                    opt.refresh_active_set()
        if packing and tokenizer is not None and train_dataset is not None \
                and not isinstance(train_dataset, (ConstantLengthDataset, torch.utils.data.Dataset)):
            train_dataset = ConstantLengthDataset(tokenizer, train_dataset, formatting_func, infinite=True,
                                                  seq_length=max_seq_length or 1024)
        super().__init__(model=model, args=args, train_dataset=train_dataset, eval_dataset=eval_dataset,
                         tokenizer=tokenizer, optimizers=optimizers, **kw)
        self.packing, self.max_seq_length = packing, max_seq_length


class AsyncSFTTrainer(NoSyncMixin, SFTTrainer):
    """reference: async_trainer.py:37-62"""
