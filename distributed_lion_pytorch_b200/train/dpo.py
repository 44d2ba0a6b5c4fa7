"""Direct Preference Optimisation trainers (stands in for TRL's ``DPOTrainer``).

reference: dpo_llama2.py:219-231 — ``DPOTrainer(model, model_ref, args, beta, train_dataset,
eval_dataset, tokenizer, peft_config, max_prompt_length, max_length)``; async_trainer.py:65-90
is its no-sync twin.  One step = concatenated chosen+rejected forward of the policy (with
grad) and of the frozen reference model (no grad), per-sequence log-prob sums over the answer
tokens, loss ``-log σ(β[(π_c-π_r)-(ref_c-ref_r)])``.  The per-token log-probs come from the
chunked N5 kernels (``model.token_log_probs``), so neither model ever materialises
[2B, L, vocab] logits.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from ..models.lora import LoraConfig, apply_lora, disable_adapters
from .trainer import NoSyncMixin, Trainer

IGNORE = -100


class DPODataCollator:
    """Builds ``[2B, L]`` batches: rows 0..B-1 = prompt+chosen, rows B..2B-1 = prompt+rejected,
    right-padded to ``max_length`` (static shapes), labels masked on prompt and pad tokens.
    Text triples (reference: dpo_llama2.py:107-116) are tokenised here when a tokenizer is given."""

    def __init__(self, tokenizer=None, max_prompt_length: int = 512, max_length: int = 1024, pad_token_id: int = 0,
                 pad_to_max: bool = True):
        self.tok, self.mp, self.ml, self.pad, self.pad_to_max = tokenizer, max_prompt_length, max_length, pad_token_id, pad_to_max

    def _ids(self, ex, key_text, key_ids):
        if key_ids in ex:
            v = ex[key_ids]
            return v.tolist() if torch.is_tensor(v) else list(v)
        return list(self.tok(ex[key_text])["input_ids"])

    def _row(self, prompt: List[int], answer: List[int]):
        prompt = prompt[-self.mp:]                               # keep the end of long prompts (TRL behaviour)
        answer = answer[: max(self.ml - len(prompt), 1)]
        ids = prompt + answer
        labels = [IGNORE] * len(prompt) + answer
        return ids[: self.ml], labels[: self.ml]

    def __call__(self, features: List[dict]) -> Dict[str, torch.Tensor]:
        rows = []
        for key_t, key_i in (("chosen", "chosen_ids"), ("rejected", "rejected_ids")):
            for ex in features:
                rows.append(self._row(self._ids(ex, "prompt", "prompt_ids"), self._ids(ex, key_t, key_i)))
        L = self.ml if self.pad_to_max else max(len(r[0]) for r in rows)
        ids = torch.full((len(rows), L), self.pad, dtype=torch.int64)
        labels = torch.full((len(rows), L), IGNORE, dtype=torch.int64)
        for i, (a, b) in enumerate(rows):
            ids[i, : len(a)] = torch.tensor(a, dtype=torch.int64)
            labels[i, : len(b)] = torch.tensor(b, dtype=torch.int64)
        return {"input_ids": ids, "labels": labels}


def dpo_loss(pi_c, pi_r, ref_c, ref_r, beta: float):
    logits = (pi_c - pi_r) - (ref_c - ref_r)
    loss = -F.logsigmoid(beta * logits)
    chosen_rewards = beta * (pi_c - ref_c).detach()
    rejected_rewards = beta * (pi_r - ref_r).detach()
    return loss.mean(), chosen_rewards, rejected_rewards


class DPOTrainer(Trainer):
    def __init__(self, model=None, ref_model=None, args=None, beta: float = 0.1, train_dataset=None, eval_dataset=None,
                 tokenizer=None, peft_config: Optional[LoraConfig] = None, max_prompt_length: int = 512,
                 max_length: int = 1024, optimizers=(None, None), data_collator=None, **kw):
        if peft_config is not None:
            model = apply_lora(model, peft_config)
        if data_collator is None:
            pad = getattr(tokenizer, "pad_token_id", None)
            data_collator = DPODataCollator(tokenizer, max_prompt_length, max_length, pad if pad is not None else 0)
        super().__init__(model=model, args=args, train_dataset=train_dataset, eval_dataset=eval_dataset,
                         tokenizer=tokenizer, optimizers=optimizers, data_collator=data_collator, **kw)
        self.beta = beta
        self.ref_model = ref_model.to(self.device) if ref_model is not None else None
        if self.ref_model is not None:
            self.ref_model.eval()
            for p in self.ref_model.parameters():
                p.requires_grad_(False)
        self._last_metrics: Dict[str, float] = {}

    def _seq_logps(self, model, ids, labels):
        return model.token_log_probs(ids, labels).sum(dim=-1)          # [2B] fp32

    def compute_loss(self, model, inputs, return_outputs: bool = False):
        ids, labels = inputs["input_ids"], inputs["labels"]
        B = ids.shape[0] // 2
        pi = self._seq_logps(model, ids, labels)
        with torch.no_grad():
            if self.ref_model is not None:
                ref = self._seq_logps(self.ref_model, ids, labels)
            elif hasattr(model, "peft_config"):     # implicit reference = the policy with its adapters switched off
                with disable_adapters(model):
                    ref = self._seq_logps(model, ids, labels)
            else:
                raise ValueError("DPO needs a reference model (or a LoRA-wrapped policy whose base serves as one)")
        loss, cr, rr = dpo_loss(pi[:B], pi[B:], ref[:B], ref[B:], self.beta)
        self._last_metrics = {"rewards/chosen": cr.mean(), "rewards/rejected": rr.mean(),
                              "rewards/accuracies": (cr > rr).float().mean(), "rewards/margins": (cr - rr).mean()}
        return (loss, None) if return_outputs else loss

    def log(self, logs):
        if "loss" in logs and self._last_metrics:
            logs = {**logs, **{k: float(v) for k, v in self._last_metrics.items()}}
        super().log(logs)

    @torch.no_grad()
    def evaluate(self, eval_dataset=None, metric_key_prefix: str = "eval"):
        loader, _ = self.get_eval_dataloader(eval_dataset)
        self.model.eval()
        tot, acc, n = 0.0, 0.0, 0
        for i, inputs in enumerate(loader):
            if self.args.max_eval_batches is not None and i >= self.args.max_eval_batches:
                break
            inputs = self._prepare_inputs(inputs)
            loss = self.compute_loss(self.model, inputs)
            tot += float(loss)
            acc += float(self._last_metrics["rewards/accuracies"])
            n += 1
        from ..parallel import runtime as rt
        m = {f"{metric_key_prefix}_loss": rt.all_reduce_scalar(tot / max(n, 1), "mean", self.device),
             f"{metric_key_prefix}_rewards/accuracies": rt.all_reduce_scalar(acc / max(n, 1), "mean", self.device)}
        super().log(m)
        self.model.train()
        return m


class AsyncDPOTrainer(NoSyncMixin, DPOTrainer):
    """reference: async_trainer.py:65-90"""
