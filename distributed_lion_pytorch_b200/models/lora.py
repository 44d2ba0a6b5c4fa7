"""Native LoRA adapters (stands in for PEFT, which is not installed on the box).

reference: ``LoraConfig(r=8, lora_alpha=16, lora_dropout=0.05, target_modules=["q_proj",
"v_proj"])`` in sft_llama2.py:44-51 and dpo_llama2.py:192-207, and
``AutoPeftModelForCausalLM.merge_and_unload`` in sft_llama2.py:195-196.  Unlike the reference
 the optimizer must be built AFTER ``apply_lora`` so it sees the adapter params.
"""
from __future__ import annotations

import contextlib
import math
from dataclasses import dataclass, field
from typing import Dict, Iterable, List

import torch
import torch.nn as nn
import torch.nn.functional as F

from .quant import Linear4bit, dequantize_base_


@dataclass
class LoraConfig:
    r: int = 8
    lora_alpha: int = 16
    lora_dropout: float = 0.05
    target_modules: List[str] = field(default_factory=lambda: ["q_proj", "v_proj"])
    bias: str = "none"
    task_type: str = "CAUSAL_LM"


class LoRALinear(nn.Module):
    """``base`` is an ``nn.Linear`` or a 4-bit :class:`Linear4bit` (the QLoRA arrangement of
    sft_llama2.py:141-154); the adapters are always 16/32-bit parameters."""

    def __init__(self, base: nn.Module, r: int, alpha: int, dropout: float):
        super().__init__()
        self.base = base
        self.r, self.scaling = r, alpha / r
        if isinstance(base, Linear4bit):
            dtype, device = base.compute_dtype, base.device
        else:
            dtype, device = base.weight.dtype, base.weight.device
            base.weight.requires_grad_(False)
        self.lora_A = nn.Parameter(torch.empty(r, base.in_features, dtype=dtype, device=device))
        self.lora_B = nn.Parameter(torch.zeros(base.out_features, r, dtype=dtype, device=device))
        nn.init.kaiming_uniform_(self.lora_A, a=math.sqrt(5))
        self.dropout = dropout
        self.enabled = True
        if base.bias is not None:
            base.bias.requires_grad_(False)

    def forward(self, x):
        y = self.base(x)
        if not self.enabled:
            return y
        xd = F.dropout(x, self.dropout, self.training) if self.dropout > 0 else x
        # the scale goes onto the [..., r] intermediate (tiny), not onto the [..., out] product.  (Folding the add into
        # the second GEMM — addmm with K = r = 8 and beta = 1 — was measured: cuBLAS runs it 25 % slower end to end.)
        h = (xd @ self.lora_A.t()) * self.scaling
        return y + h @ self.lora_B.t()

    @torch.no_grad()
    def merged(self) -> nn.Linear:
        if isinstance(self.base, Linear4bit):      # expand the 4-bit base and fold the adapters in, in fp32
            lin = self.base.to_linear(self.lora_B.float() @ self.lora_A.float() * self.scaling)
            lin.weight.requires_grad_(True)
            return lin
        self.base.weight.add_((self.lora_B.float() @ self.lora_A.float() * self.scaling).to(self.base.weight.dtype))
        self.base.weight.requires_grad_(True)
        return self.base


def apply_lora(model: nn.Module, cfg: LoraConfig) -> nn.Module:
    """Freeze the base model and wrap every ``nn.Linear`` whose name ends with a target."""
    for p in model.parameters():
        p.requires_grad_(False)
    hits = 0
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if isinstance(child, (nn.Linear, Linear4bit)) and any(name == t or name.endswith(t) for t in cfg.target_modules):
                setattr(parent, name, LoRALinear(child, cfg.r, cfg.lora_alpha, cfg.lora_dropout))
                hits += 1
    if hits == 0:
        raise ValueError(f"no module matched target_modules={cfg.target_modules}")
    model.peft_config = cfg
    return model


@contextlib.contextmanager
def disable_adapters(model: nn.Module):
    """Run the wrapped model as its frozen base (what TRL's DPOTrainer uses as the implicit reference
    model when only a PEFT policy is given)."""
    mods = [m for m in model.modules() if isinstance(m, LoRALinear)]
    old = [m.enabled for m in mods]
    for m in mods:
        m.enabled = False
    try:
        yield
    finally:
        for m, e in zip(mods, old):
            m.enabled = e


def merge_and_unload(model: nn.Module, dequantize: bool = True) -> nn.Module:
    """Fold every adapter into its base layer (reference: ``merge_and_unload`` at sft_llama2.py:195-196, done
    there on a bf16 reload of the base).  With a 4-bit base the remaining 4-bit layers are expanded too, so
    the result is a plain 16-bit model that ``save_pretrained`` writes in the ordinary HF layout."""
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if isinstance(child, LoRALinear):
                setattr(parent, name, child.merged())
    if dequantize:
        dequantize_base_(model)
    for p in model.parameters():
        p.requires_grad_(True)
    if hasattr(model, "peft_config"):
        del model.peft_config
    return model


def lora_state_dict(model: nn.Module) -> Dict[str, torch.Tensor]:
    return {k: v.detach().cpu().clone() for k, v in model.state_dict().items() if "lora_" in k}


def print_trainable_parameters(model: nn.Module) -> str:
    """reference: sft_llama2.py:78-90 (defined upstream, never called)."""
    trainable = sum(p.numel() for p in model.parameters() if p.requires_grad)
    total = sum(p.numel() for p in model.parameters())
    msg = f"trainable params: {trainable} || all params: {total} || trainable%: {100 * trainable / max(total, 1):.4f}"
    print(msg)
    return msg
