"""Model families: GPT-2 (run_clm) and Llama-2 (sft_llama2 / dpo_llama2), HF parameter names."""
from __future__ import annotations

import contextlib
import json
import os
from typing import Optional

import torch

from .common import CausalLMOutput, checkpoint_files, has_weights, load_checkpoint_into  # noqa: F401
from .gpt2 import GPT2_PRESETS, GPT2Config, GPT2LMHeadModel  # noqa: F401
from .llama import LLAMA_PRESETS, LlamaConfig, LlamaForCausalLM  # noqa: F401
from .lora import (LoraConfig, LoRALinear, apply_lora, disable_adapters, lora_state_dict,  # noqa: F401
                   merge_and_unload)
from .quant import Linear4bit, dequantize_base_, quantize_base_, quantized_bytes  # noqa: F401

_ALIASES = {
    "openai-community/gpt2": "gpt2", "gpt2": "gpt2", "gpt2-medium": "gpt2-medium", "gpt2-large": "gpt2-large",
    "gpt2-xl": "gpt2-xl", "gpt2-tiny": "gpt2-tiny",
    "meta-llama/Llama-2-7b-hf": "llama-2-7b", "llama-2-7b": "llama-2-7b", "meta-llama/Llama-2-13b-hf": "llama-2-13b",
    "llama-2-13b": "llama-2-13b", "llama-1b": "llama-1b", "llama-tiny": "llama-tiny",
}


def config_for(name_or_path: str):
    """Resolve a preset name / HF hub id / local checkpoint directory to a config object
    (stands in for ``AutoConfig.from_pretrained``, reference: run_clm.py:397-406)."""
    if os.path.isdir(name_or_path) and os.path.exists(os.path.join(name_or_path, "config.json")):
        with open(os.path.join(name_or_path, "config.json")) as f:
            mt = json.load(f).get("model_type")
        return (GPT2Config if mt == "gpt2" else LlamaConfig).from_pretrained(name_or_path)
    key = _ALIASES.get(name_or_path)
    if key is None:
        raise ValueError(f"unknown model {name_or_path!r}: give a preset ({sorted(set(_ALIASES.values()))}) "
                         "or a local checkpoint directory (no network on this box)")
    if key in GPT2_PRESETS:
        return GPT2Config(**GPT2_PRESETS[key])
    return LlamaConfig(**LLAMA_PRESETS[key])


def model_class_for(config):
    return GPT2LMHeadModel if isinstance(config, GPT2Config) else LlamaForCausalLM


def build_model(name_or_path: str, *, dtype: Optional[torch.dtype] = None, device=None, pretrained: bool = True,
                config_overrides: Optional[str] = None, load_in_4bit: bool = False, quant_type: str = "nf4"):
    """``AutoModelForCausalLM`` stand-in: load weights when ``name_or_path`` is a checkpoint
    directory, otherwise random-init the named architecture (there is no hub access here).
    ``load_in_4bit`` turns every linear layer but the LM head into a frozen 4-bit ``Linear4bit``
    (reference: ``quantization_config=bnb_config`` at sft_llama2.py:147-150, dpo_llama2.py:137)."""
    model = _build_model(name_or_path, dtype=dtype, device=device, pretrained=pretrained,
                         config_overrides=config_overrides)
    if load_in_4bit:
        quantize_base_(model, quant_type=quant_type)
    return model


def _build_model(name_or_path: str, *, dtype, device, pretrained, config_overrides):
    cfg = config_for(name_or_path)
    if config_overrides:
        cfg.update_from_string(config_overrides)
    cls = model_class_for(cfg)
    is_ckpt = has_weights(name_or_path)
    if pretrained and is_ckpt:
        return cls.from_pretrained(name_or_path, dtype=dtype, device=device)
    ctx = torch.device(device) if device is not None else contextlib.nullcontext()
    with ctx:
        model = cls(cfg)
    if dtype is not None:
        model = model.to(dtype)
        cfg.dtype_name = str(dtype).split(".")[-1]
    return model
