"""Shared model plumbing: HF-style config (de)serialisation, safetensors checkpoints, outputs."""
from __future__ import annotations

import dataclasses
import json
import os
from dataclasses import dataclass
from typing import Any, Dict, Optional

import torch


@dataclass
class CausalLMOutput:
    loss: Optional[torch.Tensor] = None
    logits: Optional[torch.Tensor] = None
    hidden_states: Optional[torch.Tensor] = None

    def __getitem__(self, k):
        if isinstance(k, int):
            return [v for v in (self.loss, self.logits) if v is not None][k]
        return getattr(self, k)


class ConfigBase:
    model_type = "base"
    architectures = ()

    def to_dict(self) -> Dict[str, Any]:
        d = dataclasses.asdict(self)
        d["model_type"] = self.model_type
        d["architectures"] = list(self.architectures)
        d["torch_dtype"] = d.pop("dtype_name", "float32")
        return d

    @classmethod
    def from_dict(cls, d: Dict[str, Any]):
        d = dict(d)
        if "torch_dtype" in d:
            d["dtype_name"] = d.pop("torch_dtype") or "float32"
        if "dtype" in d and isinstance(d["dtype"], str):      # transformers >= 5 writes "dtype"
            d["dtype_name"] = d.pop("dtype")
        names = {f.name for f in dataclasses.fields(cls)}
        return cls(**{k: v for k, v in d.items() if k in names})

    def save_pretrained(self, path: str) -> None:
        os.makedirs(path, exist_ok=True)
        with open(os.path.join(path, "config.json"), "w") as f:
            json.dump(self.to_dict(), f, indent=2, sort_keys=True)

    @classmethod
    def from_pretrained(cls, path: str):
        with open(os.path.join(path, "config.json")) as f:
            return cls.from_dict(json.load(f))

    def update_from_string(self, s: str) -> None:
        """``--config_overrides "n_embd=10,n_layer=2"`` (reference: run_clm.py:402-405)."""
        for kv in s.split(","):
            if not kv.strip():
                continue
            k, v = kv.split("=")
            k = k.strip()
            if not hasattr(self, k):
                raise ValueError(f"unknown config key {k}")
            old = getattr(self, k)
            if isinstance(old, bool):
                new = v.strip().lower() in ("1", "true", "yes")
            elif isinstance(old, int):
                new = int(v)
            elif isinstance(old, float):
                new = float(v)
            else:
                new = v.strip()
            setattr(self, k, new)


class PreTrainedMixin:
    """``save_pretrained`` / ``from_pretrained`` with the HF file layout (config.json +
    model.safetensors, HF parameter names, tied lm_head not duplicated)."""
    config_class = ConfigBase
    tied_keys = ()

    def save_pretrained(self, path: str, safe_serialization: bool = True) -> None:
        os.makedirs(path, exist_ok=True)
        self.config.save_pretrained(path)
        sd = {k: v.detach().to("cpu").contiguous().clone() for k, v in self.state_dict().items()
              if k not in self.tied_keys_present()}
        if safe_serialization:
            from safetensors.torch import save_file
            save_file(sd, os.path.join(path, "model.safetensors"), metadata={"format": "pt"})
        else:
            torch.save(sd, os.path.join(path, "pytorch_model.bin"))

    def tied_keys_present(self):
        return tuple(self.tied_keys) if getattr(self.config, "tie_word_embeddings", False) else ()

    @classmethod
    def from_pretrained(cls, path: str, dtype: Optional[torch.dtype] = None, device=None, **_):
        cfg = cls.config_class.from_pretrained(path)
        model = cls(cfg)
        st = os.path.join(path, "model.safetensors")
        if os.path.exists(st):
            from safetensors.torch import load_file
            sd = load_file(st)
        else:
            sd = torch.load(os.path.join(path, "pytorch_model.bin"), map_location="cpu")
        missing, unexpected = model.load_state_dict(sd, strict=False)
        missing = [k for k in missing if k not in cls.tied_keys]
        if missing or unexpected:
            raise RuntimeError(f"checkpoint mismatch: missing={missing} unexpected={unexpected}")
        if hasattr(model, "tie_weights"):
            model.tie_weights()
        dtype = dtype or getattr(torch, cfg.dtype_name, torch.float32)
        model = model.to(dtype=dtype)
        if device is not None:
            model = model.to(device)
        return model

    def num_parameters(self, only_trainable: bool = False) -> int:
        seen, total = set(), 0
        for p in self.parameters():
            if id(p) in seen or (only_trainable and not p.requires_grad):
                continue
            seen.add(id(p))
            total += p.numel()
        return total

    def gradient_checkpointing_enable(self, **_):
        self.gradient_checkpointing = True

    def gradient_checkpointing_disable(self):
        self.gradient_checkpointing = False


def shift_labels(labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """Next-token targets aligned with the hidden states: position t predicts labels[t+1]; the
    last position is ignored.  Avoids slicing (and copying) the [B, S, H] hidden states."""
    shifted = torch.full_like(labels, ignore_index)
    shifted[..., :-1] = labels[..., 1:]
    return shifted
