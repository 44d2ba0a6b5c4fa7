"""Shared model plumbing: HF-style config (de)serialisation, safetensors checkpoints, outputs."""
from __future__ import annotations

import dataclasses
import json
import os
import re
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence

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
        rp = d.get("rope_parameters")                        # transformers >= 5 nests rope_theta here
        if isinstance(rp, dict) and "rope_theta" in rp and "rope_theta" not in d:
            d["rope_theta"] = rp["rope_theta"]
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


def _parse_size(x) -> int:
    if isinstance(x, int):
        return x
    m = re.fullmatch(r"\s*([0-9.]+)\s*([KMGT]?)i?B?\s*", str(x), flags=re.I)
    if not m:
        raise ValueError(f"cannot parse size {x!r} (expected e.g. '5GB', '200MB' or an int of bytes)")
    return int(float(m.group(1)) * {"": 1, "K": 10 ** 3, "M": 10 ** 6, "G": 10 ** 9, "T": 10 ** 12}[m.group(2).upper()])


WEIGHT_FILES = ("model.safetensors", "model.safetensors.index.json", "pytorch_model.bin", "pytorch_model.bin.index.json")


def has_weights(path: str) -> bool:
    return os.path.isdir(path) and any(os.path.exists(os.path.join(path, f)) for f in WEIGHT_FILES)


def checkpoint_files(path: str) -> List[str]:
    """The weight files of an HF checkpoint directory, single-file or sharded (``*.index.json`` →
    ``weight_map``), safetensors preferred — what ``from_pretrained`` of transformers resolves
    (reference: ``AutoModelForCausalLM.from_pretrained`` at run_clm.py:425-440, sft_llama2.py:147)."""
    for single, index in (("model.safetensors", "model.safetensors.index.json"),
                          ("pytorch_model.bin", "pytorch_model.bin.index.json")):
        if os.path.exists(os.path.join(path, single)):
            return [os.path.join(path, single)]
        if os.path.exists(os.path.join(path, index)):
            with open(os.path.join(path, index)) as fh:
                shards = sorted(set(json.load(fh)["weight_map"].values()))
            missing = [s for s in shards if not os.path.exists(os.path.join(path, s))]
            if missing:
                raise FileNotFoundError(f"{path}: index lists shards that are not there: {missing}")
            return [os.path.join(path, s) for s in shards]
    raise FileNotFoundError(f"no model weights under {path} (looked for {', '.join(WEIGHT_FILES)})")


def _read_weight_file(file: str) -> Dict[str, torch.Tensor]:
    if file.endswith(".safetensors"):
        from safetensors.torch import load_file
        return load_file(file)
    return torch.load(file, map_location="cpu", weights_only=True)


@torch.no_grad()
def load_checkpoint_into(model: torch.nn.Module, path: str, ignore_missing: Sequence[str] = ()):
    """Copy an HF checkpoint into ``model`` one shard at a time (host memory holds one shard, never the
    whole model).  Returns ``(missing, unexpected)`` key lists; keys in ``ignore_missing`` (tied heads)
    are not reported."""
    own = dict(model.state_dict(keep_vars=True))
    seen, unexpected = set(), []
    for file in checkpoint_files(path):
        for k, v in _read_weight_file(file).items():
            t = own.get(k)
            if t is None:
                unexpected.append(k)
                continue
            if t.shape != v.shape:
                raise RuntimeError(f"{k}: checkpoint shape {tuple(v.shape)} != model shape {tuple(t.shape)}")
            t.copy_(v)                 # casts to the model's dtype and moves to its device
            seen.add(k)
    missing = [k for k in own if k not in seen and k not in ignore_missing]
    return missing, unexpected


class PreTrainedMixin:
    """``save_pretrained`` / ``from_pretrained`` with the HF file layout (config.json +
    model.safetensors or ``model-0000i-of-0000n.safetensors`` + index, HF parameter names, tied
    lm_head not duplicated)."""
    config_class = ConfigBase
    tied_keys = ()

    def save_pretrained(self, path: str, safe_serialization: bool = True, max_shard_size="5GB") -> None:
        os.makedirs(path, exist_ok=True)
        self.config.save_pretrained(path)
        skip = self.tied_keys_present()
        items = [(k, v) for k, v in self.state_dict().items() if k not in skip]
        limit = _parse_size(max_shard_size)
        shards, cur, cur_bytes = [], {}, 0
        for k, v in items:
            nb = v.numel() * v.element_size()
            if cur and cur_bytes + nb > limit:
                shards.append(cur)
                cur, cur_bytes = {}, 0
            cur[k] = v
            cur_bytes += nb
        shards.append(cur)
        stem, ext = ("model", "safetensors") if safe_serialization else ("pytorch_model", "bin")

        def write(sd, file):
            sd = {k: v.detach().to("cpu").contiguous().clone() for k, v in sd.items()}
            if safe_serialization:
                from safetensors.torch import save_file
                save_file(sd, file, metadata={"format": "pt"})
            else:
                torch.save(sd, file)
        for stale in os.listdir(path):          # a previous save with another shard count must not linger
            if re.fullmatch(rf"{stem}(-\d+-of-\d+)?\.{ext}(\.index\.json)?", stale):
                os.remove(os.path.join(path, stale))
        if len(shards) == 1:
            write(shards[0], os.path.join(path, f"{stem}.{ext}"))
            return
        weight_map, total = {}, 0
        for i, sd in enumerate(shards, 1):
            name = f"{stem}-{i:05d}-of-{len(shards):05d}.{ext}"
            write(sd, os.path.join(path, name))
            for k, v in sd.items():
                weight_map[k] = name
                total += v.numel() * v.element_size()
        with open(os.path.join(path, f"{stem}.{ext}.index.json"), "w") as fh:
            json.dump({"metadata": {"total_size": total}, "weight_map": weight_map}, fh, indent=2, sort_keys=True)

    def tied_keys_present(self):
        return tuple(self.tied_keys) if getattr(self.config, "tie_word_embeddings", False) else ()

    @classmethod
    def from_pretrained(cls, path: str, dtype: Optional[torch.dtype] = None, device=None, **_):
        """Builds the module tree on the meta device (no random initialisation, no fp32 staging copy),
        materialises it in its final dtype on its final device and streams the checkpoint in shard by shard."""
        cfg = cls.config_class.from_pretrained(path)
        dtype = dtype or getattr(torch, cfg.dtype_name, torch.float32)
        old = torch.get_default_dtype()
        try:
            if dtype.is_floating_point:
                torch.set_default_dtype(dtype)
            with torch.device("meta"):
                model = cls(cfg)
        finally:
            torch.set_default_dtype(old)
        model.to_empty(device=device if device is not None else "cpu")
        missing, unexpected = load_checkpoint_into(model, path, ignore_missing=cls.tied_keys)
        if missing or unexpected:
            raise RuntimeError(f"checkpoint mismatch: missing={missing} unexpected={unexpected}")
        if hasattr(model, "tie_weights"):
            model.tie_weights()
        cfg.dtype_name = str(dtype).split(".")[-1]
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
