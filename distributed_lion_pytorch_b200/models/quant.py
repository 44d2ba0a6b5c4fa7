"""4-bit (NF4 / FP4) frozen base weights for LoRA fine-tuning — the QLoRA arrangement.

reference: sft_llama2.py:141-154 loads the base model with ``BitsAndBytesConfig(load_in_4bit=True,
bnb_4bit_quant_type="nf4", bnb_4bit_compute_dtype=torch.bfloat16)`` and dpo_llama2.py:133-152 loads
policy and reference model with ``load_in_4bit=True``; upstream the layers are bitsandbytes'
``Linear4bit``.  Here :class:`Linear4bit` keeps ``weight_packed`` (uint8, two elements per byte) and
``absmax`` (fp32, one per block of 64) and expands them with one kernel (``ops/quant4.py`` →
``csrc/quant4.cu``) into a shared bf16 workspace right before the cuBLAS GEMM, in forward and again in
backward — so a 7 B-parameter base costs 3.6 GB of HBM instead of 13.5 GB and nothing 16-bit is kept
per layer.  The weight is frozen: only the input gradient is produced.
"""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import quant4

_WORKSPACE: Dict[Tuple[torch.device, torch.dtype], torch.Tensor] = {}


def _workspace(n: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    """One expansion buffer per (device, dtype), grown to the largest layer.  Kernels that read it are
    enqueued on the same stream as the next expansion that overwrites it, so reuse is ordered."""
    key = (device, dtype)
    buf = _WORKSPACE.get(key)
    if buf is None or buf.numel() < n:
        buf = torch.empty(n, dtype=dtype, device=device)
        _WORKSPACE[key] = buf
    return buf


class _Linear4bitFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, layer):
        ctx.layer = layer
        ctx.bias_grad = bias is not None and bias.requires_grad
        return F.linear(x, layer.dequantize(shared=True), bias)

    @staticmethod
    def backward(ctx, gy):
        layer = ctx.layer
        gx = gb = None
        if ctx.needs_input_grad[0]:
            w = layer.dequantize(shared=True)                       # expanded again: nothing 16-bit was saved
            gx = (gy.reshape(-1, gy.shape[-1]) @ w).view(*gy.shape[:-1], w.shape[1])
        if ctx.bias_grad:
            gb = gy.reshape(-1, gy.shape[-1]).sum(0)
        return gx, gb, None


class Linear4bit(nn.Module):
    """``y = x @ dequant(W4)^T + b`` with a frozen 4-bit weight."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False, *,
                 compute_dtype: torch.dtype = torch.bfloat16, blocksize: int = quant4.DEFAULT_BLOCKSIZE,
                 quant_type: str = "nf4", device=None):
        super().__init__()
        n = in_features * out_features
        quant4.code_book(quant_type)
        if n % blocksize:
            raise ValueError(f"{out_features}x{in_features} weight is not a multiple of blocksize {blocksize}")
        self.in_features, self.out_features = in_features, out_features
        self.compute_dtype, self.blocksize, self.quant_type = compute_dtype, blocksize, quant_type
        self.register_buffer("weight_packed", torch.zeros(n // 2, dtype=torch.uint8, device=device))
        self.register_buffer("absmax", torch.zeros(n // blocksize, dtype=torch.float32, device=device))
        self.bias = nn.Parameter(torch.zeros(out_features, dtype=compute_dtype, device=device),
                                 requires_grad=False) if bias else None
        self._origin_cls = nn.Linear

    @classmethod
    def from_linear(cls, lin: nn.Linear, *, blocksize: int = quant4.DEFAULT_BLOCKSIZE, quant_type: str = "nf4",
                    compute_dtype: Optional[torch.dtype] = None) -> "Linear4bit":
        w = lin.weight
        dt = compute_dtype or w.dtype           # upstream passes bnb_4bit_compute_dtype=bfloat16 for a bf16 model
        q = cls(lin.in_features, lin.out_features, lin.bias is not None, compute_dtype=dt, blocksize=blocksize,
                quant_type=quant_type, device=w.device)
        packed, absmax = quant4.quantize_4bit(w.detach(), blocksize, quant_type)
        q.weight_packed.copy_(packed)
        q.absmax.copy_(absmax)
        if lin.bias is not None:
            q.bias.data.copy_(lin.bias.detach().to(dt))
        q._origin_cls = type(lin)
        return q

    def _apply(self, fn, recurse=True):
        # module.to(dtype=...) casts every floating buffer; the block scales must stay fp32
        keep = self.absmax
        super()._apply(fn, recurse)
        if self.absmax.dtype != torch.float32:
            self.absmax = keep.to(self.absmax.device)
        return self

    @property
    def dtype(self) -> torch.dtype:
        return self.compute_dtype

    @property
    def device(self) -> torch.device:
        return self.weight_packed.device

    def dequantize(self, shared: bool = False, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
        """The [out, in] weight in ``dtype`` (default: the compute dtype).  ``shared=True`` expands into the
        per-device workspace: valid until the next 4-bit layer runs."""
        dt = dtype or self.compute_dtype
        n = self.in_features * self.out_features
        out = _workspace(n, dt, self.weight_packed.device) if shared else None
        return quant4.dequantize_4bit(self.weight_packed, self.absmax, (self.out_features, self.in_features), dt,
                                      self.blocksize, self.quant_type, out=out)

    def forward(self, x):
        if x.dtype != self.compute_dtype:        # compute in compute_dtype, hand back the caller's dtype
            return _Linear4bitFn.apply(x.to(self.compute_dtype), self.bias, self).to(x.dtype)
        return _Linear4bitFn.apply(x, self.bias, self)

    @torch.no_grad()
    def to_linear(self, delta: Optional[torch.Tensor] = None) -> nn.Linear:
        """Back to a 16-bit layer of the original class (``delta`` — e.g. a merged LoRA update — is added in fp32)."""
        w = self.dequantize(dtype=torch.float32)
        if delta is not None:
            w = w + delta.float()
        lin = self._origin_cls(self.in_features, self.out_features, bias=self.bias is not None,
                               device=self.weight_packed.device, dtype=self.compute_dtype)
        lin.weight.copy_(w.to(self.compute_dtype))
        if self.bias is not None:
            lin.bias.copy_(self.bias)
        return lin

    def extra_repr(self) -> str:
        return (f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, "
                f"{self.quant_type}, blocksize={self.blocksize}, compute_dtype={self.compute_dtype}")


def quantize_base_(model: nn.Module, *, quant_type: str = "nf4", blocksize: int = quant4.DEFAULT_BLOCKSIZE,
                   compute_dtype: Optional[torch.dtype] = None, skip: Iterable[str] = ("lm_head",)) -> nn.Module:
    """Replace every ``nn.Linear`` (except names in ``skip`` — the LM head stays 16-bit, as with
    ``load_in_4bit`` upstream) by a frozen :class:`Linear4bit`.  Call before ``apply_lora`` and before
    building the optimizer.  Returns the model; ``model.quantization`` records what was done."""
    skip = tuple(skip)
    hits = 0
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if type(child) is Linear4bit or not isinstance(child, nn.Linear) or name in skip:
                continue
            if (child.in_features * child.out_features) % blocksize:
                continue
            setattr(parent, name, Linear4bit.from_linear(child, blocksize=blocksize, quant_type=quant_type,
                                                         compute_dtype=compute_dtype))
            hits += 1
    if hits == 0:
        raise ValueError("quantize_base_ found no nn.Linear to convert")
    model.quantization = {"quant_type": quant_type, "blocksize": blocksize, "layers": hits}
    return model


def dequantize_base_(model: nn.Module) -> nn.Module:
    """Inverse of :func:`quantize_base_`: every :class:`Linear4bit` becomes a 16-bit layer of its original class
    again (values are the 4-bit ones — the rounding is not undone)."""
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if isinstance(child, Linear4bit):
                setattr(parent, name, child.to_linear())
    if hasattr(model, "quantization"):
        del model.quantization
    return model


def quantized_bytes(model: nn.Module) -> int:
    """Bytes held by the 4-bit layers (packed weights + block scales)."""
    return sum(m.weight_packed.numel() + 4 * m.absmax.numel() for m in model.modules() if isinstance(m, Linear4bit))
