"""Blockwise 4-bit weight codec (NF4 / FP4) — ``csrc/quant4.cu`` on CUDA, plain torch on CPU.

reference: ``BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_quant_type="nf4",
bnb_4bit_compute_dtype=torch.bfloat16)`` at sft_llama2.py:141-145 and ``load_in_4bit=True`` at
dpo_llama2.py:137,151.  Upstream this is the bitsandbytes package (not installed on the image); the
storage layout here is the same — element ``2i`` in the high nibble of byte ``i``, one fp32 absmax per
block of 64 — so the torch functions in this file double as the oracle of the kernels.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _cabi, require_library

# NormalFloat-4 quantiles (QLoRA, Dettmers et al. 2023, appendix E): 16 values, symmetric zero.
NF4 = (-1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453, -0.28444138169288635,
       -0.18477343022823334, -0.09105003625154495, 0.0, 0.07958029955625534, 0.16093020141124725,
       0.24611230194568634, 0.33791524171829224, 0.44070982933044434, 0.5626170039176941,
       0.7229568362236023, 1.0)
# 4-bit float, sign-magnitude: bit 3 = sign, entries scaled so the largest magnitude is 1.
FP4 = (0.0, 0.0625 / 12, 8.0 / 12, 1.0, 4.0 / 12, 6.0 / 12, 2.0 / 12, 3.0 / 12,
       -0.0, -0.0625 / 12, -8.0 / 12, -1.0, -4.0 / 12, -6.0 / 12, -2.0 / 12, -3.0 / 12)
CODE_BOOKS = {"nf4": NF4, "fp4": FP4}
DEFAULT_BLOCKSIZE = 64


def code_book(quant_type: str) -> Tuple[float, ...]:
    try:
        return CODE_BOOKS[quant_type]
    except KeyError:
        raise ValueError(f"quant_type must be one of {sorted(CODE_BOOKS)}, got {quant_type!r}") from None


_CODE_PTRS = {}


def _code_ptr(quant_type: str):
    p = _CODE_PTRS.get(quant_type)
    if p is None:
        p = _CODE_PTRS[quant_type] = (C.c_float * 16)(*code_book(quant_type))
    return p


def _check(n: int, blocksize: int) -> None:
    if blocksize < 64 or blocksize & (blocksize - 1):
        raise ValueError(f"blocksize must be a power of two >= 64 (64 ... 4096 upstream), got {blocksize}")
    if n % blocksize:
        raise ValueError(f"tensor size {n} is not a multiple of blocksize {blocksize}")


# ---------------------------------------------------------------------- torch oracle / CPU path
def quantize_4bit_reference(w: torch.Tensor, blocksize: int = DEFAULT_BLOCKSIZE, quant_type: str = "nf4"):
    n = w.numel()
    _check(n, blocksize)
    code = torch.tensor(code_book(quant_type), dtype=torch.float32, device=w.device)
    blocks = w.detach().reshape(-1, blocksize).float()
    absmax = blocks.abs().amax(dim=1)
    xn = torch.where(absmax[:, None] > 0, blocks / absmax[:, None], torch.zeros_like(blocks))
    idx = torch.empty(blocks.shape, dtype=torch.uint8, device=w.device)
    step = max(1, (1 << 22) // blocksize)            # bound the [rows, blocksize, 16] temporary
    for s in range(0, blocks.shape[0], step):
        d = (xn[s:s + step, :, None] - code).abs()
        idx[s:s + step] = d.argmin(dim=-1).to(torch.uint8)
    idx = idx.reshape(-1, 2)
    packed = (idx[:, 0] << 4) | idx[:, 1]
    return packed.contiguous(), absmax.contiguous()


def dequantize_4bit_reference(packed: torch.Tensor, absmax: torch.Tensor, shape, dtype: torch.dtype,
                              blocksize: int = DEFAULT_BLOCKSIZE, quant_type: str = "nf4") -> torch.Tensor:
    code = torch.tensor(code_book(quant_type), dtype=torch.float32, device=packed.device)
    idx = torch.stack(((packed >> 4).long(), (packed & 15).long()), dim=1).reshape(-1, blocksize)
    vals = code[idx] * absmax.float()[:, None]
    return vals.to(dtype).reshape(shape)


# ---------------------------------------------------------------------- public ops
def quantize_4bit(w: torch.Tensor, blocksize: int = DEFAULT_BLOCKSIZE, quant_type: str = "nf4"):
    """``w`` (any shape, fp32/bf16/fp16) → ``(packed uint8 [n/2], absmax fp32 [n/blocksize])``."""
    n = w.numel()
    _check(n, blocksize)
    if not (w.is_cuda and require_library()):
        return quantize_4bit_reference(w, blocksize, quant_type)
    w = w.detach().contiguous()
    packed = torch.empty(n // 2, dtype=torch.uint8, device=w.device)
    absmax = torch.empty(n // blocksize, dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device):
        _cabi.check(_cabi.lib().dlion_q4_quantize(w.data_ptr(), packed.data_ptr(), absmax.data_ptr(), n, blocksize,
                                                  _code_ptr(quant_type), _cabi.dtype_code(w.dtype),
                                                  _cabi.stream_ptr(w.device)), "q4_quantize")
    return packed, absmax


def dequantize_4bit(packed: torch.Tensor, absmax: torch.Tensor, shape, dtype: torch.dtype,
                    blocksize: int = DEFAULT_BLOCKSIZE, quant_type: str = "nf4",
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Inverse of :func:`quantize_4bit`; ``out`` (contiguous, ``dtype``, ≥ n elements) is reused when given."""
    n = packed.numel() * 2
    _check(n, blocksize)
    if absmax.dtype != torch.float32 or absmax.numel() != n // blocksize:
        raise ValueError("absmax must be fp32 with one entry per block")
    if not (packed.is_cuda and require_library()):
        res = dequantize_4bit_reference(packed, absmax, shape, dtype, blocksize, quant_type)
        if out is not None:
            out.view(-1)[:n].copy_(res.view(-1))
            return out.view(-1)[:n].view(shape)
        return res
    if out is None:
        out = torch.empty(n, dtype=dtype, device=packed.device)
    elif out.dtype != dtype or not out.is_contiguous() or out.numel() < n:
        raise ValueError("out must be a contiguous tensor of the requested dtype with at least n elements")
    with torch.cuda.device(packed.device):
        _cabi.check(_cabi.lib().dlion_q4_dequantize(packed.data_ptr(), absmax.data_ptr(), out.data_ptr(), n, blocksize,
                                                    _code_ptr(quant_type), _cabi.dtype_code(dtype),
                                                    _cabi.stream_ptr(packed.device)), "q4_dequantize")
    return out.view(-1)[:n].view(shape)
