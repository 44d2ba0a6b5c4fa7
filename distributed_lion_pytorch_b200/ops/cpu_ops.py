"""Host-side sign codec and ballot count (csrc_cpu/cpu_ops.cpp) for the CPU / gloo path.
Falls back to the pure-torch oracle (optim/semantics.py) when the library has not been built.
reference: distributed_lion.py:75-77 (encode), :84-88 (decode), :33-43 (torch.mode vote)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

import torch

from ..optim import semantics as sem

_LIB = None
_TRIED = False
_PATH = Path(__file__).resolve().parent.parent / "_lib" / "libdlion_cpu.so"


def lib():
    global _LIB, _TRIED
    if not _TRIED:
        _TRIED = True
        if _PATH.exists():
            try:
                L = C.CDLL(str(_PATH))
                L.dlion_cpu_pack_bits.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
                L.dlion_cpu_count_votes.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_int64, C.c_void_p]
                _LIB = L
            except OSError:
                _LIB = None
    return _LIB


def pack_bits(bits: torch.Tensor) -> torch.Tensor:
    L = lib()
    if L is None or bits.device.type != "cpu":
        return sem.pack_bits(bits)
    b = bits.reshape(-1).contiguous().view(torch.uint8) if bits.dtype == torch.bool else bits.reshape(-1).to(torch.uint8)
    n = b.numel()
    out = torch.empty((n + 7) // 8, dtype=torch.uint8)
    L.dlion_cpu_pack_bits(b.data_ptr(), n, out.data_ptr())
    return out


def count_votes(gathered: torch.Tensor, n: int, live_mask: Optional[int] = None) -> torch.Tensor:
    L = lib()
    if L is None or gathered.device.type != "cpu" or not gathered.is_contiguous():
        return sem.count_votes(gathered, n, live_mask)
    W, row = gathered.shape
    mask = (1 << W) - 1 if live_mask is None else int(live_mask) & ((1 << W) - 1)
    cnt = torch.empty(n, dtype=torch.int16)
    L.dlion_cpu_count_votes(gathered.data_ptr(), W, row, mask, n, cnt.data_ptr())
    return cnt
