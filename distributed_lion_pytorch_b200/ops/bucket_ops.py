"""N2 — flat-bucket kernels: multi-tensor gather/cast, sum of squares, clip scale.
reference: the per-tensor loop in distributed_lion.py:179-198 / flatten_and_pad :14-24."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from . import _cabi

_scratch = {}


def multi_tensor_gather(srcs: List[torch.Tensor], dsts: List[torch.Tensor]) -> bool:
    """Copy (and cast) ``srcs[i]`` into ``dsts[i]`` with one launch per 96 tensors.  Returns False
    when a tensor is not contiguous / dtypes are mixed so the caller can fall back."""
    if not srcs:
        return True
    sdt, ddt = srcs[0].dtype, dsts[0].dtype
    for s, d in zip(srcs, dsts):
        if s.dtype != sdt or d.dtype != ddt or not s.is_contiguous() or not d.is_contiguous() \
                or s.numel() != d.numel() or s.device != d.device:
            return False
    n = len(srcs)
    L = _cabi.lib()
    src_arr = (C.c_void_p * n)(*[s.data_ptr() for s in srcs])
    dst_arr = (C.c_void_p * n)(*[d.data_ptr() for d in dsts])
    num_arr = (C.c_int64 * n)(*[s.numel() for s in srcs])
    with torch.cuda.device(srcs[0].device):
        _cabi.check(L.dlion_multi_gather(src_arr, dst_arr, num_arr, n, _cabi.dtype_code(sdt), _cabi.dtype_code(ddt),
                                         _cabi.stream_ptr(srcs[0].device)), "multi_gather")
    return True


def _sumsq_scratch(device: torch.device) -> torch.Tensor:
    key = (device.type, device.index)
    t = _scratch.get(key)
    if t is None:
        nbytes = int(_cabi.lib().dlion_sumsq_scratch_bytes())
        t = torch.zeros((nbytes + 3) // 4, dtype=torch.int32, device=device)
        _scratch[key] = t
    return t


def sumsq(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 sum of squares of a contiguous tensor, deterministic, no host sync."""
    assert x.is_cuda and x.is_contiguous()
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _cabi.check(_cabi.lib().dlion_sumsq(x.data_ptr(), x.numel(), _cabi.dtype_code(x.dtype),
                                            _sumsq_scratch(x.device).data_ptr(), out.data_ptr(),
                                            _cabi.stream_ptr(x.device)), "sumsq")
    return out


def clip_scale_(x: torch.Tensor, sumsq_dev: torch.Tensor, max_norm: float) -> None:
    """``x *= min(1, max_norm / (sqrt(sumsq) + 1e-6))`` with the norm read on the device."""
    with torch.cuda.device(x.device):
        _cabi.check(_cabi.lib().dlion_clip_scale(x.data_ptr(), x.numel(), _cabi.dtype_code(x.dtype),
                                                 sumsq_dev.data_ptr(), float(max_norm),
                                                 _cabi.stream_ptr(x.device)), "clip_scale")


def clip_grad_norm_flat_(flat_grads: List[torch.Tensor], max_norm: float) -> torch.Tensor:
    """Global-norm clipping over flat gradient buckets; returns the (device) total norm.
    Matches ``torch.nn.utils.clip_grad_norm_`` up to fp32 summation order."""
    total = None
    parts = [sumsq(g) for g in flat_grads]
    total = parts[0] if len(parts) == 1 else torch.stack([p.view(()) for p in parts]).sum().view(1)
    for g in flat_grads:
        clip_scale_(g, total, max_norm)
    return total.sqrt()
