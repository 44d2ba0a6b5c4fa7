"""Hand-written sm_100a kernels, reached through the C ABI in ``include/dlion.h``.

``require_library()`` is the single gate the rest of the package uses: it returns True when
CUDA tensors must go through ``libdlion.so`` (and raises if the library cannot be loaded on a
machine that has a GPU — no silent eager fallback), False on CPU-only machines.
"""
from __future__ import annotations

import os

import torch

from . import _cabi

_checked = None


def require_library() -> bool:
    global _checked
    if _checked is None:
        if os.environ.get("DLION_DISABLE_KERNELS") == "1":
            _checked = False
        elif torch.cuda.is_available():
            _cabi.lib()          # raises with a build hint when missing
            _checked = True
        else:
            _checked = False
    return _checked


def library_loaded() -> bool:
    return _cabi._LIB is not None


def mark_grad_written(param) -> None:
    """Record that a backward pass produced a gradient for ``param`` in this step.  Autograd calls it through a
    post-accumulate hook; the custom backward functions that add into ``param.grad`` themselves (GEMM-epilogue
    weight gradients, the one-launch bias / LayerNorm reductions, the chunked cross-entropy) call it directly,
    because they hand autograd ``None`` and no hook fires.  ``Lion.step`` consumes the marks
    (reference: ``if p.grad is not None`` at distributed_lion.py:180)."""
    param._dlion_touched = True
