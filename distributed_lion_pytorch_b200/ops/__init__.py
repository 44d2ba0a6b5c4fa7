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
