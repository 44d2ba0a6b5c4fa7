"""Tracing helpers: NVTX ranges and device-side step timers.

reference: nothing upstream (only HF's train_samples_per_second surfaces).  Here the
trainer can wrap forward/backward/optimizer in NVTX ranges (``DLION_NVTX=1``) for Nsight, and time
phases with CUDA events without synchronising the step (``StepTimer`` resolves its events lazily).
"""
from __future__ import annotations

import contextlib
import os
from collections import defaultdict
from typing import Dict, List, Tuple

import torch

_NVTX = os.environ.get("DLION_NVTX") == "1"


@contextlib.contextmanager
def nvtx_range(name: str):
    if _NVTX and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class StepTimer:
    """Collects (start, end) CUDA-event pairs per phase; ``summary()`` synchronises once and
    returns mean milliseconds per phase.  On CPU it is a no-op that reports nothing."""

    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._pairs: Dict[str, List[Tuple[torch.cuda.Event, torch.cuda.Event]]] = defaultdict(list)

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            with nvtx_range(name):
                yield
            return
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        with nvtx_range(name):
            yield
        b.record()
        self._pairs[name].append((a, b))

    def summary(self, reset: bool = True) -> Dict[str, float]:
        if not self.enabled or not self._pairs:
            return {}
        torch.cuda.synchronize()
        out = {f"{k}_ms": sum(a.elapsed_time(b) for a, b in v) / len(v) for k, v in self._pairs.items()}
        if reset:
            self._pairs.clear()
        return out
