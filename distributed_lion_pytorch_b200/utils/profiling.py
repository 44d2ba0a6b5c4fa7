"""Tracing helpers: NVTX ranges and device-side step timers.

reference: nothing upstream (only HF's train_samples_per_second surfaces).  Here the
trainer can wrap forward/backward/optimizer in NVTX ranges (``DLION_NVTX=1``) for Nsight, and time
phases with CUDA events without synchronising the step (``StepTimer`` resolves its events lazily).
"""
from __future__ import annotations

import statistics
import threading

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


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of this rank's GPU every 200 ms during the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80)}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, v in names.items():
                    if bits & v:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}
