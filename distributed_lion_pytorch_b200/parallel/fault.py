"""Simulated worker drop-out (fault injection) for Distributed Lion.

reference: README.md:2 advertises "robust against worker drop-out" but the reference has no
code for it — ``dist.all_gather`` (distributed_lion.py:81) simply hangs on a missing rank.
BASELINE config 4 asks for a *simulated drop-out mask*: every step a W-bit live mask is
agreed by all ranks (it changes the majority threshold / averaging denominator), dropped
ranks' ballots are ignored, and everybody still applies the voted update so replicas stay
identical.  The schedule is a pure function of (seed, step), so no communication is needed
to agree on it.
"""
from __future__ import annotations

import random
from typing import Optional


class DropoutSchedule:
    """Bernoulli(drop_prob) per rank per step from a shared seed; never drops everybody."""

    def __init__(self, world_size: int, drop_prob: float = 0.0, seed: int = 0, fixed_mask: Optional[int] = None):
        if not 0.0 <= drop_prob < 1.0:
            raise ValueError("drop_prob must be in [0, 1)")
        self.world_size, self.drop_prob, self.seed, self.fixed_mask = world_size, drop_prob, seed, fixed_mask

    def mask(self, step: int) -> Optional[int]:
        W = self.world_size
        full = (1 << W) - 1
        if self.fixed_mask is not None:
            m = self.fixed_mask & full
            if m == 0:
                raise ValueError("fixed live mask excludes every worker")
            return m
        if self.drop_prob <= 0.0 or W <= 1:
            return None
        rng = random.Random((self.seed << 32) ^ (step * 0x9E3779B1 & 0xFFFFFFFF))
        m = 0
        for r in range(W):
            if rng.random() >= self.drop_prob:
                m |= 1 << r
        if m == 0:                      # keep at least one voter: the lowest rank survives
            m = 1 << rng.randrange(W)
        return m

    def live_count(self, step: int) -> int:
        m = self.mask(step)
        return self.world_size if m is None else bin(m).count("1")
