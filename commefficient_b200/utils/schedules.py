"""Learning-rate schedules and round counting.

Reference: /root/reference/CommEfficient/utils.py:26-35 (PiecewiseLinear, Exp)
and :315-321 (steps_per_epoch).
"""
from __future__ import annotations

import bisect
import math
from typing import Sequence


class PiecewiseLinear:
    """Linear interpolation through ``(knots[i], vals[i])``; clamps outside."""

    def __init__(self, knots: Sequence[float], vals: Sequence[float]):
        if len(knots) != len(vals) or len(knots) == 0:
            raise ValueError("knots and vals must be non-empty and the same length")
        self.knots = [float(k) for k in knots]
        self.vals = [float(v) for v in vals]

    def __call__(self, t: float) -> float:
        k, v = self.knots, self.vals
        if t <= k[0]:
            return v[0]
        if t >= k[-1]:
            return v[-1]
        hi = bisect.bisect_right(k, t)
        lo = hi - 1
        span = k[hi] - k[lo]
        if span == 0:
            return v[hi]
        w = (t - k[lo]) / span
        return v[lo] + w * (v[hi] - v[lo])

    def __iter__(self):  # namedtuple-like unpacking: knots, vals
        yield self.knots
        yield self.vals


class Exp:
    """Linear warm-up to ``amplitude`` then decay by 10x every ``decay_len``."""

    def __init__(self, warmup_epochs: float, amplitude: float, decay_len: float):
        self.warmup_epochs = float(warmup_epochs)
        self.amplitude = float(amplitude)
        self.decay_len = float(decay_len)

    def __call__(self, t: float) -> float:
        if t < self.warmup_epochs:
            return self.amplitude * max(t, 0.0) / self.warmup_epochs
        return self.amplitude * 10.0 ** (-(t - self.warmup_epochs) / self.decay_len)


def steps_per_epoch(local_batch_size: int, dataset, num_workers: int):
    """Rounds per epoch: whole-client rounds when ``local_batch_size == -1``,
    otherwise ceil(#samples / (local_batch_size * clients per round))."""
    if local_batch_size == -1:
        return dataset.num_clients // num_workers
    return math.ceil(len(dataset) / float(local_batch_size * num_workers))


def triangular_lr(pivot_epoch: float, num_epochs: float, peak: float) -> PiecewiseLinear:
    """cv_train schedule (cv_train.py:394-395)."""
    return PiecewiseLinear([0, pivot_epoch, num_epochs], [0, peak, 0])


def linear_decay_lr(total_steps: float, start: float) -> PiecewiseLinear:
    """gpt2_train schedule (gpt2_train.py:302-304)."""
    return PiecewiseLinear([0, total_steps], [start, 0.0])
