"""Walltime awareness: the stop-and-checkpoint policy the reference keeps inline in train.py.

Reference: train.py:163-190 (initial buffer = 10 x iter + 2 x ckpt), :223-232 (rank-0 check
`time_left < max_iter + max_ckpt + buffer`), :298-307 (running max of the iteration time, buffer
= 5 x max_iter + 1 x max_ckpt), :334-337 (running max of the checkpoint time); the job end time
comes from SLURM_JOB_END_TIME (dist_utils.py:93-101, exported by submit-training-simple.sh:46).
`get_remaining_time` / `monitor_timelimit` are the names pyrecover/__init__.py:7 advertises.
"""
from __future__ import annotations

import os
import time
from typing import Callable, Optional

INITIAL_ITER_MULT, INITIAL_CKPT_MULT = 10, 2
RUNNING_ITER_MULT, RUNNING_CKPT_MULT = 5, 1


def job_end_time() -> Optional[float]:
    val = os.environ.get("SLURM_JOB_END_TIME")
    if val is None:
        return None
    try:
        return float(val)
    except ValueError:
        return None


def get_remaining_time(now: Optional[float] = None) -> Optional[float]:
    """Seconds until the SLURM walltime limit, or None when SLURM_JOB_END_TIME is not set."""
    end = job_end_time()
    if end is None:
        return None
    return end - (time.time() if now is None else now)


class TimeAwarePolicy:
    """Decides when to stop training and write the `_final` checkpoint."""

    def __init__(self, default_iter_time: float = 1.0, default_ckpt_time: float = 10.0,
                 end_time: Optional[float] = None, clock: Callable[[], float] = time.time):
        self.max_iter_time = float(default_iter_time)
        self.max_ckpt_time = float(default_ckpt_time)
        self.buffer_time = INITIAL_ITER_MULT * self.max_iter_time + INITIAL_CKPT_MULT * self.max_ckpt_time
        self.end_time = job_end_time() if end_time is None else end_time
        self.clock = clock
        # iterations that still run after the decision (the B200 path delivers rank 0's decision inside the
        # gradient exchange, a fixed number of steps later, instead of a per-step broadcast + host sync)
        self.extra_iterations = 0

    @property
    def threshold(self) -> float:
        return (1 + self.extra_iterations) * self.max_iter_time + self.max_ckpt_time + self.buffer_time

    def time_left(self) -> Optional[float]:
        return None if self.end_time is None else self.end_time - self.clock()

    def should_stop(self) -> bool:
        left = self.time_left()
        return left is not None and left < self.threshold

    def observe_iteration(self, seconds: float) -> bool:
        """Feed one iteration time; returns True when it raised the running maximum."""
        raised = seconds > self.max_iter_time
        if raised:
            self.max_iter_time = seconds
        self.buffer_time = RUNNING_ITER_MULT * self.max_iter_time + RUNNING_CKPT_MULT * self.max_ckpt_time
        return raised

    def observe_checkpoint(self, seconds: float) -> bool:
        raised = seconds > self.max_ckpt_time
        if raised:
            self.max_ckpt_time = seconds
        return raised


def monitor_timelimit(default_iter_time: float = 1.0, default_ckpt_time: float = 10.0) -> TimeAwarePolicy:
    """Build the policy from the environment (name advertised by the reference package)."""
    return TimeAwarePolicy(default_iter_time, default_ckpt_time)
