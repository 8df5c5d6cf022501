"""Job resubmission on walltime signals (name advertised by reference pyrecover/__init__.py:6;
the reference ships no implementation -- recovery there is a manual
`sbatch submit-training-simple.sh --continue`, submit-training-simple.sh:73-76)."""
from __future__ import annotations

import os
import signal
import subprocess
from typing import Callable, Optional, Sequence


class ResubmitState:
    def __init__(self):
        self.stop_requested = False
        self.resubmitted = False


def setup_resubmission(command: Optional[Sequence[str]] = None,
                       signals: Sequence[int] = (signal.SIGUSR1, signal.SIGTERM),
                       on_signal: Optional[Callable[[int], None]] = None) -> ResubmitState:
    """Install handlers for SLURM's pre-termination signals (`#SBATCH --signal=USR1@120`).

    On the first signal `state.stop_requested` becomes True (the training loop treats it like the
    time-aware should_stop and writes the `_final` checkpoint) and, when `command` is given (e.g.
    ["sbatch", "submit-training-simple.sh", "--continue"]), the job is re-queued once."""
    state = ResubmitState()

    def handler(signum, _frame):
        state.stop_requested = True
        if on_signal is not None:
            on_signal(signum)
        if command and not state.resubmitted and int(os.environ.get("SLURM_PROCID", "0")) == 0:
            state.resubmitted = True
            try:
                subprocess.Popen(list(command))
            except OSError:
                state.resubmitted = False

    for sig in signals:
        signal.signal(sig, handler)
    return state
