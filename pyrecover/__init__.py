"""pyrecover: checkpoint / resume manager for SLURM jobs -- B200-native build.

Mirrors the reference package's public names (pyrecover/__init__.py:5-15).  The reference
imports `.resubmit` and `.timelimit`, which it does not ship (so `import pyrecover` fails
there); here both exist and the package imports cleanly.
"""
__version__ = "0.1.0"

from .checkpoint import (get_latest_checkpoint, load_ckpt_distributed, load_ckpt_vanilla,
                         save_ckpt_distributed, save_ckpt_vanilla, wait_for_checkpoints)
from .resubmit import setup_resubmission
from .timelimit import TimeAwarePolicy, get_remaining_time, monitor_timelimit

__all__ = [
    "save_ckpt_vanilla",
    "load_ckpt_vanilla",
    "save_ckpt_distributed",
    "load_ckpt_distributed",
    "get_latest_checkpoint",
    "wait_for_checkpoints",
    "monitor_timelimit",
    "get_remaining_time",
    "setup_resubmission",
    "TimeAwarePolicy",
]
