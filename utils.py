"""CLI surface and small helpers of the training driver -- the boundary the SLURM flow calls.

Every flag of the reference's `get_args` (utils.py:105-261) is kept with its name, type and
default, because submit-training-simple.sh forwards them verbatim (:63-113).  Flags below the
"additive" marker are new; their defaults reproduce the reference's behaviour.
"""
from __future__ import annotations

import argparse
import logging
from contextlib import contextmanager

import torch
from torch.optim.lr_scheduler import LambdaLR

logger = logging.getLogger()

PRECISION_STR_TO_DTYPE = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32,
                          "fp64": torch.float64}

# (flag, kwargs) -- reference utils.py:105-261, same order
_REFERENCE_FLAGS = [
    ("--dataset", dict(type=str, default="/capstor/store/cscs/ethz/large-sc/datasets/train_data.parquet",
                       help="Parquet file with a 'text' column")),
    ("--tokenizer-name-or-path", dict(type=str, default="unsloth/Mistral-Nemo-Base-2407-bnb-4bit",
                                      help="HF tokenizer id or local directory")),
    ("--sequence-length", dict(type=int, default=2048)),
    ("--batch-size", dict(type=int, default=1)),
    ("--fused-optimizer", dict(action="store_true", help="fp32 op-math AdamW (numerics of torch's fused kernel)")),
    ("--learning-rate", dict(type=float, default=1e-5)),
    ("--lr-warmup-steps", dict(type=int, default=10)),
    ("--training-steps", dict(type=int, default=1000, help="number of optimizer steps, not epochs")),
    ("--logging-frequency", dict(type=int, default=5, help="log every N steps")),
    ("--profile", dict(action="store_true", help="cudaProfilerStart/Stop + NVTX between the two steps below")),
    ("--profile-step-start", dict(type=int, default=10)),
    ("--profile-step-end", dict(type=int, default=12)),
    ("--grad-max-norm", dict(type=float, default=1)),
    ("--model-dtype", dict(type=str, default="bf16", help="dtype of parameters, gradients and optimizer state")),
    ("--compile", dict(action="store_true", help="torch.compile the model (the workload, not the hot path)")),
    ("--distributed", dict(action="store_true", help="data-parallel training over all ranks of the job")),
    ("--checkpoint-dir", dict(type=str, default="checkpoints/")),
    ("--checkpoint-frequency", dict(type=int, default=10, help="checkpoint every N steps; -1 disables")),
    ("--resume-from-checkpoint", dict(type=str, default=None, help="path, or 'latest'")),
    ("--experiment_name", dict(type=str, default="default-exp", help="sub-folder of the checkpoint dir")),
    ("--verify-checkpoints", dict(action="store_true", help="write / check checksums")),
    ("--max-kept-checkpoints", dict(type=int, default=3)),
    ("--use-torch-distributed-ckpt", dict(action="store_true", help="sharded checkpoint directories (every rank writes 1/W)")),
    ("--default-iter-time", dict(type=float, default=1.0, help="initial max_iter_time [s] for --timeaware-checkpointing")),
    ("--default-ckpt-time", dict(type=float, default=10.0, help="initial max_ckpt_time [s] for --timeaware-checkpointing")),
    ("--timeaware-checkpointing", dict(action="store_true", help="stop and checkpoint before SLURM_JOB_END_TIME")),
    ("--use_flash_attention", dict(action="store_true")),
    ("--log-loss-to-csv", dict(action="store_true")),
]

_ADDITIVE_FLAGS = [
    ("--synthetic-data", dict(action="store_true", help="tokens = f(seed, step, rank); no dataset/tokenizer needed")),
    ("--seed", dict(type=int, default=0)),
    ("--model-preset", dict(type=str, default=None, choices=[None, "tiny", "1b", "8b"],
                            help="BASELINE.json model sizes; default: the reference's hard-coded 8 B config")),
    ("--model-dim", dict(type=int, default=4096)),
    ("--model-layers", dict(type=int, default=32)),
    ("--model-heads", dict(type=int, default=32)),
    ("--model-kv-heads", dict(type=int, default=8)),
    ("--model-multiple-of", dict(type=int, default=1024)),
    ("--vocab-size", dict(type=int, default=None, help="default: tokenizer.vocab_size (131072 with --synthetic-data)")),
    ("--device", dict(type=str, default=None, choices=[None, "cuda", "cpu"], help="default: cuda when available")),
    ("--dist-backend", dict(type=str, default=None, choices=[None, "nccl", "gloo"])),
    ("--grad-sync", dict(type=str, default="replicated", choices=["replicated", "fused", "overlap"],
                         help="replicated: P2P allreduce + AdamW on every rank; fused: one reduce-scatter+AdamW+all-gather "
                              "kernel after backward; overlap: the fused kernel per bucket, under backward")),
    ("--snapshot-mode", dict(type=str, default=None, choices=[None, "staged", "direct"])),
    ("--clip-grad-norm", dict(action="store_true",
                              help="clip the global gradient norm to --grad-max-norm (the call the reference left "
                                   "commented out at train.py:271-272), folded into the optimizer kernel")),
    ("--sync-checkpoints", dict(action="store_true", help="wait for each checkpoint to be durable before continuing")),
    ("--deterministic", dict(action="store_true",
                             help="torch.use_deterministic_algorithms(True): run-to-run identical forward/backward kernels (the "
                                  "attention backward otherwise accumulates with atomics at larger sizes), so that a straight "
                                  "run and an interrupted + resumed one can be compared bit for bit at any model size")),
    ("--resubmit-command", dict(type=str, default=None,
                                help="arm pyrecover.setup_resubmission: on SIGUSR1/SIGTERM write the _final checkpoint and "
                                     "exit; a non-empty value (e.g. 'sbatch submit-training-simple.sh --continue') is run "
                                     "once by rank 0 to queue the continuation job, '' only stops")),
]


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    for flag, kw in _REFERENCE_FLAGS + _ADDITIVE_FLAGS:
        parser.add_argument(flag, **kw)
    return parser


def get_args(argv=None):
    return build_parser().parse_args(argv)


def init_logger():
    logger.setLevel(logging.INFO)
    handler = logging.StreamHandler()
    handler.setLevel(logging.INFO)
    handler.setFormatter(logging.Formatter("%(asctime)s - %(name)s - %(levelname)s - %(message)s"))
    logger.addHandler(handler)


def get_num_params(model: torch.nn.Module, exclude_embedding: bool = False) -> int:
    total = sum(p.numel() for p in model.parameters())
    if exclude_embedding:
        for child in model.children():
            if isinstance(child, torch.nn.Embedding):
                total -= sum(p.numel() for p in child.parameters())
    return total


def get_num_flop_per_token(num_params: int, model_config) -> int:
    """6 N for the dense layers + 12 l h q t for attention (reference utils.py:42-56)."""
    head_dim = model_config.dim // model_config.n_heads
    return 6 * num_params + 12 * model_config.n_layers * model_config.n_heads * head_dim * model_config.seq_len


class WarmupThenConstant:
    """lr factor min(1, (t + 1) / (warmup + 1)) -- reference utils.py:61-78."""

    def __init__(self, warmup_steps: int):
        self.warmup_steps = int(warmup_steps)

    def __call__(self, current_step: int) -> float:
        if current_step < self.warmup_steps:
            return float((current_step + 1) / (self.warmup_steps + 1))
        return 1


def build_lr_scheduler(optimizer, warmup_steps: int) -> LambdaLR:
    return LambdaLR(optimizer, WarmupThenConstant(warmup_steps))


@torch.no_grad()
def clip_grad_norm_(parameters, grad_max_norm):
    parameters = list(parameters)
    grads = [p.grad for p in parameters if p.grad is not None]
    total = torch.nn.utils.get_total_norm(grads, error_if_nonfinite=True)
    torch.nn.utils.clip_grads_with_norm_(parameters, grad_max_norm, total)
    return total


@contextmanager
def set_default_dtype(dtype: torch.dtype):
    previous = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        yield
    finally:
        torch.set_default_dtype(previous)
