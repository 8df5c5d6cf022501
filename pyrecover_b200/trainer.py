"""One data-parallel training step, assembled from the hot-path pieces.

Mirrors the body of the reference loop (train.py:245-275): fetch batch -> H2D -> zero_grad ->
forward -> sum-CE / num_items -> backward (gradient exchange fires inside) -> optimizer.step ->
lr_scheduler.step.  On CUDA the exchange is GradSync (K1) and the update is B200AdamW (K2, or the
fused K1+K2 kernel); on CPU (BASELINE config 1) it is the reference's own stack -- torch DDP on
gloo + torch.optim.AdamW -- because that IS the reference path there, not a fallback of ours.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F


def resolve_model_classes():
    """Prefer the user's `model.py` (the reference file, kept as-is next to train.py)."""
    try:
        import model as user_model  # type: ignore
        if hasattr(user_model, "Transformer") and hasattr(user_model, "TransformerModelArgs"):
            return user_model.Transformer, user_model.TransformerModelArgs
    except ImportError:
        pass
    from .workload import Transformer, TransformerModelArgs
    return Transformer, TransformerModelArgs


def model_args_from_cli(args, vocab_size: int):
    Transformer, Args = resolve_model_classes()
    if getattr(args, "model_preset", None):
        from .workload import MODEL_PRESETS
        dim, layers, heads, kv, mult, vocab = MODEL_PRESETS[args.model_preset]
        vocab_size = args.vocab_size or vocab
    else:
        dim, layers, heads, kv, mult = (args.model_dim, args.model_layers, args.model_heads,
                                        args.model_kv_heads, args.model_multiple_of)
    cfg = Args(dim=dim, n_layers=layers, n_heads=heads, n_kv_heads=kv, ffn_dim_multiplier=1.3,
               multiple_of=mult, rope_theta=500000, vocab_size=vocab_size, seq_len=args.sequence_length,
               use_flash_attention=args.use_flash_attention)      # reference train.py:88-99
    return Transformer, cfg


class Trainer:
    def __init__(self, model: torch.nn.Module, device: torch.device, *, lr: float, fused: bool,
                 warmup_steps: int, world_size: int = 1, local_rank: int = 0,
                 grad_sync_mode: str = "replicated", process_group=None, max_grad_norm: Optional[float] = None):
        from utils import build_lr_scheduler
        self.device, self.world_size = device, world_size
        self.grad_sync = None
        if device.type == "cuda":
            from .optim import B200AdamW
            if world_size > 1:
                from .gradsync import GradSync
                self.grad_sync = GradSync(process_group, mode=grad_sync_mode)
            self.model = model
            self.optimizer = B200AdamW(model.parameters(), lr=lr, fused=fused, grad_sync=self.grad_sync,
                                       max_grad_norm=max_grad_norm)
        else:
            if world_size > 1:
                from torch.nn.parallel import DistributedDataParallel as DDP
                model = DDP(model, find_unused_parameters=False)      # reference train.py:107-115
            self.model = model
            self.optimizer = torch.optim.AdamW(model.parameters(), lr=lr, fused=fused)   # train.py:120-122
        self.max_grad_norm = max_grad_norm
        self.model.train()
        self.lr_scheduler = build_lr_scheduler(self.optimizer, warmup_steps)

    def step(self, input_ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """input_ids / labels already on the device.  Returns the (device) loss tensor."""
        num_items = labels.ne(-100).sum()
        self.optimizer.zero_grad()
        logits = self.model(input_ids)
        loss = F.cross_entropy(logits.flatten(0, 1).float(), labels.flatten(0, 1), reduction="sum")
        loss = loss / num_items
        del logits
        loss.backward()
        if self.max_grad_norm and self.device.type != "cuda":
            from utils import clip_grad_norm_             # the reference's own call, train.py:271-272
            clip_grad_norm_(self.model.parameters(), self.max_grad_norm)
        self.optimizer.step()
        self.lr_scheduler.step()
        return loss

    def step_from_host(self, input_ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """The call a user makes: host batch in, device loss out (H2D inside)."""
        return self.step(input_ids.to(self.device, non_blocking=True), labels.to(self.device, non_blocking=True))

    def close(self) -> None:
        if self.grad_sync is not None:
            self.grad_sync.close()
