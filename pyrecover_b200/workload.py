"""Workload generator: a Llama-3-style decoder and a synthetic token stream.

The dense model is OUT OF SCOPE of this build ("used verbatim as the workload
generator; only its parameter list matters to the path") -- its GEMMs/attention run on library
kernels (cuBLAS, SDPA).  The reference's model.py is not redistributed here; this module builds a
decoder with the SAME parameter names and shapes (so state dicts interchange with
reference model.py:330-395: tok_embeddings, layers.<i>.attention.{wq,wk,wv,wo},
layers.<i>.feed_forward.{w1,w2,w3}, layers.<i>.{attention_norm,ffn_norm}, norm, output) and the
same SwiGLU width rule (model.py:258-262), written independently: real-valued rotary tables
instead of complex multiplication, fused rms_norm, grouped-query SDPA.

If a module named `model` exposing `Transformer` / `TransformerModelArgs` is importable (the
user's copy of the reference file next to train.py) train.py prefers it.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class TransformerModelArgs:
    dim: int = 4096
    n_layers: int = 32
    n_heads: int = 32
    n_kv_heads: Optional[int] = None
    multiple_of: int = 256
    ffn_dim_multiplier: Optional[float] = None
    norm_eps: float = 1e-5
    rope_theta: float = 10000
    norm_type: str = "rmsnorm"
    seq_len: int = 2048
    vocab_size: int = -1
    use_flash_attention: bool = False


def swiglu_width(dim: int, multiple_of: int, ffn_dim_multiplier: Optional[float]) -> int:
    h = int(2 * (4 * dim) / 3)
    if ffn_dim_multiplier is not None:
        h = int(ffn_dim_multiplier * h)
    return multiple_of * ((h + multiple_of - 1) // multiple_of)


class RMSNorm(nn.Module):
    def __init__(self, dim: int, eps: float):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))

    def forward(self, x):
        return F.rms_norm(x.float(), (x.shape[-1],), None, self.eps).type_as(x) * self.weight


def rotary_tables(head_dim: int, seq_len: int, theta: float) -> torch.Tensor:
    """(seq_len, head_dim/2, 2) float32: cos and sin of position * theta^(-2i/head_dim)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2)[: head_dim // 2].float() / head_dim))
    ang = torch.outer(torch.arange(seq_len).float(), inv)
    return torch.stack([ang.cos(), ang.sin()], dim=-1)


def rotate(x: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    """x: (B, S, H, D) with interleaved (even, odd) pairs, the layout the reference's complex
    view uses (model.py:101-127)."""
    xf = x.float().unflatten(-1, (-1, 2))
    cos, sin = table[: x.shape[1], None, :, 0], table[: x.shape[1], None, :, 1]
    re = xf[..., 0] * cos - xf[..., 1] * sin
    im = xf[..., 0] * sin + xf[..., 1] * cos
    return torch.stack([re, im], dim=-1).flatten(-2).type_as(x)


class Attention(nn.Module):
    def __init__(self, a: TransformerModelArgs):
        super().__init__()
        self.n_heads = a.n_heads
        self.n_kv_heads = a.n_heads if a.n_kv_heads is None else a.n_kv_heads
        self.head_dim = a.dim // a.n_heads
        self.wq = nn.Linear(a.dim, self.n_heads * self.head_dim, bias=False)
        self.wk = nn.Linear(a.dim, self.n_kv_heads * self.head_dim, bias=False)
        self.wv = nn.Linear(a.dim, self.n_kv_heads * self.head_dim, bias=False)
        self.wo = nn.Linear(self.n_heads * self.head_dim, a.dim, bias=False)

    def forward(self, x, table):
        b, s, _ = x.shape
        q = rotate(self.wq(x).view(b, s, self.n_heads, self.head_dim), table)
        k = rotate(self.wk(x).view(b, s, self.n_kv_heads, self.head_dim), table)
        v = self.wv(x).view(b, s, self.n_kv_heads, self.head_dim)
        o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2),
                                           is_causal=True, enable_gqa=self.n_kv_heads != self.n_heads)
        return self.wo(o.transpose(1, 2).reshape(b, s, -1))


class FeedForward(nn.Module):
    def __init__(self, dim: int, hidden: int):
        super().__init__()
        self.w1 = nn.Linear(dim, hidden, bias=False)
        self.w2 = nn.Linear(hidden, dim, bias=False)
        self.w3 = nn.Linear(dim, hidden, bias=False)

    def forward(self, x):
        return self.w2(F.silu(self.w1(x)) * self.w3(x))


class TransformerBlock(nn.Module):
    def __init__(self, a: TransformerModelArgs):
        super().__init__()
        self.attention = Attention(a)
        self.feed_forward = FeedForward(a.dim, swiglu_width(a.dim, a.multiple_of, a.ffn_dim_multiplier))
        self.attention_norm = RMSNorm(a.dim, a.norm_eps)
        self.ffn_norm = RMSNorm(a.dim, a.norm_eps)

    def forward(self, x, table):
        h = x + self.attention(self.attention_norm(x), table)
        return h + self.feed_forward(self.ffn_norm(h))


class Transformer(nn.Module):
    def __init__(self, model_args: TransformerModelArgs):
        super().__init__()
        self.model_args = model_args
        self.vocab_size = model_args.vocab_size
        self.n_layers = model_args.n_layers
        self.tok_embeddings = nn.Embedding(model_args.vocab_size, model_args.dim)
        self.register_buffer("rope", rotary_tables(model_args.dim // model_args.n_heads, model_args.seq_len,
                                                   model_args.rope_theta), persistent=False)
        self.layers = nn.ModuleDict({str(i): TransformerBlock(model_args) for i in range(model_args.n_layers)})
        self.norm = RMSNorm(model_args.dim, model_args.norm_eps)
        self.output = nn.Linear(model_args.dim, model_args.vocab_size, bias=False)

    def forward(self, tokens: torch.Tensor):
        h = self.tok_embeddings(tokens)
        table = self.rope.float()
        for layer in self.layers.values():
            h = layer(h, table)
        return self.output(self.norm(h))


# ---------------------------------------------------------------------- synthetic tokens ----
class SyntheticTokens:
    """Batches that are a pure function of (seed, step, rank) -- SURVEY.md section 8d -- so an
    interrupted + resumed run sees exactly the batches of the straight run (the reference saves no
    sampler state, pyrecover/checkpoint.py:72, and restarts its DistributedSampler at index 0)."""

    def __init__(self, vocab_size: int, seq_len: int, batch_size: int, seed: int = 0, rank: int = 0,
                 pin: bool = False):
        self.vocab_size, self.seq_len, self.batch_size = vocab_size, seq_len, batch_size
        self.seed, self.rank, self.pin = seed, rank, pin

    def batch(self, step: int) -> Tuple[torch.Tensor, torch.Tensor]:
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + step * 1_009 + self.rank)
        toks = torch.randint(0, self.vocab_size, (self.batch_size, self.seq_len + 1), generator=g)
        if self.pin:
            toks = toks.pin_memory()
        return toks[:, :-1], toks[:, 1:]


MODEL_PRESETS = {
    # name: (dim, layers, heads, kv_heads, multiple_of, vocab)  -- BASELINE.json configs
    "tiny": (256, 4, 4, 2, 64, 1024),
    "1b": (2048, 16, 16, 4, 1024, 32768),
    "8b": (4096, 32, 32, 8, 1024, 131072),
}


def preset_args(name: str, seq_len: int, use_flash_attention: bool = False) -> TransformerModelArgs:
    dim, layers, heads, kv, mult, vocab = MODEL_PRESETS[name]
    return TransformerModelArgs(dim=dim, n_layers=layers, n_heads=heads, n_kv_heads=kv, ffn_dim_multiplier=1.3,
                                multiple_of=mult, rope_theta=500000, vocab_size=vocab, seq_len=seq_len,
                                use_flash_attention=use_flash_attention)
