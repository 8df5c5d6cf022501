"""CPU port of the reference's training-step hot path -- TEST / BASELINE INFRASTRUCTURE ONLY.

The reference (Shaswat-G/PyRecover) is Python on top of torch; /root/reference does not exist on
the GPU box, so its CPU path is restated here call for call, using the same library entry points
the reference uses.  Only bench.py's `cpu_baseline` leg and `--impl reference` arm run this file;
the product (pyrecover_b200/, pyrecover/, train.py) never imports it.

Restated lines:
  model in bf16, default dtype ctx              train.py:100-101, utils.py:92-102
  DistributedDataParallel(find_unused=False)    train.py:107-115     (gloo on CPU, world 1 or 2)
  torch.optim.AdamW(lr, fused=flag)             train.py:120-122
  LambdaLR linear warm-up then constant         utils.py:59-81, train.py:123
  zero_grad / forward / sum-CE / num_items /
  backward / step / sched.step                  train.py:260-275
  torch.save({epoch, step, model, optimizer,
              lr_scheduler}) [+ md5]            pyrecover/checkpoint.py:60-84
The transformer itself is the workload generator (pyrecover_b200/workload.py, same parameter
list as reference model.py:330-395).
"""
from __future__ import annotations

import functools
import hashlib
import os
import time
from typing import Dict, Optional

import torch
import torch.nn.functional as F
from torch.optim.lr_scheduler import LambdaLR


def _warmup_constant(warmup_steps: int, current_step: int) -> float:
    if current_step < warmup_steps:
        return float((current_step + 1) / (warmup_steps + 1))
    return 1


def build(preset: str, seq_len: int, fused: bool, lr: float = 1e-5, warmup: int = 10, seed: int = 0,
          ddp: bool = False):
    from pyrecover_b200.workload import Transformer, preset_args
    torch.manual_seed(seed)
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        model = Transformer(preset_args(preset, seq_len))
    finally:
        torch.set_default_dtype(old)
    if ddp:
        model = torch.nn.parallel.DistributedDataParallel(model, find_unused_parameters=False)
    model.train()
    opt = torch.optim.AdamW(model.parameters(), lr=lr, fused=fused)
    sched = LambdaLR(opt, functools.partial(_warmup_constant, warmup))
    return model, opt, sched


def train_step(model, opt, sched, input_ids, labels):
    num_items = labels.ne(-100).sum()
    opt.zero_grad()
    logits = model(input_ids)
    loss = F.cross_entropy(logits.flatten(0, 1).float(), labels.flatten(0, 1), reduction="sum")
    loss = loss / num_items
    del logits
    loss.backward()
    opt.step()
    sched.step()
    return loss


def save_vanilla(model, opt, sched, step: int, path: str, verify: bool) -> float:
    t0 = time.perf_counter()
    inner = model.module if hasattr(model, "module") else model
    state = {"epoch": 1, "step": step, "model": inner.state_dict(), "optimizer": opt.state_dict(),
             "lr_scheduler": sched.state_dict()}
    torch.save(state, path)
    if verify:
        with open(path, "rb") as f:
            digest = hashlib.md5(f.read()).hexdigest()
        with open(path + ".md5", "w") as f:
            f.write(digest)
    return time.perf_counter() - t0


def run(preset: str, full_seq: int, steps: int, warmup: int, fused: bool = True, budget_s_per_step: float = 10.0,
        sample_seq: Optional[int] = None, ckpt_path: Optional[str] = None, threads: Optional[int] = None,
        max_seconds: float = 90.0) -> Dict:
    """Time `steps` CPU training steps on a bounded token sample of the named workload.

    The per-step optimizer / checkpoint cost does not depend on the token count, so tokens/s on a
    short sample UNDER-states the CPU's full-sequence throughput only through that fixed cost --
    which is exactly the hot path being compared.  The sample length is chosen so one step takes
    about `budget_s_per_step` seconds (calibrated on a 16-token probe)."""
    from pyrecover_b200.workload import MODEL_PRESETS, SyntheticTokens
    if threads:
        torch.set_num_threads(threads)
    cores = torch.get_num_threads()
    vocab = MODEL_PRESETS[preset][5]
    model, opt, sched = build(preset, full_seq, fused)
    probe = SyntheticTokens(vocab, 16, 1)
    x, y = probe.batch(0)
    t0 = time.perf_counter()
    train_step(model, opt, sched, x, y)
    t_probe = time.perf_counter() - t0          # includes lazy optimizer-state allocation
    t0 = time.perf_counter()
    train_step(model, opt, sched, x, y)
    t16 = time.perf_counter() - t0
    if sample_seq is None:
        x64, y64 = SyntheticTokens(vocab, 144, 1).batch(0)
        train_step(model, opt, sched, x64, y64)          # first touch of the new shape
        t0 = time.perf_counter()
        train_step(model, opt, sched, x64, y64)
        t64 = time.perf_counter() - t0
        per_tok = max((t64 - t16) / 128.0, 1e-6)
        fixed = max(t16 - 16 * per_tok, 0.0)
        sample_seq = int(max(16, min(full_seq, (budget_s_per_step - fixed) / per_tok)))
        sample_seq = max(16, sample_seq // 16 * 16)
    data = SyntheticTokens(vocab, sample_seq, 1)
    for s in range(warmup):
        train_step(model, opt, sched, *data.batch(s))
    t0 = time.perf_counter()
    done = 0
    for s in range(steps):
        loss = train_step(model, opt, sched, *data.batch(warmup + s))
        float(loss.detach())
        done += 1
        if time.perf_counter() - t0 > max_seconds:      # keep the whole arm within minutes
            break
    steps = done
    dt = time.perf_counter() - t0
    out = {"tokens_per_s": sample_seq * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": cores,
           "sample_seq": sample_seq, "steps": steps, "probe_first_step_s": t_probe,
           "sample": f"{steps} steps of {sample_seq} tokens (of {full_seq}) on the {preset} model, "
                     f"torch CPU fwd/bwd + torch.optim.AdamW(fused={fused}), {cores} threads"}
    if ckpt_path:
        out["ckpt_store_s"] = save_vanilla(model, opt, sched, steps, ckpt_path, verify=False)
        out["ckpt_bytes"] = os.path.getsize(ckpt_path)
        os.remove(ckpt_path)
    return out
