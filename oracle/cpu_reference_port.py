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

When oracle/_ref/ holds the reference's own files (staged by oracle/stage_ref.py from /root/reference
at build time; git-ignored, shipped to the GPU box), `build()` and `save_vanilla()` use THOSE instead:
model.Transformer / TransformerModelArgs (model.py:9-395), utils.build_lr_scheduler (utils.py:59-81) and
pyrecover/checkpoint.py:save_ckpt_vanilla (:25-105) -- kind "reference" in bench.py's lines; this file
is then only the loop around them.  `run_ddp()` is the W > 1 arm: W gloo ranks on this host's cores,
DistributedDataParallel as at train.py:107-115, the cores split evenly between the ranks.
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


def have_ref() -> bool:
    from oracle import stage_ref
    return stage_ref.available()


def build(preset: str, seq_len: int, fused: bool, lr: float = 1e-5, warmup: int = 10, seed: int = 0,
          ddp: bool = False, use_ref: Optional[bool] = None):
    """(model, optimizer, lr_scheduler, kind) -- kind "reference" when the staged reference files were used."""
    from pyrecover_b200.workload import MODEL_PRESETS, Transformer, preset_args
    use_ref = have_ref() if use_ref is None else use_ref
    torch.manual_seed(seed)
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        if use_ref:
            from oracle import stage_ref
            ref_model = stage_ref.load("model")
            dim, layers, heads, kv, mult, vocab = MODEL_PRESETS[preset]
            cfg = ref_model.TransformerModelArgs(dim=dim, n_layers=layers, n_heads=heads, n_kv_heads=kv,
                                                 ffn_dim_multiplier=1.3, multiple_of=mult, rope_theta=500000,
                                                 vocab_size=vocab, seq_len=seq_len)     # train.py:88-99
            model = ref_model.Transformer(cfg)
        else:
            model = Transformer(preset_args(preset, seq_len))
    finally:
        torch.set_default_dtype(old)
    if ddp:
        model = torch.nn.parallel.DistributedDataParallel(model, find_unused_parameters=False)
    model.train()
    opt = torch.optim.AdamW(model.parameters(), lr=lr, fused=fused)
    if use_ref:
        from oracle import stage_ref
        sched = stage_ref.load("utils").build_lr_scheduler(opt, warmup)
    else:
        sched = LambdaLR(opt, functools.partial(_warmup_constant, warmup))
    return model, opt, sched, ("reference" if use_ref else "port")


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


def save_vanilla(model, opt, sched, step: int, path: str, verify: bool, kind: str = "port") -> float:
    t0 = time.perf_counter()
    if kind == "reference":
        from oracle import stage_ref
        stage_ref.load("checkpoint").save_ckpt_vanilla(model, opt, sched, None, step, 1, path, max_keep=0, verify=verify)
        return time.perf_counter() - t0
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


def _calibrate(model, opt, sched, vocab: int, full_seq: int, budget_s_per_step: float):
    """Sample length (a multiple of 16 tokens, <= full_seq) at which one step takes about the budget."""
    from pyrecover_b200.workload import SyntheticTokens
    x, y = SyntheticTokens(vocab, 16, 1).batch(0)
    t0 = time.perf_counter()
    train_step(model, opt, sched, x, y)
    t_probe = time.perf_counter() - t0          # includes lazy optimizer-state allocation
    t0 = time.perf_counter()
    train_step(model, opt, sched, x, y)
    t16 = time.perf_counter() - t0
    x64, y64 = SyntheticTokens(vocab, 144, 1).batch(0)
    train_step(model, opt, sched, x64, y64)          # first touch of the new shape
    t0 = time.perf_counter()
    train_step(model, opt, sched, x64, y64)
    t64 = time.perf_counter() - t0
    per_tok = max((t64 - t16) / 128.0, 1e-6)
    fixed = max(t16 - 16 * per_tok, 0.0)
    sample_seq = int(max(16, min(full_seq, (budget_s_per_step - fixed) / per_tok)))
    return max(16, sample_seq // 16 * 16), t_probe


def run(preset: str, full_seq: int, steps: int, warmup: int, fused: bool = True, budget_s_per_step: float = 10.0,
        sample_seq: Optional[int] = None, ckpt_path: Optional[str] = None, threads: Optional[int] = None,
        max_seconds: float = 90.0) -> Dict:
    """Time `steps` CPU training steps on a bounded token sample of the named workload.

    The per-step optimizer / checkpoint cost does not depend on the token count, so tokens/s on a
    short sample UNDER-states the CPU's full-sequence throughput only through that fixed cost --
    which is exactly the hot path being compared.  The sample length is chosen so one step takes
    about `budget_s_per_step` seconds (calibrated on a 16-token probe)."""
    from pyrecover_b200.workload import MODEL_PRESETS, SyntheticTokens
    torch.set_num_threads(threads or os.cpu_count() or 1)    # torchrun exports OMP_NUM_THREADS=1: say it explicitly
    cores = torch.get_num_threads()
    vocab = MODEL_PRESETS[preset][5]
    model, opt, sched, kind = build(preset, full_seq, fused)
    t_probe = None
    if sample_seq is None:
        sample_seq, t_probe = _calibrate(model, opt, sched, vocab, full_seq, budget_s_per_step)
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
    stack = ("the reference's own model.py + utils.build_lr_scheduler (staged in oracle/_ref)" if kind == "reference"
             else "port of the reference step (oracle/cpu_reference_port.py)")
    out = {"tokens_per_s": sample_seq * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": cores, "kind": kind,
           "sample_seq": sample_seq, "steps": steps, "probe_first_step_s": t_probe, "world": 1,
           "sample": f"{steps} steps of {sample_seq} tokens (of {full_seq}) on the {preset} model, {stack}, "
                     f"torch CPU fwd/bwd + torch.optim.AdamW(fused={fused}), {cores} threads"}
    if ckpt_path:
        out["ckpt_store_s"] = save_vanilla(model, opt, sched, steps, ckpt_path, verify=False, kind=kind)
        out["ckpt_bytes"] = os.path.getsize(ckpt_path)
        os.remove(ckpt_path)
    return out


# ------------------------------------------------------------------------- W > 1: DDP on gloo --
def _free_port() -> int:
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _ddp_worker(rank: int, world: int, port: int, threads: int, preset: str, full_seq: int, steps: int, warmup: int,
                fused: bool, budget_s_per_step: float, max_seconds: float, result_path: str) -> None:
    import json

    import torch.distributed as dist
    from pyrecover_b200.workload import MODEL_PRESETS, SyntheticTokens
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    vocab = MODEL_PRESETS[preset][5]
    model, opt, sched, kind = build(preset, full_seq, fused, ddp=True)       # train.py:107-115
    with model.no_sync():       # calibrate the compute only: a gloo all-reduce of the whole model takes seconds by itself
        sample_seq, _ = _calibrate(model, opt, sched, vocab, full_seq, budget_s_per_step)
    t = torch.tensor([sample_seq], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)                                 # every rank the same shape
    sample_seq = int(t.item())
    data = SyntheticTokens(vocab, sample_seq, 1, rank=rank)
    for s in range(warmup):
        train_step(model, opt, sched, *data.batch(s))
    dist.barrier()
    t0 = time.perf_counter()
    done = 0
    for s in range(steps):
        loss = train_step(model, opt, sched, *data.batch(warmup + s))
        float(loss.detach())
        done += 1
        stop = torch.tensor([int(time.perf_counter() - t0 > max_seconds)])
        dist.all_reduce(stop, op=dist.ReduceOp.MAX)                          # all ranks leave together
        if stop.item():
            break
    dist.barrier()
    dt = time.perf_counter() - t0
    if rank == 0:
        stack = ("the reference's own model.py + utils.build_lr_scheduler (staged in oracle/_ref)" if kind == "reference"
                 else "port of the reference step (oracle/cpu_reference_port.py)")
        with open(result_path, "w") as f:
            json.dump({"tokens_per_s": world * sample_seq * done / dt, "ms_per_step": 1e3 * dt / done,
                       "cores": threads * world, "kind": kind, "sample_seq": sample_seq, "steps": done, "world": world,
                       "sample": f"{done} steps of {world} x {sample_seq} tokens (of {world} x {full_seq}) on the {preset} "
                                 f"model, {stack}, DistributedDataParallel on gloo with {world} ranks x {threads} "
                                 f"threads, torch.optim.AdamW(fused={fused})"}, f)
    dist.destroy_process_group()


def run_ddp(preset: str, full_seq: int, steps: int, warmup: int, world: int, fused: bool = True,
            budget_s_per_step: float = 8.0, max_seconds: float = 60.0) -> Dict:
    """The reference's CPU path at world size W: W processes on this host, DDP over gloo, the host's cores
    split evenly (the reference is launched one task per GPU, train.py:107-115)."""
    import json
    import subprocess
    import sys
    import tempfile
    cores = os.cpu_count() or world
    threads = max(1, cores // world)
    port = _free_port()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # the children are a job of their own: none of torchrun's variables may leak into them
    # (TORCHELASTIC_USE_AGENT_STORE alone makes every rank look for a store that the agent hosts)
    env = {k: v for k, v in os.environ.items()
           if not k.startswith("TORCHELASTIC_") and
           k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS", "GROUP_RANK",
                     "GROUP_WORLD_SIZE", "ROLE_RANK", "ROLE_NAME", "ROLE_WORLD_SIZE", "LOCAL_WORLD_SIZE")}
    env["CUDA_VISIBLE_DEVICES"] = ""                        # the CPU arm must not touch a GPU
    env["OMP_NUM_THREADS"] = str(threads)
    with tempfile.TemporaryDirectory(prefix="pr_cpu_ddp_") as tmp:
        result_path = os.path.join(tmp, "result.json")
        procs = [subprocess.Popen([sys.executable, "-m", "oracle.cpu_reference_port", str(r), str(world), str(port),
                                   str(threads), preset, str(full_seq), str(steps), str(warmup), str(int(fused)),
                                   str(budget_s_per_step), str(max_seconds), result_path], cwd=root, env=env)
                 for r in range(world)]
        deadline = time.perf_counter() + max(600.0, 6 * max_seconds)      # never hang the caller
        codes = []
        for p in procs:
            try:
                codes.append(p.wait(timeout=max(1.0, deadline - time.perf_counter())))
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise RuntimeError("CPU DDP arm: workers did not finish in time")
        if any(codes):
            raise RuntimeError(f"CPU DDP arm: worker exit codes {codes}")
        with open(result_path) as f:
            return json.load(f)


if __name__ == "__main__":          # one gloo rank of run_ddp
    import sys
    a = sys.argv[1:]
    _ddp_worker(int(a[0]), int(a[1]), int(a[2]), int(a[3]), a[4], int(a[5]), int(a[6]), int(a[7]), bool(int(a[8])),
                float(a[9]), float(a[10]), a[11])
