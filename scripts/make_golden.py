#!/usr/bin/env python
"""Generate tests/golden/*.npz|*.pt from the reference stack itself.

Run in the BUILD container, where /root/reference (PyRecover) and torch 2.11 are importable:
    python scripts/make_golden.py
The GPU box has neither /root/reference nor network access, so the vectors are committed.

What is pinned (the reference holds no golden vectors of its own, so
these are outputs of the reference's call sites run here):
  adamw_bf16.npz / adamw_f32.npz   torch.optim.AdamW(lr, fused=False|True) exactly as
                                   reference train.py:120-122 builds it, stepped with the
                                   reference's LR schedule (utils.py:59-81, imported from
                                   /root/reference/utils.py).
  ddp_mean_w{2,4}.npz              torch DistributedDataParallel on gloo, world 2 and 4
                                   (reference train.py:110-115): the averaged bf16 gradient.
  ref_vanilla.pt(+.md5)            a checkpoint written by the reference's own
                                   save_ckpt_vanilla (pyrecover/checkpoint.py:25-105).
  clip_grad_norm.npz               the reference's clip_grad_norm_ (utils.py:84-89, imported from
                                   /root/reference/utils.py) on fp32 and bf16 gradient lists:
                                   total norm and the clipped gradients.
  checksum_kat.npz                 PR-SUM64 known answers from an independent pure-python
                                   big-int evaluation of the definition.
"""
import importlib.util
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
OUT = ROOT / "tests" / "golden"
REF = Path("/root/reference")


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def bits(t):
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


def gen_adamw(dtype, name, n=4096, steps=12):
    sys.path.insert(0, str(REF))
    ref_utils = _load(REF / "utils.py", "ref_utils")
    out = {}
    g = torch.Generator().manual_seed(1234)
    p0 = (torch.randn(n, generator=g) * 0.02).to(dtype)
    # gradient magnitudes sweep 1e-6 .. 1 so both the eps-dominated and the normalised regime
    # are exercised; a few exact zeros and one huge value as edge cases
    grads = []
    for s in range(steps):
        scale = 10.0 ** (-6 + 6 * s / (steps - 1))
        gr = (torch.randn(n, generator=g) * scale).to(dtype)
        gr[:8] = 0
        gr[8] = 1e4
        grads.append(gr)
    out["p0"] = bits(p0)
    out["grads"] = np.stack([bits(x) for x in grads])
    lrs = []
    for fused in (False, True):
        P = torch.nn.Parameter(p0.clone())
        opt = torch.optim.AdamW([P], lr=1e-5, fused=fused)        # train.py:120-122
        sched = ref_utils.build_lr_scheduler(opt, 10)              # train.py:123
        tag = "fused" if fused else "single"
        for s in range(steps):
            lrs.append(opt.param_groups[0]["lr"]) if not fused else None
            P.grad = grads[s].clone()
            opt.step()
            sched.step()
            st = opt.state[P]
            out[f"{tag}_p_{s + 1}"] = bits(P.data)
            out[f"{tag}_m_{s + 1}"] = bits(st["exp_avg"])
            out[f"{tag}_v_{s + 1}"] = bits(st["exp_avg_sq"])
    out["lrs"] = np.array(lrs, dtype=np.float64)
    out["hyper"] = np.array([0.9, 0.999, 1e-8, 0.01], dtype=np.float64)
    np.savez_compressed(OUT / name, **out)
    print("wrote", name, "lrs", lrs[:3], "...")


def _ddp_worker(rank, world, port, n, ret):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    lin = torch.nn.Linear(n // 64, 64, bias=False).to(torch.bfloat16)
    ddp = torch.nn.parallel.DistributedDataParallel(lin, find_unused_parameters=False)  # train.py:110-115
    g = torch.Generator().manual_seed(100 + rank)
    local = (torch.randn(64, n // 64, generator=g) * 1e-3).to(torch.bfloat16)
    # make backward produce exactly `local` as this rank's weight gradient
    x = torch.eye(n // 64, dtype=torch.bfloat16)
    y = ddp(x)                      # (n/64, 64) = x @ W^T
    y.backward(local.t().contiguous())
    ret[rank] = (bits(local.flatten()), bits(lin.weight.grad.flatten()))
    dist.barrier()
    dist.destroy_process_group()


def gen_ddp(world, n=4096):
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29650 + world
    mp.spawn(_ddp_worker, args=(world, port, n, ret), nprocs=world, join=True)
    locals_ = np.stack([ret[r][0] for r in range(world)])
    outs = np.stack([ret[r][1] for r in range(world)])
    assert all((outs[r] == outs[0]).all() for r in range(world))
    np.savez_compressed(OUT / f"ddp_mean_w{world}.npz", local=locals_, mean=outs[0])
    print("wrote ddp_mean", world)


def gen_ref_ckpt():
    ck = _load(REF / "pyrecover" / "checkpoint.py", "ref_checkpoint")
    torch.manual_seed(7)
    model = torch.nn.Sequential(torch.nn.Linear(24, 16, bias=False), torch.nn.Linear(16, 8, bias=True)).to(torch.bfloat16)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, fused=False)
    sys.path.insert(0, str(REF))
    ref_utils = sys.modules.get("ref_utils") or _load(REF / "utils.py", "ref_utils")
    sched = ref_utils.build_lr_scheduler(opt, 10)
    for s in range(3):
        opt.zero_grad()
        x = torch.randn(4, 24).to(torch.bfloat16)
        model(x).float().pow(2).sum().backward()
        opt.step()
        sched.step()
    path = OUT / "ref_vanilla.pt"
    ck.save_ckpt_vanilla(model, opt, sched, None, 3, 1, str(path), max_keep=0, verify=True)
    print("wrote", path, path.stat().st_size, "bytes")


def gen_checksum_kat():
    rng = np.random.default_rng(99)
    cases = {}
    for i, nbytes in enumerate([0, 1, 3, 4, 15, 16, 17, 63, 64, 1000, 4099, 65536 + 5]):
        data = rng.integers(0, 256, size=nbytes, dtype=np.uint8)
        base = [0, 0, 7, 2**33 + 5][i % 4]
        padded = bytes(data) + b"\0" * ((-nbytes) % 4)
        A = B = 0
        for w in range(len(padded) // 4):
            word = int.from_bytes(padded[4 * w:4 * w + 4], "little")
            A += word
            B += (base + w + 1) * word
        cases[f"data_{i}"] = data
        cases[f"meta_{i}"] = np.array([base, A % 2**64, B % 2**64], dtype=np.uint64)
    cases["count"] = np.array([12])
    np.savez_compressed(OUT / "checksum_kat.npz", **cases)
    print("wrote checksum_kat")


def gen_clip():
    """Reference clip_grad_norm_ on seeded gradient lists, fp32 and bf16, above and below the threshold."""
    sys.path.insert(0, str(REF))
    ref_utils = sys.modules.get("ref_utils") or _load(REF / "utils.py", "ref_utils")
    out = {}
    shapes = [(64, 48), (48,), (300, 17), (5,), (4096,)]
    cases = [("big", 0.3, 1.0), ("small", 1e-4, 1.0), ("tight", 0.3, 1e-2)]
    for tag, scale, max_norm in cases:
        for dtype, dname in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):
            gen = torch.Generator().manual_seed(hash(tag) % 1000 if False else len(tag) * 17)
            params = [torch.nn.Parameter(torch.zeros(s, dtype=dtype)) for s in shapes]
            grads = [(torch.randn(s, generator=gen) * scale).to(dtype) for s in shapes]
            for q, g in zip(params, grads):
                q.grad = g.clone()
            total = ref_utils.clip_grad_norm_(params, max_norm)
            key = f"{tag}_{dname}"
            out[key + "_grads"] = np.concatenate([bits(g).reshape(-1) for g in grads])
            out[key + "_clipped"] = np.concatenate([bits(q.grad).reshape(-1) for q in params])
            out[key + "_total"] = np.array([float(total)], dtype=np.float64)
            out[key + "_max_norm"] = np.array([max_norm], dtype=np.float64)
    np.savez_compressed(OUT / "clip_grad_norm.npz", **out)
    print("wrote clip_grad_norm.npz")


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    gen_clip()
    gen_adamw(torch.bfloat16, "adamw_bf16.npz")
    gen_adamw(torch.float32, "adamw_f32.npz")
    gen_ddp(2)
    gen_ddp(4)
    gen_ref_ckpt()
    gen_checksum_kat()
    gen_boundary()


def gen_boundary():
    """Dump the reference's CLI defaults and checkpoint-function signatures (the boundary)."""
    import inspect
    import json
    sys.path.insert(0, str(REF))
    old = sys.argv
    sys.argv = ["train.py"]
    try:
        ref_utils = sys.modules.get("ref_utils") or _load(REF / "utils.py", "ref_utils")
        flags = vars(ref_utils.get_args())
    finally:
        sys.argv = old
    (OUT / "reference_flags.json").write_text(json.dumps(flags, indent=1, sort_keys=True))
    ck = _load(REF / "pyrecover" / "checkpoint.py", "ref_checkpoint2")
    sigs = {}
    for name in ("save_ckpt_vanilla", "load_ckpt_vanilla", "save_ckpt_distributed", "load_ckpt_distributed",
                 "get_latest_checkpoint"):
        sig = inspect.signature(getattr(ck, name))
        sigs[name] = [[p.name, None if p.default is inspect._empty else p.default] for p in sig.parameters.values()]
    (OUT / "reference_signatures.json").write_text(json.dumps(sigs, indent=1))
    print("wrote boundary fixtures")


if __name__ == "__main__":
    if "--boundary-only" in sys.argv:
        gen_boundary()
    elif "--clip-only" in sys.argv:
        gen_clip()
    else:
        main()
