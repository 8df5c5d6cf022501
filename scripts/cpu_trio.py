#!/usr/bin/env python
"""CPU context numbers for the three pieces of the hot path ("microbench trio"):
the reference's own library calls on the host cores of the box -- gloo all_reduce (W = 2), torch.optim.AdamW.step
on one flat bf16 tensor, torch.save + md5 of the same state.  A baseline for context, not a target.

    python scripts/cpu_trio.py [--json profiles/r1_cpu_trio.json]
"""
import argparse
import hashlib
import json
import os
import tempfile
import time
from pathlib import Path

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _median(ts):
    ts = sorted(ts)
    return ts[len(ts) // 2]


def _gloo_worker(rank, world, port, sizes, threads, ret):
    torch.set_num_threads(threads)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rows = []
    for nbytes in sizes:
        t = (torch.randn(nbytes // 2) * 1e-3).bfloat16()
        for _ in range(2):
            dist.all_reduce(t)
        ts = []
        for _ in range(5):
            dist.barrier()
            t0 = time.perf_counter()
            dist.all_reduce(t)
            ts.append(time.perf_counter() - t0)
        rows.append({"bytes": nbytes, "ms": _median(ts) * 1e3, "bus_gbs": nbytes / _median(ts) / 1e9 * 2 * (world - 1) / world})
    if rank == 0:
        ret.put(rows)
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    ap.add_argument("--adamw-log2", type=int, default=26)
    a = ap.parse_args()
    cores = len(os.sched_getaffinity(0))
    out = {"cores": cores, "torch": torch.__version__}
    # 1. gloo all_reduce, W = 2 (the reference's CPU collective: DDP on gloo)
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    sizes = [1 << 20, 1 << 24, 1 << 27]
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, 29650 + os.getpid() % 100, sizes, max(1, cores // 2), ret)) for r in range(2)]
    for p in procs:
        p.start()
    out["gloo_allreduce_w2"] = ret.get(timeout=600)
    for p in procs:
        p.join()
    # 2. torch.optim.AdamW.step on one flat bf16 tensor (reference train.py:120-122, :274)
    torch.set_num_threads(cores)
    n = 1 << a.adamw_log2
    for fused in (False, True):
        p = torch.nn.Parameter((torch.randn(n) * 0.02).bfloat16())
        p.grad = (torch.randn(n) * 1e-3).bfloat16()
        opt = torch.optim.AdamW([p], lr=1e-5, fused=fused)
        opt.step()
        ts = []
        for _ in range(5):
            t0 = time.perf_counter()
            opt.step()
            ts.append(time.perf_counter() - t0)
        out[f"adamw_step_fused_{fused}"] = {"elements": n, "ms": _median(ts) * 1e3, "gbs_at_14B_per_elt": 14 * n / _median(ts) / 1e9}
    # 3. torch.save + md5 of {model, optimizer} for the same tensor (reference pyrecover/checkpoint.py:74-83)
    state = {"model": {"w": p.detach()}, "optimizer": opt.state_dict()}
    d = tempfile.mkdtemp(prefix="pr_trio_")
    path = os.path.join(d, "ckpt.pt")
    t0 = time.perf_counter()
    torch.save(state, path)
    t1 = time.perf_counter()
    h = hashlib.md5()
    with open(path, "rb") as f:
        h.update(f.read())            # the reference reads the whole file at once (checkpoint.py:79-80)
    t2 = time.perf_counter()
    size = os.path.getsize(path)
    out["torch_save_md5"] = {"bytes": size, "save_s": t1 - t0, "md5_s": t2 - t1, "gbs_total": size / (t2 - t0) / 1e9}
    os.remove(path)
    os.rmdir(d)
    print(json.dumps(out, indent=1))
    if a.json:
        Path(a.json).parent.mkdir(parents=True, exist_ok=True)
        Path(a.json).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
