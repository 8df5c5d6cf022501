#!/usr/bin/env python
"""Where does the checkpoint stall come from?  Per-step CUDA-event timeline and host-side timing
of one save_ckpt_vanilla inside a run of training steps (1 GPU, BASELINE configs[1] workload).

    python scripts/ckpt_stall_probe.py [--preset 1b] [--json gpurun_out/stall_probe.json]
"""
import argparse
import cProfile
import io
import json
import pstats
import sys
import tempfile
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from pyrecover import checkpoint as ckpt  # noqa: E402
from pyrecover_b200.trainer import Trainer  # noqa: E402
from pyrecover_b200.workload import MODEL_PRESETS, SyntheticTokens, Transformer, preset_args  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="1b")
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--json", default=None)
    ap.add_argument("--profile", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = Transformer(preset_args(a.preset, a.seq))
    torch.set_default_dtype(torch.float32)
    tr = Trainer(model, dev, lr=1e-5, fused=True, warmup_steps=10)
    data = SyntheticTokens(MODEL_PRESETS[a.preset][5], a.seq, 1, seed=0, rank=0, pin=True)
    batches = [tuple(t.to(dev) for t in data.batch(s)) for s in range(4)]
    tmp = Path(tempfile.mkdtemp(prefix="pr_stall_"))
    snap = ckpt._snapshotter(tr.model, tr.optimizer)
    snap.full()
    for s in range(6):
        tr.step(*batches[s % 4])
    torch.cuda.synchronize()
    out = {}

    def run(tag, save_at, n=12, what="full"):
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        host = []
        (tmp / tag).mkdir(parents=True, exist_ok=True)
        torch.cuda.synchronize()
        evs[0].record()
        t_begin = time.perf_counter()
        save_host = None
        for i in range(n):
            h0 = time.perf_counter()
            tr.step(*batches[i % 4])
            if i == save_at:
                s0 = time.perf_counter()
                if what == "full":
                    ckpt.save_ckpt_vanilla(tr.model, tr.optimizer, tr.lr_scheduler, None, 100 + i, 1,
                                           tmp / tag / f"ckpt_{100 + i}.pt", max_keep=1, verify=False)
                elif what == "state_only":
                    ckpt._collect_state(tr.model, tr.optimizer, tr.lr_scheduler, None, 1, 1)
                elif what == "sleep20":
                    time.sleep(0.020)
                save_host = time.perf_counter() - s0
            evs[i + 1].record()
            host.append(time.perf_counter() - h0)
        t_issue = time.perf_counter() - t_begin
        torch.cuda.synchronize()
        t_total = time.perf_counter() - t_begin
        gpu = [evs[i].elapsed_time(evs[i + 1]) for i in range(n)]
        ckpt.wait_for_checkpoints()
        out[tag] = {"gpu_ms_per_step": [round(x, 2) for x in gpu], "host_ms_per_step": [round(x * 1e3, 2) for x in host],
                    "save_host_ms": None if save_host is None else round(save_host * 1e3, 2),
                    "host_issue_ms": round(t_issue * 1e3, 1), "wall_ms": round(t_total * 1e3, 1),
                    "sum_gpu_ms": round(sum(gpu), 1), "stats": dict(snap.last_stats) if what == "full" and save_at >= 0 else None}
        print(tag, json.dumps(out[tag]), flush=True)

    (tmp / "a").mkdir(parents=True, exist_ok=True)
    run("baseline1", -1)
    run("baseline2", -1)
    run("save_full", 3)
    run("baseline3", -1)
    run("state_only", 3, what="state_only")
    run("sleep20", 3, what="sleep20")
    run("save_full2", 3)
    if a.profile:
        pr = cProfile.Profile()
        (tmp / "p").mkdir(parents=True, exist_ok=True)
        torch.cuda.synchronize()
        pr.enable()
        ckpt.save_ckpt_vanilla(tr.model, tr.optimizer, tr.lr_scheduler, None, 999, 1, tmp / "p" / "ckpt_999.pt",
                               max_keep=1, verify=False)
        pr.disable()
        ckpt.wait_for_checkpoints()
        s = io.StringIO()
        pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(25)
        print(s.getvalue())
    if a.json:
        Path(a.json).parent.mkdir(parents=True, exist_ok=True)
        Path(a.json).write_text(json.dumps(out, indent=1))
    tr.close()
    ckpt.release_checkpoint_engines()
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
