#!/usr/bin/env python
"""Single-GPU micro-benchmarks of K2 (AdamW), K3 (checksum / pack) against the HBM roofline
and the torch incumbents.  CUDA-event timing, L2 flushed by size (all arenas >> 126 MB).
Usage: python scripts/microbench.py [--n-log2 30] [--iters 10] [--json out.json]"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from pyrecover_b200 import kernels as K  # noqa: E402


def timeit(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def sweep(a, hbm):
    """BASELINE configs[4], single-GPU legs: AdamW and checkpoint pack over buckets of 1 MB - 16 GB
    (bf16 bucket bytes), torch's fused AdamW beside it.  Buckets whose working set fits the 126 MB
    L2 are timed with an L2 flush (a 512 MB memset) before every launch."""
    dev = "cuda"
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for lg in range(a.sweep_min_log2, a.sweep_max_log2 + 1):
        nbytes = 1 << lg
        n = nbytes // 2
        p = (torch.randn(n, device=dev) * 0.02).bfloat16()
        g = (torch.randn(n, device=dev) * 1e-3).bfloat16()
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        need_flush = 4 * nbytes < (512 << 20)

        def t(fn, iters=a.iters):
            for _ in range(3):
                fn()
            ts = []
            for _ in range(iters):
                if need_flush:
                    flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            return ts[len(ts) // 2]
        step = [0]

        def k2():
            step[0] += 1
            K.adamw_(p, g, m, v, lr=1e-5, step=step[0], numerics=0)
        row = {"bucket_bytes": nbytes, "l2_flushed": need_flush}
        ms = t(k2)
        row["adamw_ms"], row["adamw_gbs"] = ms, 14 * n / ms / 1e6
        P = torch.nn.Parameter(p)
        P.grad = g
        opt = torch.optim.AdamW([P], lr=1e-5, fused=True)
        ms = t(opt.step)
        row["torch_fused_adamw_ms"], row["torch_fused_adamw_gbs"] = ms, 14 * n / ms / 1e6
        del opt, P
        if lg <= a.sweep_pack_max_log2:
            s = torch.zeros(2, dtype=torch.int64, device=dev)
            image = torch.zeros(3 * nbytes, dtype=torch.uint8, device=dev)
            plan = K.PackPlan([p, m, v], [0, nbytes, 2 * nbytes], 3 * nbytes)
            ms = t(lambda: plan.pack(image.data_ptr(), s))
            row["pack_ms"], row["pack_gbs"] = ms, 2 * 3 * nbytes / ms / 1e6
            ms = t(lambda: K.checksum_into(p, s))
            row["checksum_ms"], row["checksum_gbs"] = ms, nbytes / ms / 1e6
            plan.close()
            del image
        for k in ("adamw_gbs", "torch_fused_adamw_gbs", "pack_gbs", "checksum_gbs"):
            if k in row:
                row[k.replace("_gbs", "_frac")] = row[k] / hbm
        rows.append(row)
        print(f"{nbytes >> 20:6d} MiB bucket | AdamW {row['adamw_ms']:9.4f} ms {row['adamw_gbs']:7.0f} GB/s ({row['adamw_frac']:.2f}) | "
              f"torch fused {row['torch_fused_adamw_ms']:9.4f} ms {row['torch_fused_adamw_gbs']:7.0f} GB/s | "
              + (f"pack {row['pack_ms']:9.4f} ms {row['pack_gbs']:7.0f} GB/s ({row['pack_frac']:.2f}) | checksum {row['checksum_gbs']:7.0f} GB/s"
                 if "pack_ms" in row else "pack: skipped (image would not fit)"), flush=True)
        del p, g, m, v
        torch.cuda.empty_cache()
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true", help="bucket sweep 1 MB - 16 GB instead of the single-size table")
    ap.add_argument("--sweep-min-log2", type=int, default=20)
    ap.add_argument("--sweep-max-log2", type=int, default=34)
    ap.add_argument("--sweep-pack-max-log2", type=int, default=33)
    ap.add_argument("--n-log2", type=int, default=30)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", type=str, default=None)
    a = ap.parse_args()
    n = 1 << a.n_log2
    peaks = {}
    pk = Path(__file__).resolve().parents[1] / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    hbm = peaks.get("hbm_gbs", 6650.0)
    res = {"n": n, "hbm_peak_gbs": hbm, "device": torch.cuda.get_device_name()}
    if a.sweep:
        res = {"hbm_peak_gbs": hbm, "device": torch.cuda.get_device_name(), "rows": sweep(a, hbm)}
        if a.json:
            Path(a.json).parent.mkdir(parents=True, exist_ok=True)
            Path(a.json).write_text(json.dumps(res, indent=1))
        return
    dev = "cuda"
    p = (torch.randn(n, device=dev) * 0.02).bfloat16()
    g = (torch.randn(n, device=dev) * 1e-3).bfloat16()
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)

    def row(name, ms, best, nbytes):
        gbs = nbytes / ms / 1e6
        res[name] = {"ms_median": ms, "ms_best": best, "gbs": gbs, "frac_of_measured_hbm": gbs / hbm}
        print(f"{name:34s} {ms:9.3f} ms (best {best:9.3f})  {gbs:8.1f} GB/s  {gbs / hbm:6.3f} of measured HBM copy")

    # plain copy for reference (the denominators' own method)
    ms, best = timeit(lambda: m.copy_(p), a.iters)
    row("torch copy_ (2 B/elt r+w)", ms, best, 4 * n)
    m.zero_()
    for numerics in (0, 1):
        step = [0]

        def f():
            step[0] += 1
            K.adamw_(p, g, m, v, lr=1e-5, step=step[0], numerics=numerics)
        ms, best = timeit(f, a.iters)
        row(f"pr_adamw bf16 numerics={numerics}", ms, best, 14 * n)
    # torch incumbents on the same flat tensor and on a 291-tensor split
    P = torch.nn.Parameter(p.clone())
    P.grad = g
    for fused in (True, False):
        opt = torch.optim.AdamW([P], lr=1e-5, fused=fused)
        ms, best = timeit(opt.step, a.iters)
        row(f"torch AdamW flat fused={fused}", ms, best, 14 * n)
        del opt
    del P
    torch.cuda.empty_cache()
    chunks = list(torch.split(p.clone(), n // 291 + 1))
    params = [torch.nn.Parameter(c) for c in chunks]
    for q, gg in zip(params, torch.split(g, n // 291 + 1)):
        q.grad = gg
    opt = torch.optim.AdamW(params, lr=1e-5, fused=True)
    ms, best = timeit(opt.step, a.iters)
    row("torch AdamW 291 tensors fused=True", ms, best, 14 * n)
    del opt, params, chunks
    torch.cuda.empty_cache()

    # checksum / pack
    s = torch.zeros(2, dtype=torch.int64, device=dev)
    ms, best = timeit(lambda: K.checksum_into(p, s), a.iters)
    row("pr_checksum (1 B/B read)", ms, best, 2 * n)
    chunk = (n // 291 + 127) // 128 * 128      # 256-B aligned tensor starts, like the flat arena
    views = list(torch.split(p, chunk)) + list(torch.split(m, chunk)) + list(torch.split(v, chunk))
    offs, cur = [], 0
    for t in views:
        offs.append(cur)
        cur += (t.numel() * 2 + 255) // 256 * 256
    image = torch.zeros(cur, dtype=torch.uint8, device=dev)
    plan = K.PackPlan(views, offs, cur)
    from pyrecover_b200 import _lib
    ms, best = timeit(lambda: plan.pack(image.data_ptr(), s, algo=_lib.PR_PACK_LDG), a.iters)
    row("pr_ckpt_pack LDG/STG p+m+v (+sum)", ms, best, 2 * 6 * n)
    ms, best = timeit(lambda: plan.pack(image.data_ptr(), None, algo=_lib.PR_PACK_LDG), a.iters)
    row("pr_ckpt_pack LDG/STG (no sum)", ms, best, 2 * 6 * n)
    ms, best = timeit(lambda: plan.unpack(image.data_ptr(), s, algo=_lib.PR_PACK_LDG), a.iters)
    row("pr_ckpt_unpack LDG/STG (+sum)", ms, best, 2 * 6 * n)
    # bulk-copy (TMA-staged) kernel: grid sweep -- how few CTAs carry the copy
    for blocks in (8, 16, 24, 32, 48, 64, 96, 148):
        ms, best = timeit(lambda: plan.pack(image.data_ptr(), s, algo=_lib.PR_PACK_BULK, max_blocks=blocks), a.iters)
        row(f"pr_ckpt_pack BULK {blocks:3d} CTAs (+sum)", ms, best, 2 * 6 * n)
    for blocks in (32, 64, 148):
        ms, best = timeit(lambda: plan.pack(image.data_ptr(), None, algo=_lib.PR_PACK_BULK, max_blocks=blocks), a.iters)
        row(f"pr_ckpt_pack BULK {blocks:3d} CTAs (no sum)", ms, best, 2 * 6 * n)
        ms, best = timeit(lambda: plan.unpack(image.data_ptr(), s, algo=_lib.PR_PACK_BULK, max_blocks=blocks), a.iters)
        row(f"pr_ckpt_unpack BULK {blocks:3d} CTAs (+sum)", ms, best, 2 * 6 * n)
    # gather: scattered gradients -> flat arena (pr_gather_grads), 4 B per element
    gchunks = [c.clone() for c in torch.split(g, chunk)]
    aplan = K.AdamWPlan([i * chunk for i in range(len(gchunks))], [c.numel() for c in gchunks], chunk * len(gchunks))
    garena = torch.zeros(chunk * len(gchunks), dtype=torch.bfloat16, device=dev)
    ms, best = timeit(lambda: aplan.gather(gchunks, garena), a.iters)
    row("pr_gather_grads (2 B r + 2 B w)", ms, best, 4 * n)
    gviews = [garena[i * chunk:i * chunk + c.numel()] for i, c in enumerate(gchunks)]
    ms, best = timeit(lambda: torch._foreach_copy_(gviews, gchunks), a.iters)
    row("torch._foreach_copy_ same tensors", ms, best, 4 * n)
    del gchunks, garena
    # D2H of the image into the pinned arena (PCIe leg)
    arena = K.HostArena(cur)
    ms, best = timeit(lambda: K.copy_d2h_async(arena.ptr, image, cur, 64 << 20), 3, warmup=1)
    res["d2h_pinned"] = {"ms_median": ms, "gbs": cur / ms / 1e6}
    print(f"{'D2H image -> pinned arena':34s} {ms:9.3f} ms  {cur / ms / 1e6:8.1f} GB/s (PCIe)")
    if a.json:
        Path(a.json).parent.mkdir(parents=True, exist_ok=True)
        Path(a.json).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
