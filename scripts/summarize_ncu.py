#!/usr/bin/env python
"""Summaries of ncu output for profiles/ (run here, no GPU needed):

    python scripts/summarize_ncu.py launches gpurun_out/r2_ncu_launches_bench_n1.csv profiles/r2_ncu_launch_summary.json
    ncu -i gpurun_out/r2_full.ncu-rep --page raw --csv > /tmp/raw.csv
    python scripts/summarize_ncu.py full /tmp/raw.csv profiles/r2_ncu_full_summary.json
"""
import csv
import json
import sys

OURS = ("adamw", "pack_", "gather_grads", "checksum", "sumsq", "norm_finish", "allreduce", "allgather", "signal_post")


def launches(src, dst):
    rows = [r for r in csv.reader(line for line in open(src) if not line.startswith("==")) if r]
    hdr = rows[0]
    k, m, v = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    unit = hdr.index("Metric Unit")
    per = {}
    total = 0.0
    n = 0
    for r in rows[1:]:
        if len(r) <= v or r[m] != "gpu__time_duration.sum":
            continue
        ms = float(r[v].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[unit], 1e-6)
        e = per.setdefault(r[k], {"ms": 0.0, "launches": 0})
        e["ms"] += ms
        e["launches"] += 1
        total += ms
        n += 1
    for e in per.values():
        e["share"] = e["ms"] / total
    top = sorted(({"kernel": name[:140], **e} for name, e in per.items()), key=lambda x: -x["ms"])[:12]
    ours = {name: e for name, e in per.items() if any(t in name for t in OURS) and "at::" not in name}
    json.dump({"source": src, "total_ms": total, "launches": n, "ours": ours, "top": top,
               "note": "per-launch times under ncu are cold-cache and serialised: the SHARE of the step is what must agree "
                       "with bench.py's live CUDA-event figure (roofline.share_of_step)"}, open(dst, "w"), indent=1)
    print(f"{n} launches, {total:.2f} ms;", {k[:50]: round(e["share"], 4) for k, e in ours.items()})


def full(src, dst):
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct"]
    out = []
    for r in rows[2:]:
        d = {k: r[idx[k]] for k in want if k in idx}
        d["units"] = {k: units[idx[k]] for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum") if k in idx}
        out.append(d)
    json.dump(out, open(dst, "w"), indent=1)
    print(len(out), "kernels")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
