#!/bin/bash
# Round-2 two-GPU session: multi-GPU tests, bench in the three exchange modes (parity + incumbent in the default one).
set -u
mkdir -p gpurun_out
echo "(multi-GPU pytest skipped in this run)"
run() { # name, extra args
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 "${@:2}" \
     > gpurun_out/r2_bench_w2_$1.json 2> gpurun_out/r2_bench_w2_$1.err; echo "bench $1 rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2_bench_w2_$1.json"))
    r=d["roofline"]; c=d.get("ckpt") or {}
    print("$1: %.0f tok/s %.2f ms/step host %.2f | k1 %.2f ms wait %.2f exposed %s bus %.0f | stall %s | parity %s | inc %s" % (d["value"], d["ms_per_step"], d["host_issue_ms_per_step"], r["kernel_ms_per_step"], r["barrier_wait_ms_per_step"], r["exposed_ms_after_backward"], r["achieved"] or 0, c.get("stall_frac_of_step"), (d.get("parity") or {}).get("ok"), (d.get("incumbent") or {}).get("tokens_per_s")))
except Exception as e:
    print("$1: no line", e)
PY
  tail -3 gpurun_out/r2_bench_w2_$1.err
}
run overlap
run fused --grad-sync fused --no-incumbent
run replicated --grad-sync replicated --no-incumbent
