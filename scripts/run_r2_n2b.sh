#!/bin/bash
# Round-2 two-GPU session B: new multi-GPU tests, overlap with learned trigger hooks (32 / 64 / 148 CTAs under backward).
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "many_buckets or timeaware" > gpurun_out/r2_pytest_multi_b.log 2>&1; echo "pytest multi rc=$?"; tail -15 gpurun_out/r2_pytest_multi_b.log
run() { # name, env, extra args
  env $2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 --no-incumbent "${@:3}" \
     > gpurun_out/r2_bench_w2_$1.json 2> gpurun_out/r2_bench_w2_$1.err; echo "bench $1 rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_bench_w2_$1.json").read().strip().splitlines()[-1])
    r=d["roofline"]; c=d.get("ckpt") or {}
    print("$1: %.0f tok/s %.2f ms/step host %.2f | k1 %.2f ms wait %.2f exposed %s bus %.0f | stall %s | parity %s" % (d["value"], d["ms_per_step"], d["host_issue_ms_per_step"], r["kernel_ms_per_step"], r["barrier_wait_ms_per_step"], r["exposed_ms_after_backward"], r["achieved"] or 0, c.get("stall_frac_of_step"), (d.get("parity") or {}).get("ok")))
except Exception as e:
    print("$1: no line", e)
PY
  grep -v "^W11\|\*\*\*\|OMP_NUM" gpurun_out/r2_bench_w2_$1.err | tail -3
}
run overlap32 PYRECOVER_B200_OVERLAP_BLOCKS=32
run overlap64 PYRECOVER_B200_OVERLAP_BLOCKS=64
run overlap148 PYRECOVER_B200_OVERLAP_BLOCKS=148
run overlap64_mc "PYRECOVER_B200_OVERLAP_BLOCKS=32 PR_COMM_MULTICAST=1"
