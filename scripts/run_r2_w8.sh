#!/bin/bash
# Round-2 eight-GPU session A: the training bench at N = 8 (default = overlap, 32 CTAs under backward; 16 / 64; fused) and N = 4.
set -u
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_w8.txt 2>&1; free -g | head -2; df -h /tmp /dev/shm | tail -2; nproc
run() { # name, ngpus, env, extra args
  env $3 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $2 --steps 20 --warmup 5 "${@:4}" \
     > gpurun_out/r2_bench_$1.json 2> gpurun_out/r2_bench_$1.err; echo "bench $1 rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_bench_$1.json").read().strip().splitlines()[-1])
    r=d["roofline"]; c=d.get("ckpt") or {}
    print("$1: %.0f tok/s %.2f ms/step host %.2f | k1 %.2f ms wait %.2f exposed %s bus %.0f transfer %s | stall %s | parity %s | inc %s" % (d["value"], d["ms_per_step"], d["host_issue_ms_per_step"], r["kernel_ms_per_step"], r["barrier_wait_ms_per_step"], r["exposed_ms_after_backward"], r["achieved"] or 0, r["transfer_only"]["bus_gbs"], c.get("stall_frac_of_step"), (d.get("parity") or {}).get("ok"), (d.get("incumbent") or {}).get("tokens_per_s")))
except Exception as e:
    print("$1: no line", e)
PY
  grep -v "^W11\|\*\*\*\|OMP_NUM\|^\[W" gpurun_out/r2_bench_$1.err | tail -3
}
LIGHT="--no-incumbent --no-parity --stall-blocks 2 --no-cpu-baseline"
run w8_1b 8 X=1
run w8_1b_overlap16 8 PYRECOVER_B200_OVERLAP_BLOCKS=16 $LIGHT
run w8_1b_overlap64 8 PYRECOVER_B200_OVERLAP_BLOCKS=64 $LIGHT
run w8_1b_fused 8 X=1 --grad-sync fused $LIGHT
run w4_1b 4 CUDA_VISIBLE_DEVICES=0,1,2,3
run w4_1b_fused 4 CUDA_VISIBLE_DEVICES=0,1,2,3 --grad-sync fused $LIGHT
