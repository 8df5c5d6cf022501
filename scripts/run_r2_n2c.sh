#!/bin/bash
set -u
mkdir -p gpurun_out
for mode in fused overlap; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 5 \
     --grad-sync $mode --no-incumbent --no-parity --stall-blocks 7 > gpurun_out/r2_stall_w2_$mode.json 2> gpurun_out/r2_stall_w2_$mode.err; echo "rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_stall_w2_$mode.json").read().strip().splitlines()[-1]); c=d["ckpt"]
print("$mode: %.2f ms/step | stall %.2f ms = %.3f pairs %s" % (d["ms_per_step"], c["stall_ms"], c["stall_frac_of_step"], [round(x,2) for x in c["stall_ms_per_pair"]]))
PY
done
