#!/bin/bash
# Checkpoint-stall sensitivity at N=1: grid of the bulk-copy pack kernel that runs beside the next forward.
set -u
mkdir -p gpurun_out
for b in 8 16 32 64 148; do
  PYRECOVER_B200_PACK_BLOCKS=$b python bench.py --steps 6 --warmup 3 --no-incumbent --no-cpu-baseline --stall-blocks 7 \
     > gpurun_out/r2_stall_b$b.json 2> gpurun_out/r2_stall_b$b.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2_stall_b$b.json"))
c=d["ckpt"]
print("blocks $b: stall %.2f ms = %.3f of step, pairs %s, enqueue %.2f ms" % (c["stall_ms"], c["stall_frac_of_step"], [round(x,2) for x in c["stall_ms_per_pair"]], c["enqueue_ms"]))
PY
done
