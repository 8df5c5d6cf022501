#!/bin/bash
# Round-2 eight-GPU session B: BASELINE configs[3] at 8 B (time-aware stop, sharded DCP save, resume both ways, equality with an
# uninterrupted run) and configs[4] (K1 sweep 1 MiB - 16 GiB with NVLink byte counters).
set -u
mkdir -p gpurun_out
EXTRA_FLAGS=--deterministic bash scripts/run_config3.sh 8 8b 100 r2_config3_w8_8b > gpurun_out/r2_config3_w8_8b.log 2>&1
cat gpurun_out/r2_config3_w8_8b/summary.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 scripts/allreduce_bench.py \
   --min-log2 20 --max-log2 34 --mc-blocks 16,32 --skip-nccl-above-log2 32 --nvlink-counters --json gpurun_out/r2_allreduce_w8.json \
   > gpurun_out/r2_allreduce_w8.log 2>&1; echo "allreduce sweep rc=$?"
grep -v "^W11\|\*\*\*\|OMP_NUM\|^\[W" gpurun_out/r2_allreduce_w8.log | tail -28
