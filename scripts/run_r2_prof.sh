#!/bin/bash
# Round-2 ncu session (1 GPU): launch list of the bench step, then full captures of the hot kernels.
set -u
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --profile-range --no-cpu-baseline --no-incumbent --stall-blocks 1"
$BENCH > gpurun_out/r2_prof_plain_bench.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_ncu_launches_bench_n1.csv \
    $BENCH > gpurun_out/r2_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python scripts/profile_kernels.py 1 > gpurun_out/r2_prof_plain_kernels.log 2>&1 &&
ncu --set full --clock-control none --import-source on -f -o gpurun_out/r2_full -k regex:'adamw_bf16_scattered|pack_bulk|gather_grads' \
    python scripts/profile_kernels.py 1 > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/r2_ncu_launches_bench_n1.csv 2>/dev/null; tail -3 gpurun_out/r2_ncu_full.log
