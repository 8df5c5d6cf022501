#!/bin/bash
# One 1-GPU session: GPU tests, bench (own arm + reference arm), microbenches, ncu launch list and full capture.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu3.log
python bench.py > gpurun_out/bench_n1_b.json 2> gpurun_out/bench_n1_b.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_n1_b.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_n1_ref.json 2> gpurun_out/bench_n1_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_n1_ref.json
python scripts/microbench.py --json gpurun_out/microbench_n1.json > gpurun_out/microbench_n1.log 2>&1; echo "microbench rc=$?"; cat gpurun_out/microbench_n1.log
python scripts/microbench.py --sweep --json gpurun_out/microbench_sweep.json > gpurun_out/microbench_sweep.log 2>&1; echo "sweep rc=$?"; cat gpurun_out/microbench_sweep.log
python scripts/profile_kernels.py 2 > gpurun_out/profile_kernels.log 2>&1; echo "profile_kernels rc=$?"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1b_ncu_launches_bench_n1.csv \
    python bench.py --steps 2 --warmup 3 --profile-range --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -f -o gpurun_out/r1b_full -k regex:'adamw|pack_kernel|checksum|sumsq|norm_finish' \
    python scripts/profile_kernels.py 1 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null
