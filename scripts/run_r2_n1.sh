#!/bin/bash
# Round-2 single-GPU session: GPU tests, smoke, micro-benchmarks, bench.  Logs land in gpurun_out/.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest_gpu.log
tail -5 gpurun_out/r2_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2_smoke.log
python scripts/microbench.py --n-log2 30 --json gpurun_out/r2_microbench_n1.json > gpurun_out/r2_microbench_n1.log 2>&1; echo "microbench rc=$?"; tail -40 gpurun_out/r2_microbench_n1.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/r2_bench_n1.json; tail -5 gpurun_out/r2_bench_n1.err
