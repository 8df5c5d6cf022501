#!/bin/bash
# Round-2 final single-GPU check: GPU tests, smoke, the default bench line, the reference arm.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_final_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_final_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_final_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_final_smoke.log
python bench.py > gpurun_out/r2_final_bench_n1.json 2> gpurun_out/r2_final_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2_final_bench_n1.json").read().strip().splitlines()[-1]); c=d["ckpt"]; r=d["roofline"]
print("N=1: %.0f tok/s (e2e %.0f) %.2f ms | K2 frac %.3f | stall %.2f ms = %.3f pairs %s blocks %s | inc %s | cpu %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], r["frac"], c["stall_ms"], c["stall_frac_of_step"], [round(x,2) for x in c["stall_ms_per_pair"]], None, d["incumbent"]["tokens_per_s"], d["cpu_baseline"]))
PY
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_final_ref_n1.json 2> gpurun_out/r2_final_ref_n1.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/r2_final_ref_n1.json
