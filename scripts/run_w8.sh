#!/bin/bash
# One 8-GPU session: K1 sweep (P2P, NVSwitch multicast, NCCL), then the training bench on
# BASELINE configs[1] (~1B) and configs[2] (8B).  Outputs under gpurun_out/.
set -u
export PR_COMM_TIMEOUT_S=20
N=${1:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_w$N.txt 2>&1
timeout 420 $RUN --master-port 29511 scripts/allreduce_bench.py --max-log2 31 --mc-blocks 16,32,64,148 \
    --json gpurun_out/ar_w$N.json > gpurun_out/ar_w$N.log 2>&1
rc=$?
echo "allreduce_bench rc=$rc"
grep -E "MiB|multimem|world" gpurun_out/ar_w$N.log | tail -30
if [ $rc -ne 0 ]; then
    echo "multicast sweep failed: training benches run with PR_COMM_MULTICAST=0"
    export PR_COMM_MULTICAST=0
    tail -20 gpurun_out/ar_w$N.log
fi
timeout 300 $RUN --master-port 29512 bench.py --gpus $N > gpurun_out/bench_w${N}_1b.json 2> gpurun_out/bench_w${N}_1b.err
echo "bench 1b rc=$?"
cut -c1-250 gpurun_out/bench_w${N}_1b.json
timeout 600 $RUN --master-port 29513 bench.py --gpus $N --preset 8b --steps 8 --warmup 3 > gpurun_out/bench_w${N}_8b.json 2> gpurun_out/bench_w${N}_8b.err
echo "bench 8b rc=$?"
cut -c1-250 gpurun_out/bench_w${N}_8b.json
tail -5 gpurun_out/bench_w${N}_8b.err
df -h /tmp | tail -1
