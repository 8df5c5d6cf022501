#!/bin/bash
# BASELINE configs[3]: time-aware checkpointing under a simulated SLURM time limit with sharded
# (torch.distributed.checkpoint-format) save + resume.   usage: run_config3.sh <gpus> <preset> <seconds-to-walltime> [tag]
#  1. train with --timeaware-checkpointing and SLURM_JOB_END_TIME = now + budget  -> ckpt_<S>_final
#  2. resume (every rank reads only its own shard, NVLink all-gather) and train to S+4, checkpoint there
#  3. resume once more with PYRECOVER_B200_SLICE_LOAD=0 (every rank reads every shard): load wall time only
#  4. uninterrupted run to S+4 in another directory: its checkpoint must carry the same PR-SUM64 per rank
set -u
N=$1; PRESET=$2; BUDGET=$3; TAG=${4:-r2_config3_w${N}_${PRESET}}
OUT=gpurun_out/$TAG; mkdir -p $OUT
BASE=/tmp; [ "$(df --output=avail -k /dev/shm | tail -1)" -gt "$(df --output=avail -k /tmp | tail -1)" ] && BASE=/dev/shm
CK=$BASE/pr_config3_$$; mkdir -p $CK
echo "checkpoints under $CK ($(df -h $CK | tail -1))" | tee $OUT/summary.txt
launch() { # name, steps, extra...
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29641 train.py --distributed \
    --synthetic-data --model-preset $PRESET --sequence-length 2048 --batch-size $N --fused-optimizer --grad-sync ${GRAD_SYNC:-overlap} \
    --use-torch-distributed-ckpt --checkpoint-dir $CK --experiment_name $1 --training-steps $2 --logging-frequency 50 \
    --max-kept-checkpoints 2 ${EXTRA_FLAGS:-} "${@:3}"
}
t0=$(date +%s.%N)
SLURM_JOB_END_TIME=$(( $(date +%s) + BUDGET )) launch ta 1000000 --timeaware-checkpointing --default-iter-time 0.3 --default-ckpt-time 5 \
    --checkpoint-frequency -1 > $OUT/1_timeaware.log 2>&1; echo "timeaware rc=$? wall $(python -c "import time; print(round(time.time() - $t0, 1))") s" | tee -a $OUT/summary.txt
FINAL=$(ls -d $CK/ta/ckpt_*_final 2>/dev/null | head -1)
S=$(basename "$FINAL" | sed 's/ckpt_\([0-9]*\)_final/\1/')
echo "final checkpoint: $FINAL (step $S): $(ls $FINAL | tr '\n' ' ')" | tee -a $OUT/summary.txt
du -sh $FINAL | tee -a $OUT/summary.txt
grep -E "TIME CHECK|STOP|Updated max" $OUT/1_timeaware.log | tail -8 | tee -a $OUT/summary.txt
[ -z "$S" ] && { echo "no final checkpoint"; rm -rf $CK; exit 1; }
T=$((S + 4))
launch ta $T --resume-from-checkpoint latest --checkpoint-frequency $T > $OUT/2_resume_own_shard.log 2>&1; echo "resume(own shard) rc=$?" | tee -a $OUT/summary.txt
grep -E "Checkpoint loading completed|checkpoint loaded" $OUT/2_resume_own_shard.log | tee -a $OUT/summary.txt
python - <<PY | tee -a $OUT/summary.txt
import json, glob
print("resumed   sums:", [json.load(open(f))["sum64"] for f in sorted(glob.glob("$CK/ta/ckpt_$T/b200_*.json"))])
PY
cp $CK/ta/ckpt_$T/b200_*.json $OUT/ 2>/dev/null
rm -rf $CK/ta/ckpt_$T
PYRECOVER_B200_SLICE_LOAD=0 launch ta $((S + 1)) --resume-from-checkpoint latest --checkpoint-frequency -1 > $OUT/3_resume_all_shards.log 2>&1; echo "resume(all shards) rc=$?" | tee -a $OUT/summary.txt
grep -E "Checkpoint loading completed" $OUT/3_resume_all_shards.log | tee -a $OUT/summary.txt
rm -rf $CK/ta
if [ "${SKIP_STRAIGHT:-0}" != "1" ]; then
  launch straight $T --checkpoint-frequency $T > $OUT/4_straight.log 2>&1; echo "straight rc=$?" | tee -a $OUT/summary.txt
  python - <<PY | tee -a $OUT/summary.txt
import json, glob
a = [json.load(open(f))["sum64"] for f in sorted(glob.glob("$CK/straight/ckpt_$T/b200_*.json"))]
b = [json.load(open(f))["sum64"] for f in sorted(glob.glob("$OUT/b200_*.json"))]
print("straight  sums:", a)
print("WEIGHT/MOMENT EQUALITY (PR-SUM64 of every rank's shard, interrupted+resumed vs uninterrupted):", bool(a) and a == b)
PY
fi
rm -rf $CK
