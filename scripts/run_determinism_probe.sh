#!/bin/bash
# Is a straight 1 B run reproducible run to run on one GPU?  (with and without --deterministic)
set -u
mkdir -p gpurun_out; CK=/dev/shm/pr_det_$$
run() { python train.py --synthetic-data --model-preset 1b --sequence-length 2048 --fused-optimizer --training-steps 12 --checkpoint-frequency 12 \
   --checkpoint-dir $CK --experiment_name $1 --logging-frequency 6 "${@:2}" > gpurun_out/r2_det_$1.log 2>&1; echo "$1 rc=$?"; grep "Step: 12" gpurun_out/r2_det_$1.log | sed 's/.*Step: 12/Step: 12/'; cat $CK/$1/ckpt_12.pt.b200sum | cut -c1-80; echo; }
run da --deterministic; run db --deterministic
rm -rf $CK
