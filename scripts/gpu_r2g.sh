#!/bin/bash
# K7t: one full ncu capture of k_score_mma at 65k x 100k x K16 (one batch).
mkdir -p gpurun_out
CMD="python scripts/profile_kernels.py --variants 65000 --samples 100000 --ops score16 --score-variants 65000 --reps 1"
$CMD > gpurun_out/plain_score.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base function -k regex:k_score_mma -s 1 -c 1 -o gpurun_out/r02_score_mma $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_full.log
