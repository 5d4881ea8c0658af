#!/bin/bash
# int8 MMA peak + per-kernel breakdown of one plink_score batch (launch list of our kernels only).
mkdir -p gpurun_out
timeout 120 scripts/build/i8_mma_peak > gpurun_out/i8_mma_peak.json 2> gpurun_out/i8_mma_peak.err; echo "peak rc=$?"; cat gpurun_out/i8_mma_peak.json
CMD="python scripts/profile_kernels.py --variants 65000 --samples 100000 --ops score16 --score-variants 65000 --reps 1"
$CMD > gpurun_out/plain_score.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base function -k regex:^k_ -c 60 --csv --log-file gpurun_out/launches_score.csv $CMD > gpurun_out/ncu_ll.log 2>&1
echo "ll rc=$?"; tail -2 gpurun_out/plain_score.log
