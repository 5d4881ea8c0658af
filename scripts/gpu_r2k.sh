#!/bin/bash
# Round 2, kernel tables with the final K7t: score tests, 200k x 100k table of every op, 1M x 500k (125 GB resident)
# counts / per-sample / plink_score K=16 over all 1M variants (BASELINE configs[4] at N = 1), ncu full of k_score_mma.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dosage.py -m gpu -x -q -k "score" > gpurun_out/pytest_score.log 2>&1; echo "score tests rc=$?"; tail -2 gpurun_out/pytest_score.log
python scripts/profile_kernels.py --variants 200000 --samples 100000 --score-variants 200000 > gpurun_out/kernels.jsonl 2> gpurun_out/kernels.err; echo "table rc=$?"; cat gpurun_out/kernels.jsonl
python scripts/profile_kernels.py --variants 1000000 --samples 500000 --ops counts,missing_by_sample,sample_counts,score16 --score-variants 1000000 --reps 2 > gpurun_out/kernels_1Mx500k.jsonl 2> gpurun_out/kernels_1Mx500k.err; echo "1Mx500k rc=$?"; cat gpurun_out/kernels_1Mx500k.jsonl; tail -2 gpurun_out/kernels_1Mx500k.err
CMD="python scripts/profile_kernels.py --variants 65000 --samples 100000 --ops score16 --score-variants 65000 --reps 1"
$CMD > gpurun_out/plain_score.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base function -k regex:k_score_mma -s 1 -c 1 -o gpurun_out/r02_score_mma $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_full.log
