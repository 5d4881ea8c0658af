#!/bin/bash
# K7t side kernels: score tests, then one full ncu capture of k_tc_finish + k_tc_plan (second batch of a 65k x 100k x K16 call).
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dosage.py -m gpu -x -q -k "score" > gpurun_out/pytest_score.log 2>&1; echo "score tests rc=$?"; tail -2 gpurun_out/pytest_score.log
CMD="python scripts/profile_kernels.py --variants 65000 --samples 100000 --ops score16 --score-variants 65000 --reps 1"
$CMD > gpurun_out/plain_score.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base function -k regex:"k_tc_finish|k_tc_plan" -s 2 -c 2 -o gpurun_out/r02_score_side $CMD > gpurun_out/ncu_side.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_side.log
