#!/bin/bash
# ncu launch list (per-kernel device time) of profile_kernels.py: gpu_launches.sh <out-name> <profile args...>
OUT=$1; shift
mkdir -p gpurun_out
python scripts/profile_kernels.py "$@" > gpurun_out/plain_$OUT.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 400 --csv --log-file gpurun_out/launches_$OUT.csv python scripts/profile_kernels.py "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$OUT.log
