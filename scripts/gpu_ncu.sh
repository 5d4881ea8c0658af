#!/bin/bash
# ncu --set full of selected kernels: bash scripts/gpu_ncu.sh <kernel-regex> <out-name> <skip> <count> -- <profile_kernels.py args>
KRE=$1; OUT=$2; SKIP=$3; CNT=$4; shift 5
mkdir -p gpurun_out
CMD="python scripts/profile_kernels.py $*"
$CMD > gpurun_out/ncu_plain_$OUT.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c $CNT -f -o gpurun_out/$OUT $CMD > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$OUT.log
