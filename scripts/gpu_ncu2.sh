#!/bin/bash
# ncu --set full on selected kernels of profile_kernels.py: gpu_ncu2.sh <kernel-regex> <out-name> <skip> <count> <profile args...>
KRE=$1; OUT=$2; SKIP=$3; CNT=$4; shift 4
mkdir -p gpurun_out
python scripts/profile_kernels.py "$@" > gpurun_out/ncu_plain_$OUT.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c $CNT -f -o gpurun_out/$OUT python scripts/profile_kernels.py "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$OUT.log
