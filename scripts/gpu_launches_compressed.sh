#!/bin/bash
# ncu launch list (per-kernel device time) of scripts/bench_compressed.py: gpu_launches_compressed.sh <out-name> <args...>
OUT=$1; shift
mkdir -p gpurun_out
timeout 600 python scripts/bench_compressed.py "$@" > gpurun_out/plain_$OUT.jsonl 2> gpurun_out/plain_$OUT.err || { echo "plain failed"; tail -5 gpurun_out/plain_$OUT.err; exit 1; }
cut -c1-300 gpurun_out/plain_$OUT.jsonl
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 400 --csv --log-file gpurun_out/launches_$OUT.csv python scripts/bench_compressed.py "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$OUT.log
