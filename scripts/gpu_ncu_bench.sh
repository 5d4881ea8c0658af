#!/bin/bash
# ncu launch list of the bench command itself (after the same command ran clean without ncu):
#   gpu_ncu_bench.sh <out-name> <bench args...>
OUT=$1; shift
mkdir -p gpurun_out
timeout 600 python bench.py "$@" > gpurun_out/bench_plain_$OUT.json 2> gpurun_out/bench_plain_$OUT.err; RC=$?
echo "plain rc=$RC"; cat gpurun_out/bench_plain_$OUT.json | cut -c1-300
[ $RC -eq 0 ] || exit $RC
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 600 --csv --log-file gpurun_out/launches_$OUT.csv python bench.py "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$OUT.log; wc -l gpurun_out/launches_$OUT.csv
