#!/bin/bash
# ncu --set full of the sparse k_densify<1> launch of missing_by_sample on the configs[3] fileset (65k variants).
mkdir -p gpurun_out
CMD="python scripts/bench_compressed.py --variants 65000 --samples 100000 --reps 1"
$CMD > gpurun_out/plain_c4.jsonl 2> gpurun_out/plain_c4.err && \
ncu --set full --clock-control none --import-source on --kernel-name-base function -k regex:k_densify -s 1 -c 1 -o gpurun_out/r02_densify $CMD > gpurun_out/ncu_densify.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_densify.log
