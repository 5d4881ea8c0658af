#!/bin/bash
# Quick loop: selected GPU tests + selected kernel table rows.  usage: gpu_quick.sh "<pytest -k expr>" "<ops>" [profile args]
mkdir -p gpurun_out
K="$1"; OPS="$2"; shift 2
timeout 150 python -m pytest tests -m gpu -x -q -k "$K" > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_quick.log
timeout 600 python scripts/profile_kernels.py --ops "$OPS" "$@" > gpurun_out/kernels_quick.jsonl 2> gpurun_out/kernels_quick.err; echo "profile rc=$?"; cat gpurun_out/kernels_quick.jsonl; tail -3 gpurun_out/kernels_quick.err
