#!/bin/bash
# Fileset tests, difflist tests with the buffered lane walker, compressed bench.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fileset.py -m gpu -x -q > gpurun_out/pytest_fileset.log 2>&1; echo "fileset rc=$?"; tail -15 gpurun_out/pytest_fileset.log
bash scripts/gpu_r2n.sh 200000
