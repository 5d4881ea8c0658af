#!/bin/bash
# f4: multiallelic decode (pd_decode_alleles vs PgrGetM) + the other f4 tests.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_f4.py -m gpu -x -q > gpurun_out/pytest_f4.log 2>&1; echo "f4 rc=$?"; tail -25 gpurun_out/pytest_f4.log
