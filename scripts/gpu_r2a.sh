#!/bin/bash
# Round 2, trip A: HWE parity far from equilibrium + whole GPU suite, the box's H2D ceiling, a 1-GPU bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/box.txt; nproc >> gpurun_out/box.txt; free -g >> gpurun_out/box.txt; df -h /dev/shm >> gpurun_out/box.txt
python -m pytest tests/test_gpu_hwe.py -m gpu -x -q > gpurun_out/pytest_hwe.log 2>&1; echo "hwe rc=$?"; tail -15 gpurun_out/pytest_hwe.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 scripts/build/h2d_probe 4 8 > gpurun_out/h2d_probe.json 2> gpurun_out/h2d_probe.err; echo "probe rc=$?"; cat gpurun_out/h2d_probe.json; tail -3 gpurun_out/h2d_probe.err
python bench.py --steps 5 --warmup 3 --skip-cpu > gpurun_out/bench_r2a.json 2> gpurun_out/bench_r2a.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r2a.err
cat gpurun_out/bench_r2a.json
