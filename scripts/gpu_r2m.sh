#!/bin/bash
# Round 2: the allocation-free entry point's test alone in a fresh process, then the whole GPU suite (which ends with
# it), then both bench arms.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -3 gpurun_out/pytest_no_alloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_ref.json
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_full.err; cut -c1-1500 gpurun_out/bench_full.json
python scripts/profile_kernels.py --variants 200000 --samples 100000 --score-variants 200000 --cpu-variants 200000 > gpurun_out/kernels.jsonl 2> gpurun_out/kernels.err; echo "table rc=$?"; cut -c1-330 gpurun_out/kernels.jsonl; tail -2 gpurun_out/kernels.err
