#!/bin/bash
# Sparse walk A/B: parity tests that walk difflists, then scripts/bench_compressed.py with the SWAR walker (the library)
# and with the round-1 lane-per-group walker (scripts/build/libplinkgpu_lanes.so, -DPD_WALK_LANES).
mkdir -p gpurun_out
V=${1:-200000}
python -m pytest tests/test_gpu_counts.py tests/test_gpu_sample.py tests/test_gpu_fuzz.py tests/test_gpu_dosage.py tests/test_gpu_f4.py -m gpu -x -q > gpurun_out/pytest_walk.log 2>&1; echo "walk tests rc=$?"; tail -3 gpurun_out/pytest_walk.log
python scripts/bench_compressed.py --variants $V --samples 100000 > gpurun_out/compressed_swar.jsonl 2> gpurun_out/compressed_swar.err; echo "swar rc=$?"; tail -2 gpurun_out/compressed_swar.err; cut -c1-420 gpurun_out/compressed_swar.jsonl
if [ -f scripts/build/libplinkgpu_lanes.so ]; then
PLINKGPU_LIB=scripts/build/libplinkgpu_lanes.so python scripts/bench_compressed.py --variants $V --samples 100000 > gpurun_out/compressed_lanes.jsonl 2> gpurun_out/compressed_lanes.err; echo "lanes rc=$?"; tail -2 gpurun_out/compressed_lanes.err; cut -c1-420 gpurun_out/compressed_lanes.jsonl
fi
