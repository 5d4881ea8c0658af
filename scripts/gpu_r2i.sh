#!/bin/bash
# Round 2, second 2-GPU trip: pinned-file group contexts (piece-crossing DMAs) + bench at N=2 as the driver launches it.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_group.py -k pinned -m gpu -x -q > gpurun_out/pytest_group_pinned.log 2>&1; echo "group pinned rc=$?"; tail -4 gpurun_out/pytest_group_pinned.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/bench_n2.err; cat gpurun_out/bench_n2.json
