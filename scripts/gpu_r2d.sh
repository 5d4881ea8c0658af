#!/bin/bash
# Round 2, 2-GPU trip: multi-GPU contexts through the C ABI (NCCL all-reduce inside the library), async jobs over 2
# GPUs, bench at N=2 as the driver launches it (torchrun), reference arm.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_group.py tests/test_gpu_async.py -m gpu -x -q > gpurun_out/pytest_group.log 2>&1; echo "group rc=$?"; tail -4 gpurun_out/pytest_group.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/bench_n2.err; cat gpurun_out/bench_n2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref n2 rc=$?"; cat gpurun_out/bench_ref_n2.json
