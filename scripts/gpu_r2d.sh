#!/bin/bash
# Round 2, trip D (2 GPUs): multi-GPU contexts through the C ABI (NCCL all-reduce inside the library), bench at N=2
# both as the driver launches it (torchrun) and plain.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/box2.txt; nproc >> gpurun_out/box2.txt; free -g >> gpurun_out/box2.txt; nvidia-smi topo -m >> gpurun_out/box2.txt 2>&1
python -m pytest tests/test_gpu_group.py tests/test_gpu_ops.py -m gpu -x -q -s > gpurun_out/pytest_group.log 2>&1; echo "group rc=$?"; tail -8 gpurun_out/pytest_group.log
timeout 300 scripts/build/h2d_probe 4 4 > gpurun_out/h2d_probe_2gpu.json 2> gpurun_out/h2d_probe.err; echo "probe rc=$?"; cat gpurun_out/h2d_probe_2gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/bench_n2.err; cat gpurun_out/bench_n2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref n2 rc=$?"; cat gpurun_out/bench_ref_n2.json
