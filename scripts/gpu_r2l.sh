#!/bin/bash
# Round 2, 8-GPU trip: one pd_ctx over 8 GPUs (NCCL all-reduce inside the library): a small slice of the group parity
# tests at k = 8, then bench.py at N = 8 as the driver launches it (torchrun).
mkdir -p gpurun_out
{ nvidia-smi -L; free -g; nproc; df -h /dev/shm; } > gpurun_out/box8.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_group.py -m gpu -x -q -k "8] and (pgen_example or syn_dosage)" > gpurun_out/pytest_group8.log 2>&1; echo "group8 rc=$?"; tail -2 gpurun_out/pytest_group8.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench n8 rc=$?"; grep -v "^$" gpurun_out/bench_n8.err | tail -5; cat gpurun_out/bench_n8.json
