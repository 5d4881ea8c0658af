#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): N-rank parity check over NCCL, then the bench at N ranks.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/multi_box.txt; nvidia-smi topo -m >> gpurun_out/multi_box.txt 2>&1; free -g >> gpurun_out/multi_box.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_rank_check.py > gpurun_out/multi_check_$N.log 2>&1; echo "multi check rc=$?"; tail -3 gpurun_out/multi_check_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; cat gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
