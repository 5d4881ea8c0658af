#!/bin/bash
# First GPU pass of a session: box facts, GPU parity tests, bench (small + full), ncu launch list + full capture.
mkdir -p gpurun_out
{ free -g | head -2; nproc; lscpu | grep "Model name"; nvidia-smi -L; df -h /dev/shm | tail -1; } > gpurun_out/box.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 3 --warmup 3 --variants 200000 > gpurun_out/bench_200k.json 2> gpurun_out/bench_200k.err; echo "bench200k rc=$?"
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "benchfull rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "benchref rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --variants 100000 --skip-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -c 60 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_count_fixed -s 3 -c 2 -o gpurun_out/prof_count_fixed $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
cat gpurun_out/bench_full.json
