#!/bin/bash
# K7t iteration: score parity tests + timing at 200k x 100k x K16 (and K1) + per-kernel launch list of one batch.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dosage.py -m gpu -x -q -k "score" > gpurun_out/pytest_score.log 2>&1; echo "score tests rc=$?"; tail -4 gpurun_out/pytest_score.log
python scripts/profile_kernels.py --variants 200000 --samples 100000 --ops score1,score16 --score-variants 200000 > gpurun_out/score_table.jsonl 2> gpurun_out/score_table.err; echo "table rc=$?"; cat gpurun_out/score_table.jsonl; tail -3 gpurun_out/score_table.err
CMD="python scripts/profile_kernels.py --variants 65000 --samples 100000 --ops score16 --score-variants 65000 --reps 1"
$CMD > gpurun_out/plain_score.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base function -k regex:^k_ -c 14 --csv --log-file gpurun_out/launches_score.csv $CMD > gpurun_out/ncu_ll.log 2>&1
echo "ll rc=$?"
python - <<'PY'
import csv
lines=[l for l in open('gpurun_out/launches_score.csv') if not l.startswith('==')]
rows=[(r['Kernel Name'][:44], float(r['Metric Value'].replace(',',''))) for r in csv.DictReader(lines)]
for n,v in rows[7:]: print(f"{n:44s} {v/1e3:9.1f} us")
PY
