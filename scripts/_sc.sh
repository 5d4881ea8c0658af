python scripts/profile_kernels.py --ops score1,score16 --variants 200000 --score-variants 200000 2>&1 | tail -2 | cut -c1-400
PD_SCORE_PATH=fp64 python scripts/profile_kernels.py --ops score16 --variants 200000 --score-variants 200000 --reps 1 2>&1 | tail -1 | cut -c1-400
