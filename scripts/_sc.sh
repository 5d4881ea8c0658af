python scripts/profile_kernels.py --ops score1,score16 --variants 100000 --score-variants 20000 2>&1 | tail -2 | cut -c1-330
PD_SCORE_PATH=fp64 python scripts/profile_kernels.py --ops score1,score16 --variants 100000 --score-variants 20000 2>&1 | tail -2 | cut -c1-330
