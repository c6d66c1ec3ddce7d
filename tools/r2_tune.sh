for b in 3 4 5 6; do echo BLOCKS_U=$b; SU2B200_SWEEP_BLOCKS_U=$b python tools/bench_prec.py 185 5 2>&1 | grep "sys=0"; done
for w in 32 64; do echo WIDE=$w; SU2B200_SWEEP_WIDE=$w SU2B200_FACTOR_WIDE=$w python tools/bench_prec.py 185 5 2>&1 | grep "sys=0"; done
