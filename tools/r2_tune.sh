for w in 128 256 512 768; do echo WIDE=$w; SU2B200_SWEEP_WIDE=$w SU2B200_FACTOR_WIDE=$w python tools/bench_prec.py 185 5 2>&1 | grep "sys=0"; done
echo "sweep 512 factor 1024"; SU2B200_SWEEP_WIDE=512 SU2B200_FACTOR_WIDE=1024 python tools/bench_prec.py 185 5 2>&1 | grep "sys=0"
echo "129 cube"; for w in 512 1024 4096; do echo WIDE=$w; SU2B200_SWEEP_WIDE=$w SU2B200_FACTOR_WIDE=$w python tools/bench_prec.py 129 5 2>&1 | grep "sys=0"; done
