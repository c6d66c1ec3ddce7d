export TETS=1
echo base; python tools/bench_prec.py 129 3 2>&1 | grep "sys="
for w in 512 1024 2048; do echo SWEEP_WIDE=$w FACTOR_WIDE=$w; SU2B200_SWEEP_WIDE=$w SU2B200_FACTOR_WIDE=$w python tools/bench_prec.py 129 3 2>&1 | grep "sys="; done
for w in 4096 1024 256; do echo own16 FACTOR_WIDE=$w; SU2B200_LIB=$PWD/su2_b200/variants/libsu2b200_own16.so SU2B200_FACTOR_WIDE=$w SU2B200_SWEEP_WIDE=1024 python tools/bench_prec.py 129 3 2>&1 | grep "sys="; done
