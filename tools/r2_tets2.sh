python -m pytest tests/test_linsys_gpu.py tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -3
echo hex; python tools/bench_prec.py 185 5 2>&1 | grep "sys=0"
export TETS=1
for w in 4096 2048 1024 512; do echo tets SWEEP_WIDE=$w; SU2B200_SWEEP_WIDE=$w python tools/bench_prec.py 129 3 2>&1 | grep "sys=0"; done
