set -x
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2_zz_alone.log 2>&1; tail -3 gpurun_out/r2_zz_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_full.log 2>&1; tail -5 gpurun_out/r2_gpu_full.log
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 400 gpurun_out/r2_bench_n1.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_launches_rans185.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_bench.log 2>&1; tail -2 gpurun_out/r2_ncu_bench.log
