set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_fin3_smoke.log 2>&1; tail -2 gpurun_out/r2_fin3_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2_fin3_gpu_tests.log 2>&1; tail -3 gpurun_out/r2_fin3_gpu_tests.log
timeout 600 python bench.py > gpurun_out/r2_fin3_bench_rans185.json 2> gpurun_out/r2_fin3_bench.err; tail -c 300 gpurun_out/r2_fin3_bench_rans185.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_fin3_launches_rans185.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_fin3_ncu.log 2>&1
