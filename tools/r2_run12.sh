set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_full12.log 2>&1; tail -5 gpurun_out/r2_gpu_full12.log
for f in 1 0; do SU2B200_FUSED_UPPER=$f python bench.py --no-cpu-baseline > gpurun_out/r2_bench_fused$f.json 2> gpurun_out/r2_bench_fused$f.err; tail -c 300 gpurun_out/r2_bench_fused$f.json; done
