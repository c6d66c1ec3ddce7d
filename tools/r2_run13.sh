for rep in 1 2; do for f in 1 0; do SU2B200_FUSED_UPPER=$f python bench.py --no-cpu-baseline > gpurun_out/r2_bench_fused${f}_$rep.json 2> gpurun_out/r2_bench_fused${f}_$rep.err; done; done
