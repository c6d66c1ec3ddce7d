for rep in 1 2; do for iv in 0.25 0.05; do SU2B200_CLOCK_SAMPLE_S=$iv python bench.py --no-cpu-baseline > gpurun_out/r2_bench_iv${iv}_$rep.json 2> gpurun_out/r2_bench_iv${iv}_$rep.err; done; done
