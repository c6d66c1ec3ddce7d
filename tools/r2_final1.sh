set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_final_gpu_tests.log 2>&1; tail -3 gpurun_out/r2_final_gpu_tests.log
timeout 600 python bench.py > gpurun_out/r2_final_bench_rans185.json 2> gpurun_out/r2_final_bench_rans185.err
timeout 600 python bench.py --no-cpu-baseline --solver NAVIER_STOKES > gpurun_out/r2_final_bench_ns185.json 2>/dev/null
timeout 600 python bench.py --no-cpu-baseline --solver EULER > gpurun_out/r2_final_bench_euler185.json 2>/dev/null
timeout 600 python bench.py --no-cpu-baseline --no-rcm > gpurun_out/r2_final_bench_rans185_lexicographic.json 2>/dev/null
timeout 600 python bench.py --no-cpu-baseline --tets --n 129,129,129 > gpurun_out/r2_final_bench_rans129_tets.json 2> gpurun_out/r2_final_bench_tets.err
timeout 900 python bench.py --no-cpu-baseline --solver EULER --n 292,292,292 --steps 10 --warmup 3 > gpurun_out/r2_final_bench_euler292_config5.json 2> gpurun_out/r2_final_bench_config5.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_final_launches_rans185.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_final_ncu_bench.log 2>&1
timeout 900 python bench.py --impl reference > gpurun_out/r2_final_bench_reference_arm.json 2> gpurun_out/r2_final_ref.err
for f in gpurun_out/r2_final_bench_*.json; do echo $f; head -c 260 $f; echo; done
