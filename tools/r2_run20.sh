timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_factor_rows -c 1 -o /tmp/fr python tools/bench_prec.py 185 1 > gpurun_out/r2_ncu_fr.log 2>&1
ncu -i /tmp/fr.ncu-rep --page raw --csv > gpurun_out/r2_ncu_factor_rows_raw.csv
ncu -i /tmp/fr.ncu-rep --page source --csv > gpurun_out/r2_ncu_factor_rows_source.csv
ncu -i /tmp/fr.ncu-rep --page source --csv --print-source cuda > gpurun_out/r2_ncu_factor_rows_cuda.csv 2> gpurun_out/r2_ncu_cuda.err
ncu -i /tmp/fr.ncu-rep --page details > gpurun_out/r2_ncu_factor_rows_details.txt
