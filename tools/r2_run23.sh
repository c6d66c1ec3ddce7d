timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_poll_sweep_rowsU -c 2 -o /tmp/sw python tools/bench_prec.py 185 1 > gpurun_out/r2_ncu_sw.log 2>&1
ncu -i /tmp/sw.ncu-rep --page raw --csv > gpurun_out/r2_ncu_sweep_raw.csv
ncu -i /tmp/sw.ncu-rep --page source --csv > gpurun_out/r2_ncu_sweep_source.csv
ncu -i /tmp/sw.ncu-rep --page details > gpurun_out/r2_ncu_sweep_details.txt
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_edge_roe -c 1 -o /tmp/ed python tools/bench_edge.py 185 > gpurun_out/r2_ncu_ed.log 2>&1
ncu -i /tmp/ed.ncu-rep --page raw --csv > gpurun_out/r2_ncu_edge_raw.csv
ncu -i /tmp/ed.ncu-rep --page source --csv > gpurun_out/r2_ncu_edge_source.csv
ncu -i /tmp/ed.ncu-rep --page details > gpurun_out/r2_ncu_edge_details.txt
tail -3 gpurun_out/r2_ncu_ed.log
