# round 2: ncu --set full of the six math kernels of one step + the launch list (after the same command exited 0 without ncu)
CMD="python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-eager --skip-other-workloads"
$CMD > gpurun_out/ncu_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:chain2_kernel|dw_kernel|knn_cull_kernel' -s 12 -c 6 -o gpurun_out/r2_prof_math $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu_full=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo ncu_list=$?
ls -la gpurun_out/r2_prof_math.ncu-rep gpurun_out/r2_launches.csv
