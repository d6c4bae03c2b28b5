# round 2, final build: launch list (after the same command exited 0 without ncu)
CMD="python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-eager --skip-other-workloads"
$CMD > gpurun_out/ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches_final.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo ncu_list=$?
ls -la gpurun_out/r2_launches_final.csv
