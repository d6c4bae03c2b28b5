# round 2, final build: ncu --set full of one step's math kernels (7 launches: trunk_fwd, head_fwd, knn, head_bwd, dw early, trunk_bwd, dw late)
CMD="python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-eager --skip-other-workloads"
$CMD > gpurun_out/ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:chain2_kernel|dw_kernel|knn_cull_kernel' -s 14 -c 7 -f -o gpurun_out/r2_prof_math_final $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu_full=$?
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/r2_prof_math_final.ncu-rep
