# final-build profile: default bench (no ncu), then the ncu launch list and one --set full capture of dw_kernel + trunk_fwd
timeout 300 python bench.py --skip-cpu-baseline > gpurun_out/pf_bench.json 2> gpurun_out/pf_bench.err; echo bench=$?
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/pf_launches.csv python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/pf_ncu.log 2>&1; echo ncu=$?
timeout 400 ncu --set full --clock-control none --import-source on -k 'regex:^dw_kernel|^chain_kernel' -s 10 -c 5 -o gpurun_out/pf_math python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/pf_ncu2.log 2>&1; echo ncu2=$?
ls -la gpurun_out/pf_math.ncu-rep
