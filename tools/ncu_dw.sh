timeout 200 python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/nd_plain.json 2> gpurun_out/nd_plain.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:^dw_kernel' -s 4 -c 2 -f -o gpurun_out/prof_dw3 python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/nd_ncu.log 2>&1; echo ncu=$?; tail -2 gpurun_out/nd_ncu.log
