# default bench (no profiler) then one `ncu --set full` pass over the six math kernels of one warm step
timeout 300 python bench.py --skip-cpu-baseline > gpurun_out/nf_bench.json 2> gpurun_out/nf_bench.err; echo bench=$?; tail -2 gpurun_out/nf_bench.err; python tools/gt_show.py gpurun_out/nf_bench.json
timeout 200 python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/nf_plain.json 2> gpurun_out/nf_plain.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:dw_kernel|chain_kernel|knn_cull_kernel' -s 18 -c 6 -f -o gpurun_out/prof_math python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/nf_ncu.log 2>&1; echo ncu=$?; tail -3 gpurun_out/nf_ncu.log
