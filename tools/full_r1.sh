# Round regression + profile: gpu tests, smoke, default bench, reference arm, ncu launch list, other workloads at N=1
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/fr_tests.log 2>&1; echo tests=$?; tail -3 gpurun_out/fr_tests.log
timeout 200 python __graft_entry__.py --smoke > gpurun_out/fr_smoke.log 2>&1; echo smoke=$?; tail -2 gpurun_out/fr_smoke.log
timeout 300 python bench.py > gpurun_out/fr_bench.json 2> gpurun_out/fr_bench.err; echo bench=$?; tail -2 gpurun_out/fr_bench.err; cat gpurun_out/fr_bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/fr_ref.json 2> gpurun_out/fr_ref.err; echo ref=$?; cat gpurun_out/fr_ref.json
for w in cfg3 cfg4 cfg5; do timeout 300 python bench.py --workload $w --skip-cpu-baseline --steps 20 --warmup 3 > gpurun_out/fr_bench_$w.json 2> gpurun_out/fr_bench_$w.err; echo $w=$?; cat gpurun_out/fr_bench_$w.json; done
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/fr_launches.csv python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/fr_ncu.log 2>&1; echo ncu=$?
