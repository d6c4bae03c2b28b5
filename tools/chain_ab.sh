timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/ca_tests.log 2>&1; echo tests=$?; tail -3 gpurun_out/ca_tests.log
timeout 100 python tools/chain_trace.py
timeout 200 python bench.py --skip-cpu-baseline > gpurun_out/ca_bench.json 2>gpurun_out/ca_bench.err; echo bench=$?; python tools/gt_show.py gpurun_out/ca_bench.json | sed -n 1p; python tools/gt_show.py gpurun_out/ca_bench.json | sed -n 3p
for w in cfg4 cfg5; do timeout 200 python bench.py --skip-cpu-baseline --workload $w --steps 30 > gpurun_out/ca_$w.json 2>/dev/null; python tools/gt_show.py gpurun_out/ca_$w.json | sed -n 1p; done
