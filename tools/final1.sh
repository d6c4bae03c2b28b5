timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/f1_tests.log 2>&1; echo tests=$?; tail -2 gpurun_out/f1_tests.log
timeout 100 python tools/chain_trace.py > gpurun_out/f1_trace.txt 2>&1; echo trace=$?; tail -22 gpurun_out/f1_trace.txt
timeout 300 python bench.py > gpurun_out/f1_bench.json 2>gpurun_out/f1_bench.err; echo bench=$?; python tools/gt_show.py gpurun_out/f1_bench.json | sed -n 1p
