# GPU regression: all gpu tests then the default bench; every step under its own timeout
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/gt_tests.log 2>&1; echo tests=$?; tail -3 gpurun_out/gt_tests.log
timeout 200 python bench.py "$@" > gpurun_out/gt_bench.json 2> gpurun_out/gt_bench.err; echo bench=$?; tail -2 gpurun_out/gt_bench.err
