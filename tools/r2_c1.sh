# the no-allocation entry point: alone in a fresh process, then after the whole suite
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c1_tests.log 2>&1; echo tests=$?; tail -4 gpurun_out/c1_tests.log
