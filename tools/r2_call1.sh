# round 2, call 1: 2-CTA self-test + probe, full GPU suite (incl. the new full-size parity tests), smoke, bench line
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 120 python -m pytest tests/test_umma_gpu.py -m gpu -x -q > gpurun_out/c1_umma.log 2>&1; echo umma=$?; tail -5 gpurun_out/c1_umma.log
timeout 120 python tools/cta2_probe.py > gpurun_out/c1_probe.json 2> gpurun_out/c1_probe.err; echo probe=$?; cat gpurun_out/c1_probe.json; tail -3 gpurun_out/c1_probe.err
timeout 900 python -m pytest tests -m gpu -q --durations=12 > gpurun_out/c1_tests.log 2>&1; echo tests=$?; tail -40 gpurun_out/c1_tests.log
timeout 200 python __graft_entry__.py --smoke > gpurun_out/c1_smoke.log 2>&1; echo smoke=$?; tail -2 gpurun_out/c1_smoke.log
timeout 400 python bench.py > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo bench=$?; tail -3 gpurun_out/c1_bench.err; cut -c1-1500 gpurun_out/c1_bench.json
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/c1_ref.json 2> gpurun_out/c1_ref.err; echo ref=$?; cut -c1-600 gpurun_out/c1_ref.json
