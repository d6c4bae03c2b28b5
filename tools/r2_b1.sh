# round 2: early / late launch of the dW kernel beside the trunk's backward chain — correctness, then A/B over the early weight
timeout 900 python -m pytest tests/test_field_gpu.py tests/test_step_gpu.py tests/test_parity_full_gpu.py -m gpu -q -x > gpurun_out/b1_tests.log 2>&1; echo tests=$?; tail -6 gpurun_out/b1_tests.log
for w in 0 1.9 1.5 2.3 1.0; do
  FINDB200_DW_SPLIT=$w timeout 300 python bench.py --steps 200 --warmup 20 --skip-cpu-baseline --skip-eager --skip-other-workloads > gpurun_out/b1_bench_$w.json 2> gpurun_out/b1_bench_$w.err; echo "w=$w rc=$?"
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/b1_bench_$w.json'))
    print('  ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    print('  ', [(k['name'], k['ms_per_step']) for k in d['kernels']])
except Exception as e: print('  fail', e); print(open('gpurun_out/b1_bench_$w.err').read()[-1500:])
PY
done
