# round 2: presorted training step (target sort + template order on a side stream beside the forward) — correctness, then A/B
timeout 900 python -m pytest tests/test_chamfer_gpu.py tests/test_step_gpu.py tests/test_parity_full_gpu.py tests/test_properties.py -m gpu -q -x > gpurun_out/b8_tests.log 2>&1; echo tests=$?; tail -5 gpurun_out/b8_tests.log
for v in 0 1; do
  FINDB200_PRESORT=$v timeout 300 python bench.py --steps 300 --warmup 20 --skip-cpu-baseline --skip-eager --skip-other-workloads > gpurun_out/b8_bench_$v.json 2> gpurun_out/b8_bench_$v.err; echo "presort=$v rc=$?"
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/b8_bench_$v.json'))
    print('  ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'launches', d['gpu_launches_per_step'])
    print('  ', [(k['name'], k['ms_per_step']) for k in d['kernels']])
    print('  knn pairs', [ (k.get('pairs_executed'), k.get('pairs_algorithmic')) for k in d['kernels'] if k['name']=='knn_main'])
except Exception as e: print('  fail', e); print(open('gpurun_out/b8_bench_$v.err').read()[-1500:])
PY
done
