# round 2, call 2: trunk NT=2 fix on the single-CTA kernels, then the CTA-pair chain kernels (default)
FINDB200_CHAIN=1 timeout 300 python -m pytest tests/test_field_gpu.py tests/test_parity_full_gpu.py -m gpu -q -k "two_per_cta or cfg4 or trajectory" > gpurun_out/c2_v1.log 2>&1; echo v1=$?; tail -4 gpurun_out/c2_v1.log
timeout 300 python -m pytest tests/test_field_gpu.py -m gpu -x -q > gpurun_out/c2_field.log 2>&1; echo field=$?; tail -15 gpurun_out/c2_field.log
timeout 600 python -m pytest tests/test_step_gpu.py tests/test_parity_full_gpu.py -m gpu -x -q > gpurun_out/c2_step.log 2>&1; echo step=$?; tail -15 gpurun_out/c2_step.log
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --steps 50 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err; echo bench=$?; tail -3 gpurun_out/c2_bench.err; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/c2_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    for k in d['kernels']: print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print({k:v['ms_per_step'] for k,v in d['other_workloads'].items()})
except Exception as e: print('no bench', e)
PY
FINDB200_CHAIN=1 timeout 300 python bench.py --skip-cpu-baseline --skip-eager --skip-other-workloads --steps 50 > gpurun_out/c2_bench_v1.json 2> gpurun_out/c2_bench_v1.err; echo bench_v1=$?; python -c "
import json; d=json.load(open('gpurun_out/c2_bench_v1.json')); print('v1 ms_per_step', d['ms_per_step'])"
