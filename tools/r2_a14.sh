timeout 600 python -m pytest tests/test_field_gpu.py tests/test_step_gpu.py tests/test_parity_full_gpu.py tests/test_chamfer_gpu.py -m gpu -x -q > gpurun_out/a14_tests.log 2>&1; echo tests=$?; tail -5 gpurun_out/a14_tests.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -1; timeout 200 python tools/chain_stress.py 200 2>&1 | tail -3
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --steps 50 > gpurun_out/a14_bench.json 2> gpurun_out/a14_bench.err; echo bench=$?; tail -3 gpurun_out/a14_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a14_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'launches', d['gpu_launches_per_step'])
    for k in d['kernels']: print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print({k:v.get('ms_per_step') for k,v in d.get('other_workloads',{}).items()})
    print(d['loss'])
except Exception as e: print('no bench', e)
PY
