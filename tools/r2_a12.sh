timeout 400 python -m pytest tests/test_field_gpu.py tests/test_step_gpu.py tests/test_parity_full_gpu.py -m gpu -x -q > gpurun_out/a12_tests.log 2>&1; echo tests=$?; tail -5 gpurun_out/a12_tests.log
timeout 120 python tools/dw_trace.py > gpurun_out/a12_dw_trace.txt 2>&1; grep -v Warn gpurun_out/a12_dw_trace.txt | tail -38
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --steps 50 > gpurun_out/a12_bench.json 2> gpurun_out/a12_bench.err; echo bench=$?; tail -3 gpurun_out/a12_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a12_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    for k in d['kernels']:
        if k.get('frac'): print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print({k:v.get('ms_per_step') for k,v in d.get('other_workloads',{}).items()})
except Exception as e: print('no bench', e)
PY
