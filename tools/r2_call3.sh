timeout 200 python -m pytest tests/test_umma_gpu.py tests/test_field_gpu.py -m gpu -x -q > gpurun_out/c3_field.log 2>&1; echo field=$?; tail -3 gpurun_out/c3_field.log
timeout 600 python -m pytest tests/test_step_gpu.py tests/test_parity_full_gpu.py -m gpu -x -q > gpurun_out/c3_step.log 2>&1; echo step=$?; tail -3 gpurun_out/c3_step.log
timeout 120 python tools/chain2_trace.py > gpurun_out/c3_trace.txt 2>&1; echo trace=$?; cat gpurun_out/c3_trace.txt
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --steps 50 > gpurun_out/c3_bench.json 2> gpurun_out/c3_bench.err; echo bench=$?; tail -3 gpurun_out/c3_bench.err; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/c3_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    for k in d['kernels']:
        if k.get('frac'): print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print({k:v['ms_per_step'] for k,v in d['other_workloads'].items()})
except Exception as e: print('no bench', e)
PY
