# round 2 call A3: uniform-register MMA issue path (elect.sync guard): rate probe, field/step tests, bench both chains, trace
timeout 120 python tools/umma_rate.py > gpurun_out/a3_umma_rate.json 2> gpurun_out/a3_umma_rate.err; echo rate=$?; cat gpurun_out/a3_umma_rate.json
timeout 400 python -m pytest tests/test_umma_gpu.py tests/test_field_gpu.py tests/test_step_gpu.py tests/test_parity_full_gpu.py -m gpu -x -q > gpurun_out/a3_tests.log 2>&1; echo tests=$?; tail -5 gpurun_out/a3_tests.log
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --steps 50 > gpurun_out/a3_bench.json 2> gpurun_out/a3_bench.err; echo bench=$?; tail -3 gpurun_out/a3_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a3_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    for k in d['kernels']:
        if k.get('frac'): print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print({k:v.get('ms_per_step') for k,v in d.get('other_workloads',{}).items()})
except Exception as e: print('no bench', e)
PY
FINDB200_CHAIN=1 timeout 300 python bench.py --skip-cpu-baseline --skip-eager --skip-other-workloads --steps 50 > gpurun_out/a3_bench_v1.json 2> gpurun_out/a3_bench_v1.err; echo bench_v1=$?
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a3_bench_v1.json'))
    print('single-CTA chain ms_per_step', d['ms_per_step'])
    for k in d['kernels']:
        if k.get('frac'): print(' ', k['name'], k['ms_per_step'], k.get('frac'))
except Exception as e: print('no bench', e)
PY
timeout 100 python tools/chain2_trace.py > gpurun_out/a3_trace.txt 2>&1; head -30 gpurun_out/a3_trace.txt
