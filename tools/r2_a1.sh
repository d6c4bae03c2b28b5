# round 2 re-entry, call A1: full GPU suite, smoke, bench (pair chain default + single-CTA A/B), ncu launch list
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -q --durations=10 > gpurun_out/a1_tests.log 2>&1; echo tests=$?; tail -30 gpurun_out/a1_tests.log
timeout 200 python __graft_entry__.py --smoke > gpurun_out/a1_smoke.log 2>&1; echo smoke=$?; tail -2 gpurun_out/a1_smoke.log
timeout 600 python bench.py > gpurun_out/a1_bench.json 2> gpurun_out/a1_bench.err; echo bench=$?; tail -3 gpurun_out/a1_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a1_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'])
    print('roofline', d['roofline'])
    for k in d['kernels']: print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print({k:v.get('ms_per_step') for k,v in d.get('other_workloads',{}).items()})
    print('eager', d.get('gpu_eager_baseline'))
    print('cpu', d.get('cpu_baseline'))
except Exception as e: print('no bench', e)
PY
FINDB200_CHAIN=1 timeout 300 python bench.py --skip-cpu-baseline --skip-eager --skip-other-workloads --steps 50 > gpurun_out/a1_bench_v1.json 2> gpurun_out/a1_bench_v1.err; echo bench_v1=$?
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a1_bench_v1.json'))
    print('single-CTA chain ms_per_step', d['ms_per_step'])
    for k in d['kernels']:
        if k.get('frac'): print(' ', k['name'], k['ms_per_step'], k.get('frac'))
except Exception as e: print('no bench', e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/a1_launches.csv python bench.py --skip-cpu-baseline --skip-eager --skip-other-workloads --steps 2 --warmup 1 > gpurun_out/a1_ncu.log 2>&1; echo ncu=$?
