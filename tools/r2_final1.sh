# round 2: full GPU suite, smoke, default bench line (+ CPU baseline, eager comparators, other workloads), reference arm
timeout 1200 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/f1_tests.log 2>&1; echo tests=$?; tail -14 gpurun_out/f1_tests.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/f1_bench.json 2> gpurun_out/f1_bench.err; echo bench=$?; tail -2 gpurun_out/f1_bench.err
timeout 400 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/f1_ref.json 2> gpurun_out/f1_ref.err; echo ref=$?; cut -c1-700 gpurun_out/f1_ref.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/f1_bench.json'))
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['value'], 'launches', d['gpu_launches_per_step'])
print('roofline', json.dumps(d['roofline']))
for k in d['kernels']: print(' ', k['name'], k['ms_per_step'], k.get('frac'))
print({k:(v.get('ms_per_step'), v.get('value')) for k,v in d.get('other_workloads',{}).items()})
print('eager', json.dumps(d.get('gpu_eager_baseline'))[:600])
print('cpu', d.get('cpu_baseline'))
print('shipped', d.get('shipped_3d_step'))
print('clocks', d.get('clocks'))
PY
