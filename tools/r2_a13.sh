for c in "0.45,0.8,0.5" "0.45,0.8,0.75" "0.45,0.8,0.95" "0.6,0.7,0.8" "0.3,0.9,0.9"; do echo "== cost $c"; FINDB200_DW_COST=$c timeout 120 python tools/dw_trace.py 2>&1 | grep "kernel"; done
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --skip-other-workloads --steps 50 > gpurun_out/a13_bench.json 2> gpurun_out/a13_bench.err; echo bench=$?; tail -3 gpurun_out/a13_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/a13_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d.get('kernel_timing'))
    tot=0
    for k in d['kernels']:
        tot+=k['ms_per_step']; print(' ', k['name'], k['ms_per_step'], k.get('frac'))
    print('sum', tot)
except Exception as e: print('no bench', e)
PY
