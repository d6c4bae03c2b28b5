timeout 200 python -m pytest tests/test_field_gpu.py -m gpu -x -q > gpurun_out/c8_field.log 2>&1; echo field=$?; tail -3 gpurun_out/c8_field.log
timeout 100 python tools/chain2_trace.py > gpurun_out/c8_trace.txt 2>&1; head -16 gpurun_out/c8_trace.txt
for f in 2 4 8 16 30; do echo "== dbg flags $f"; FINDB200_CHAIN_DBG=$f TRACE_ACCOUNT_ONLY=1 timeout 100 python tools/chain2_trace.py 2>&1 | grep head_fwd; done
timeout 300 python bench.py --skip-cpu-baseline --skip-eager --skip-other-workloads --steps 50 > gpurun_out/c8_bench.json 2> gpurun_out/c8_bench.err; echo bench=$?; tail -3 gpurun_out/c8_bench.err; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/c8_bench.json'))
    print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    for k in d['kernels']:
        if k.get('frac'): print(' ', k['name'], k['ms_per_step'], k.get('frac'))
except Exception as e: print('no bench', e)
PY
