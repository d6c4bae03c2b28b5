# round 2: dW plan with a per-segment cost x early weight
timeout 900 python -m pytest tests/test_field_gpu.py tests/test_step_gpu.py -m gpu -q -x > gpurun_out/b6_tests.log 2>&1; echo tests=$?; tail -3 gpurun_out/b6_tests.log
for cfg in "0 6" "1.35 6" "1.65 6" "1.65 4" "1.65 9" "1.9 6"; do
  set -- $cfg
  FINDB200_DW_SPLIT=$1 FINDB200_DW_COST=0.70,0.55,0.45,$2 timeout 300 python bench.py --steps 300 --warmup 20 --skip-cpu-baseline --skip-eager --skip-other-workloads > gpurun_out/b6_bench.json 2> gpurun_out/b6_bench.err; echo "w=$1 seg=$2 rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/b6_bench.json'))
print('  ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], [(k['name'], k['ms_per_step']) for k in d['kernels'] if k['name'] in ('sum_feet','trunk_bwd','dw_gemms')])
PY
done
for w in 1.65 0; do echo "== trace w=$w"; FINDB200_DW_SPLIT=$w timeout 200 python tools/dw_trace.py 2>&1 | grep -A14 "^latest\|^CTA start" | grep -v "cloud_sort\|^ *[0-9]*:" ; done
