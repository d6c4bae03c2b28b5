# round 2: early / late dW launch — CTA lifetimes and a finer sweep of the early weight
for w in 1.5 1.0; do echo "== trace w=$w"; FINDB200_DW_SPLIT=$w timeout 200 python tools/dw_trace.py 2>/dev/null | grep -v "^dw gemm" | head -24; done
for w in 1.2 1.35 1.5 1.65; do
  FINDB200_DW_SPLIT=$w timeout 300 python bench.py --steps 300 --warmup 20 --skip-cpu-baseline --skip-eager --skip-other-workloads > gpurun_out/b2_bench_$w.json 2> gpurun_out/b2_bench_$w.err; echo "w=$w rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/b2_bench_$w.json'))
print('  ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], [(k['name'], k['ms_per_step']) for k in d['kernels'] if k['name'] in ('sum_feet','trunk_bwd','dw_gemms')])
PY
done
