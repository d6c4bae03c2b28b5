# round 2: dW cost model refit (fixed 0.70 us, 0.55 us / 64 KB HBM, 0.45 us / 64 KB L2) x early weight
for w in 0 1.35 1.65 1.9; do
  FINDB200_DW_SPLIT=$w timeout 300 python bench.py --steps 300 --warmup 20 --skip-cpu-baseline --skip-eager --skip-other-workloads > gpurun_out/b4_bench_$w.json 2> gpurun_out/b4_bench_$w.err; echo "w=$w rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/b4_bench_$w.json'))
print('  ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], [(k['name'], k['ms_per_step']) for k in d['kernels'] if k['name'] in ('sum_feet','trunk_bwd','dw_gemms')])
PY
done
for w in 1.65 0; do echo "== trace w=$w"; FINDB200_DW_SPLIT=$w timeout 200 python tools/dw_trace.py 2>&1 | grep -A40 "single-GEMM\|^CTA start" | grep -v "cloud_sort\|^ *[0-9]*:" ; done
