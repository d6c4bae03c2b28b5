timeout 300 python -m pytest tests/test_field_gpu.py tests/test_step_gpu.py -x -q > gpurun_out/k2_tests.log 2>&1; echo tests=$?; tail -2 gpurun_out/k2_tests.log
timeout 200 python bench.py --skip-cpu-baseline > gpurun_out/k2_bench.json 2>gpurun_out/k2_bench.err; echo bench=$?; python tools/gt_show.py gpurun_out/k2_bench.json | sed -n 1p; python tools/gt_show.py gpurun_out/k2_bench.json | sed -n 3p
KB_MODES=0:1,0:2,0:4 timeout 200 python tools/knn_bench.py 8 10242 10000 > gpurun_out/k2_knn.jsonl 2>gpurun_out/k2_knn.err; echo kb=$?
python -c "
import json
for l in open('gpurun_out/k2_knn.jsonl'):
    d=json.loads(l); print(d['kq'], d['ms_median'], d['pairs_exec'], d['max_evals'], d['groups'], {k:round(v,4) for k,v in d['kernels_ms'].items()})"
